// GetVTable.h -- the plugin ABI of a netsDB user type (objectModel/headers/GetVTable.h:8-58): every UDF / record type is
// its own shared library built from a .cc file that holds nothing but GET_V_TABLE(Type), exporting
//     extern "C" char *getObjectTypeName();
//     extern "C" void *getObjectVTable();
//     extern "C" void  setAllGlobalVariables(Allocator *, VTableMap *, void *stackBase, void *stackEnd);
// which PDBClient::registerType ships to the catalog and every worker dlopen()s on first use of the type
// (objectModel/source/VTableMapCatalogLookup.cc:77-128).  Same three symbols here, so the reference's one-line .cc files
// (src/FF/source/FFTransposeMult.cc, ...) compile unmodified into libraries/lib<Type>.so against the shim classes.
#pragma once
#include <cxxabi.h>

#include "pdb_shim.h"

namespace pdb {
class Allocator;   // opaque here: the shim's objects live on the C++ heap
class VTableMap;
// TypeName.h:19: the demangled name of the type
template <class T>
std::string getTypeName() {
    int status = 0;
    char *d = abi::__cxa_demangle(typeid(T).name(), nullptr, nullptr, &status);
    std::string s = (status == 0 && d) ? d : typeid(T).name();
    free(d);
    return s;
}
}  // namespace pdb

#undef GET_V_TABLE
#define GET_V_TABLE(TYPE_NAME)                                                                         \
    namespace pdb {                                                                                    \
    char typeName[300];                                                                                \
    Allocator *pluginAllocator = nullptr;                                                              \
    VTableMap *pluginVTable = nullptr;                                                                 \
    void *pluginStackBase = nullptr, *pluginStackEnd = nullptr;                                        \
    bool inSharedLibrary = false;                                                                      \
    extern "C" {                                                                                       \
    char *getObjectTypeName() {                                                                        \
        std::string typeString = getTypeName<TYPE_NAME>();                                             \
        memmove(typeName, typeString.c_str(), typeString.size() + 1);                                  \
        return typeName;                                                                               \
    }                                                                                                  \
    void *getObjectVTable() {                                                                          \
        TYPE_NAME temp;                                                                                \
        return *reinterpret_cast<void **>(&temp); /* the vptr of a default-constructed instance */     \
    }                                                                                                  \
    void setAllGlobalVariables(Allocator *newAllocator, VTableMap *theVTableIn, void *stackBaseIn, void *stackEndIn) { \
        pluginStackBase = stackBaseIn;                                                                 \
        pluginAllocator = newAllocator;                                                                \
        pluginStackEnd = stackEndIn;                                                                   \
        pluginVTable = theVTableIn;                                                                    \
        inSharedLibrary = true;                                                                        \
    }                                                                                                  \
    }                                                                                                  \
    }
