// PDBDebug.h -- PDB_COUT as in pdbServer/headers/PDBDebug.h:20-33: std::cout with -DPDB_DEBUG, a sink otherwise
#pragma once
#include <iostream>
struct NullStream {
    template <typename T> NullStream &operator<<(const T &) { return *this; }
    NullStream &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
#ifdef PDB_DEBUG
#define PDB_COUT std::cout
#else
#define PDB_COUT NullStream()
#endif
