// PDBClient.h -- the client facade the drivers use (mainClient/headers/PDBClient.h:40-259), in process:
// sets live on the device behind the C ABI, executeComputations hands the graph to the fused FF stage.
// Calls return bool + errMsg like the reference's; nothing here computes on the host.
#pragma once
// the real PDBClient.h pulls these in transitively and the reference drivers rely on that
#include <dlfcn.h>
#include <unistd.h>

#include <chrono>
#include <fstream>

#include "ff_stage.h"

namespace pdb {

class CatalogClient {
public:
    CatalogClient() {}
    CatalogClient(int, std::string, PDBLoggerPtr) {}
    bool setExists(const std::string &db, const std::string &set) {
        nb200_set *s = nullptr;
        return nb200_set_lookup(db.c_str(), set.c_str(), &s) == NB200_OK;
    }
};

// getSetIterator<T>(db, set): range-for over Handle<FFMatrixBlock>, one download per block
// (queries/headers/SetIterator.h:12-68)
template <class T>
class SetIterator {
public:
    class It {
    public:
        It(nb200_set *s, int idx, int n, int nbc) : s(s), idx(idx), n(n), nbc(nbc) {}
        bool operator!=(const It &o) const { return idx != o.idx; }
        void operator++() { idx++; }
        It(const std::vector<Handle<T>> *host, int idx) : s(nullptr), idx(idx), n(0), nbc(1), host(host) {}
        It(const std::pair<std::string, std::string> *names, int idx, int nbc) : s(nullptr), idx(idx), n(0), nbc(nbc), names(names) {}
        Handle<T> operator*() const {
            if constexpr (std::is_base_of<conv2d_memory_fusion::Matrix3D, T>::value) return (*host)[idx];
            else if (names) return nb200_host::fetch_block_sharded<T>(names->first, names->second, idx / nbc, idx % nbc);
            else return nb200_host::fetch_block<T>(s, idx / nbc, idx % nbc);
        }

    private:
        nb200_set *s;
        int idx, n, nbc;
        const std::vector<Handle<T>> *host = nullptr;
        const std::pair<std::string, std::string> *names = nullptr;   // a set whose rows are spread over several devices
    };
    SetIterator() {}
    explicit SetIterator(const std::vector<Handle<T>> *records) : host(records) {}
    explicit SetIterator(nb200_set *s) : s(s) {
        if (s) {
            int tr, tc, br, bc;
            nb200_set_info(s, &tr, &tc, &br, &bc);
            nbr = (tr + br - 1) / br;
            nbc = (tc + bc - 1) / bc;
        }
    }
    // a row-sharded set (NB200_DEVICES > 1): the block grid spans the rows of all devices
    SetIterator(const std::string &db, const std::string &set, int total_rows, int total_cols, int br, int bc)
        : names(db, set), sharded(true) {
        nbr = (total_rows + br - 1) / br;
        nbc = (total_cols + bc - 1) / bc;
    }
    It begin() {
        if (sharded) return It(&names, 0, nbc > 0 ? nbc : 1);
        return host ? It(host, 0) : It(s, 0, nbr * nbc, nbc > 0 ? nbc : 1);
    }
    It end() {
        if (sharded) return It(&names, nbr * nbc, nbc > 0 ? nbc : 1);
        return host ? It(host, static_cast<int>(host->size())) : It(s, nbr * nbc, nbr * nbc, nbc > 0 ? nbc : 1);
    }

private:
    nb200_set *s = nullptr;
    int nbr = 0, nbc = 0;
    const std::vector<Handle<T>> *host = nullptr;
    std::pair<std::string, std::string> names;
    bool sharded = false;
};

class PDBClient {
public:
    PDBClient() { nb200_host::ensure_init(); }
    PDBClient(int /*port*/, std::string /*address*/, PDBLoggerPtr /*logger*/, bool /*usePangea*/ = false,
              bool /*useQueryScheduler*/ = false) { nb200_host::ensure_init(); }

    bool createDatabase(const std::string &, std::string &) { return true; }
    bool removeDatabase(const std::string &, std::string &) { return true; }
    // registerType (mainClient/headers/PDBClient.h:176-180): the reference ships the UDF's shared library to the catalog.
    // In process there is nothing to ship; when the library exists (as given, under $NB200_LIBRARIES, or next to the
    // driver: <exe dir>/../<path>) it is loaded and its plugin ABI checked -- getObjectTypeName / getObjectVTable /
    // setAllGlobalVariables, GetVTable.h -- and a broken library fails the call.  A path that does not exist is accepted:
    // the shim's types are compiled into the driver.
    bool registerType(const std::string &fileContainingSharedLib, std::string &errMsg) {
        std::vector<std::string> candidates{fileContainingSharedLib};
        const std::string base = fileContainingSharedLib.substr(fileContainingSharedLib.find_last_of('/') + 1);
        if (const char *e = getenv("NB200_LIBRARIES")) candidates.push_back(std::string(e) + "/" + base);
        char self[4096];
        const ssize_t len = readlink("/proc/self/exe", self, sizeof(self) - 1);
        if (len > 0) {
            std::string exe(self, len);
            candidates.push_back(exe.substr(0, exe.find_last_of('/')) + "/../" + fileContainingSharedLib);
        }
        for (const std::string &path : candidates) {
            if (access(path.c_str(), R_OK) != 0) continue;
            void *h = dlopen(path.c_str(), RTLD_LOCAL | RTLD_LAZY);  // the flags of VTableMapCatalogLookup.cc:92
            if (!h) { errMsg = std::string("registerType: ") + dlerror(); return false; }
            typedef char *(*name_fn)();
            typedef void *(*vtable_fn)();
            typedef void (*globals_fn)(void *, void *, void *, void *);  // (Allocator *, VTableMap *, stackBase, stackEnd)
            name_fn getName = reinterpret_cast<name_fn>(dlsym(h, "getObjectTypeName"));
            vtable_fn getVTable = reinterpret_cast<vtable_fn>(dlsym(h, "getObjectVTable"));
            globals_fn setGlobals = reinterpret_cast<globals_fn>(dlsym(h, "setAllGlobalVariables"));
            if (!getName || !getVTable || !setGlobals) { errMsg = "registerType: " + path + " does not export the plugin ABI"; return false; }
            setGlobals(nullptr, nullptr, nullptr, nullptr);
            if (!getVTable()) { errMsg = "registerType: " + path + ": no vtable"; return false; }
            registered[getName()] = path;
            return true;
        }
        return true;
    }
    const std::map<std::string, std::string> &registeredTypes() const { return registered; }
    // the device set is sized when the first block arrives (sendData), as the reference's set is filled by sendData
    template <class T>
    bool createSet(const std::string &db, const std::string &set, std::string &, size_t = 0, const std::string & = "",
                   Handle<Computation> = nullptr, Handle<Computation> = nullptr) {
        declared.insert(std::make_pair(db, set));
        return true;
    }
    bool createSet(const std::string &db, const std::string &set, const std::string & /*typeName*/, std::string &,
                   size_t = 0, const std::string & = "") {
        declared.insert(std::make_pair(db, set));
        return true;
    }
    bool removeSet(const std::string &db, const std::string &set, std::string &) { return drop(db, set); }
    bool deleteSet(std::string db, std::string set) { return drop(db, set); }  // mainClient/headers/PDBClient.h:253
    bool clearSet(const std::string &db, const std::string &set, const std::string &, std::string &) { return drop(db, set); }
    bool flushData(std::string &errMsg) { return sync_all(errMsg); }

    // NOTE the (set, db) order of the pair, as in the reference (FFMatrixUtil.cc:110-111)
    template <class T>
    bool sendData(std::pair<std::string, std::string> setAndDatabase, Handle<Vector<Handle<T>>> data, std::string &errMsg) {
        const std::string &set = setAndDatabase.first, &db = setAndDatabase.second;
        if constexpr (std::is_base_of<conv2d_memory_fusion::Matrix3D, T>::value) {
            // Image / Kernel records stay on the host until a job scans them (conv_stage.h)
            auto &dst = nb200_host::host_sets<T>()[std::make_pair(db, set)];
            for (size_t i = 0; i < data->size(); i++) dst.push_back((*data)[i]);
            (void)errMsg;
            return true;
        } else {
        int tr = 0, tc = 0;  // extents of the block grid, for blocks whose meta carry no totals
        for (size_t i = 0; i < data->size(); i++) {
            Handle<T> &b = (*data)[i];
            tr = std::max(tr, (b->getBlockRowIndex() + 1) * b->getRowNums());
            tc = std::max(tc, (b->getBlockColIndex() + 1) * b->getColNums());
        }
        for (size_t i = 0; i < data->size(); i++)
            if (!ok(nb200_host::ship_block(db, set, (*data)[i], tr, tc), errMsg)) return false;
        return sync_all(errMsg);  // the caller reuses its allocation block after sendData
        }
    }

    template <class... Sinks>
    bool executeComputations(std::string &errMsg, const std::string &jobName, bool /*preCompile*/, Handle<Computation> first,
                             Sinks... rest) {
        std::vector<Handle<Computation>> sinks{first, Handle<Computation>(rest)...};
        return stage.run(jobName, sinks, errMsg);
    }
    template <class... Sinks>
    bool executeComputations(std::string &errMsg, const std::string &jobName, Handle<Computation> first, Sinks... rest) {
        return executeComputations(errMsg, jobName, false, first, rest...);
    }
    template <class... Sinks>
    bool executeComputations(std::string &errMsg, Handle<Computation> first, Sinks... rest) {
        return executeComputations(errMsg, std::string("job"), false, first, rest...);
    }

    template <class T>
    SetIterator<T> getSetIterator(const std::string &db, const std::string &set) {
        if constexpr (std::is_base_of<conv2d_memory_fusion::Matrix3D, T>::value) {
            auto &m = nb200_host::host_sets<T>();
            auto it = m.find(std::make_pair(db, set));
            return it == m.end() ? SetIterator<T>() : SetIterator<T>(&it->second);
        }
        nb200_set *s = nullptr;
        if (nb200_host::num_devices() > 1 && nb200_host::sharded_sets().count(std::make_pair(db, set))) {
            int tr = 0, tc = 0, br = 1, bc = 1;
            for (int d = 0; d < nb200_host::num_devices(); d++) {
                nb200_use_device(d);
                int r = 0;
                if (nb200_set_lookup(db.c_str(), set.c_str(), &s) != NB200_OK || nb200_set_info(s, &r, &tc, &br, &bc) != NB200_OK) {
                    nb200_use_device(0);
                    return SetIterator<T>();
                }
                tr += r;
            }
            nb200_use_device(0);
            return SetIterator<T>(db, set, tr, tc, br, bc);
        }
        if (nb200_set_lookup(db.c_str(), set.c_str(), &s) != NB200_OK) return SetIterator<T>();
        return SetIterator<T>(s);
    }

    // extension: labels of the last fused inference (first row argmax of the logits, equal to the fp64 pipeline's)
    const std::vector<int32_t> &lastLabels() const { return stage.labels; }
    const nb200_ff_stats &lastStats() const { return stage.stats; }
    nb200_host::FusedStage stage;

private:
    static bool ok(int st, std::string &errMsg) {
        if (st == NB200_OK) return true;
        errMsg = nb200_last_error();
        return false;
    }
    static bool sync_all(std::string &errMsg) {
        bool good = true;
        for (int d = 0; d < nb200_host::num_devices(); d++) {
            if (nb200_host::num_devices() > 1) nb200_use_device(d);
            good = ok(nb200_synchronize(), errMsg) && good;
        }
        if (nb200_host::num_devices() > 1) nb200_use_device(0);
        return good;
    }
    bool drop(const std::string &db, const std::string &set) {
        nb200_host::host_set_drop(db, set);
        for (int d = 0; d < nb200_host::num_devices(); d++) {
            if (nb200_host::num_devices() > 1) nb200_use_device(d);
            nb200_set *s = nullptr;
            if (nb200_set_lookup(db.c_str(), set.c_str(), &s) == NB200_OK) nb200_set_destroy(s);
        }
        if (nb200_host::num_devices() > 1) nb200_use_device(0);
        nb200_host::sharded_sets().erase(std::make_pair(db, set));
        return true;
    }
    std::set<std::pair<std::string, std::string>> declared;
    std::map<std::string, std::string> registered;  // type name -> library it was loaded from
};

}  // namespace pdb
