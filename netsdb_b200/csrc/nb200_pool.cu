// nb200_pool.cu -- caching device allocator behind every cudaMalloc / cudaFree of the library.
//
// netsDB allocates every intermediate block out of its page pool (Pangea); the operator-by-operator entry points here
// create a result set, operand planes and temporaries per call, and a raw cudaFree is a device-wide synchronisation.
// Freed blocks of up to kMaxPooled bytes are kept (up to kMaxCached in total) and handed out again for requests of
// nearly the same size, so a steady sequence of calls (one LSTM step, the jobs of an inference driver) stops touching
// the CUDA allocator.
//
// Why immediate reuse is safe: all kernels and device-side copies of the library run on ONE compute stream, so a block
// "freed" while work that reads it is still queued is only ever rewritten by work queued later on the same stream.
// The two upload streams are ordered against the compute stream in both directions by the entry points
// (wait_uploads / wait_set_uploads before compute, main_ev before an upload overwrites a set), and a set with uploads
// in flight is waited for before its storage is released (free_set_storage).
#include "nb200_internal.h"

#include <map>
#include <unordered_map>

namespace nb200 {

namespace {
constexpr size_t kMaxPooled = size_t(2) << 30;   // larger blocks (the 47.8 GB input sets) go straight to the driver
constexpr size_t kMaxCached = size_t(16) << 30;  // bytes kept in the free lists, per device
constexpr int kMaxDev = 16;
struct Pool {                                    // one per device: a block is only ever handed out on the device it lives on
    std::mutex mu;
    std::multimap<size_t, void *> free_blocks;   // size -> block
    std::unordered_map<void *, size_t> size;     // every live or cached block of the pool -> its true size
    size_t cached = 0;
};
Pool g_pool[kMaxDev];

Pool &current_pool() {
    int d = 0;
    cudaGetDevice(&d);
    return g_pool[d >= 0 && d < kMaxDev ? d : 0];
}

void trim_locked(Pool &P) {
    for (auto &kv : P.free_blocks) {
        P.size.erase(kv.second);
        cudaFree(kv.second);
    }
    P.free_blocks.clear();
    P.cached = 0;
}
}  // namespace

cudaError_t pool_malloc(void **p, size_t bytes) {
    const size_t sz = bytes < 512 ? 512 : (bytes + 511) / 512 * 512;
    Pool &P = current_pool();
    std::lock_guard<std::mutex> lk(P.mu);
    if (sz <= kMaxPooled) {
        auto it = P.free_blocks.lower_bound(sz);
        if (it != P.free_blocks.end() && it->first <= sz + sz / 8) {  // at most 12.5 % larger than asked for
            *p = it->second;
            P.cached -= it->first;
            P.free_blocks.erase(it);
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMalloc(p, sz);
    if (e == cudaErrorMemoryAllocation) {  // give the cached blocks back and try once more
        cudaGetLastError();
        trim_locked(P);
        e = cudaMalloc(p, sz);
    }
    if (e == cudaSuccess) P.size[*p] = sz;
    return e;
}

cudaError_t pool_free(void *p) {
    if (!p) return cudaSuccess;
    // the owning pool: the current device's in all but exotic cases (a set destroyed from a thread working elsewhere)
    Pool *owner = nullptr;
    {
        Pool &P = current_pool();
        std::lock_guard<std::mutex> lk(P.mu);
        if (P.size.count(p)) owner = &P;
    }
    for (int d = 0; d < kMaxDev && !owner; d++) {
        std::lock_guard<std::mutex> lk(g_pool[d].mu);
        if (g_pool[d].size.count(p)) owner = &g_pool[d];
    }
    if (!owner) return cudaFree(p);  // not ours (never happens for library allocations)
    std::lock_guard<std::mutex> lk(owner->mu);
    auto it = owner->size.find(p);
    if (it == owner->size.end()) return cudaFree(p);
    const size_t sz = it->second;
    if (sz <= kMaxPooled && owner->cached + sz <= kMaxCached) {
        owner->free_blocks.emplace(sz, p);
        owner->cached += sz;
        return cudaSuccess;
    }
    owner->size.erase(it);
    return cudaFree(p);
}

void pool_release_all() {   // the current device's cached blocks
    Pool &P = current_pool();
    std::lock_guard<std::mutex> lk(P.mu);
    trim_locked(P);
}

}  // namespace nb200
