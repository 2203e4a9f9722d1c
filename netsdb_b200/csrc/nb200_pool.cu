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
constexpr size_t kMaxCached = size_t(16) << 30;  // bytes kept in the free lists
std::mutex g_pool_mu;
std::multimap<size_t, void *> g_free;            // size -> block
std::unordered_map<void *, size_t> g_size;       // every live or cached block of the pool -> its true size
size_t g_cached = 0;

void trim_locked() {
    for (auto &kv : g_free) {
        g_size.erase(kv.second);
        cudaFree(kv.second);
    }
    g_free.clear();
    g_cached = 0;
}
}  // namespace

cudaError_t pool_malloc(void **p, size_t bytes) {
    const size_t sz = bytes < 512 ? 512 : (bytes + 511) / 512 * 512;
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (sz <= kMaxPooled) {
        auto it = g_free.lower_bound(sz);
        if (it != g_free.end() && it->first <= sz + sz / 8) {  // at most 12.5 % larger than asked for
            *p = it->second;
            g_cached -= it->first;
            g_free.erase(it);
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMalloc(p, sz);
    if (e == cudaErrorMemoryAllocation) {  // give the cached blocks back and try once more
        cudaGetLastError();
        trim_locked();
        e = cudaMalloc(p, sz);
    }
    if (e == cudaSuccess) g_size[*p] = sz;
    return e;
}

cudaError_t pool_free(void *p) {
    if (!p) return cudaSuccess;
    std::lock_guard<std::mutex> lk(g_pool_mu);
    auto it = g_size.find(p);
    if (it == g_size.end()) return cudaFree(p);  // not ours (never happens for library allocations)
    const size_t sz = it->second;
    if (sz <= kMaxPooled && g_cached + sz <= kMaxCached) {
        g_free.emplace(sz, p);
        g_cached += sz;
        return cudaSuccess;
    }
    g_size.erase(it);
    return cudaFree(p);
}

void pool_release_all() {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    trim_locked();
}

}  // namespace nb200
