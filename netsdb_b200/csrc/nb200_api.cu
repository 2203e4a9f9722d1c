// nb200_api.cu -- the extern "C" boundary of libnetsdb_b200.so (see include/netsdb_b200.h):
// context, device block sets, block upload/download (de-blocking on the way in, re-blocking on
// the way out), the fused operators and the fused ff::inference_unit stage.
#include "nb200_internal.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <initializer_list>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace nb200 {

static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- devices --------------------------------------------------------------------------------------
// One Context per device.  netsDB runs `numThreads` pipeline threads inside ONE backend process (conf/headers/
// Configuration.h:52-53, work/source/PDBWorkerQueue.cc:163): every thread picks the device it works on (nb200_init or
// nb200_use_device, thread-local), and each device has its own context, sets, scratch, allocator pool and lock, so
// threads on different devices never wait for one another.  A thread that never chose uses the first device initialised.
constexpr int kMaxDev = 16;
struct DeviceState {
    Context *ctx = nullptr;
    std::recursive_mutex mu;   // serialises the entry points of THIS device (they compose: recursive)
    std::map<std::pair<std::string, std::string>, nb200_set *> sets;
    void *ff = nullptr, *lstm = nullptr;   // FFScratch / LstmScratch, created on first use
};
static DeviceState g_dev[kMaxDev];
static int g_default_dev = -1;
static thread_local int tls_dev = -1;
static int cur_dev() { return tls_dev >= 0 ? tls_dev : g_default_dev; }
static DeviceState &cur_state() { const int d = cur_dev(); return g_dev[d >= 0 ? d : 0]; }
#define g_sets (cur_state().sets)
static int g_opt_fused = -1;  // -1: from the environment (NB200_FUSED, default 2); 0 separate kernels, 1 ring kernel, 2 stream kernel
static int g_opt_tau_x10 = 25;  // near-tie band in units of 0.1 eps_r (nb200_set_option "near_tie_tau_x10")
static int fused_mode() {
    if (g_opt_fused >= 0) return g_opt_fused;
    static const int env = getenv("NB200_FUSED") ? atoi(getenv("NB200_FUSED")) : 2;
    return env < 0 ? 0 : (env > 2 ? 2 : env);
}
static std::mutex g_mu;   // the registry of devices and the per-device set maps
// Every entry point holds its device's lock for its whole duration: a device keeps one compute stream, one split-K workspace
// and one set of scratch buffers, so calls from several pipeline threads onto the SAME device run one at a time.
static std::atomic<uint64_t> g_version{0};  // source of set versions: unique across sets, so a cache keyed on
                                            // (pointer, version) cannot be fooled by a recycled address

Context *ctx() { const int d = cur_dev(); return d >= 0 ? g_dev[d].ctx : nullptr; }
std::recursive_mutex &api_mutex() { return cur_state().mu; }

int ensure_ws(Context *c, size_t bytes) {
    if (bytes <= c->ws_bytes) return NB200_OK;
    if (c->ws) {
        NB_CUDA(cudaStreamSynchronize(c->stream));
        nb_free(c->ws);
        c->ws = nullptr;
        c->ws_bytes = 0;
    }
    NB_CUDA(nb_malloc(&c->ws, bytes));
    c->ws_bytes = bytes;
    return NB200_OK;
}

static int ensure_stage(Context *c, size_t bytes) {
    if (bytes <= c->stage_bytes) return NB200_OK;
    if (c->stage) {
        NB_CUDA(cudaStreamSynchronize(c->stream));
        nb_free(c->stage);
        c->stage = nullptr;
        c->stage_bytes = 0;
    }
    NB_CUDA(nb_malloc(&c->stage, bytes));
    c->stage_bytes = bytes;
    return NB200_OK;
}

int wait_uploads(Context *c) {
    if (!c->uploads_pending) return NB200_OK;
    for (int i = 0; i < 2; i++) {
        NB_CUDA(cudaEventRecord(c->up_ev[i], c->copy[i]));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->up_ev[i], 0));
    }
    c->uploads_pending = false;
    c->up_epoch++;
    return NB200_OK;
}

// Order the compute stream after the uploads of ONE set.  With a commit fence (nb200_upload_commit) only that set's
// copies are waited for, so uploads queued behind them -- the next batch into another set -- keep running under the
// computation; without one this falls back to waiting for everything.
int wait_set_uploads(Context *c, nb200_set *s) {
    if (!s) return NB200_OK;
    if (s->view_of) s = s->view_of;
    if (s->up_epoch != c->up_epoch) return NB200_OK;  // nothing uploaded since the compute stream last waited
    if (!s->up_fenced) return wait_uploads(c);
    for (int i = 0; i < 2; i++) NB_CUDA(cudaStreamWaitEvent(c->stream, s->up_fence[i], 0));
    s->up_epoch = 0;
    return NB200_OK;
}

int check_kernel_fault(Context *c, const char *what) {
    NB_CUDA(cudaMemcpyAsync(c->err_flag_host, c->err_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    if (*c->err_flag_host != 0) {
        const int code = *c->err_flag_host;
        cudaMemsetAsync(c->err_flag, 0, sizeof(int), c->stream);
        set_error("%s: kernel fault %d (101 TMA producer / 102 MMA issuer / 103 epilogue timed out on a pipeline barrier)",
                  what, code);
        return NB200_ERR_KERNEL;
    }
    return NB200_OK;
}

static void free_planes(Planes *p) {
    nb_free(p->hi); nb_free(p->lo); nb_free(p->row_sumsq); nb_free(p->maxabs);
    nb_free(p->f16); nb_free(p->f16_exp); nb_free(p->f16_maxscaled);
    *p = Planes();
}

static void free_set_storage(nb200_set *s) {
    // uploads still in flight into this set must land before its storage can be handed to someone else
    if (Context *c0 = ctx()) { if (s->up_epoch == c0->up_epoch && c0->uploads_pending) wait_uploads(c0); }
    for (auto &e : s->up_fence) { if (e) cudaEventDestroy(e); e = nullptr; }
    s->up_fenced = false; s->up_epoch = 0;
    nb_free(s->f64); s->f64 = nullptr;
    nb_free(s->f32); s->f32 = nullptr;
    nb_free(s->wo_rownorm_max); s->wo_rownorm_max = nullptr;
    nb_free(s->hb_dev); s->hb_dev = nullptr; s->hb_host.clear(); s->side_format = false;
    free_planes(&s->rowp);
    free_planes(&s->colp);
    nb_free(s->img_hi); nb_free(s->img_lo); s->img_hi = s->img_lo = nullptr; s->img_n = 0;
}

#define NEED_CTX_RAW()                                          \
    std::lock_guard<std::recursive_mutex> _api_lock(::nb200::api_mutex()); \
    Context *c = ::nb200::ctx();                                \
    if (!c) {                                                   \
        ::nb200::set_error("nb200_init has not been called");   \
        return NB200_ERR_STATE;                                 \
    }                                                           \
    cudaSetDevice(c->device);   /* the calling thread may have worked on another device before */
// every entry point that reads or writes set data on the compute stream first orders itself after queued uploads
#define NEED_CTX()                                              \
    NEED_CTX_RAW()                                              \
    {                                                           \
        int _w = ::nb200::wait_uploads(c);                      \
        if (_w != NB200_OK) return _w;                          \
        c->main_dirty = true;                                   \
    }

// a set can only be used by a thread that works on the set's device
static int same_device(Context *c, std::initializer_list<const nb200_set *> sets) {
    for (const nb200_set *s : sets)
        if (s && s->device != c->device) {
            set_error("set %s:%s lives on device %d but the calling thread works on device %d (nb200_use_device)", s->db.c_str(),
                      s->name.c_str(), s->device, c->device);
            return NB200_ERR_STATE;
        }
    return NB200_OK;
}

// ---- derived representations ------------------------------------------------------------------
static int ensure_f64(Context *c, nb200_set *s) {
    if (s->f64) return NB200_OK;
    const size_t n = static_cast<size_t>(s->total_rows) * s->total_cols;
    NB_CUDA(nb_malloc(&s->f64, (n ? n : 1) * sizeof(double)));
    NB_CUDA(cudaMemsetAsync(s->f64, 0, (n ? n : 1) * sizeof(double), c->stream));
    return NB200_OK;
}
static int ensure_f32(Context *c, nb200_set *s) {
    if (s->f32) return NB200_OK;
    if (!s->f64) { set_error("set %s:%s holds no data", s->db.c_str(), s->name.c_str()); return NB200_ERR_STATE; }
    const size_t n = static_cast<size_t>(s->total_rows) * s->total_cols;
    NB_CUDA(nb_malloc(&s->f32, (n ? n : 1) * sizeof(float)));
    s->f32_ld = s->total_cols;
    return launch_f64_to_f32(c, s->f64, static_cast<int64_t>(n), s->f32);
}
static int ensure_rowp(Context *c, nb200_set *s, bool want_lo) {
    if (s->rowp.valid && (!want_lo || s->rowp.lo_valid)) return NB200_OK;
    if (s->f64) return launch_convert_planes(c, s->f64, true, s->total_rows, s->total_cols, s->total_cols, &s->rowp, want_lo);
    if (s->f32) return launch_convert_planes(c, s->f32, false, s->total_rows, s->total_cols, s->f32_ld, &s->rowp, want_lo);
    set_error("set %s:%s holds no data", s->db.c_str(), s->name.c_str());
    return NB200_ERR_STATE;
}
static int ensure_colp(Context *c, nb200_set *s, bool want_lo) {
    if (s->colp.valid && (!want_lo || s->colp.lo_valid)) return NB200_OK;
    if (s->f64) return launch_convert_planes_t(c, s->f64, true, s->total_rows, s->total_cols, s->total_cols, &s->colp, want_lo);
    if (s->f32) return launch_convert_planes_t(c, s->f32, false, s->total_rows, s->total_cols, s->f32_ld, &s->colp, want_lo);
    set_error("set %s:%s holds no data", s->db.c_str(), s->name.c_str());
    return NB200_ERR_STATE;
}

static int make_set(const char *db, const char *name, int tr, int tc, int br, int bc, nb200_set **out) {
    if (!db || !name || tr < 0 || tc < 0 || br <= 0 || bc <= 0) {
        set_error("set_create: invalid arguments (total %d x %d, block %d x %d)", tr, tc, br, bc);
        return NB200_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto key = std::make_pair(std::string(db), std::string(name));
    auto it = g_sets.find(key);
    nb200_set *s;
    if (it != g_sets.end()) {  // createSet on an existing set: replace its contents
        s = it->second;
        free_set_storage(s);
    } else {
        s = new nb200_set();
        g_sets[key] = s;
    }
    s->db = db; s->name = name;
    s->device = cur_dev();
    s->total_rows = tr; s->total_cols = tc; s->block_rows = br; s->block_cols = bc;
    s->view_exp = false; s->view_of = nullptr;
    s->version = ++g_version;
    *out = s;
    return NB200_OK;
}

static float prec_unit(int precision) {
    // unit of the relative rounding error of one product (DESIGN.md "error model")
    if (precision == NB200_PREC_BF16) return 0.00390625f;      // 2^-8
    if (precision == NB200_PREC_F16) return 0.00048828125f;    // 2^-11
    return 1.52587890625e-05f;                                 // 2^-16 (bf16x3)
}

// fp16 planes: exact scale for model-side sets (stride 1), sampled scale for per-query inputs
static int ensure_f16_row(Context *c, nb200_set *s, int sample_stride) {
    if (s->rowp.f16_valid) return NB200_OK;
    if (s->f64) return launch_convert_f16(c, s->f64, true, s->total_rows, s->total_cols, s->total_cols, &s->rowp, sample_stride, nullptr);
    if (s->f32) return launch_convert_f16(c, s->f32, false, s->total_rows, s->total_cols, s->f32_ld, &s->rowp, sample_stride, nullptr);
    set_error("set %s:%s holds no data", s->db.c_str(), s->name.c_str());
    return NB200_ERR_STATE;
}
static int ensure_f16_col(Context *c, nb200_set *s) {
    if (s->colp.f16_valid) return NB200_OK;
    NB_TRY(ensure_f32(c, s));
    float *t = nullptr;  // transposed fp32 copy, then a plain conversion
    const size_t n = static_cast<size_t>(s->total_rows) * s->total_cols;
    NB_CUDA(nb_malloc(&t, (n ? n : 1) * sizeof(float)));
    int st = launch_transpose_f32(c, s->f32, s->total_rows, s->total_cols, s->f32_ld, t, s->total_rows);
    if (st == NB200_OK) st = launch_convert_f16(c, t, false, s->total_cols, s->total_rows, s->total_rows, &s->colp, 1, nullptr);
    cudaStreamSynchronize(c->stream);
    nb_free(t);
    return st;
}

}  // namespace nb200

using namespace nb200;

extern "C" {

const char *nb200_last_error(void) { return g_err; }

int nb200_device_count(int *count) {
    if (!count) return NB200_ERR_INVALID;
    NB_CUDA(cudaGetDeviceCount(count));
    return NB200_OK;
}

int nb200_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    int n = 0;
    NB_CUDA(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n || device >= kMaxDev) {
        set_error("nb200_init: device %d out of range (%d visible); there is no CPU fallback", device, n);
        return NB200_ERR_INVALID;
    }
    if (g_dev[device].ctx) {   // idempotent; the calling thread now works on this device
        tls_dev = device;
        NB_CUDA(cudaSetDevice(device));
        return NB200_OK;
    }
    NB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    NB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("nb200_init: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
        return NB200_ERR_STATE;
    }
    Context *c = new Context();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    NB_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    NB_CUDA(nb_malloc(&c->err_flag, sizeof(int)));
    NB_CUDA(cudaMemset(c->err_flag, 0, sizeof(int)));
    NB_CUDA(cudaMallocHost(&c->err_flag_host, sizeof(int)));
    for (auto &e : c->ev) NB_CUDA(cudaEventCreate(&e));
    {   // the pipelined GEMM launches must win SM slots over the conversion they run beside
        int lo_prio = 0, hi_prio = 0;
        NB_CUDA(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
        NB_CUDA(cudaStreamCreateWithPriority(&c->aux[0], cudaStreamNonBlocking, lo_prio));
        NB_CUDA(cudaStreamCreateWithPriority(&c->aux[1], cudaStreamNonBlocking, hi_prio));
        NB_CUDA(cudaStreamCreateWithPriority(&c->aux[2], cudaStreamNonBlocking, hi_prio));
    }
    for (auto &e : c->chunk_ev) NB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : c->join_ev) NB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &st : c->copy) NB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto &e : c->up_ev) NB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    NB_CUDA(cudaEventCreateWithFlags(&c->main_ev, cudaEventDisableTiming));
    g_dev[device].ctx = c;
    if (g_default_dev < 0) g_default_dev = device;
    tls_dev = device;
    return NB200_OK;
}

int nb200_init_devices(const int *devices, int n) {
    if (!devices || n <= 0) { set_error("nb200_init_devices: invalid arguments"); return NB200_ERR_INVALID; }
    for (int i = 0; i < n; i++) NB_TRY(nb200_init(devices[i]));
    return nb200_use_device(devices[0]);
}

int nb200_use_device(int device) {
    if (device < 0 || device >= kMaxDev || !g_dev[device].ctx) {
        set_error("nb200_use_device: device %d has not been initialised (nb200_init / nb200_init_devices)", device);
        return NB200_ERR_STATE;
    }
    tls_dev = device;
    NB_CUDA(cudaSetDevice(device));
    return NB200_OK;
}

int nb200_current_device(int *device) {
    if (!device) return NB200_ERR_INVALID;
    *device = cur_dev();
    return NB200_OK;
}

static void release_lstm_scratch(void);
static void release_ff_scratch(void);

static void shutdown_device(int d) {
    DeviceState &D = g_dev[d];
    Context *c = D.ctx;
    if (!c) return;
    std::lock_guard<std::recursive_mutex> api_lock(D.mu);
    const int saved = tls_dev;
    tls_dev = d;                       // the scratch / set accessors below address THIS device
    cudaSetDevice(d);
    cudaStreamSynchronize(c->stream);
    release_lstm_scratch();
    release_ff_scratch();
    for (auto &kv : D.sets) {
        free_set_storage(kv.second);
        delete kv.second;
    }
    D.sets.clear();
    nb_free(c->ws);
    nb_free(c->stage);
    nb_free(c->err_flag);
    cudaFreeHost(c->err_flag_host);
    for (auto &e : c->ev) cudaEventDestroy(e);
    for (auto &e : c->chunk_ev) cudaEventDestroy(e);
    for (auto &e : c->join_ev) cudaEventDestroy(e);
    for (auto &st : c->aux) cudaStreamDestroy(st);
    for (auto &st : c->copy) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    for (auto &e : c->up_ev) cudaEventDestroy(e);
    cudaEventDestroy(c->main_ev);
    nb_free(c->up_stage[0]); nb_free(c->up_stage[1]);
    cudaStreamDestroy(c->own_stream);
    delete c;
    D.ctx = nullptr;
    pool_release_all();                // this device's cached blocks
    tls_dev = saved;
}

int nb200_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int d = 0; d < kMaxDev; d++) shutdown_device(d);
    g_default_dev = -1;
    tls_dev = -1;
    return NB200_OK;
}

int nb200_set_stream(void *cuda_stream) {
    NEED_CTX();
    NB_CUDA(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : c->own_stream;
    return NB200_OK;
}

int nb200_synchronize(void) {
    NEED_CTX();
    NB_CUDA(cudaStreamSynchronize(c->stream));  // NEED_CTX made it wait for the copy streams
    return NB200_OK;
}

int64_t nb200_launch_count(void) { Context *c = ctx(); return c ? c->launches : 0; }

int nb200_set_option(const char *key, int value) {
    if (key && !strcmp(key, "fused_layer1")) { g_opt_fused = value < 0 ? -1 : (value > 2 ? 2 : value); return NB200_OK; }
    if (key && !strcmp(key, "near_tie_tau_x10")) { g_opt_tau_x10 = value < 5 ? 5 : (value > 160 ? 160 : value); return NB200_OK; }
    if (key && !strcmp(key, "l1_stream_mt")) { g_l1_stream_mt = (value == 1 || value == 2) ? value : 0; return NB200_OK; }
    set_error("set_option: unknown key");
    return NB200_ERR_INVALID;
}

int nb200_pin_pool(void *base, size_t bytes) {
    NEED_CTX();
    (void)c;
    NB_CUDA(cudaHostRegister(base, bytes, cudaHostRegisterPortable));
    return NB200_OK;
}
int nb200_unpin_pool(void *base) {
    NEED_CTX();
    NB_CUDA(cudaStreamSynchronize(c->stream));
    NB_CUDA(cudaHostUnregister(base));
    return NB200_OK;
}

// ---- sets ---------------------------------------------------------------------------------------
int nb200_set_create(const char *db, const char *set, int total_rows, int total_cols, int block_rows, int block_cols,
                     nb200_set **out) {
    NEED_CTX();
    (void)c;
    if (!out) return NB200_ERR_INVALID;
    return make_set(db, set, total_rows, total_cols, block_rows, block_cols, out);
}

int nb200_set_lookup(const char *db, const char *set, nb200_set **out) {
    if (!db || !set || !out) return NB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(std::make_pair(std::string(db), std::string(set)));
    if (it == g_sets.end()) {
        set_error("set %s:%s does not exist", db, set);
        return NB200_ERR_STATE;
    }
    *out = it->second;
    return NB200_OK;
}

int nb200_set_destroy(nb200_set *s) {
    NEED_CTX();
    if (!s) return NB200_ERR_INVALID;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto &kv : g_sets)
        if (kv.second->view_of == s) kv.second->view_of = nullptr;
    g_sets.erase(std::make_pair(s->db, s->name));
    free_set_storage(s);
    delete s;
    return NB200_OK;
}

int nb200_set_info(const nb200_set *s, int *total_rows, int *total_cols, int *block_rows, int *block_cols) {
    if (!s) return NB200_ERR_INVALID;
    if (total_rows) *total_rows = s->total_rows;
    if (total_cols) *total_cols = s->total_cols;
    if (block_rows) *block_rows = s->block_rows;
    if (block_cols) *block_cols = s->block_cols;
    return NB200_OK;
}

static void drop_derived(nb200_set *s) {
    s->version = ++g_version;
    s->rowp.valid = s->rowp.lo_valid = s->rowp.f16_valid = false;
    s->colp.valid = s->colp.lo_valid = s->colp.f16_valid = false;
    if (s->f64 && s->f32) { nb_free(s->f32); s->f32 = nullptr; }
}

int nb200_upload_rect(nb200_set *s, int row0, int col0, const double *host, int rows, int cols, int64_t ld) {
    NEED_CTX_RAW();
    if (!s || !host || row0 < 0 || col0 < 0 || rows < 0 || cols < 0 || ld < cols) {
        set_error("upload: invalid arguments");
        return NB200_ERR_INVALID;
    }
    if (s->view_of) { set_error("upload: %s:%s is a read-only view", s->db.c_str(), s->name.c_str()); return NB200_ERR_STATE; }
    NB_TRY(same_device(c, {s}));
    if (!s->f64) {  // first block of the set: allocate + zero on the compute stream
        NB_TRY(ensure_f64(c, s));
        c->main_dirty = true;
    }
    const int vr = std::min(rows, s->total_rows - row0), vc = std::min(cols, s->total_cols - col0);
    drop_derived(s);
    if (vr <= 0 || vc <= 0) return NB200_OK;  // a block that is all padding
    if (c->main_dirty) {  // the copy streams must not overtake compute-stream work that touches the set
        NB_CUDA(cudaEventRecord(c->main_ev, c->stream));
        for (auto &st : c->copy) NB_CUDA(cudaStreamWaitEvent(st, c->main_ev, 0));
        c->main_dirty = false;
    }
    const int slot = c->up_next;
    c->up_next ^= 1;
    cudaStream_t cs = c->copy[slot];
    double *dst = s->f64 + static_cast<size_t>(row0) * s->total_cols + col0;
    const size_t dpitch = static_cast<size_t>(s->total_cols) * sizeof(double);
    // (the rectangle's last row only guarantees vc elements: never read the ld - vc doubles after it)
    const size_t block_bytes = (static_cast<size_t>(vr - 1) * ld + vc) * sizeof(double);
    if (block_bytes >= (1u << 20) && dpitch != static_cast<size_t>(ld) * sizeof(double)) {
        // contiguous H2D of the host rows into a staging slot, then de-block on the device
        if (c->up_bytes[slot] < block_bytes) {
            NB_CUDA(cudaStreamSynchronize(cs));
            nb_free(c->up_stage[slot]);
            c->up_stage[slot] = nullptr;
            c->up_bytes[slot] = 0;
            NB_CUDA(nb_malloc(&c->up_stage[slot], block_bytes));
            c->up_bytes[slot] = block_bytes;
        }
        NB_CUDA(cudaMemcpyAsync(c->up_stage[slot], host, block_bytes, cudaMemcpyHostToDevice, cs));
        NB_CUDA(cudaMemcpy2DAsync(dst, dpitch, c->up_stage[slot], static_cast<size_t>(ld) * sizeof(double),
                                  static_cast<size_t>(vc) * sizeof(double), vr, cudaMemcpyDeviceToDevice, cs));
    } else {
        NB_CUDA(cudaMemcpy2DAsync(dst, dpitch, host, static_cast<size_t>(ld) * sizeof(double),
                                  static_cast<size_t>(vc) * sizeof(double), vr, cudaMemcpyHostToDevice, cs));
    }
    c->uploads_pending = true;
    s->up_epoch = c->up_epoch;
    s->up_fenced = false;
    return NB200_OK;
}

int nb200_upload_commit(nb200_set *s) {
    NEED_CTX_RAW();
    if (!s || s->view_of) { set_error("upload_commit: invalid set"); return NB200_ERR_INVALID; }
    for (int i = 0; i < 2; i++) {
        if (!s->up_fence[i]) NB_CUDA(cudaEventCreateWithFlags(&s->up_fence[i], cudaEventDisableTiming));
        NB_CUDA(cudaEventRecord(s->up_fence[i], c->copy[i]));
    }
    s->up_fenced = true;
    return NB200_OK;
}

// ---- fp16 side format ---------------------------------------------------------------------------
int nb200_set_export_f16(nb200_set *s, void *plane_host, int64_t ldk, int *exp_out, float *row_sumsq_host) {
    NEED_CTX();
    if (!s || !plane_host || !exp_out || !row_sumsq_host || s->view_of) { set_error("set_export_f16: invalid arguments"); return NB200_ERR_INVALID; }
    if (ldk != round_up(s->total_cols, 64)) {
        set_error("set_export_f16: ldk must be %lld (columns rounded up to a multiple of 64)", (long long)round_up(s->total_cols, 64));
        return NB200_ERR_INVALID;
    }
    s->rowp.f16_valid = false;            // always from the exact maximum: this is the persistent copy
    NB_TRY(ensure_f16_row(c, s, 1));
    const size_t rows = static_cast<size_t>(s->total_rows);
    NB_CUDA(cudaMemcpyAsync(plane_host, s->rowp.f16, rows * ldk * sizeof(__half), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaMemcpyAsync(exp_out, s->rowp.f16_exp, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaMemcpyAsync(row_sumsq_host, s->rowp.row_sumsq, rows * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB200_OK;
}

int nb200_set_import_f16(nb200_set *s, const void *plane_host, int64_t ldk, int exp, const float *row_sumsq_host,
                         const double *const *host_blocks, int edge_ld) {
    NEED_CTX_RAW();
    if (!s || !plane_host || !row_sumsq_host || !host_blocks || s->view_of || edge_ld <= 0 || edge_ld > s->block_cols) {
        set_error("set_import_f16: invalid arguments");
        return NB200_ERR_INVALID;
    }
    if (ldk != round_up(s->total_cols, 64)) {
        set_error("set_import_f16: ldk must be %lld (columns rounded up to a multiple of 64)", (long long)round_up(s->total_cols, 64));
        return NB200_ERR_INVALID;
    }
    const int n_rb = (s->total_rows + s->block_rows - 1) / s->block_rows, n_cb = (s->total_cols + s->block_cols - 1) / s->block_cols;
    // the cascade's kernels dereference these pointers: they must be pinned host memory the device can see
    for (int i = 0; i < n_rb * n_cb; i += (n_rb * n_cb > 1 ? n_rb * n_cb - 1 : 1)) {
        cudaPointerAttributes at{};
        if (!host_blocks[i] || cudaPointerGetAttributes(&at, host_blocks[i]) != cudaSuccess || at.type != cudaMemoryTypeHost ||
            at.devicePointer != static_cast<const void *>(host_blocks[i])) {
            cudaGetLastError();
            set_error("set_import_f16: the fp64 host blocks must lie in pinned, device-visible memory (nb200_pin_pool / cudaHostAlloc)");
            return NB200_ERR_INVALID;
        }
    }
    if (s->f64) {  // the set now lives on the device in its side format only
        if (s->up_epoch == c->up_epoch && c->uploads_pending) NB_TRY(wait_uploads(c));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        nb_free(s->f64); s->f64 = nullptr;
    }
    if (s->f32) { nb_free(s->f32); s->f32 = nullptr; }
    drop_derived(s);
    NB_TRY(alloc_f16_plane(c, &s->rowp, s->total_rows, s->total_cols));
    if (!s->hb_dev) NB_CUDA(nb_malloc(&s->hb_dev, static_cast<size_t>(n_rb) * n_cb * sizeof(double *)));
    s->hb_host.assign(host_blocks, host_blocks + static_cast<size_t>(n_rb) * n_cb);
    s->hb_edge_ld = edge_ld;
    s->hb_exp = exp;
    c->main_dirty = true;  // (the plane may have been handed out by the allocator a moment ago)
    NB_CUDA(cudaEventRecord(c->main_ev, c->stream));
    for (auto &st : c->copy) NB_CUDA(cudaStreamWaitEvent(st, c->main_ev, 0));
    c->main_dirty = false;
    // the plane in pieces of 256 MB, alternating between the two upload streams like block uploads do
    const size_t total = static_cast<size_t>(s->total_rows) * ldk * sizeof(__half), piece = size_t(256) << 20;
    for (size_t off = 0; off < total; off += piece) {
        const size_t n = total - off < piece ? total - off : piece;
        NB_CUDA(cudaMemcpyAsync(reinterpret_cast<char *>(s->rowp.f16) + off, static_cast<const char *>(plane_host) + off, n,
                                cudaMemcpyHostToDevice, c->copy[c->up_next]));
        c->up_next ^= 1;
    }
    cudaStream_t cs = c->copy[0];
    NB_CUDA(cudaMemcpyAsync(s->rowp.row_sumsq, row_sumsq_host, static_cast<size_t>(s->total_rows) * sizeof(float), cudaMemcpyHostToDevice, cs));
    NB_CUDA(cudaMemcpyAsync(s->rowp.f16_exp, &s->hb_exp, sizeof(int), cudaMemcpyHostToDevice, cs));
    NB_CUDA(cudaMemsetAsync(s->rowp.f16_maxscaled, 0, sizeof(float), cs));
    NB_CUDA(cudaMemcpyAsync(s->hb_dev, s->hb_host.data(), s->hb_host.size() * sizeof(double *), cudaMemcpyHostToDevice, cs));
    s->rowp.f16_valid = true;
    s->side_format = true;
    c->uploads_pending = true;
    s->up_epoch = c->up_epoch;
    s->up_fenced = false;
    return NB200_OK;
}

int nb200_upload_block(nb200_set *s, int block_row, int block_col, const double *host, int rows, int cols) {
    if (!s || block_row < 0 || block_col < 0) {
        set_error("upload_block: invalid arguments");
        return NB200_ERR_INVALID;
    }
    return nb200_upload_rect(s, block_row * s->block_rows, block_col * s->block_cols, host, rows, cols, cols);
}

int nb200_download_rect(nb200_set *s, int row0, int col0, double *host, int rows, int cols, int64_t ld) {
    NEED_CTX_RAW();
    NB_TRY(same_device(c, {s}));
    NB_TRY(wait_set_uploads(c, s));  // only this set's uploads: a download of results must not wait for the next batch's
    c->main_dirty = true;
    if (!s || !host || row0 < 0 || col0 < 0 || rows < 0 || cols < 0 || ld < cols) {
        set_error("download: invalid arguments");
        return NB200_ERR_INVALID;
    }
    nb200_set *src = s->view_of ? s->view_of : s;
    const int vr = std::max(0, std::min(rows, s->total_rows - row0)), vc = std::max(0, std::min(cols, s->total_cols - col0));
    if (rows == 0 || cols == 0) return NB200_OK;
    if (src->f64 && !s->view_exp) {
        if (vr < rows || vc < cols)
            for (int r = 0; r < rows; r++) memset(host + static_cast<size_t>(r) * ld, 0, static_cast<size_t>(cols) * sizeof(double));
        if (vr > 0 && vc > 0)
            NB_CUDA(cudaMemcpy2DAsync(host, static_cast<size_t>(ld) * sizeof(double),
                                      src->f64 + static_cast<size_t>(row0) * src->total_cols + col0,
                                      static_cast<size_t>(src->total_cols) * sizeof(double),
                                      static_cast<size_t>(vc) * sizeof(double), vr, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        return NB200_OK;
    }
    if (!src->f32 && !src->f64) { set_error("download: set %s:%s holds no data", s->db.c_str(), s->name.c_str()); return NB200_ERR_STATE; }
    NB_TRY(ensure_stage(c, static_cast<size_t>(rows) * cols * sizeof(double)));
    if (src->f64) NB_TRY(launch_extract_to_f64(c, src->f64, true, src->total_cols, row0, col0, rows, cols, vr, vc, c->stage, cols, s->view_exp ? 1 : 0));
    else NB_TRY(launch_extract_to_f64(c, src->f32, false, src->f32_ld, row0, col0, rows, cols, vr, vc, c->stage, cols, s->view_exp ? 1 : 0));
    NB_CUDA(cudaMemcpy2DAsync(host, static_cast<size_t>(ld) * sizeof(double), c->stage, static_cast<size_t>(cols) * sizeof(double),
                              static_cast<size_t>(cols) * sizeof(double), rows, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB200_OK;
}

int nb200_download_block(nb200_set *s, int block_row, int block_col, double *host, int rows, int cols) {
    if (!s || block_row < 0 || block_col < 0) {
        set_error("download_block: invalid arguments");
        return NB200_ERR_INVALID;
    }
    return nb200_download_rect(s, block_row * s->block_rows, block_col * s->block_cols, host, rows, cols, cols);
}

int nb200_set_fill_synth(nb200_set *s, uint64_t seed, int64_t row_offset) {
    NEED_CTX();
    if (!s) return NB200_ERR_INVALID;
    NB_TRY(ensure_f64(c, s));
    drop_derived(s);
    return launch_synth_fill(c, s->f64, s->total_rows, s->total_cols, s->total_cols, seed, row_offset, s->total_cols);
}

int nb200_set_exp_view(nb200_set *src, const char *db, const char *set, nb200_set **out) {
    NEED_CTX();
    (void)c;
    if (!src || !db || !set || !out) return NB200_ERR_INVALID;
    nb200_set *base = src->view_of ? src->view_of : src;
    nb200_set *v = nullptr;
    NB_TRY(make_set(db, set, base->total_rows, base->total_cols, base->block_rows, base->block_cols, &v));
    v->view_of = base;
    v->view_exp = true;
    *out = v;
    return NB200_OK;
}

int nb200_set_rename(nb200_set *s, const char *db, const char *set) {
    NEED_CTX();
    if (!s || !db || !set) return NB200_ERR_INVALID;
    if (s->db == db && s->name == set) return NB200_OK;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    std::lock_guard<std::mutex> lk(g_mu);
    auto key = std::make_pair(std::string(db), std::string(set));
    auto it = g_sets.find(key);
    if (it != g_sets.end() && it->second != s) {
        nb200_set *old = it->second;
        for (auto &kv : g_sets)
            if (kv.second->view_of == old) kv.second->view_of = nullptr;
        free_set_storage(old);
        delete old;
        g_sets.erase(it);
    }
    g_sets.erase(std::make_pair(s->db, s->name));
    s->db = db;
    s->name = set;
    g_sets[key] = s;
    return NB200_OK;
}

int nb200_set_maxabs(nb200_set *s, float *out) {
    NEED_CTX();
    if (!s || !out) return NB200_ERR_INVALID;
    if (!s->rowp.valid && !s->rowp.f16_valid && !s->colp.valid) NB_TRY(ensure_rowp(c, s, false));
    const Planes &p = (s->rowp.valid || s->rowp.f16_valid) ? s->rowp : s->colp;
    NB_CUDA(cudaMemcpyAsync(out, p.maxabs, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB200_OK;
}

int nb200_set_invalidate_planes(nb200_set *s) {
    if (!s) return NB200_ERR_INVALID;
    s->rowp.valid = s->rowp.lo_valid = s->rowp.f16_valid = false;
    s->colp.valid = s->colp.lo_valid = s->colp.f16_valid = false;
    return NB200_OK;
}

// ---- operators ------------------------------------------------------------------------------------
static int bias_check(nb200_set *bias, int rows, const char *who) {
    // FFReluBiasSum.h:49-56 / FFTransposeBiasSum.h:56-60: in2 must be (rows of in1) x 1
    if (!bias || bias->total_cols != 1 || bias->total_rows != rows) {
        set_error("[%s] Block dimemsions mismatch! bias must be %d x 1", who, rows);
        return NB200_ERR_SHAPE;
    }
    return NB200_OK;
}

static int gemm_blocks_impl(int op, nb200_set *a, nb200_set *b, int epilogue, nb200_set *bias, int precision,
                            const char *db, const char *out_set, nb200_set **out, int extra_relu) {
    NEED_CTX();
    if (!a || !b || !db || !out_set || !out || (op != NB200_OP_NT && op != NB200_OP_NN && op != NB200_OP_TN) || (op == NB200_OP_TN && epilogue != NB200_EPI_NONE) || epilogue < 0 || epilogue > 2 ||
        precision < 0 || precision > 3) {
        set_error("gemm_blocks: invalid arguments");
        return NB200_ERR_INVALID;
    }
    NB_TRY(same_device(c, {a, b, bias}));
    // join predicate + per-pair dimension check (FFTransposeMult.h:38-39,61-65; FFInputLayerJoin.h:30-31,57-60)
    // (LASillyTransposeMultiply1Join.h:24-33,44-47: there the rows of both operands are the contraction index)
    const bool a_t = op == NB200_OP_TN;  // A enters transposed
    const int Ka = a_t ? a->total_rows : a->total_cols;
    const int Kb = (op == NB200_OP_NT) ? b->total_cols : b->total_rows;
    const int bkA = a_t ? a->block_rows : a->block_cols, bkB = (op == NB200_OP_NT) ? b->block_cols : b->block_rows;
    if (bkA != bkB) {  // the per-pair check of the projection lambda
        set_error("Block dimemsions mismatch! (block K %d vs %d)", bkA, bkB);
        return NB200_ERR_SHAPE;
    }
    // Total K extents may differ: the join simply finds no partner for the surplus column blocks of the longer
    // operand (src/tests/source/LSTMTest.cc:470-480 multiplies a 128 x 500 set with a 128 x 500 set this way), and
    // inside the last matched block the shorter operand's zero padding cancels the rest: the contraction runs over
    // k < min(Ka, Kb).
    const int K = Ka < Kb ? Ka : Kb;
    const int I = a_t ? a->total_cols : a->total_rows, J = (op == NB200_OP_NT) ? b->total_rows : b->total_cols;
    const int bI = a_t ? a->block_cols : a->block_rows, bJ = (op == NB200_OP_NT) ? b->block_rows : b->block_cols;
    if (epilogue != NB200_EPI_NONE) NB_TRY(bias_check(bias, I, epilogue == NB200_EPI_BIAS_RELU ? "ReluBiasSum" : "TransposeBiasSum"));
    const bool transpose_out = (epilogue == NB200_EPI_BIAS_TRANSPOSE);
    // output meta: FFTransposeMult.h:68-71 / FFInputLayerJoin.h:64-67 / FFTransposeBiasSum.h:64-67
    nb200_set *o = nullptr;
    NB_TRY(make_set(db, out_set, transpose_out ? J : I, transpose_out ? I : J, transpose_out ? bJ : bI,
                    transpose_out ? bI : bJ, &o));
    const int relu = (epilogue == NB200_EPI_BIAS_RELU) || extra_relu;
    const size_t on = static_cast<size_t>(o->total_rows) * o->total_cols;

    if (precision == NB200_PREC_FP64) {
        if (!a->f64 || !b->f64 || (bias && !bias->f64)) { set_error("gemm_blocks(FP64): inputs hold no fp64 data"); return NB200_ERR_STATE; }
        NB_CUDA(nb_malloc(&o->f64, (on ? on : 1) * sizeof(double)));
        F64Epilogue e;
        e.relu = relu;
        // element strides of A(i,k) and B'(j,k)
        const int64_t a_sm = a_t ? 1 : a->total_cols, a_sk = a_t ? a->total_cols : 1;
        const int64_t b_sn = (op == NB200_OP_NT) ? b->total_cols : 1, b_sk = (op == NB200_OP_NT) ? 1 : b->total_cols;
        if (!transpose_out) {
            e.bias_m = bias ? bias->f64 : nullptr;
            NB_TRY(launch_gemm_f64(c, a->f64, a_sm, a_sk, b->f64, b_sn, b_sk, I, J, K, o->f64, o->total_cols, e));
        } else {
            e.bias_n = bias->f64;
            NB_TRY(launch_gemm_f64(c, b->f64, b_sn, b_sk, a->f64, a_sm, a_sk, J, I, K, o->f64, o->total_cols, e));
        }
        *out = o;
        return NB200_OK;
    }

    const int nterms = (precision == NB200_PREC_BF16X3) ? 3 : 1;
    const bool lo = nterms == 3;
    const int fmt = (precision == NB200_PREC_F16) ? kFmtF16 : kFmtBF16;
    if (fmt == kFmtF16) {
        if (a_t) NB_TRY(ensure_f16_col(c, a));
        else NB_TRY(ensure_f16_row(c, a, 1));
        if (op == NB200_OP_NT) NB_TRY(ensure_f16_row(c, b, 1));
        else NB_TRY(ensure_f16_col(c, b));
    } else {
        if (a_t) NB_TRY(ensure_colp(c, a, lo));
        else NB_TRY(ensure_rowp(c, a, lo));
        if (op == NB200_OP_NT) NB_TRY(ensure_rowp(c, b, lo));
        else NB_TRY(ensure_colp(c, b, lo));
    }
    const Planes &pa = a_t ? a->colp : a->rowp;
    const Planes &pb = (op == NB200_OP_NT) ? b->rowp : b->colp;
    const float *bias32 = nullptr;
    if (bias) {
        NB_TRY(ensure_f32(c, bias));
        bias32 = bias->f32;
    }
    NB_CUDA(nb_malloc(&o->f32, (on ? on : 1) * sizeof(float)));
    o->f32_ld = o->total_cols;
    // P is always the logical output: roles are swapped for the transposing epilogue
    const Planes &U = transpose_out ? pb : pa;
    const Planes &V = transpose_out ? pa : pb;
    const int M = o->total_rows, N = o->total_cols;
    if (M == 0 || N == 0) { *out = o; return NB200_OK; }
    const int splits = choose_splits(c, M, N, U.ldk < V.ldk ? U.ldk : V.ldk, nterms);
    int Mpad = 0, Npad = 0;
    if (splits == 1) {
        GemmEpilogue ge;
        ge.mode = 1; ge.out = o->f32; ge.out_ld = o->f32_ld; ge.relu = relu;
        if (transpose_out) ge.bias_n = bias32; else ge.bias_m = bias32;
        NB_TRY(launch_gemm_tc(c, U, V, M, N, nterms, 1, ge, &Mpad, &Npad, nullptr, fmt));
    } else {
        GemmEpilogue ge;
        ge.mode = 0;
        int real_splits = splits;
        NB_TRY(launch_gemm_tc(c, U, V, M, N, nterms, splits, ge, &Mpad, &Npad, &real_splits, fmt));
        ReduceOut ro;
        ro.f32 = o->f32; ro.f32_ld = o->f32_ld;
        if (fmt == kFmtF16) { ro.exp_u = U.f16_exp; ro.exp_v = V.f16_exp; }
        NB_TRY(launch_splitk_reduce(c, M, N, real_splits, Mpad, Npad, transpose_out ? nullptr : bias32,
                                    transpose_out ? bias32 : nullptr, relu, ro));
    }
    NB_TRY(check_kernel_fault(c, "gemm_blocks"));
    *out = o;
    return NB200_OK;
}

int nb200_gemm_blocks(int op, nb200_set *a, nb200_set *b, int epilogue, nb200_set *bias, int precision, const char *db,
                      const char *out_set, nb200_set **out) {
    return gemm_blocks_impl(op, a, b, epilogue, bias, precision, db, out_set, out, 0);
}

int nb200_sum_blocks(nb200_set *a, nb200_set *b, nb200_set *cc, int activation, const char *db, const char *out_set,
                     nb200_set **out) {
    NEED_CTX();
    if (!a || (!b && cc) || !db || !out_set || !out || activation < NB200_ACT_NONE || activation > NB200_ACT_TANH) {
        set_error("sum_blocks: invalid arguments");
        return NB200_ERR_INVALID;
    }
    nb200_set *in[3] = {a, b, cc};
    EwiseSrc src[3];
    int nsrc = 0;
    for (int i = 0; i < 3 && in[i]; i++, nsrc++) {
        nb200_set *s = in[i];
        if (s->view_of) { set_error("sum_blocks: %s:%s is a view", s->db.c_str(), s->name.c_str()); return NB200_ERR_INVALID; }
        if (s->total_rows != a->total_rows || s->total_cols != a->total_cols || s->block_rows != a->block_rows ||
            s->block_cols != a->block_cols) {
            set_error("Block dimemsions mismatch! (%d x %d in blocks of %d x %d vs %d x %d in blocks of %d x %d)", a->total_rows,
                      a->total_cols, a->block_rows, a->block_cols, s->total_rows, s->total_cols, s->block_rows, s->block_cols);
            return NB200_ERR_SHAPE;
        }
        if (s->f64) src[i] = EwiseSrc{s->f64, nullptr, s->total_cols};
        else if (s->f32) src[i] = EwiseSrc{nullptr, s->f32, s->f32_ld};
        else { set_error("sum_blocks: set %s:%s holds no data", s->db.c_str(), s->name.c_str()); return NB200_ERR_STATE; }
    }
    // the output may replace one of the inputs (same db:name): compute into fresh storage, then swap it in
    const size_t n = static_cast<size_t>(a->total_rows) * a->total_cols;
    double *res = nullptr;
    NB_CUDA(nb_malloc(&res, (n ? n : 1) * sizeof(double)));
    const int tr = a->total_rows, tc = a->total_cols, br = a->block_rows, bc = a->block_cols;
    int st = launch_sum_blocks(c, src, nsrc, activation, tr, tc, res);
    if (st != NB200_OK) { nb_free(res); return st; }
    NB_CUDA(cudaStreamSynchronize(c->stream));  // make_set below may free an input that is being replaced
    nb200_set *o = nullptr;
    st = make_set(db, out_set, tr, tc, br, bc, &o);
    if (st != NB200_OK) { nb_free(res); return st; }
    o->f64 = res;
    *out = o;
    return NB200_OK;
}

// ---- the networks that live inside one SelectionComp ------------------------------------------------
// FullyConnectedNetwork (FF_proj/headers/FullyConnectedNetwork.h:62-123, driver tests/source/FCProjTest.cc:88-97) and
// SemanticClassifier (word2vec/headers/SemanticClassifier.h:43-141) map every FFMatrixBlock of a set through a two-layer
// network held by the computation object: x0 = the block's raw buffer viewed as features x batch,
// y0 = relu(w0 x0 + b0), out = outLabel(sigmoid(w1 y0 + b1)), a labels x batch block with the input's block indices.
// Here all blocks go through two fp64 tensor-core GEMMs at once (the blocks side by side along the batch axis):
// fp64 because the result is a thresholded sign and the layers are tiny (word2vec: 500 -> 16 -> 1).
int nb200_udf_dense2(nb200_set *in, int features, int batch, int hidden, int labels, const double *w0_host, const double *b0_host,
                     const double *w1_host, const double *b1_host, const char *db, const char *out_set, nb200_set **out) {
    NEED_CTX();
    if (!in || features <= 0 || batch <= 0 || hidden <= 0 || labels <= 0 || !w0_host || !b0_host || !w1_host || !b1_host || !db || !out_set || !out) {
        set_error("udf_dense2: invalid arguments");
        return NB200_ERR_INVALID;
    }
    NB_TRY(same_device(c, {in}));
    if (in->view_of || (!in->f64 && !in->f32)) { set_error("udf_dense2: the input set holds no data (or is a view)"); return NB200_ERR_STATE; }
    if (static_cast<int64_t>(in->block_rows) * in->block_cols < static_cast<int64_t>(features) * batch) {
        set_error("udf_dense2: a %d x %d block holds fewer than features x batch = %d x %d values", in->block_rows, in->block_cols, features, batch);
        return NB200_ERR_SHAPE;
    }
    NB_TRY(wait_set_uploads(c, in));
    const int nbr = static_cast<int>(ceil_div(in->total_rows, in->block_rows)), nbc = static_cast<int>(ceil_div(in->total_cols, in->block_cols));
    const int64_t M = static_cast<int64_t>(nbr) * nbc * batch;
    if (M > (1 << 30)) { set_error("udf_dense2: too many rows"); return NB200_ERR_SHAPE; }
    double *xt = nullptr, *y0 = nullptr, *z = nullptr, *w0 = nullptr, *w1 = nullptr, *b0 = nullptr, *b1 = nullptr, *res = nullptr;
    const size_t on = static_cast<size_t>(nbr) * labels * nbc * batch;
    int st = NB200_OK;
    auto fail = [&](int code) { nb_free(xt); nb_free(y0); nb_free(z); nb_free(w0); nb_free(w1); nb_free(b0); nb_free(b1); nb_free(res); return code; };
    if (nb_malloc(&xt, (M ? M : 1) * features * sizeof(double)) != cudaSuccess || nb_malloc(&y0, (M ? M : 1) * hidden * sizeof(double)) != cudaSuccess ||
        nb_malloc(&z, (M ? M : 1) * labels * sizeof(double)) != cudaSuccess || nb_malloc(&w0, static_cast<size_t>(hidden) * features * sizeof(double)) != cudaSuccess ||
        nb_malloc(&w1, static_cast<size_t>(labels) * hidden * sizeof(double)) != cudaSuccess || nb_malloc(&b0, hidden * sizeof(double)) != cudaSuccess ||
        nb_malloc(&b1, labels * sizeof(double)) != cudaSuccess || nb_malloc(&res, (on ? on : 1) * sizeof(double)) != cudaSuccess) {
        set_error("udf_dense2: out of device memory");
        return fail(NB200_ERR_NOMEM);
    }
    if (cudaMemcpyAsync(w0, w0_host, static_cast<size_t>(hidden) * features * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(b0, b0_host, hidden * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(w1, w1_host, static_cast<size_t>(labels) * hidden * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(b1, b1_host, labels * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
        set_error("udf_dense2: weight upload failed");
        return fail(NB200_ERR_CUDA);
    }
    st = launch_udf_gather(c, in->f64 ? static_cast<const void *>(in->f64) : in->f32, in->f64 != nullptr, in->f64 ? in->total_cols : in->f32_ld, in->total_rows, in->total_cols, in->block_rows, in->block_cols, nbr, nbc, features, batch, xt);
    F64Epilogue e0; e0.bias_n = b0; e0.relu = 1;
    if (st == NB200_OK) st = launch_gemm_f64(c, xt, features, 1, w0, features, 1, static_cast<int>(M), hidden, features, y0, hidden, e0);
    F64Epilogue e1; e1.bias_n = b1;
    if (st == NB200_OK) st = launch_gemm_f64(c, y0, hidden, 1, w1, hidden, 1, static_cast<int>(M), labels, hidden, z, labels, e1);
    if (st == NB200_OK) st = launch_udf_labels(c, z, labels, batch, nbr, nbc, res, static_cast<int64_t>(nbc) * batch);
    if (st == NB200_OK) st = check_kernel_fault(c, "udf_dense2");
    if (cudaStreamSynchronize(c->stream) != cudaSuccess && st == NB200_OK) { set_error("udf_dense2: stream synchronisation failed"); st = NB200_ERR_CUDA; }
    if (st != NB200_OK) return fail(st);
    nb200_set *o = nullptr;
    st = make_set(db, out_set, nbr * labels, nbc * batch, labels, batch, &o);   // may replace (and free) the input
    if (st != NB200_OK) return fail(st);
    o->f64 = res;
    res = nullptr;
    fail(NB200_OK);
    *out = o;
    return NB200_OK;
}

// ---- LSTM step -----------------------------------------------------------------------------------
namespace {
struct LstmScratch {
    nb200_set *w[4] = {}, *u[4] = {};
    uint64_t ver[8] = {};
    bool wu_lo = false, wu_valid = false;
    Planes WU, XH, FI, CC, OP, TH;
    float *G = nullptr, *C32 = nullptr;
    double *T64 = nullptr;
    size_t G_cap = 0, C_cap = 0, T_cap = 0;
};
static LstmScratch &lstm_ref() { DeviceState &D = cur_state(); if (!D.lstm) D.lstm = new LstmScratch(); return *static_cast<LstmScratch *>(D.lstm); }
#define g_lstm (lstm_ref())

int grow_bytes(void **p, size_t *have, size_t need) {
    if (*have >= need) return NB200_OK;
    nb_free(*p);
    *p = nullptr; *have = 0;
    NB_CUDA(nb_malloc(p, need ? need : 1));
    *have = need;
    return NB200_OK;
}

// P = U . V^T over K-major bf16 planes into an fp32 matrix (split-K + deterministic reduce when that is faster)
int gemm_planes_f32(Context *c, const Planes &U, const Planes &V, int M, int N, int nterms, float *out, int64_t ld) {
    if (M == 0 || N == 0) return NB200_OK;
    const int splits = choose_splits(c, M, N, U.ldk < V.ldk ? U.ldk : V.ldk, nterms);
    int Mpad = 0, Npad = 0;
    GemmEpilogue ge;
    if (splits == 1) {
        ge.mode = 1; ge.out = out; ge.out_ld = ld;
        return launch_gemm_tc(c, U, V, M, N, nterms, 1, ge, &Mpad, &Npad, nullptr, kFmtBF16);
    }
    int real_splits = splits;
    NB_TRY(launch_gemm_tc(c, U, V, M, N, nterms, splits, ge, &Mpad, &Npad, &real_splits, kFmtBF16));
    ReduceOut ro;
    ro.f32 = out; ro.f32_ld = ld;
    return launch_splitk_reduce(c, M, N, real_splits, Mpad, Npad, nullptr, nullptr, 0, ro);
}

// result set holding fp64: storage of an existing set of the same shape is reused (no cudaMalloc / cudaFree per step)
int out_set_f64(const char *db, const char *name, int tr, int tc, int br, int bc, nb200_set **out) {
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_sets.find(std::make_pair(std::string(db), std::string(name)));
        if (it != g_sets.end()) {
            nb200_set *s = it->second;
            if (s->f64 && !s->f32 && !s->view_of && s->total_rows == tr && s->total_cols == tc && s->block_rows == br &&
                s->block_cols == bc) {
                s->version = ++g_version;
                s->rowp.valid = s->rowp.lo_valid = s->rowp.f16_valid = false;
                s->colp.valid = s->colp.lo_valid = s->colp.f16_valid = false;
                *out = s;
                return NB200_OK;
            }
        }
    }
    NB_TRY(make_set(db, name, tr, tc, br, bc, out));
    const size_t n = static_cast<size_t>(tr) * tc;
    NB_CUDA(nb_malloc(&(*out)->f64, (n ? n : 1) * sizeof(double)));
    return NB200_OK;
}
}  // namespace

static void release_lstm_scratch(void) {
    for (Planes *p : {&g_lstm.WU, &g_lstm.XH, &g_lstm.FI, &g_lstm.CC, &g_lstm.OP, &g_lstm.TH}) free_planes(p);
    nb_free(g_lstm.G); nb_free(g_lstm.C32); nb_free(g_lstm.T64);
    g_lstm = LstmScratch();
}

static int lstm_cell_composed(const nb200_lstm_sets *in, int precision, const char *db) {
    static const char *gate_names[4] = {"f_t", "i_t", "o_t", "c_t_temp"};
    nb200_set *gate[4] = {};
    for (int g = 0; g < 4; g++) {
        nb200_set *wx = nullptr, *uh = nullptr;
        NB_TRY(gemm_blocks_impl(NB200_OP_NN, in->w[g], in->x_t, NB200_EPI_NONE, nullptr, precision, db, "__lstm_wx", &wx, 0));
        NB_TRY(gemm_blocks_impl(NB200_OP_NN, in->u[g], in->h_t_1, NB200_EPI_NONE, nullptr, precision, db, "__lstm_uh", &uh, 0));
        NB_TRY(nb200_sum_blocks(wx, uh, in->b[g], g == 3 ? NB200_ACT_TANH : NB200_ACT_SIGMOID, db, gate_names[g], &gate[g]));
    }
    nb200_set *fc = nullptr, *ic = nullptr, *c_t = nullptr, *th = nullptr, *h_t = nullptr;
    NB_TRY(gemm_blocks_impl(NB200_OP_NN, gate[0], in->c_t_1, NB200_EPI_NONE, nullptr, precision, db, "__lstm_wx", &fc, 0));
    NB_TRY(gemm_blocks_impl(NB200_OP_NN, gate[1], gate[3], NB200_EPI_NONE, nullptr, precision, db, "__lstm_uh", &ic, 0));
    NB_TRY(nb200_sum_blocks(fc, ic, nullptr, NB200_ACT_NONE, db, "c_t", &c_t));
    NB_TRY(nb200_sum_blocks(c_t, nullptr, nullptr, NB200_ACT_TANH, db, "__lstm_tanh_c_t", &th));
    NB_TRY(gemm_blocks_impl(NB200_OP_NN, gate[2], th, NB200_EPI_NONE, nullptr, precision, db, "h_t", &h_t, 0));
    return NB200_OK;
}

int nb200_lstm_cell(const nb200_lstm_sets *in, int precision, const char *db) {
    NEED_CTX();
    if (!in || !db || precision < 0 || precision > 3 || !in->x_t || !in->h_t_1 || !in->c_t_1) { set_error("lstm_cell: invalid arguments"); return NB200_ERR_INVALID; }
    for (int g = 0; g < 4; g++)
        if (!in->w[g] || !in->u[g] || !in->b[g]) { set_error("lstm_cell: invalid arguments"); return NB200_ERR_INVALID; }
    const int L = in->w[0]->total_rows, D = in->w[0]->total_cols, B = in->x_t->total_cols;
    const int br = in->w[0]->block_rows, bk = in->w[0]->block_cols, bc = in->x_t->block_cols;
    // the fused path wants the textbook shapes; anything else (the driver's own ragged joins) goes operator by operator
    bool regular = (precision == NB200_PREC_BF16 || precision == NB200_PREC_BF16X3) && L > 0 && D > 0 && B > 0 &&
                   in->x_t->total_rows == D && in->x_t->f64 && in->x_t->block_rows == bk &&
                   in->h_t_1->total_rows == L && in->h_t_1->total_cols == B && in->h_t_1->f64 && in->h_t_1->block_cols == bc &&
                   in->c_t_1->total_rows == L && in->c_t_1->total_cols == B && in->c_t_1->f64 && in->c_t_1->block_cols == bc &&
                   in->c_t_1->block_rows == bc &&  // f_t's column blocks join c_t_1's row blocks
                   br == bc;  // ... and i_t's / o_t's join c_t_temp's / c_t's, whose row blocks are br tall (else the reference exits: composed path)
    for (int g = 0; g < 4 && regular; g++) {
        nb200_set *w = in->w[g], *u = in->u[g], *b = in->b[g];
        regular = w->f64 && u->f64 && b->f64 && w->total_rows == L && w->total_cols == D && w->block_rows == br && w->block_cols == bk &&
                  u->total_rows == L && u->total_cols == L && u->block_rows == br && u->block_cols == in->h_t_1->block_rows &&
                  b->total_rows == L && b->total_cols == B && b->block_rows == br && b->block_cols == bc;
    }
    if (!regular) return lstm_cell_composed(in, precision, db);

    LstmScratch &S = g_lstm;
    const bool lo = precision == NB200_PREC_BF16X3;
    const int nterms = lo ? 3 : 1;
    const int64_t Dp = round_up(D, 64), Lp = round_up(L, 64);
    // 1. stacked weight operand [w_g | u_g] for g = f, i, o, c: rebuilt only when a weight set changed
    bool cached = S.wu_valid && (S.wu_lo || !lo) && S.WU.rows == 4ll * L && S.WU.ldk == Dp + Lp;
    for (int g = 0; g < 4 && cached; g++)
        cached = S.w[g] == in->w[g] && S.u[g] == in->u[g] && S.ver[g] == in->w[g]->version && S.ver[4 + g] == in->u[g]->version;
    if (!cached) {
        S.wu_valid = false;
        NB_TRY(alloc_planes_raw(&S.WU, 4ll * L, Dp + Lp, lo));
        for (int g = 0; g < 4; g++) {
            NB_TRY(launch_convert_sub(c, in->w[g]->f64, D, L, D, false, &S.WU, static_cast<int64_t>(g) * L, 0, lo));
            NB_TRY(launch_convert_sub(c, in->u[g]->f64, L, L, L, false, &S.WU, static_cast<int64_t>(g) * L, Dp, lo));
            S.w[g] = in->w[g]; S.u[g] = in->u[g];
            S.ver[g] = in->w[g]->version; S.ver[4 + g] = in->u[g]->version;
        }
        S.wu_lo = lo;
        S.wu_valid = true;
    }
    // 2. [x_t; h_t_1] as the transposed operand, then every gate's W.x + U.h in one GEMM
    NB_TRY(alloc_planes_raw(&S.XH, B, Dp + Lp, lo));
    NB_TRY(launch_convert_sub(c, in->x_t->f64, B, D, B, true, &S.XH, 0, 0, lo));
    NB_TRY(launch_convert_sub(c, in->h_t_1->f64, B, L, B, true, &S.XH, 0, Dp, lo));
    NB_TRY(grow_bytes(reinterpret_cast<void **>(&S.G), &S.G_cap, static_cast<size_t>(4) * L * B * sizeof(float)));
    NB_TRY(gemm_planes_f32(c, S.WU, S.XH, 4 * L, B, nterms, S.G, B));
    // 3. the four LSTMThreeWaySum joins
    nb200_set *f_t, *i_t, *o_t, *c_tmp, *c_t, *h_t;
    NB_TRY(out_set_f64(db, "f_t", L, B, br, bc, &f_t));
    NB_TRY(out_set_f64(db, "i_t", L, B, br, bc, &i_t));
    NB_TRY(out_set_f64(db, "o_t", L, B, br, bc, &o_t));
    NB_TRY(out_set_f64(db, "c_t_temp", L, B, br, bc, &c_tmp));
    NB_TRY(launch_lstm_gates(c, S.G, B, L, B, in->b[0]->f64, in->b[1]->f64, in->b[2]->f64, in->b[3]->f64, f_t->f64, i_t->f64,
                             o_t->f64, c_tmp->f64));
    // 4. c_t = f_t . c_t_1 + i_t . c_t_temp: the joins match column blocks of the left with row blocks of the right,
    //    i.e. contract over k < Kc = min(B, L); one GEMM over [f_t | i_t] and [c_t_1; c_t_temp]
    const int Kc = B < L ? B : L;
    const int64_t Kcp = round_up(Kc, 64);
    NB_TRY(alloc_planes_raw(&S.FI, L, 2 * Kcp, lo));
    NB_TRY(launch_convert_sub(c, f_t->f64, B, L, Kc, false, &S.FI, 0, 0, lo));
    NB_TRY(launch_convert_sub(c, i_t->f64, B, L, Kc, false, &S.FI, 0, Kcp, lo));
    NB_TRY(alloc_planes_raw(&S.CC, B, 2 * Kcp, lo));
    NB_TRY(launch_convert_sub(c, in->c_t_1->f64, B, Kc, B, true, &S.CC, 0, 0, lo));
    NB_TRY(launch_convert_sub(c, c_tmp->f64, B, Kc, B, true, &S.CC, 0, Kcp, lo));
    NB_TRY(grow_bytes(reinterpret_cast<void **>(&S.C32), &S.C_cap, static_cast<size_t>(L) * B * sizeof(float)));
    NB_TRY(gemm_planes_f32(c, S.FI, S.CC, L, B, nterms, S.C32, B));
    NB_TRY(out_set_f64(db, "c_t", L, B, br, bc, &c_t));
    EwiseSrc src{nullptr, S.C32, B};
    NB_TRY(launch_sum_blocks(c, &src, 1, NB200_ACT_NONE, L, B, c_t->f64));
    // 5. h_t = o_t . tanh(c_t)
    NB_TRY(grow_bytes(reinterpret_cast<void **>(&S.T64), &S.T_cap, static_cast<size_t>(L) * B * sizeof(double)));
    NB_TRY(launch_sum_blocks(c, &src, 1, NB200_ACT_TANH, L, B, S.T64));
    NB_TRY(alloc_planes_raw(&S.OP, L, Kcp, lo));
    NB_TRY(launch_convert_sub(c, o_t->f64, B, L, Kc, false, &S.OP, 0, 0, lo));
    NB_TRY(alloc_planes_raw(&S.TH, B, Kcp, lo));
    NB_TRY(launch_convert_sub(c, S.T64, B, Kc, B, true, &S.TH, 0, 0, lo));
    NB_TRY(gemm_planes_f32(c, S.OP, S.TH, L, B, nterms, S.C32, B));
    NB_TRY(out_set_f64(db, "h_t", L, B, br, bc, &h_t));
    NB_TRY(launch_sum_blocks(c, &src, 1, NB200_ACT_NONE, L, B, h_t->f64));
    return check_kernel_fault(c, "lstm_cell");
}

int nb200_softmax_rows(nb200_set *logits, const char *db, const char *out_set, nb200_set **out, int32_t *labels_host) {
    NEED_CTX();
    if (!logits || !db || !out_set || !out) return NB200_ERR_INVALID;
    nb200_set *src = logits->view_of ? logits->view_of : logits;
    nb200_set *o = nullptr;
    NB_TRY(make_set(db, out_set, logits->total_rows, logits->total_cols, logits->block_rows, logits->block_cols, &o));
    const int64_t rows = o->total_rows, cols = o->total_cols;
    const size_t on = static_cast<size_t>(rows) * cols;
    NB_CUDA(nb_malloc(&o->f64, (on ? on : 1) * sizeof(double)));
    int32_t *labels = nullptr;
    NB_CUDA(nb_malloc(&labels, (rows ? rows : 1) * sizeof(int32_t)));
    int st;
    if (src->f64) st = launch_softmax_rows(c, src->f64, true, rows, cols, src->total_cols, o->f64, cols, labels, nullptr, nullptr, 0, 0);
    else if (src->f32) st = launch_softmax_rows(c, src->f32, false, rows, cols, src->f32_ld, o->f64, cols, labels, nullptr, nullptr, 0, 0);
    else { set_error("softmax_rows: set %s:%s holds no data", logits->db.c_str(), logits->name.c_str()); st = NB200_ERR_STATE; }
    if (st == NB200_OK && labels_host && rows)
        if (cudaMemcpyAsync(labels_host, labels, rows * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
    cudaStreamSynchronize(c->stream);
    nb_free(labels);
    if (st != NB200_OK) return st;
    *out = o;
    return NB200_OK;
}

// ---- fused ff::inference_unit -------------------------------------------------------------------
// One tensor-core forward pass over `rows` rows whose bf16 planes are X:
//   y1 = relu(X W1^T + b1)  (split-K block GEMM + fused reduce, written as operand planes for layer 2)
//   logits = y1 Wo^T + bo   (block GEMM with the bias in its epilogue)
// When `pipe` is given the fp16 planes of X do not exist yet: their conversion from pipe->src is chunked along K
// (one chunk per split) on a side stream and each split's GEMM starts as soon as its chunk is converted, on two
// alternating streams so that one launch's tail overlaps the next one's head.
struct PipeSrc {
    const void *src;
    bool is_f64;
    int64_t ld, cols;
    Planes *planes;
    FusedScratch *fused;  // non-null: use the fused conversion + GEMM kernel (fp64 source, fp16 operands)
    L1StreamScratch *stream = nullptr;  // non-null: prefer the cluster kernel of nb200_l1stream.cu where the shape suits it
};
static int tc_forward(Context *c, const Planes &X, int rows, int nterms, int fmt, nb200_set *w1, nb200_set *wo, const float *b1,
                      const float *bo, Planes *y1, float **y1_f32, size_t *y1_f32_bytes, float *y1_max, float *logits,
                      int64_t logits_ld, int H, int Lb, int *splits1, cudaEvent_t k0, cudaEvent_t k1, cudaEvent_t after_l1,
                      const PipeSrc *pipe = nullptr) {
    const bool lo = nterms == 3;
    const bool f16 = fmt == kFmtF16;
    int Mpad = 0, Npad = 0, used1 = 1;
    GemmEpilogue g1;
    g1.mode = 0;
    if (k0) NB_CUDA(cudaEventRecord(k0, c->stream));
    const int want_splits = choose_splits(c, rows, H, X.ldk, nterms);
    bool fused_done = false;
    int fused_slots = 1;
    if (pipe && pipe->fused && pipe->stream && f16 && pipe->is_f64 && X.ldk >= 4096 &&
        l1_stream_supported(static_cast<const double *>(pipe->src), pipe->ld, pipe->cols, H, w1->rowp)) {
        // Streamed layer 1: the whole batch through the cluster kernel (conversion inside, K split in waves), then one
        // pass that sums the partial tiles and applies FFReluBiasSum.
        const size_t need = static_cast<size_t>(rows) * H * sizeof(float);
        if (need > *y1_f32_bytes) {
            nb_free(*y1_f32);
            *y1_f32 = nullptr;
            *y1_f32_bytes = 0;
            NB_CUDA(nb_malloc(y1_f32, need));
            *y1_f32_bytes = need;
        }
        if (k0) NB_CUDA(cudaEventRecord(k0, c->stream));
        NB_TRY(launch_l1_stream(c, static_cast<const double *>(pipe->src), pipe->ld, rows, static_cast<int>(pipe->cols),
                                pipe->planes->f16_exp, pipe->planes->row_sumsq, pipe->planes->f16_maxscaled, w1->rowp, H, pipe->stream));
        if (k1) NB_CUDA(cudaEventRecord(k1, c->stream));
        NB_TRY(launch_l1_stream_reduce(c, *pipe->stream, rows, H, b1, 1, pipe->planes->f16_exp, w1->rowp.f16_exp, *y1_f32, H, y1_max));
        pipe->planes->f16_valid = false;  // the fp16 plane of the inputs was never materialised
        fused_done = true;
        fused_slots = pipe->stream->max_slots();
        k1 = nullptr;                     // recorded above: the kernel-only window excludes the reduction
    } else if (pipe && pipe->fused && f16 && pipe->is_f64) {
        // Fused layer 1: as many whole 256-row tiles as there are SM pairs go through the persistent kernel that
        // converts fp64 -> fp16 inside the GEMM; the left-over rows take the two-kernel path first.
        const int sms = c->sm_count > 0 ? c->sm_count : 148;
        const int tiles_n = static_cast<int>(ceil_div(H, 512));
        int n_mtiles = (sms / 2) / tiles_n;
        if (n_mtiles > rows / 256) n_mtiles = rows / 256;
        if (n_mtiles >= 8 && X.ldk >= 4096 && tiles_n <= 2) {  // wider hidden layers take the two-kernel path
            const int rows_f = n_mtiles * 256;
            const size_t need = static_cast<size_t>(rows) * H * sizeof(float);
            if (need > *y1_f32_bytes) {
                nb_free(*y1_f32);
                *y1_f32 = nullptr;
                *y1_f32_bytes = 0;
                NB_CUDA(nb_malloc(y1_f32, need));
                *y1_f32_bytes = need;
            }
            NB_CUDA(cudaMemsetAsync(y1_max, 0, sizeof(float), c->stream));
            if (rows_f < rows) {
                Planes sub = *pipe->planes;  // view on the tail rows of the fp16 plane
                sub.f16 += static_cast<int64_t>(rows_f) * sub.ldk;
                sub.row_sumsq += rows_f;
                sub.rows = rows - rows_f;
                NB_TRY(launch_convert_f16_cols(c, c->stream, static_cast<const double *>(pipe->src) + static_cast<int64_t>(rows_f) * pipe->ld,
                                               true, rows - rows_f, pipe->cols, pipe->ld, &sub, 0, sub.ldk));
                int mp = 0, np = 0, used = 1;
                GemmEpilogue gl;
                gl.mode = 0;
                NB_TRY(launch_gemm_tc(c, sub, w1->rowp, rows - rows_f, H, 1, choose_splits(c, rows - rows_f, H, sub.ldk, 1), gl, &mp, &np, &used, fmt));
                ReduceOut rl;
                rl.f32 = *y1_f32 + static_cast<int64_t>(rows_f) * H; rl.f32_ld = H;
                rl.exp_u = sub.f16_exp; rl.exp_v = w1->rowp.f16_exp;
                rl.out_maxabs = y1_max;  // cleared and set here; the fused kernel folds its own maximum in afterwards
                NB_TRY(launch_splitk_reduce(c, rows - rows_f, H, used, mp, np, nullptr, b1, 1, rl));
            }
            if (k0) NB_CUDA(cudaEventRecord(k0, c->stream));  // the kernel-only window (ms_gemm1_kernel) is the fused kernel's
            NB_TRY(launch_l1_fused(c, static_cast<const double *>(pipe->src), pipe->ld, rows, static_cast<int>(pipe->cols), 0, n_mtiles,
                                   *pipe->planes, w1->rowp, H, b1, *y1_f32, H, y1_max, pipe->fused));
            pipe->planes->f16_valid = false;  // the full fp16 plane of the inputs was never materialised
            fused_done = true;
            used1 = 1;
        }
    }
    if (fused_done) {
        if (k1) NB_CUDA(cudaEventRecord(k1, c->stream));
        *splits1 = fused_slots;
        NB_TRY(launch_convert_f16(c, *y1_f32, false, rows, H, H, y1, 1, y1_max));
        if (after_l1) NB_CUDA(cudaEventRecord(after_l1, c->stream));
    } else if (!pipe || pipe->fused) {
        if (pipe)  // fused path requested but the shape does not suit it: materialise the fp16 plane now
            NB_TRY(launch_convert_f16_cols(c, c->stream, pipe->src, pipe->is_f64, rows, pipe->cols, pipe->ld, pipe->planes, 0, pipe->planes->ldk));
        NB_TRY(launch_gemm_tc(c, X, w1->rowp, rows, H, nterms, want_splits, g1, &Mpad, &Npad, &used1, fmt));
    } else {
        const int64_t num_kb = X.ldk / 64;
        const int64_t per = ceil_div(num_kb, want_splits > 0 ? want_splits : 1);
        const int nsplit = static_cast<int>(ceil_div(num_kb, per));
        NB_CUDA(cudaEventRecord(c->join_ev[0], c->stream));
        for (auto &st : c->aux) NB_CUDA(cudaStreamWaitEvent(st, c->join_ev[0], 0));
        for (int j = 0; j < nsplit; j++) {
            NB_TRY(launch_convert_f16_cols(c, c->aux[0], pipe->src, pipe->is_f64, rows, pipe->cols, pipe->ld, pipe->planes,
                                           j * per * 64, per * 64));
            cudaEvent_t ce = c->chunk_ev[j % 64];
            NB_CUDA(cudaEventRecord(ce, c->aux[0]));
            cudaStream_t sg = c->aux[1 + (j & 1)];
            NB_CUDA(cudaStreamWaitEvent(sg, ce, 0));
            g1.split_begin = j; g1.split_count = 1; g1.stream = sg;
            NB_TRY(launch_gemm_tc(c, X, w1->rowp, rows, H, nterms, want_splits, g1, &Mpad, &Npad, &used1, fmt));
        }
        NB_CUDA(cudaEventRecord(c->join_ev[1], c->aux[1]));
        NB_CUDA(cudaEventRecord(c->join_ev[2], c->aux[2]));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->join_ev[1], 0));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->join_ev[2], 0));
    }
    if (!fused_done) {
    if (k1) NB_CUDA(cudaEventRecord(k1, c->stream));
    *splits1 = used1;
    ReduceOut r1;
    if (!f16) {
        r1.planes = y1;
        r1.want_lo = lo;
        NB_TRY(launch_splitk_reduce(c, rows, H, used1, Mpad, Npad, nullptr, b1, 1, r1));
    } else {
        // fp16: un-scale the partial sums, keep y1 in fp32 for a moment to learn its maximum, then pick its scale
        const size_t need = static_cast<size_t>(rows) * H * sizeof(float);
        if (need > *y1_f32_bytes) {
            nb_free(*y1_f32);
            *y1_f32 = nullptr;
            *y1_f32_bytes = 0;
            NB_CUDA(nb_malloc(y1_f32, need));
            *y1_f32_bytes = need;
        }
        r1.f32 = *y1_f32; r1.f32_ld = H;
        r1.exp_u = X.f16_exp; r1.exp_v = w1->rowp.f16_exp;
        r1.out_maxabs = y1_max;
        NB_TRY(launch_splitk_reduce(c, rows, H, used1, Mpad, Npad, nullptr, b1, 1, r1));
        NB_TRY(launch_convert_f16(c, *y1_f32, false, rows, H, H, y1, 1, y1_max));
    }
    if (after_l1) NB_CUDA(cudaEventRecord(after_l1, c->stream));
    }  // !fused_done
    const int splits2 = choose_splits(c, rows, Lb, y1->ldk, nterms);
    if (splits2 == 1) {
        GemmEpilogue g2;
        g2.mode = 1; g2.out = logits; g2.out_ld = logits_ld; g2.bias_n = bo;
        NB_TRY(launch_gemm_tc(c, *y1, wo->rowp, rows, Lb, nterms, 1, g2, &Mpad, &Npad, nullptr, fmt));
    } else {
        GemmEpilogue g2;
        g2.mode = 0;
        int used2 = splits2;
        NB_TRY(launch_gemm_tc(c, *y1, wo->rowp, rows, Lb, nterms, splits2, g2, &Mpad, &Npad, &used2, fmt));
        ReduceOut r2;
        r2.f32 = logits; r2.f32_ld = logits_ld;
        if (f16) { r2.exp_u = y1->f16_exp; r2.exp_v = wo->rowp.f16_exp; }
        NB_TRY(launch_splitk_reduce(c, rows, Lb, used2, Mpad, Npad, nullptr, bo, 0, r2));
    }
    return NB200_OK;
}

struct FFScratch {  // reused across calls
    Planes y1;                 // relu(X W1^T + b1) as operand planes, rows = batch, K = hidden
    float *logits = nullptr;   // batch x ld fp32
    int64_t logits_rows = 0, logits_ld = 0;
    int32_t *labels = nullptr;
    float *gap = nullptr;
    int32_t *list = nullptr;
    int *count = nullptr;
    int *count_host = nullptr;
    int64_t cap_rows = 0;
    double *xg = nullptr, *y1g = nullptr, *zg = nullptr;  // fp64 refinement buffers
    size_t xg_bytes = 0, y1g_bytes = 0, zg_bytes = 0;
    // second cascade level (bf16x3 on the rows the single-pass result left undecided)
    Planes xsub, y1sub;
    float *logits_sub = nullptr;
    size_t logits_sub_bytes = 0;
    int32_t *list2 = nullptr;
    int64_t list2_cap = 0;
    float *y1_f32 = nullptr;   // fp16 path: layer-1 result before its scale is known
    size_t y1_f32_bytes = 0;
    float *y1_max = nullptr;
    float *host_scalars = nullptr;  // pinned: [0] = largest scaled |x| of the inputs' fp16 conversion
    FusedScratch fused;             // ring + flags of the fused layer-1 kernel
    L1StreamScratch l1s;            // wave plan of the streamed layer-1 kernel
};
static FFScratch &ff_ref() { DeviceState &D = cur_state(); if (!D.ff) D.ff = new FFScratch(); return *static_cast<FFScratch *>(D.ff); }
#define g_ff (ff_ref())

// nb200_shutdown: nothing of the fused stage's scratch may outlive the context (a later nb200_init may bind another device)
static void release_ff_scratch(void) {
    FFScratch &S = g_ff;
    for (Planes *p : {&S.y1, &S.xsub, &S.y1sub}) free_planes(p);
    nb_free(S.logits); nb_free(S.labels); nb_free(S.gap); nb_free(S.list); nb_free(S.count); nb_free(S.xg); nb_free(S.y1g);
    nb_free(S.zg); nb_free(S.logits_sub); nb_free(S.list2); nb_free(S.y1_f32); nb_free(S.y1_max);
    nb_free(S.fused.ring); nb_free(S.fused.flags); nb_free(S.l1s.part[0].dev); nb_free(S.l1s.part[1].dev);
    if (S.count_host) cudaFreeHost(S.count_host);
    if (S.host_scalars) cudaFreeHost(S.host_scalars);
    S = FFScratch();
}

static int grow(double **p, size_t *have, size_t need) {
    if (need <= *have) return NB200_OK;
    nb_free(*p);
    *p = nullptr;
    *have = 0;
    NB_CUDA(nb_malloc(p, need));
    *have = need;
    return NB200_OK;
}

int nb200_ff_inference_unit(nb200_set *w1, nb200_set *wo, nb200_set *inputs, nb200_set *b1, nb200_set *bo, int precision,
                            const char *db, const char *out_set, nb200_set **out, int32_t *labels_host,
                            nb200_ff_stats *stats) {
    NEED_CTX_RAW();
    if (!w1 || !wo || !inputs || !b1 || !bo || !db || !out_set || !out || precision < 0 || precision > 3) {
        set_error("ff_inference_unit: invalid arguments");
        return NB200_ERR_INVALID;
    }
    NB_TRY(same_device(c, {w1, wo, inputs, b1, bo}));
    // wait for the uploads of the five operand sets only (see nb200_upload_commit): the next batch may already be on its
    // way into another set
    for (nb200_set *s : {w1, wo, inputs, b1, bo}) NB_TRY(wait_set_uploads(c, s));
    c->main_dirty = true;
    const int B = inputs->total_rows, F = inputs->total_cols, H = w1->total_rows, Lb = wo->total_rows;
    // FFTransposeMult predicate/dimension check (w1 x inputs), FFInputLayerJoin (wo x y1)
    if (w1->total_cols != F || w1->block_cols != inputs->block_cols) { set_error("Block dimemsions mismatch! (w1 vs inputs)"); return NB200_ERR_SHAPE; }
    if (wo->total_cols != H || wo->block_cols != w1->block_rows) { set_error("Block dimemsions mismatch! (wo vs y1)"); return NB200_ERR_SHAPE; }
    NB_TRY(bias_check(b1, H, "ReluBiasSum"));
    NB_TRY(bias_check(bo, Lb, "TransposeBiasSum"));
    const bool side = inputs->side_format && inputs->rowp.f16_valid && inputs->hb_dev;  // resident in its fp16 side format only
    if (!w1->f64 || !wo->f64 || (!inputs->f64 && !side) || !b1->f64 || !bo->f64) { set_error("ff_inference_unit: an input set holds no data"); return NB200_ERR_STATE; }
    if (side && !inputs->f64 && precision != NB200_PREC_F16) {
        set_error("ff_inference_unit: a set imported in the fp16 side format can only be evaluated with NB200_PREC_F16");
        return NB200_ERR_INVALID;
    }
    const HostBlocks hb{inputs->hb_dev, inputs->block_rows, inputs->block_cols,
                        (inputs->total_cols + inputs->block_cols - 1) / inputs->block_cols, inputs->hb_edge_ld};
    const bool from_host = side && !inputs->f64;

    const int64_t launches0 = c->launches;
    // a repeated call replaces its result sets: take over their buffers instead of freeing and re-allocating
    auto take = [&](const char *nm, bool want_f64) -> void * {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_sets.find(std::make_pair(std::string(db), std::string(nm)));
        if (it == g_sets.end() || it->second->total_rows != B || it->second->total_cols != Lb) return nullptr;
        void *p = nullptr;
        if (want_f64) { p = it->second->f64; it->second->f64 = nullptr; }
        else if (it->second->f32_ld == Lb) { p = it->second->f32; it->second->f32 = nullptr; }
        return p;
    };
    float *reuse_logits = static_cast<float *>(take("__logits", false));
    double *reuse_out = static_cast<double *>(take(out_set, true));
    // output sets: yo (view: exp of the logits, FFTransposeBiasSum.h:64-67 meta) and `output`
    nb200_set *zset = nullptr, *yo = nullptr, *o = nullptr;
    NB_TRY(make_set(db, "__logits", B, Lb, inputs->block_rows, wo->block_rows, &zset));
    NB_TRY(make_set(db, "yo", B, Lb, inputs->block_rows, wo->block_rows, &yo));
    yo->view_of = zset;
    yo->view_exp = true;
    NB_TRY(make_set(db, out_set, B, Lb, inputs->block_rows, wo->block_rows, &o));
    const size_t on = static_cast<size_t>(B) * Lb;
    if (reuse_out) o->f64 = reuse_out;
    else NB_CUDA(nb_malloc(&o->f64, (on ? on : 1) * sizeof(double)));
    if (reuse_logits) { zset->f32 = reuse_logits; zset->f32_ld = Lb; }

    FFScratch &S = g_ff;
    if (S.cap_rows < B) {
        nb_free(S.labels); nb_free(S.gap); nb_free(S.list);
        NB_CUDA(nb_malloc(&S.labels, (B ? B : 1) * sizeof(int32_t)));
        NB_CUDA(nb_malloc(&S.gap, (B ? B : 1) * sizeof(float)));
        NB_CUDA(nb_malloc(&S.list, (B ? B : 1) * sizeof(int32_t)));
        S.cap_rows = B;
    }
    if (!S.count) {
        NB_CUDA(nb_malloc(&S.count, sizeof(int)));
        NB_CUDA(cudaMallocHost(&S.count_host, sizeof(int)));
        NB_CUDA(nb_malloc(&S.y1_max, sizeof(float)));
        NB_CUDA(cudaMallocHost(&S.host_scalars, 4 * sizeof(float)));
    }
    cudaEvent_t *ev = c->ev;
    NB_CUDA(cudaEventRecord(ev[0], c->stream));
    int refined = 0, splits1 = 1, relevel = 0;
    float ms_k1 = 0.f;

    if (B == 0 || Lb == 0) {
        *out = o;
        if (stats) memset(stats, 0, sizeof(*stats));
        return NB200_OK;
    }

    if (precision == NB200_PREC_FP64) {
        // exact path: everything in fp64 DFMA
        NB_TRY(grow(&S.y1g, &S.y1g_bytes, static_cast<size_t>(B) * H * sizeof(double)));
        if (zset->f32) { nb_free(zset->f32); zset->f32 = nullptr; }
        NB_CUDA(nb_malloc(&zset->f64, on * sizeof(double)));
        NB_CUDA(cudaEventRecord(ev[1], c->stream));
        F64Epilogue e1; e1.bias_n = b1->f64; e1.relu = 1;
        NB_TRY(launch_gemm_f64(c, inputs->f64, F, 1, w1->f64, F, 1, B, H, F, S.y1g, H, e1));
        NB_CUDA(cudaEventRecord(ev[2], c->stream));
        F64Epilogue e2; e2.bias_n = bo->f64;
        NB_TRY(launch_gemm_f64(c, S.y1g, H, 1, wo->f64, H, 1, B, Lb, H, zset->f64, Lb, e2));
        NB_CUDA(cudaEventRecord(ev[3], c->stream));
        NB_TRY(launch_softmax_rows(c, zset->f64, true, B, Lb, Lb, o->f64, Lb, S.labels, nullptr, nullptr, 0, 0));
        NB_CUDA(cudaEventRecord(ev[4], c->stream));
        NB_CUDA(cudaEventRecord(ev[5], c->stream));
    } else {
        const int nterms = (precision == NB200_PREC_BF16X3) ? 3 : 1;
        const bool lo = nterms == 3;
        const int fmt = (precision == NB200_PREC_F16) ? kFmtF16 : kFmtBF16;
        // model side (cached across calls, like a materialised model): weight planes, fp32 biases.
        // The single-pass variants also keep the residual planes: their undecided rows are re-run in bf16x3.
        NB_TRY(ensure_rowp(c, w1, true));
        NB_TRY(ensure_rowp(c, wo, true));
        if (fmt == kFmtF16) {
            NB_TRY(ensure_f16_row(c, w1, 1));
            NB_TRY(ensure_f16_row(c, wo, 1));
        }
        NB_TRY(ensure_f32(c, b1));
        NB_TRY(ensure_f32(c, bo));
        if (!wo->wo_rownorm_max) NB_CUDA(nb_malloc(&wo->wo_rownorm_max, sizeof(float)));
        NB_TRY(launch_max_rownorm(c, wo->rowp.row_sumsq, Lb, wo->wo_rownorm_max));
        NB_CUDA(cudaEventRecord(ev[0], c->stream));
        // query side: de-blocked fp64 inputs -> 16-bit operand planes.  With NB200_OVERLAP=1 the fp16 conversion is
        // pipelined with the layer-1 GEMM chunk by chunk (tc_forward) and only the scale is picked here; measured on
        // B200 the two kernels do not overlap usefully (phase time = sum of both), so it is off by default.
        static const bool overlap = getenv("NB200_OVERLAP") && atoi(getenv("NB200_OVERLAP")) != 0;  // off by default, see DESIGN.md
        const bool fused = fused_mode() != 0;
        PipeSrc pipe{inputs->f64, true, F, F, &inputs->rowp, fused ? &S.fused : nullptr, fused_mode() == 2 ? &S.l1s : nullptr};
        bool pipelined = false;
        if (fmt == kFmtF16) {
            if (!inputs->rowp.f16_valid && (overlap || fused) && static_cast<int64_t>(F) >= 8192) {
                NB_TRY(prepare_convert_f16(c, inputs->f64, true, B, F, F, &inputs->rowp, 128, nullptr));
                pipelined = true;
            } else {
                NB_TRY(ensure_f16_row(c, inputs, 128));
            }
        } else {
            NB_TRY(ensure_rowp(c, inputs, lo));
        }
        NB_CUDA(cudaEventRecord(ev[1], c->stream));
        if (!zset->f32) NB_CUDA(nb_malloc(&zset->f32, on * sizeof(float)));
        zset->f32_ld = Lb;
        // level 1 on every row
        NB_TRY(tc_forward(c, inputs->rowp, B, nterms, fmt, w1, wo, b1->f32, bo->f32, &S.y1, &S.y1_f32, &S.y1_f32_bytes, S.y1_max,
                          zset->f32, Lb, H, Lb, &splits1, ev[6], ev[7], ev[2], pipelined ? &pipe : nullptr));
        NB_CUDA(cudaEventRecord(ev[3], c->stream));
        // FFRowAggregate + FFOutputLayer + label, and the near-tie scan
        NB_TRY(launch_softmax_rows(c, zset->f32, false, B, Lb, Lb, o->f64, Lb, S.labels, S.gap, nullptr, 0, 0));
        NB_CUDA(cudaMemsetAsync(S.count, 0, sizeof(int), c->stream));
        NB_TRY(launch_flag_rows(c, B, S.gap, inputs->rowp.row_sumsq, S.y1.row_sumsq, w1->rowp.maxabs, wo->rowp.maxabs,
                                wo->wo_rownorm_max, prec_unit(precision), 0.1f * g_opt_tau_x10, S.list, S.count, B, nullptr));
        NB_CUDA(cudaMemcpyAsync(S.count_host, S.count, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        if (fmt == kFmtF16)
            NB_CUDA(cudaMemcpyAsync(S.host_scalars, inputs->rowp.f16_maxscaled, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaEventRecord(ev[4], c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        if (fmt == kFmtF16 && !(S.host_scalars[0] < 65504.f)) {
            // the sampled scale let an input overflow fp16: redo this call in bf16 (full exponent range)
            inputs->rowp.f16_valid = false;
            return nb200_ff_inference_unit(w1, wo, inputs, b1, bo, NB200_PREC_BF16, db, out_set, out, labels_host, stats);
        }
        int n1 = std::min(*S.count_host, B);
        const int32_t *final_list = S.list;
        refined = n1;
        // (a handful of undecided rows go straight to the fp64 level: both levels stream all of W1 once, and the fp64 kernel
        // handles up to 16 rows in that one pass)
        if ((precision == NB200_PREC_BF16 || precision == NB200_PREC_F16) && n1 > 16) {
            // level 2: bf16x3 on the undecided rows only (gathered straight from the fp64 inputs)
            if (S.list2_cap < B) {
                nb_free(S.list2);
                NB_CUDA(nb_malloc(&S.list2, static_cast<size_t>(B) * sizeof(int32_t)));
                S.list2_cap = B;
            }
            NB_CUDA(cudaMemsetAsync(S.count, 0, sizeof(int), c->stream));
            const int chunk2 = 8192;
            for (int r0 = 0; r0 < n1; r0 += chunk2) {
                const int R = std::min(chunk2, n1 - r0);
                if (from_host) NB_TRY(launch_gather_convert_rows_hb(c, hb, F, S.list + r0, R, &S.xsub));
                else NB_TRY(launch_gather_convert_rows(c, inputs->f64, F, F, S.list + r0, R, &S.xsub));
                const size_t need = static_cast<size_t>(R) * Lb * sizeof(float);
                if (need > S.logits_sub_bytes) {
                    nb_free(S.logits_sub);
                    S.logits_sub = nullptr;
                    S.logits_sub_bytes = 0;
                    NB_CUDA(nb_malloc(&S.logits_sub, need));
                    S.logits_sub_bytes = need;
                }
                int dummy = 0;
                NB_TRY(tc_forward(c, S.xsub, R, 3, kFmtBF16, w1, wo, b1->f32, bo->f32, &S.y1sub, nullptr, nullptr, nullptr,
                                  S.logits_sub, Lb, H, Lb, &dummy, nullptr, nullptr, nullptr));
                NB_TRY(launch_scatter_rows_f32(c, S.logits_sub, Lb, Lb, S.list + r0, R, zset->f32, Lb));
                NB_TRY(launch_softmax_rows(c, S.logits_sub, false, B, Lb, Lb, o->f64, Lb, S.labels, S.gap, S.list + r0, R, 1));
                NB_TRY(launch_flag_rows(c, R, S.gap, inputs->rowp.row_sumsq, S.y1.row_sumsq, w1->rowp.maxabs,
                                        wo->rowp.maxabs, wo->wo_rownorm_max, prec_unit(NB200_PREC_BF16X3), 0.1f * g_opt_tau_x10, S.list2,
                                        S.count, B, S.list + r0));
            }
            NB_CUDA(cudaMemcpyAsync(S.count_host, S.count, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
            NB_CUDA(cudaStreamSynchronize(c->stream));
            refined = std::min(*S.count_host, B);
            final_list = S.list2;
            relevel = n1;
        }
        // last level: fp64 re-evaluation of what is still undecided (K9), in chunks that bound the gather buffer
        const int chunk = 256;
        for (int r0 = 0; r0 < refined; r0 += chunk) {
            const int R = std::min(chunk, refined - r0);
            NB_TRY(grow(&S.xg, &S.xg_bytes, static_cast<size_t>(chunk) * F * sizeof(double)));
            NB_TRY(grow(&S.y1g, &S.y1g_bytes, static_cast<size_t>(chunk) * H * sizeof(double)));
            NB_TRY(grow(&S.zg, &S.zg_bytes, static_cast<size_t>(chunk) * Lb * sizeof(double)));
            if (from_host) NB_TRY(launch_gather_rows_f64_hb(c, hb, F, final_list + r0, R, S.xg));
            else NB_TRY(launch_gather_rows_f64(c, inputs->f64, F, F, final_list + r0, R, S.xg));
            F64Epilogue e1; e1.bias_n = b1->f64; e1.relu = 1;
            NB_TRY(launch_gemm_f64(c, S.xg, F, 1, w1->f64, F, 1, R, H, F, S.y1g, H, e1));
            F64Epilogue e2; e2.bias_n = bo->f64;
            NB_TRY(launch_gemm_f64(c, S.y1g, H, 1, wo->f64, H, 1, R, Lb, H, S.zg, Lb, e2));
            NB_TRY(launch_scatter_rows(c, S.zg, Lb, final_list + r0, R, zset->f32, Lb));
            NB_TRY(launch_softmax_rows(c, S.zg, true, B, Lb, Lb, o->f64, Lb, S.labels, nullptr, final_list + r0, R, 1));
        }
        NB_CUDA(cudaEventRecord(ev[5], c->stream));
    }
    if (labels_host) NB_CUDA(cudaMemcpyAsync(labels_host, S.labels, static_cast<size_t>(B) * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    NB_TRY(check_kernel_fault(c, "ff_inference_unit"));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        cudaEventElapsedTime(&stats->ms_convert, ev[0], ev[1]);
        cudaEventElapsedTime(&stats->ms_gemm1, ev[1], ev[2]);
        cudaEventElapsedTime(&stats->ms_gemm2, ev[2], ev[3]);
        cudaEventElapsedTime(&stats->ms_softmax, ev[3], ev[4]);
        cudaEventElapsedTime(&stats->ms_refine, ev[4], ev[5]);
        cudaEventElapsedTime(&stats->ms_total, ev[0], ev[5]);
        if (precision != NB200_PREC_FP64) cudaEventElapsedTime(&ms_k1, ev[6], ev[7]);
        stats->ms_gemm1_kernel = ms_k1;
        stats->refined_rows = refined;
        stats->relevel_rows = relevel;
        stats->kernel_launches = static_cast<int>(c->launches - launches0);
        stats->splits1 = splits1;
    }
    *out = o;
    return NB200_OK;
}


// ---- K-sharded variant ----------------------------------------------------------------------------
int nb200_ff_layer1_partial(nb200_set *w1, nb200_set *inputs, int precision, float *partial_dev, int64_t ld) {
    NEED_CTX();
    if (!w1 || !inputs || !partial_dev || precision == NB200_PREC_FP64 || precision < 0 || precision > 3) {
        set_error("ff_layer1_partial: invalid arguments (tensor-core precisions only)");
        return NB200_ERR_INVALID;
    }
    const int B = inputs->total_rows, Fs = inputs->total_cols, H = w1->total_rows;
    if (w1->total_cols != Fs || w1->block_cols != inputs->block_cols || ld < H) {
        set_error("Block dimemsions mismatch! (w1 shard vs inputs shard)");
        return NB200_ERR_SHAPE;
    }
    if (B == 0 || H == 0) return NB200_OK;
    const int nterms = precision == NB200_PREC_BF16X3 ? 3 : 1;
    const int fmt = precision == NB200_PREC_F16 ? kFmtF16 : kFmtBF16;
    if (fmt == kFmtF16) {
        NB_TRY(ensure_f16_row(c, w1, 1));
        if (fused_mode() == 2 && inputs->f64 && !inputs->rowp.f16_valid && Fs >= 4096 &&
            l1_stream_supported(inputs->f64, Fs, Fs, H, w1->rowp)) {
            // the shard's fp64 blocks go through the cluster kernel as they are (no fp16 plane of the inputs)
            NB_TRY(prepare_convert_f16(c, inputs->f64, true, B, Fs, Fs, &inputs->rowp, 64, nullptr));
            NB_TRY(launch_l1_stream(c, inputs->f64, Fs, B, Fs, inputs->rowp.f16_exp, inputs->rowp.row_sumsq,
                                    inputs->rowp.f16_maxscaled, w1->rowp, H, &g_ff.l1s));
            NB_TRY(launch_l1_stream_reduce(c, g_ff.l1s, B, H, nullptr, 0, inputs->rowp.f16_exp, w1->rowp.f16_exp, partial_dev, ld, nullptr));
            inputs->rowp.f16_valid = false;
            if (ld > H)
                NB_CUDA(cudaMemcpy2DAsync(partial_dev + H, static_cast<size_t>(ld) * sizeof(float), inputs->rowp.row_sumsq, sizeof(float),
                                          sizeof(float), B, cudaMemcpyDeviceToDevice, c->stream));
            float mx = 0.f;
            NB_CUDA(cudaMemcpyAsync(&mx, inputs->rowp.f16_maxscaled, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
            NB_CUDA(cudaStreamSynchronize(c->stream));
            if (!(mx < 65504.f)) {
                set_error("ff_layer1_partial: an input overflowed the sampled fp16 scale; use NB200_PREC_BF16 or BF16X3");
                return NB200_ERR_STATE;
            }
            return check_kernel_fault(c, "ff_layer1_partial");
        }
        NB_TRY(ensure_f16_row(c, inputs, 64));
    } else {
        NB_TRY(ensure_rowp(c, w1, nterms == 3));
        NB_TRY(ensure_rowp(c, inputs, nterms == 3));
    }
    int Mpad = 0, Npad = 0, used = 1;
    GemmEpilogue g;
    g.mode = 0;
    NB_TRY(launch_gemm_tc(c, inputs->rowp, w1->rowp, B, H, nterms, choose_splits(c, B, H, inputs->rowp.ldk, nterms), g, &Mpad,
                          &Npad, &used, fmt));
    ReduceOut r;
    r.f32 = partial_dev; r.f32_ld = ld;
    if (fmt == kFmtF16) { r.exp_u = inputs->rowp.f16_exp; r.exp_v = w1->rowp.f16_exp; }
    NB_TRY(launch_splitk_reduce(c, B, H, used, Mpad, Npad, nullptr, nullptr, 0, r));
    if (ld > H)  // extra column: this shard's per-row sum of squares of the inputs (summed by the caller's reduction)
        NB_CUDA(cudaMemcpy2DAsync(partial_dev + H, static_cast<size_t>(ld) * sizeof(float), inputs->rowp.row_sumsq, sizeof(float),
                                  sizeof(float), B, cudaMemcpyDeviceToDevice, c->stream));
    if (fmt == kFmtF16) {  // a sampled scale that overflowed would have produced inf: report instead of returning garbage
        float mx = 0.f;
        NB_CUDA(cudaMemcpyAsync(&mx, inputs->rowp.f16_maxscaled, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        if (!(mx < 65504.f)) {
            inputs->rowp.f16_valid = false;
            set_error("ff_layer1_partial: an input overflowed the sampled fp16 scale; use NB200_PREC_BF16 or BF16X3");
            return NB200_ERR_STATE;
        }
    }
    return check_kernel_fault(c, "ff_layer1_partial");
}

int nb200_ff_finish_from_layer1(const float *y1pre_dev, int64_t ld, int rows, nb200_set *b1, nb200_set *wo, nb200_set *bo,
                                int precision, float w1_maxabs, const char *db, const char *out_set, nb200_set **out,
                                int32_t *labels_host, int *undecided_rows) {
    NEED_CTX();
    if (!y1pre_dev || !b1 || !wo || !bo || !db || !out_set || !out || rows < 0 || precision == NB200_PREC_FP64 ||
        precision < 0 || precision > 3) {
        set_error("ff_finish_from_layer1: invalid arguments (tensor-core precisions only)");
        return NB200_ERR_INVALID;
    }
    const int H = wo->total_cols, Lb = wo->total_rows;
    NB_TRY(bias_check(b1, H, "ReluBiasSum"));
    NB_TRY(bias_check(bo, Lb, "TransposeBiasSum"));
    if (ld < H) { set_error("ff_finish_from_layer1: pitch smaller than the hidden width"); return NB200_ERR_INVALID; }
    nb200_set *zset = nullptr, *yo = nullptr, *o = nullptr;
    NB_TRY(make_set(db, "__logits", rows, Lb, b1->block_rows, wo->block_rows, &zset));
    NB_TRY(make_set(db, "yo", rows, Lb, b1->block_rows, wo->block_rows, &yo));
    yo->view_of = zset;
    yo->view_exp = true;
    NB_TRY(make_set(db, out_set, rows, Lb, b1->block_rows, wo->block_rows, &o));
    const size_t on = static_cast<size_t>(rows) * Lb;
    NB_CUDA(nb_malloc(&o->f64, (on ? on : 1) * sizeof(double)));
    NB_CUDA(nb_malloc(&zset->f32, (on ? on : 1) * sizeof(float)));
    zset->f32_ld = Lb;
    *out = o;
    if (undecided_rows) *undecided_rows = 0;
    if (rows == 0 || Lb == 0) return NB200_OK;
    const int nterms = precision == NB200_PREC_BF16X3 ? 3 : 1;
    const int fmt = precision == NB200_PREC_F16 ? kFmtF16 : kFmtBF16;
    NB_TRY(ensure_f32(c, b1));
    NB_TRY(ensure_f32(c, bo));
    NB_TRY(ensure_rowp(c, wo, nterms == 3));
    if (fmt == kFmtF16) NB_TRY(ensure_f16_row(c, wo, 1));
    // y1 = relu(sum of partials + b1)  [FFReluBiasSum], then operand planes for layer 2
    FFScratch &S = g_ff;
    const size_t need = static_cast<size_t>(rows) * H * sizeof(float);
    if (need > S.y1_f32_bytes) {
        nb_free(S.y1_f32);
        S.y1_f32 = nullptr;
        S.y1_f32_bytes = 0;
        NB_CUDA(nb_malloc(&S.y1_f32, need));
        S.y1_f32_bytes = need;
    }
    NB_CUDA(cudaMemcpy2DAsync(S.y1_f32, static_cast<size_t>(H) * sizeof(float), y1pre_dev, static_cast<size_t>(ld) * sizeof(float),
                              static_cast<size_t>(H) * sizeof(float), rows, cudaMemcpyDeviceToDevice, c->stream));
    NB_TRY(launch_bias_relu_f32(c, S.y1_f32, rows, H, H, b1->f32));
    if (fmt == kFmtF16) NB_TRY(launch_convert_f16(c, S.y1_f32, false, rows, H, H, &S.y1, 1, nullptr));
    else NB_TRY(launch_convert_planes(c, S.y1_f32, false, rows, H, H, &S.y1, nterms == 3));
    int Mpad = 0, Npad = 0, used = 1;
    const int splits2 = choose_splits(c, rows, Lb, S.y1.ldk, nterms);
    GemmEpilogue g2;
    if (splits2 == 1) {
        g2.mode = 1; g2.out = zset->f32; g2.out_ld = Lb; g2.bias_n = bo->f32;
        NB_TRY(launch_gemm_tc(c, S.y1, wo->rowp, rows, Lb, nterms, 1, g2, &Mpad, &Npad, nullptr, fmt));
    } else {
        g2.mode = 0;
        NB_TRY(launch_gemm_tc(c, S.y1, wo->rowp, rows, Lb, nterms, splits2, g2, &Mpad, &Npad, &used, fmt));
        ReduceOut r2;
        r2.f32 = zset->f32; r2.f32_ld = Lb;
        if (fmt == kFmtF16) { r2.exp_u = S.y1.f16_exp; r2.exp_v = wo->rowp.f16_exp; }
        NB_TRY(launch_splitk_reduce(c, rows, Lb, used, Mpad, Npad, nullptr, bo->f32, 0, r2));
    }
    if (S.cap_rows < rows) {
        nb_free(S.labels); nb_free(S.gap); nb_free(S.list);
        NB_CUDA(nb_malloc(&S.labels, static_cast<size_t>(rows) * sizeof(int32_t)));
        NB_CUDA(nb_malloc(&S.gap, static_cast<size_t>(rows) * sizeof(float)));
        NB_CUDA(nb_malloc(&S.list, static_cast<size_t>(rows) * sizeof(int32_t)));
        S.cap_rows = rows;
    }
    NB_TRY(launch_softmax_rows(c, zset->f32, false, rows, Lb, Lb, o->f64, Lb, S.labels, S.gap, nullptr, 0, 0));
    if (labels_host) NB_CUDA(cudaMemcpyAsync(labels_host, S.labels, static_cast<size_t>(rows) * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    if (undecided_rows && ld > H && w1_maxabs > 0.f) {
        // same near-tie criterion as the fused stage, with the globally reduced ||x_r||^2 and max |w1|
        if (!S.count) {
            NB_CUDA(nb_malloc(&S.count, sizeof(int)));
            NB_CUDA(cudaMallocHost(&S.count_host, sizeof(int)));
            NB_CUDA(nb_malloc(&S.y1_max, sizeof(float)));
            NB_CUDA(cudaMallocHost(&S.host_scalars, 4 * sizeof(float)));
        }
        float *xs = nullptr;
        NB_CUDA(nb_malloc(&xs, static_cast<size_t>(rows + 1) * sizeof(float)));
        int st = NB200_OK;
        if (cudaMemcpy2DAsync(xs, sizeof(float), y1pre_dev + H, static_cast<size_t>(ld) * sizeof(float), sizeof(float), rows,
                              cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
        if (st == NB200_OK && cudaMemcpyAsync(xs + rows, &w1_maxabs, sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
        if (!wo->wo_rownorm_max && st == NB200_OK && nb_malloc(&wo->wo_rownorm_max, sizeof(float)) != cudaSuccess) st = NB200_ERR_CUDA;
        if (st == NB200_OK) st = launch_max_rownorm(c, wo->rowp.row_sumsq, Lb, wo->wo_rownorm_max);
        if (st == NB200_OK && cudaMemsetAsync(S.count, 0, sizeof(int), c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
        if (st == NB200_OK)
            st = launch_flag_rows(c, rows, S.gap, xs, S.y1.row_sumsq, xs + rows, wo->rowp.maxabs, wo->wo_rownorm_max,
                                  prec_unit(precision), 0.1f * g_opt_tau_x10, S.list, S.count, rows, nullptr);
        if (st == NB200_OK && cudaMemcpyAsync(S.count_host, S.count, sizeof(int), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
        cudaStreamSynchronize(c->stream);
        nb_free(xs);
        if (st != NB200_OK) return st;
        *undecided_rows = *S.count_host;
    }
    return check_kernel_fault(c, "ff_finish_from_layer1");
}

// ---- conv2d_memory_fusion -------------------------------------------------------------------------
int nb200_conv2d_image_flat(const double *images_host, int n, int ch, int h, int w, int k, int stride, int padding,
                            int block_rows, int block_cols, const char *db, const char *out_set, nb200_set **out) {
    NEED_CTX();
    if (!images_host || n <= 0 || ch <= 0 || h <= 0 || w <= 0 || k <= 0 || stride <= 0 || padding < 0 || !out) {
        set_error("conv2d_image_flat: invalid arguments");
        return NB200_ERR_INVALID;
    }
    const int vs = (h + 2 * padding - k) / stride + 1, hs = (w + 2 * padding - k) / stride + 1;
    if (vs <= 0 || hs <= 0) { set_error("conv2d_image_flat: kernel larger than image"); return NB200_ERR_SHAPE; }
    nb200_set *o = nullptr;
    NB_TRY(make_set(db, out_set, n * vs * hs, ch * k * k + 1, block_rows, block_cols, &o));
    NB_TRY(ensure_f64(c, o));
    const size_t ib = static_cast<size_t>(n) * ch * h * w * sizeof(double);
    double *dimg = nullptr;
    NB_CUDA(nb_malloc(&dimg, ib));
    int st = NB200_OK;
    if (cudaMemcpyAsync(dimg, images_host, ib, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) st = NB200_ERR_CUDA;
    if (st == NB200_OK) st = launch_im2col(c, dimg, n, ch, h, w, k, stride, padding, o->f64, o->total_cols);
    cudaStreamSynchronize(c->stream);
    nb_free(dimg);
    if (st != NB200_OK) return st;
    *out = o;
    return NB200_OK;
}

int nb200_conv2d_kernel_flat(const double *kernels_host, const double *bias_host, int kout, int ch, int k, int block_rows,
                             int block_cols, const char *db, const char *out_set, nb200_set **out) {
    NEED_CTX();
    if (!kernels_host || !bias_host || kout <= 0 || ch <= 0 || k <= 0 || !out) {
        set_error("conv2d_kernel_flat: invalid arguments");
        return NB200_ERR_INVALID;
    }
    // tiny one-off pack on the host (KernelToChunks.h:39-83 column order, KernelBiasJoin.h:42-46 bias column)
    const int cols = ch * k * k + 1;
    std::vector<double> flat(static_cast<size_t>(kout) * cols);
    for (int r = 0; r < kout; r++) {
        memcpy(&flat[static_cast<size_t>(r) * cols], kernels_host + static_cast<size_t>(r) * (cols - 1), (cols - 1) * sizeof(double));
        flat[static_cast<size_t>(r) * cols + cols - 1] = bias_host[r];
    }
    nb200_set *o = nullptr;
    NB_TRY(make_set(db, out_set, kout, cols, block_rows, block_cols, &o));
    NB_TRY(ensure_f64(c, o));
    NB_CUDA(cudaMemcpyAsync(o->f64, flat.data(), flat.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    *out = o;
    return NB200_OK;
}

int nb200_conv2d_result_to_images(nb200_set *result, int num_images, int x, int y, int z, double *images_host) {
    NEED_CTX();
    if (!result || !images_host || num_images <= 0 || x <= 0 || y <= 0 || z <= 0) {
        set_error("conv2d_result_to_images: invalid arguments");
        return NB200_ERR_INVALID;
    }
    const int group = result->total_rows / num_images;
    if (result->total_rows % num_images != 0 || static_cast<long long>(x) * y != group || result->total_cols > z) {
        set_error("conv2d_result_to_images: %d rows do not split into %d images of %d x %d windows (or %d columns > %d channels)",
                  result->total_rows, num_images, x, y, result->total_cols, z);
        return NB200_ERR_SHAPE;
    }
    if (!result->f64 && !result->f32) { set_error("conv2d_result_to_images: set %s:%s holds no data", result->db.c_str(), result->name.c_str()); return NB200_ERR_STATE; }
    const size_t n = static_cast<size_t>(num_images) * z * group;
    double *dev = nullptr;
    NB_CUDA(nb_malloc(&dev, n * sizeof(double)));
    int st = result->f64 ? launch_result_to_images(c, result->f64, true, result->total_cols, num_images, group, result->total_cols, z, dev)
                         : launch_result_to_images(c, result->f32, false, result->f32_ld, num_images, group, result->total_cols, z, dev);
    if (st == NB200_OK && (cudaMemcpyAsync(images_host, dev, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
                           cudaStreamSynchronize(c->stream) != cudaSuccess)) st = NB200_ERR_CUDA;
    nb_free(dev);
    if (st == NB200_ERR_CUDA) set_error("conv2d_result_to_images: CUDA error %s", cudaGetErrorString(cudaGetLastError()));
    return st;
}

// the conv2d job on NHWC bf16 image planes: kernel planes, bias, the implicit-GEMM launch; result set (n * hout * wout) x kout
static int conv2d_fused_planes(Context *c, const __nv_bfloat16 *ihi, const __nv_bfloat16 *ilo, int n, int ch, int h, int w, int cp,
                               const double *kernels_host, const double *bias_host, int kout, int k, int padding, int relu, int nterms,
                               int block_rows, int block_cols, const char *db, const char *out_set, nb200_set **out, float *kernel_ms) {
    const int hout = h + 2 * padding - k + 1, wout = w + 2 * padding - k + 1;
    double *dker = nullptr;
    float *dbias = nullptr;
    Planes kp;
    nb200_set *o = nullptr;
    int st = make_set(db, out_set, n * hout * wout, kout, block_rows, block_cols, &o);
    const size_t on = static_cast<size_t>(n) * hout * wout * kout;
    const size_t kelems = static_cast<size_t>(kout) * ch * k * k;
    std::vector<float> bias32(kout);
    for (int i = 0; i < kout; i++) bias32[i] = static_cast<float>(bias_host[i]);
    if (st == NB200_OK && (nb_malloc(&dker, kelems * sizeof(double)) != cudaSuccess || nb_malloc(&dbias, static_cast<size_t>(kout) * sizeof(float)) != cudaSuccess ||
                           nb_malloc(&o->f32, (on ? on : 1) * sizeof(float)) != cudaSuccess)) {
        set_error("conv2d: out of device memory");
        st = NB200_ERR_NOMEM;
    }
    if (st == NB200_OK) {
        o->f32_ld = kout;
        if (cudaMemcpyAsync(dker, kernels_host, kelems * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
            cudaMemcpyAsync(dbias, bias32.data(), static_cast<size_t>(kout) * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
            set_error("conv2d: kernel upload failed");
            st = NB200_ERR_CUDA;
        }
    }
    if (st == NB200_OK) st = launch_kernel_to_planes(c, dker, kout, ch, k, cp, &kp, nterms == 3);
    if (kernel_ms) { c->time_ev0 = c->ev[6]; c->time_ev1 = c->ev[7]; }   // recorded right around the kernel launch
    if (st == NB200_OK) st = launch_conv_implicit(c, ihi, ilo, n, h, w, cp, kp, kout, k, padding, nterms, dbias, relu, o->f32);
    c->time_ev0 = c->time_ev1 = nullptr;
    if (st == NB200_OK) st = check_kernel_fault(c, "conv2d_fused");
    cudaStreamSynchronize(c->stream);   // bias32 / the caller's kernels were read asynchronously
    if (st == NB200_OK && kernel_ms) cudaEventElapsedTime(kernel_ms, c->ev[6], c->ev[7]);
    nb_free(dker); nb_free(dbias);
    free_planes(&kp);
    if (st != NB200_OK) return st;
    *out = o;
    return NB200_OK;
}

int nb200_conv2d_fused(const double *images_host, int n, int ch, int h, int w, const double *kernels_host,
                       const double *bias_host, int kout, int k, int padding, int relu, int precision, int block_rows,
                       int block_cols, const char *db, const char *out_set, nb200_set **out) {
    NEED_CTX();
    if (!images_host || !kernels_host || !bias_host || n <= 0 || ch <= 0 || h <= 0 || w <= 0 || kout <= 0 || k <= 0 ||
        padding < 0 || !db || !out_set || !out || precision < 0 || precision > 3) {
        set_error("conv2d_fused: invalid arguments");
        return NB200_ERR_INVALID;
    }
    const int hout = h + 2 * padding - k + 1, wout = w + 2 * padding - k + 1;
    if (hout <= 0 || wout <= 0) { set_error("conv2d_fused: kernel larger than image"); return NB200_ERR_SHAPE; }
    if (precision == NB200_PREC_FP64 || kout > 256) {
        // exact path (and very wide filters): materialised im2col on the device + the block GEMM
        nb200_set *img = nullptr, *ker = nullptr, *res = nullptr;
        NB_TRY(nb200_conv2d_image_flat(images_host, n, ch, h, w, k, 1, padding, block_rows, block_cols, db, "__image_flat", &img));
        NB_TRY(nb200_conv2d_kernel_flat(kernels_host, bias_host, kout, ch, k, block_rows, block_cols, db, "__kernel_flat", &ker));
        int st = gemm_blocks_impl(NB200_OP_NT, img, ker, NB200_EPI_NONE, nullptr, precision, db, out_set, &res, relu ? 1 : 0);
        nb200_set_destroy(img);
        nb200_set_destroy(ker);
        if (st != NB200_OK) return st;
        *out = res;
        return NB200_OK;
    }
    // implicit GEMM: image_flat is never built (K6).  bf16 (fp16 requests use bf16 planes here: the operands are
    // tiny and the kernel is HBM-bound on the result) or bf16x3.
    const int nterms = precision == NB200_PREC_BF16X3 ? 3 : 1;
    const int cp = static_cast<int>(round_up(ch, 64));
    const size_t img_elems = static_cast<size_t>(n) * ch * h * w, nhwc_elems = static_cast<size_t>(n) * h * w * cp;
    double *dimg = nullptr;
    __nv_bfloat16 *ihi = nullptr, *ilo = nullptr;
    int st = NB200_OK;
    if (nb_malloc(&dimg, img_elems * sizeof(double)) != cudaSuccess || nb_malloc(&ihi, nhwc_elems * sizeof(__nv_bfloat16)) != cudaSuccess ||
        (nterms == 3 && nb_malloc(&ilo, nhwc_elems * sizeof(__nv_bfloat16)) != cudaSuccess)) {
        set_error("conv2d_fused: out of device memory");
        st = NB200_ERR_NOMEM;
    }
    if (st == NB200_OK && cudaMemcpyAsync(dimg, images_host, img_elems * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
        set_error("conv2d_fused: image upload failed");
        st = NB200_ERR_CUDA;
    }
    if (st == NB200_OK) st = launch_nchw_to_nhwc_planes(c, dimg, n, ch, h, w, cp, ihi, ilo);
    if (st == NB200_OK)
        st = conv2d_fused_planes(c, ihi, ilo, n, ch, h, w, cp, kernels_host, bias_host, kout, k, padding, relu, nterms, block_rows, block_cols, db,
                                 out_set, out, nullptr);
    cudaStreamSynchronize(c->stream);
    nb_free(dimg); nb_free(ihi); nb_free(ilo);
    return st;
}

// The reference keeps the images as a set of Image records (conv2d_memory_fusion/headers/Image.h) that the `image_ops` job
// scans (PipelinedConv2dMemFuseTest.cc:136-180).  nb200_conv2d_images_create is that set on the device: the NCHW fp64 data
// as they were sent, plus -- derived on first use and kept until the set changes -- the NHWC bf16 operand planes the
// implicit GEMM reads.  A conv2d job on a resident image set then moves nothing but the result.
int nb200_conv2d_images_create(const double *images_host, int n, int ch, int h, int w, const char *db, const char *set, nb200_set **out) {
    NEED_CTX();
    if (!images_host || n <= 0 || ch <= 0 || h <= 0 || w <= 0 || !db || !set || !out) { set_error("conv2d_images_create: invalid arguments"); return NB200_ERR_INVALID; }
    nb200_set *s = nullptr;
    NB_TRY(make_set(db, set, n * ch * h, w, h, w, &s));   // as a matrix: one (h x w) block per (image, channel)
    const size_t elems = static_cast<size_t>(n) * ch * h * w;
    NB_CUDA(nb_malloc(&s->f64, elems * sizeof(double)));
    NB_CUDA(cudaMemcpyAsync(s->f64, images_host, elems * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));             // the caller's buffer is free again
    s->img_n = n; s->img_c = ch; s->img_h = h; s->img_w = w;
    *out = s;
    return NB200_OK;
}

int nb200_conv2d_fused_from_set(nb200_set *images, const double *kernels_host, const double *bias_host, int kout, int k, int padding,
                                int relu, int precision, int block_rows, int block_cols, const char *db, const char *out_set,
                                nb200_set **out, float *kernel_ms) {
    NEED_CTX();
    if (!images || images->img_n <= 0 || !images->f64 || !kernels_host || !bias_host || kout <= 0 || kout > 256 || k <= 0 || padding < 0 || !db ||
        !out_set || !out || precision == NB200_PREC_FP64 || precision < 0 || precision > 3) {
        set_error("conv2d_fused_from_set: invalid arguments (an image set of nb200_conv2d_images_create, kout <= 256, a tensor-core precision)");
        return NB200_ERR_INVALID;
    }
    NB_TRY(same_device(c, {images}));
    const int n = images->img_n, ch = images->img_c, h = images->img_h, w = images->img_w;
    if (h + 2 * padding - k + 1 <= 0 || w + 2 * padding - k + 1 <= 0) { set_error("conv2d_fused_from_set: kernel larger than image"); return NB200_ERR_SHAPE; }
    const int nterms = precision == NB200_PREC_BF16X3 ? 3 : 1;
    const int cp = static_cast<int>(round_up(ch, 64));
    const size_t nhwc = static_cast<size_t>(n) * h * w * cp;
    if (!images->img_hi || (nterms == 3 && !images->img_lo)) {   // derive the operand planes once
        nb_free(images->img_hi); nb_free(images->img_lo);
        images->img_hi = images->img_lo = nullptr;
        NB_CUDA(nb_malloc(&images->img_hi, nhwc * sizeof(__nv_bfloat16)));
        NB_CUDA(nb_malloc(&images->img_lo, nhwc * sizeof(__nv_bfloat16)));
        NB_TRY(launch_nchw_to_nhwc_planes(c, images->f64, n, ch, h, w, cp, images->img_hi, images->img_lo));
    }
    return conv2d_fused_planes(c, images->img_hi, nterms == 3 ? images->img_lo : nullptr, n, ch, h, w, cp, kernels_host, bias_host, kout, k, padding,
                               relu, nterms, block_rows, block_cols, db, out_set, out, kernel_ms);
}

}  // extern "C"
