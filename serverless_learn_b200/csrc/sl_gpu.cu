// sl_gpu.cu — the C ABI of include/sl_gpu.h: handle, staging, launches.
// There is no CPU fallback anywhere in this file: every entry point runs on the GPU
// or returns an error.
#include "../../include/sl_gpu.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "kernels_wide.cuh"
#include "kernels_multi.cuh"
#include "kernels_cluster.cuh"

namespace {

thread_local std::string g_create_error;

constexpr int kStageSlots = 4;
constexpr size_t kStageBytes = 8u << 20;
constexpr int kMaxStepsPerLaunch = 1 << 15;
constexpr size_t kDefaultShardCap = 256u << 20;

// ---- NCCL, resolved at run time (sl_comm_init_rank) -------------------------------
typedef struct ncclComm* ncclComm_t;
struct NcclUniqueId { char internal[128]; };
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, NcclUniqueId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclFloat64 = 8;  // ncclDataType_t::ncclFloat64
constexpr int kNcclSum = 0;      // ncclRedOp_t::ncclSum
NcclApi g_nccl;
std::mutex g_nccl_mu;

const char* load_nccl() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib) return nullptr;
    // a process that already imported torch has its bundled libnccl.so.2 mapped;
    // dlopen by soname returns that copy instead of loading a second NCCL.
    // SL_NCCL_LIB names the copy to use when the process will ALSO load another library that
    // bundles its own NCCL later (PyTorch does): two different libnccl.so.2 cannot coexist, so
    // whoever comes first must pick the bundled one.
    const char* names[] = {getenv("SL_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* n : names) {
        if (!n || !*n) continue;
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) return "libnccl.so.2 not found (dlopen)";
    NcclApi api;
    api.lib = lib;
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(lib, "ncclCommInitRank");
    api.AllReduce = (decltype(api.AllReduce))dlsym(lib, "ncclAllReduce");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(lib, "ncclCommDestroy");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy)
        return "NCCL symbols missing";
    g_nccl = api;
    return nullptr;
}

// ---- persistent-kernel variants -----------------------------------------------------
struct Variant {
    int ch, r, rs, ncw, xb, ws, kind;  // kind 0: lr_persist_kernel, 1: lr_wide_kernel
    const void* fn;
};
#define SL_VARIANT(CH, R, RS, NCW, XB, WS) \
    Variant { CH, R, RS, NCW, XB, WS, 0, \
              (const void*)slk::lr_persist_kernel<CH, R, RS, NCW, (XB) != 0, WS> }
// wide-row kernel: `r` holds the number of row groups, `rs` the rows a group takes per pass
#define SL_WIDE(CH, NCW, NG, R) \
    Variant { CH, NG, R, NCW, 1, 0, 1, (const void*)slk::lr_wide_kernel<CH, NCW, NG, R> }
const Variant kVariants[] = {
    SL_VARIANT(1, 16, 4, 8, 1, 0),    // 1: F <= 1024; measured best at F=1024, B/grid = 28
    SL_VARIANT(1, 16, 4, 16, 1, 0),   // 2: F <= 2048; two row groups of 16 rows
    SL_VARIANT(1, 32, 4, 8, 1, 0),    // 3: F <= 1024; a CTA's 28 rows of a step in one sub-batch
    SL_VARIANT(1, 8, 4, 16, 1, 0),    // 4: F <= 2048
    SL_VARIANT(2, 8, 4, 8, 1, 0),     // 5: F <= 2048
    SL_VARIANT(4, 4, 2, 8, 1, 0),     // 6: F <= 4096
    Variant{4, 4, 2, 16, 0, 0, 0, nullptr},   // 7-9: retired (the in-kernel wide modes that lr_wide_kernel replaced;
    Variant{8, 1, 1, 16, 0, 1, 0, nullptr},   //      numbers kept so that the later entries keep theirs)
    Variant{8, 2, 1, 8, 0, 1, 0, nullptr},
    SL_VARIANT(4, 4, 4, 14, 1, 2),    // 10: F <= 1024 as 7 row groups of 2 warps x 4 rows: 28 rows in one pass
    SL_WIDE(10, 14, 1, 1),            // 11: F <= 17920, wide-row kernel, 14+1 warps = 128 registers/thread
    SL_WIDE(5, 14, 1, 1),             // 12: F <= 8960,  wide-row kernel
    SL_WIDE(2, 16, 1, 1),             // 13: F <= 4096,  wide-row kernel
    SL_WIDE(10, 14, 2, 1),            // 14: F <= 8960,  two row groups of 7 warps
    SL_WIDE(5, 14, 2, 1),             // 15: F <= 4480,  two row groups of 7 warps
    SL_WIDE(4, 16, 2, 1),             // 16: F <= 4096,  two row groups of 8 warps
    SL_WIDE(4, 16, 4, 1),             // 17: F <= 2048,  four row groups of 4 warps
    SL_WIDE(8, 12, 4, 1),             // 18: F <= 3072,  four row groups of 3 warps
    SL_WIDE(2, 16, 4, 1),             // 19: F <= 1024,  four row groups of 4 warps
    SL_WIDE(5, 14, 1, 2),             // 20: F <= 8960,  one group, two rows per pass
    SL_WIDE(10, 14, 2, 2),            // 21: F <= 8960,  two groups, two rows per pass
    SL_WIDE(5, 14, 2, 2),             // 22: F <= 4480,  two groups, two rows per pass
    SL_WIDE(3, 14, 1, 2),             // 23: F <= 5376,  one group, two rows per pass
    SL_WIDE(4, 16, 4, 2),             // 24: F <= 2048,  four groups, two rows per pass
    SL_WIDE(2, 16, 2, 2),             // 25: F <= 2048,  two groups, two rows per pass
    SL_WIDE(10, 14, 1, 2),            // 26: F <= 17920, one group, two rows per pass
    SL_VARIANT(1, 16, 8, 8, 1, 0),    // 27: F <= 1024, 8-row ring slots
    SL_VARIANT(1, 16, 16, 8, 1, 0),   // 28: F <= 1024, one 16-row ring slot per pass: fewer barrier operations per
                                      //     row; measured best once a CTA has >= ~40 rows per step
    SL_VARIANT(2, 8, 8, 8, 1, 0),     // 29: F <= 2048, one 8-row ring slot per pass
};
// auto-selection: the first of these that is wide enough for F
const int kAutoOrder[] = {1, 29, 22, 14, 11};  // measured best per width (tools/sweep.py, profiles/r1_summary.md)
constexpr int kNumVariants = sizeof(kVariants) / sizeof(kVariants[0]);

// ---- cluster kernel (K1c) variants ---------------------------------------------------
struct ClVariant {
    int wpg, rg, r, rs, cs, warps;
    const void* fn;
    const void* fn_dbg;   // the same kernel with the timeline stamps compiled in (option debug_timeline)
};
#define SL_CLUSTER_P(WPG, RG, R, RS, CS, XW, CREG, XREG, PW) \
    ClVariant { WPG, RG, R, RS, CS, slk::cl_warps<WPG, RG, XW, CREG, PW>(), \
                (const void*)slk::lr_cluster_kernel<WPG, RG, R, RS, CS, XW, CREG, XREG, PW, false>, \
                (const void*)slk::lr_cluster_kernel<WPG, RG, R, RS, CS, XW, CREG, XREG, PW, true> }
#define SL_CLUSTER(WPG, RG, R, RS, CS, XW, CREG, XREG) SL_CLUSTER_P(WPG, RG, R, RS, CS, XW, CREG, XREG, 1)
const ClVariant kClVariants[] = {
    SL_CLUSTER(8, 2, 16, 4, 8, 2, 104, 64),  // 1: F <= 1024, two row groups of 16 rows: a CTA's 32 rows of config 3 in
                                             //    one pass; 20 warps, the compute warpgroups take 104 registers
    SL_CLUSTER(8, 2, 8, 4, 8, 4, 0, 0),      // 2: F <= 1024, two row groups of 8 rows
    SL_CLUSTER(8, 1, 16, 4, 8, 4, 0, 0),     // 3: F <= 1024, one row group (8 compute warps)
    SL_CLUSTER(4, 4, 8, 4, 8, 4, 0, 0),      // 4: F <= 512, four row groups of 8 rows
    SL_CLUSTER(8, 2, 16, 8, 4, 3, 104, 64),  // 5: as 1 in clusters of 4 (more SMs take part where 8 do not tile the GPCs), 8-row slots
    SL_CLUSTER(8, 2, 8, 4, 4, 4, 0, 0),      // 6: as 2 in clusters of 4
    SL_CLUSTER(8, 1, 16, 4, 4, 4, 0, 0),     // 7: as 3 in clusters of 4
    SL_CLUSTER(8, 2, 16, 8, 4, 4, 104, 32),  // 8: as 5 with four exchange warps (24 warps, 32 registers for the exchange)
    SL_CLUSTER(8, 1, 16, 16, 4, 4, 0, 0),    // 9: as 7 with one 16-row ring slot per pass
    SL_CLUSTER(8, 1, 16, 8, 4, 4, 0, 0),     // 10: as 7 with 8-row ring slots
    SL_CLUSTER(8, 2, 8, 8, 4, 4, 0, 0),      // 11: as 6 with one 8-row ring slot per pass
    SL_CLUSTER(8, 1, 16, 16, 8, 4, 0, 0),    // 12: as 3 with one 16-row ring slot per pass
    SL_CLUSTER(8, 2, 16, 16, 4, 3, 104, 64), // 13: as 5 with one 16-row slot per pass
    SL_CLUSTER(8, 2, 16, 8, 8, 2, 104, 64),  // 14: as 1 with 8-row slots
    SL_CLUSTER_P(8, 2, 16, 8, 4, 4, 104, 64, 0),  // 15: as 5 with four exchange warps, one of which also issues the row copies
    SL_CLUSTER_P(8, 2, 16, 8, 8, 4, 104, 64, 0),  // 16: the same in clusters of 8
    SL_CLUSTER(16, 1, 16, 4, 4, 3, 104, 64),      // 17: F <= 2048: one row group of 16 warps, 16 rows per pass
    SL_CLUSTER(16, 1, 16, 4, 4, 4, 0, 0),         // 18: the same without the register re-split, four exchange warps
};
constexpr int kNumClVariants = sizeof(kClVariants) / sizeof(kClVariants[0]);

}  // namespace

constexpr int kMaxRemoteStates = 16;
struct sl_handle {
    int device = 0;
    int D = 0, F = 0, B = 0;
    double eta = 0, lambda = 0;
    int num_sms = 0, coop = 0;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;

    double *model = nullptr, *old = nullptr;  // device fp64 [D]
    // exchange staging: pinned + mapped, the kernels read / write it directly
    double *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    int *h_err = nullptr, *d_err = nullptr;
    unsigned long long *h_sum = nullptr, *d_sum = nullptr;

    // shard
    unsigned char* shard = nullptr;
    size_t shard_cap = 0, shard_len = 0;
    int npart = 0;                       // >1 while sl_train_steps_multi runs: independent replicas, one grid partition each
    double* pmodel[8] = {};
    int vw = 0;                          // >1 while sl_train_steps_group runs: worker replicas trained as one job
    bool forbid_simple = false;          // group / multi calls: the three-launch path has no multi-shard mode
    const float* vX[8] = {};
    const float* vy[8] = {};
    size_t shard_off = 0;                // 64 when the sealed stream starts with an sl_shard_header
    unsigned long long shard_row0 = 0;   // from the header (0 for raw streams)
    // incoming stream (sl_shard_begin .. sl_shard_end) lands here while `shard` stays trainable
    unsigned char* in_buf = nullptr;
    size_t in_cap = 0, in_len = 0;
    long long n_rows = 0;
    bool shard_open = false, shard_sealed = false;
    unsigned char* stage[kStageSlots] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t stage_ev[kStageSlots] = {nullptr, nullptr, nullptr, nullptr};
    int stage_idx = 0;

    // training scratch
    long long step = 0;
    unsigned long long* acc = nullptr;   // [3][krep][Fp] gradient accumulators
    size_t acc_bytes = 0;
    unsigned* counter = nullptr;
    double* wmaster = nullptr;           // [grid][F] (WM=1 variants)
    float* w32buf = nullptr;             // [2][F] (K1w)
    size_t wmaster_bytes = 0;
    double* d_loss = nullptr;
    double* h_loss = nullptr;
    double* h_loss_dev = nullptr;        // device alias of h_loss (mapped): K1c writes its losses straight there
    size_t loss_cap = 0;
    float *resid = nullptr, *rowloss = nullptr, *gpart = nullptr;
    double* gsum = nullptr;
    int nsplit = 0, rows_per_split = 0;

    // options
    long long opt_grid = 0, opt_variant = 0, opt_path = 0, opt_steps_per_launch = 0;
    long long opt_dbg_step = -1, opt_ring = 0, opt_krep = 0, opt_dbg_skip = 0, opt_tma_chunk = 0, opt_block_rows = 0, opt_multi_mode = 0, opt_multi_lpw = 0;
    unsigned long long* loss_fix = nullptr;
    unsigned long long* dbg = nullptr;
    size_t dbg_words = 0;
    // info
    long long k1_launches = 0;
    long long launches = 0, path_used = 0, variant_used = 0, grid_used = 0, ns_used = 0,
              smem_used = 0;

    // NCCL
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
    // peer-memory exchange of the fused multi-GPU kernel
    unsigned char* xblock = nullptr;     // [inbox 2*8*F packets][fin 3*F u64]
    size_t xblock_bytes = 0;
    void* peer_base[8] = {nullptr};      // xblock of every rank as mapped here (own = xblock)
    int peer_n = 0, peer_rank = 0;       // 0 = not attached
    unsigned xflag_base = 0;
    // states of workers in OTHER processes, mapped here for the one-kernel gossip round (sl_state_attach)
    struct RemoteState { double* model = nullptr; double* old = nullptr; double lambda = 0; bool used = false; };
    RemoteState remote[kMaxRemoteStates];
    std::thread::id lent_by;             // thread that holds mu between sl_state_lend and sl_state_reclaim
    std::atomic<bool> lent{false};
    long long opt_sync_path = 0;         // 0 auto (fused when attached), 1 NCCL per step
    // cluster kernel (K1c)
    unsigned long long* cl_scratch = nullptr;  // [gacc 2F][gbase 2F][lacc 3 + pad][lbase 3 + pad]
    int cl_scratch_F = 0;
    bool cl_dirty = false;               // a launch aborted: the accumulators are re-zeroed before the next one
    long long cl_xstep = 0;              // steps taken through K1c so far (accumulator parity continues across launches)
    long long cl_gstep = 0;              // ... of them together with the attached group (cross-GPU buffer and tag)
    int cl_max_clusters[32] = {};        // per variant, for the shared-memory size in cl_smem_set
    size_t cl_smem_set[32] = {};
    int cl_coop = -1;                    // cooperative + cluster launch accepted? (-1 = not tried)
    // [record ev0] -> [kernel] -> [record ev1] as one instantiated graph: the three reach the GPU together, so
    // the device does not sit between the first time stamp and the kernel waiting for the host's launch call
    cudaGraph_t cl_graph = nullptr;
    cudaGraphExec_t cl_gexec = nullptr;
    cudaGraphNode_t cl_gnode = nullptr;
    int cl_graph_ok = -1;                // -1 not tried, 0 unusable here, 1 in use
    int cl_graph_variant = -1, cl_graph_grid = 0, cl_graph_threads = 0;
    size_t cl_graph_smem = 0;
    long long opt_cl_variant = 0, opt_clusters = 0, opt_poll_delay = -1, opt_poll_gap = -1, opt_xpoll_delay = -1, opt_xpoll_gap = -1, opt_cl_graph = 1, opt_wait_all = 0;
    // caller-provided symmetric allocation for K1c's all-rank words (sl_peer_attach_symmetric)
    unsigned long long* sym_buf[8] = {};
    unsigned long long* mc_gx = nullptr; // its multicast alias (one store reaches every rank), or nullptr
    int sym_n = 0;
    bool ipc_attached = false;

    std::mutex mu;       // replica state, shard, training stream: every entry point except the ingest ones below
    std::mutex in_mu;    // the incoming stream (in_buf, in_len, staging ring, copy_stream): sl_shard_begin / append, and
                         // sl_shard_end, which takes `mu` after it only for the swap — a ReceiveFile thread streams
                         // while the training thread is inside a launch (src/worker.cc:242-246)
    std::mutex err_mu;
    std::string err;
};

namespace {

int fail(sl_handle* h, int code, const char* fmt, ...);
// what the kernels left in the watchdog word: a range error is the caller's data, anything else a wait that gave up
int fail_watchdog(sl_handle* h, int code) {
    if (code == slk::kErrRange)
        return fail(h, SL_EINVAL, "a gradient partial left the fixed-point range of the exchange (features too large for "
                    "the 64-bit sums, or not finite): the model is not usable");
    return fail(h, SL_ETIMEOUT, "in-kernel wait timed out (watchdog code %d)", code);
}

int fail(sl_handle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) {
        std::lock_guard<std::mutex> lk(h->err_mu);
        h->err = buf;
    } else {
        g_create_error = buf;
    }
    return code;
}

#define CK(h, call)                                                                      \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return fail((h), SL_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                             \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int ensure_buf_cap(sl_handle* h, unsigned char** buf, size_t* cap_io, size_t len, size_t need,
                   bool exact = false) {
    if (need <= *cap_io) return SL_OK;
    size_t cap = std::max(need, std::max(*cap_io * 2, kDefaultShardCap));
    if (exact) cap = std::max<size_t>(need, 1u << 20);  // caller told us the size
    unsigned char* nb = nullptr;
    cudaError_t e = cudaMalloc(&nb, cap);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(h, SL_ENOMEM, "cudaMalloc(%zu) for the shard failed: %s", cap,
                    cudaGetErrorString(e));
    }
    if (*buf && len) {
        CK(h, cudaMemcpyAsync(nb, *buf, len, cudaMemcpyDeviceToDevice, h->copy_stream));
    }
    CK(h, cudaStreamSynchronize(h->copy_stream));  // also drains DMAs still aimed at the old buffer
    if (*buf) cudaFree(*buf);
    *buf = nb;
    *cap_io = cap;
    return SL_OK;
}

// geometry of the simple path
int ensure_simple_scratch(sl_handle* h) {
    if (h->resid) return SL_OK;
    const int col_blocks = (h->F + 127) / 128;
    int want = std::max(1, (4 * h->num_sms + col_blocks - 1) / col_blocks);
    want = std::min(want, h->B);
    h->rows_per_split = (h->B + want - 1) / want;
    h->nsplit = (h->B + h->rows_per_split - 1) / h->rows_per_split;
    CK(h, cudaMalloc(&h->resid, sizeof(float) * h->B));
    CK(h, cudaMalloc(&h->rowloss, sizeof(float) * h->B));
    CK(h, cudaMalloc(&h->gpart, sizeof(float) * (size_t)h->nsplit * h->F));
    CK(h, cudaMalloc(&h->gsum, sizeof(double) * (h->F + 1)));
    return SL_OK;
}

int ensure_loss_cap(sl_handle* h, size_t n) {
    if (n <= h->loss_cap) return SL_OK;
    if (h->d_loss) cudaFree(h->d_loss);
    if (h->loss_fix) cudaFree(h->loss_fix);
    h->loss_fix = nullptr;
    if (h->h_loss) cudaFreeHost(h->h_loss);
    h->d_loss = nullptr;
    h->h_loss = nullptr;
    h->loss_cap = 0;
    size_t cap = std::max<size_t>(n, 1024);
    CK(h, cudaMalloc(&h->d_loss, sizeof(double) * cap));
    CK(h, cudaMalloc(&h->loss_fix, 8 * cap));
    CK(h, cudaHostAlloc(&h->h_loss, sizeof(double) * cap, cudaHostAllocMapped));
    h->h_loss_dev = nullptr;
    if (cudaHostGetDevicePointer(&h->h_loss_dev, h->h_loss, 0) != cudaSuccess) {
        cudaGetLastError();
        h->h_loss_dev = nullptr;
    }
    h->loss_cap = cap;
    return SL_OK;
}

struct PersistPlan {
    int variant = -1;  // index into kVariants
    int grid = 0;
    slk::LrArgs args;
    size_t smem = 0;
};

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// float4 columns a variant can hold in registers (the wide kernel splits its warps into row groups)
int variant_granules(const Variant& v) { return v.ch * 32 * (v.kind == 1 ? v.ncw / v.r : v.ncw); }

// Returns false (with reason) when the persistent kernel cannot run this shape.
bool plan_persist(sl_handle* h, PersistPlan* p, std::string* why) {
    const int F = h->F;
    if (!h->coop) { *why = "device lacks cooperative launch"; return false; }
    if (F % 4 != 0) { *why = "F is not a multiple of 4 (16-byte rows needed)"; return false; }
    if (((uintptr_t)h->shard & 15) != 0) { *why = "shard not 16-byte aligned"; return false; }
    const int granules = F / 4;
    int vi = -1;
    if (h->opt_variant > 0 && h->opt_variant <= kNumVariants) {
        vi = (int)h->opt_variant - 1;
        if (!kVariants[vi].fn) { *why = "this variant number is retired"; return false; }
        if (granules > variant_granules(kVariants[vi])) {
            *why = "requested variant too narrow for F";
            return false;
        }
    } else {
        for (int k : kAutoOrder) {
            const int i = k - 1;
            if (granules <= variant_granules(kVariants[i])) { vi = i; break; }
        }
        if (vi < 0) { *why = "F too wide for the persistent kernel"; return false; }
        // many rows per CTA and step: every ring slot costs each warp a barrier wait and an arrival, so
        // one slot per pass wins (F=1024, B=16384: 14.3 vs 16.5 us/step); at config 3's 28 rows the
        // finer slots refill earlier and win (7.3 vs 7.75)
        {
            int g = h->num_sms;
            if (h->opt_grid > 0) g = (int)std::min<long long>(h->opt_grid, h->num_sms);
            if (h->npart > 1) g = std::max(1, g / h->npart);
            const long long beff = (long long)h->B * (h->vw > 1 ? h->vw : 1);
            if (vi == 0 && (beff + g - 1) / g >= 40) vi = 27;  // variant 28
        }
    }
    const Variant& v = kVariants[vi];
    int grid = h->num_sms;
    if (h->opt_grid > 0) grid = (int)std::min<long long>(h->opt_grid, h->num_sms);
    const int npart = h->npart > 1 ? h->npart : 1;
    if (npart > 1) {
        if (v.kind != 0 || v.ws != 0 || !v.xb) { *why = "replica partitions run on the register-row kernel only (F <= 2048)"; return false; }
        grid = grid / npart;  // CTAs per partition; the launch has npart of them
        if (grid < 1) { *why = "fewer SMs than replicas"; return false; }
    }
    slk::LrArgs& a = p->args;
    memset(&a, 0, sizeof(a));
    a.F = F;
    const long long Beff = (long long)h->B * (h->vw > 1 ? h->vw : 1);  // positions per step
    a.B = (int)Beff;
    a.vB = h->B;
    a.nvw = h->vw > 1 ? h->vw : 0;
    a.npart = npart > 1 ? npart : 0;
    a.Fp = F;
    a.zy_rows = (int)((((Beff + grid - 1) / grid + v.r - 1) / v.r) * v.r);
    a.krep = !v.xb ? 1 : (F <= 1024 ? 4 : (F <= 2048 ? 2 : 1));
    a.xslice = 32;
    if (v.kind == 1) {
        a.krep = 1;
        a.row_bytes = 4 * F;
        a.zy_rows = (int)((Beff + grid - 1) / grid);
        a.xslice = 4 * ((F + 4 * grid - 1) / (4 * grid));  // columns a slice owner sums and updates
        const size_t red_w = sizeof(float) * 2 * v.ncw * v.rs;
        const size_t zy_w = sizeof(float) * 2 * 2 * (size_t)a.zy_rows;
        // owner partial sums: groups x slice (a slice wider than the CTA's threads: one group)
        const size_t gout_w = std::max<size_t>(8 * (size_t)v.ncw * 32 * 4, 8 * (size_t)a.xslice) + 64;
        const size_t fixed_w = red_w + zy_w + gout_w + 512;
        if (h->smem_optin < fixed_w + 16 + (size_t)a.row_bytes) { *why = "row does not fit in shared memory"; return false; }
        size_t nsw = (h->smem_optin - fixed_w) / ((size_t)a.row_bytes + 16);
        nsw = std::min<size_t>(nsw, 32);  // one producer lane per slot
        if (h->opt_ring > 0) nsw = std::min<size_t>(nsw, (size_t)h->opt_ring);
        // a slot must always be consumed by the same row group (the parity waits on its barriers
        // cannot tell two rounds apart otherwise): ring length = a multiple of NG*R rows
        nsw -= nsw % ((size_t)v.r * v.rs);
        if (nsw < (size_t)v.r * v.rs) { *why = "ring too short for this variant's row groups"; return false; }
        a.ns = (int)nsw;
        size_t offw = (size_t)a.ns * a.row_bytes;
        offw = align_up(offw, 16); a.off_bar = (int)offw;   offw += 8 * (2 * nsw + 3);
        offw = align_up(offw, 16); a.off_red = (int)offw;   offw += red_w;
        offw = align_up(offw, 16); a.off_zy = (int)offw;    offw += zy_w;
        offw = align_up(offw, 128); a.off_gout = (int)offw; offw += gout_w;
        p->smem = offw;
        p->variant = vi;
        p->grid = grid;
        return true;
    }
    if (v.xb && h->opt_krep > 0) a.krep = (int)h->opt_krep;
    a.wpg = (granules + v.ch * 32 - 1) / (v.ch * 32);
    a.row_bytes = 4 * F;
    const int rg = v.ncw / a.wpg;
    // fixed shared-memory regions after the ring
    const size_t red_b = sizeof(float) * 2 * v.ncw * v.r;
    const size_t zy_b = sizeof(float) * 2 * 2 * (size_t)a.zy_rows;
    const size_t gsm_b = rg > 1 ? sizeof(float) * (size_t)rg * F : 0;
    const size_t gout_b = v.xb ? 8 * (size_t)a.Fp : 0;
    const size_t gin_b = v.xb ? 8 * (size_t)a.Fp * a.krep : 0;
    const size_t wsm_b = v.ws ? sizeof(float) * (size_t)F : 0;
    const size_t msm_b = v.ws == 2 ? sizeof(double) * (size_t)F : 0;
    const size_t fixed = red_b + zy_b + gsm_b + gout_b + gin_b + wsm_b + msm_b + 512;
    const size_t slot_b = (size_t)v.rs * a.row_bytes;
    if (slot_b > (1u << 20) - 16 || h->smem_optin < fixed + 16 + slot_b) {
        *why = "a ring slot does not fit in shared memory";
        return false;
    }
    size_t ns = (h->smem_optin - fixed) / (slot_b + 16);
    ns = std::min<size_t>(ns, 256);
    if (h->opt_ring > 0) ns = std::min<size_t>(ns, (size_t)h->opt_ring);
    if (ns < 1) { *why = "no ring slot fits"; return false; }
    {   // with several row groups a slot should always be consumed by the same group (see K1w)
        const size_t per_round = (size_t)rg * (size_t)(v.r / v.rs);
        if (rg > 1 && ns >= per_round) ns -= ns % per_round;
    }
    a.ns = (int)ns;
    size_t off = (size_t)a.ns * slot_b;
    off = align_up(off, 16); a.off_bar = (int)off;   off += 8 * (2 * ns + 3);
    off = align_up(off, 16); a.off_red = (int)off;   off += red_b;
    off = align_up(off, 16); a.off_zy = (int)off;    off += zy_b;
    off = align_up(off, 16); a.off_gsm = (int)off;   off += gsm_b;
    off = align_up(off, 128); a.off_gout = (int)off; off += gout_b;
    off = align_up(off, 128); a.off_gin = (int)off;  off += gin_b;
    off = align_up(off, 16); a.off_wsm = (int)off;   off += wsm_b;
    off = align_up(off, 16); a.off_msm = (int)off;   off += msm_b;
    p->smem = off;
    p->variant = vi;
    p->grid = grid * npart;  // geometry above is per partition
    return true;
}

int ensure_exchange(sl_handle* h, const PersistPlan& plan) {
    const Variant& v = kVariants[plan.variant];
    const size_t need = v.kind == 1 ? 4 * (size_t)plan.grid * v.r * plan.args.Fp   // K1w: one fp32 partial vector per row group
                                    : 8 * (size_t)3 * plan.args.krep * plan.args.Fp * (plan.args.npart > 1 ? plan.args.npart : 1);
    if (h->acc_bytes < need) {
        if (h->acc) cudaFree(h->acc);
        h->acc = nullptr;
        h->acc_bytes = 0;
        CK(h, cudaMalloc(&h->acc, need));
        h->acc_bytes = need;
    }
    if (!h->counter) CK(h, cudaMalloc(&h->counter, 256));
    if (v.kind == 1 && !h->w32buf) CK(h, cudaMalloc(&h->w32buf, sizeof(float) * 2 * (size_t)h->F));
    if (v.ws == 1) {
        const size_t wb = sizeof(double) * (size_t)plan.grid * h->F;
        if (h->wmaster_bytes < wb) {
            if (h->wmaster) cudaFree(h->wmaster);
            h->wmaster = nullptr;
            h->wmaster_bytes = 0;
            CK(h, cudaMalloc(&h->wmaster, wb));
            h->wmaster_bytes = wb;
        }
    }
    return SL_OK;
}

int run_persist(sl_handle* h, const PersistPlan& plan, int n_steps, bool want_loss,
                bool fused_sync = false) {
    const Variant& v = kVariants[plan.variant];
    int rc = ensure_exchange(h, plan);
    if (rc) return rc;
    CK(h, cudaFuncSetAttribute(v.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    int occ = 0;
    CK(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, v.fn, (v.ncw + 2 - v.kind) * 32, plan.smem));
    if (occ < 1) return fail(h, SL_ECUDA, "persistent kernel does not fit on an SM");
    int max_per_launch =
        h->opt_steps_per_launch > 0 ? (int)h->opt_steps_per_launch : kMaxStepsPerLaunch;
    {   // the kernel counts its streamed rows in 32 bits
        const long long rows_per_cta = (plan.args.B + plan.grid - 1) / plan.grid;  // >= stages per CTA
        const long long cap = ((1ll << 31) - 1 - plan.args.ns) / std::max(1ll, rows_per_cta);
        max_per_launch = (int)std::max(1ll, std::min<long long>(max_per_launch, cap));
    }
    int done = 0;
    *h->h_err = 0;
    const int np = plan.args.npart > 1 ? plan.args.npart : 1;
    if (want_loss) CK(h, cudaMemsetAsync(h->loss_fix, 0, 8 * (size_t)n_steps * np, h->stream));
    while (done < n_steps) {
        const int n = std::min(max_per_launch, n_steps - done);
        slk::LrArgs a = plan.args;
        a.X = reinterpret_cast<const float*>(h->shard + h->shard_off);
        a.y = a.X + (size_t)h->n_rows * h->F;
        a.n_rows = h->n_rows;
        a.step0 = h->step;
        a.n_steps = n;
        a.model = h->model;
        a.acc = h->acc;
        a.counter = h->counter;
        a.wmaster = h->wmaster;
        a.w32buf = h->w32buf;
        if (kVariants[plan.variant].kind == 1) {  // the wide kernel always runs its owner stage
            a.counter2 = h->counter + 16;
            a.n_own = (h->F + plan.args.xslice - 1) / plan.args.xslice;
        }
        a.loss_fix = want_loss ? h->loss_fix + done : nullptr;
        // accumulators and the arrival counter start every launch at zero
        if (kVariants[plan.variant].kind == 0)
            CK(h, cudaMemsetAsync(h->acc, 0, 8 * (size_t)3 * a.krep * a.Fp * np, h->stream));
        CK(h, cudaMemsetAsync(h->counter, 0, 256, h->stream));
        a.eta_over_b = h->eta / (double)a.B;
        a.fix_limit = 134217728.0f / (float)((long long)plan.grid * (kVariants[plan.variant].kind == 1 ? kVariants[plan.variant].r : 1) * (fused_sync ? std::max(1, h->peer_n) : 1));
        for (int k = 0; k < 8; ++k) { a.vX[k] = h->vX[k]; a.vy[k] = h->vy[k]; a.pmodel[k] = h->pmodel[k]; }
        a.loss_stride = n_steps;
        a.block_rows = h->opt_block_rows != 0 ? 1 : 0;
        a.dbg_skip = h->opt_dbg_skip != 0 ? 1 : 0;
        a.tma_chunk = h->opt_tma_chunk > 0 ? (int)((h->opt_tma_chunk + 15) / 16 * 16) : (1 << 20);
        if (fused_sync) {
            a.nranks = h->peer_n;
            a.rank = h->peer_rank;
            for (int p = 0; p < h->peer_n; ++p) a.peer_x[p] = reinterpret_cast<uint4*>(h->peer_base[p]);
            a.fin = reinterpret_cast<unsigned long long*>(h->xblock + sizeof(uint4) * 2 * 8 * (size_t)h->F);
            a.counter2 = h->counter + 16;
            a.xflag0 = h->xflag_base;
            a.n_own = (h->F + plan.args.xslice - 1) / plan.args.xslice;
            a.eta_over_b = h->eta / ((double)h->B * (double)h->peer_n);
        }
        a.err = h->d_err;
        a.dbg = nullptr;
        a.dbg_step = 0;
        if (h->opt_dbg_step >= 0 && done == 0) {
            const size_t words = (size_t)4 * plan.grid * 16;
            if (h->dbg_words < words) {
                if (h->dbg) cudaFree(h->dbg);
                h->dbg = nullptr;
                CK(h, cudaMalloc(&h->dbg, words * 8));
                h->dbg_words = words;
            }
            CK(h, cudaMemsetAsync(h->dbg, 0, words * 8, h->stream));
            a.dbg = h->dbg;
            a.dbg_step = (int)h->opt_dbg_step;
        }
        void* params[] = {&a};
        CK(h, cudaLaunchCooperativeKernel(v.fn, dim3(plan.grid), dim3((v.ncw + 2 - v.kind) * 32), params,
                                          plan.smem, h->stream));
        if (fused_sync) h->xflag_base += (unsigned)n;  // only a launch that went out consumes packet flags
        h->launches++;
        h->k1_launches++;
        h->step += n;
        done += n;
    }
    if (want_loss) {
        slk::loss_finish_kernel<<<(n_steps * np + 255) / 256, 256, 0, h->stream>>>(
            h->loss_fix, h->d_loss, n_steps * np, 1.0 / (double)plan.args.B);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    h->path_used = fused_sync ? 4 : 1;
    h->variant_used = plan.variant + 1;
    h->grid_used = plan.grid;
    h->ns_used = plan.args.ns;
    h->smem_used = (long long)plan.smem;
    return SL_OK;
}


// ---- K1c: cluster kernel (kernels_cluster.cuh) -----------------------------------------
struct ClusterPlan {
    int variant = -1;  // index into kClVariants
    int grid = 0, threads = 0;
    slk::ClArgs args;
    size_t smem = 0;
};

// Largest number of co-resident clusters of variant vi (one CTA per SM), cached on the handle.
int cluster_capacity(sl_handle* h, int vi) {
    if (h->cl_smem_set[vi] == h->smem_optin) return h->cl_max_clusters[vi];
    const ClVariant& v = kClVariants[vi];
    h->cl_max_clusters[vi] = 0;
    h->cl_smem_set[vi] = h->smem_optin;
    for (const void* fn : {v.fn, v.fn_dbg}) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_optin) != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
        if (v.cs > 8 && cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(v.cs * h->num_sms));
    cfg.blockDim = dim3((unsigned)(v.warps * 32));
    cfg.dynamicSmemBytes = h->smem_optin;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)v.cs;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, v.fn, &cfg) != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    h->cl_max_clusters[vi] = n;
    return n;
}

// Returns false (with reason) when the cluster kernel cannot run this shape.
bool plan_cluster(sl_handle* h, ClusterPlan* p, std::string* why, bool fused) {
    const int F = h->F;
    if (h->npart > 1) { *why = "independent replicas run on the interleaved kernel"; return false; }
    if (((uintptr_t)(h->shard + h->shard_off) & 15) != 0) { *why = "shard not 16-byte aligned"; return false; }
    if (fused && F > 1024) { *why = "rows wider than 1024 floats exchange across GPUs in the older kernels"; return false; }
    int vi = -1;
    auto fits = [&](const ClVariant& v) { return F % (4 * v.cs) == 0 && F <= 128 * v.wpg; };
    if (h->opt_cl_variant > 0 && h->opt_cl_variant <= kNumClVariants) {
        vi = (int)h->opt_cl_variant - 1;
        if (!fits(kClVariants[vi])) { *why = "requested cluster variant does not hold this F"; return false; }
    } else {
        // measured at config 3 (tools/k1c_bench.py, profiles/r2_k1c_timeline.md): clusters of 4 put 132 of the 148 SMs to
        // work (clusters of 8 only tile 120), and 16 compute warps take a CTA's 32 rows of a step in one pass.
        // Variant 15 (row copies issued by an exchange warp, 20 warps) measures within a few percent of 5 either
        // way from box to box; 5 also serves shapes with several passes per step.
        vi = F <= 512 ? 3 : 4;
        if (!fits(kClVariants[vi]) || cluster_capacity(h, vi) < 1) {
            vi = -1;
            for (int k = 0; k < kNumClVariants; ++k)
                if (fits(kClVariants[k]) && cluster_capacity(h, k) >= 1) { vi = k; break; }
        }
        if (vi < 0) { *why = "F must be a multiple of 16 and at most 2048 for the cluster kernel"; return false; }
        // rows wider than 1024 floats (variant 17: 16 rows per pass): measured ahead of the older kernel up to three
        // passes per step (F = 2048: 8.5 against 10.7 us at B = 4096), behind it from four on (15.5 against 15.0 at 8192)
        if (F > 1024 && h->opt_path == 0) {
            const long long Bq = (long long)h->B * (h->vw > 1 ? h->vw : 1);
            const long long per_cta = (Bq + (long long)cluster_capacity(h, vi) * kClVariants[vi].cs - 1) /
                                      ((long long)cluster_capacity(h, vi) * kClVariants[vi].cs);
            if (per_cta > 48) { *why = "wide rows with more than three passes per step run on the older kernel"; return false; }
        }
    }
    const ClVariant& v = kClVariants[vi];
    int ncl = cluster_capacity(h, vi);
    if (ncl < 1) { *why = "no cluster of this kernel fits on the device"; return false; }
    if (h->opt_clusters > 0) ncl = (int)std::min<long long>(ncl, h->opt_clusters);
    else if (h->opt_grid > 0) ncl = (int)std::max<long long>(1, std::min<long long>(ncl, h->opt_grid / v.cs));
    ncl = std::min(ncl, (int)slk::kClCntMask);
    slk::ClArgs& a = p->args;
    memset(&a, 0, sizeof(a));
    const long long Beff = (long long)h->B * (h->vw > 1 ? h->vw : 1);  // positions per step
    const int grid = ncl * v.cs;
    a.F = F;
    a.B = (int)Beff;
    a.vB = h->B;
    a.nvw = h->vw > 1 ? h->vw : 0;
    a.ncl = ncl;
    a.fc = F / v.cs;
    a.row_bytes = 4 * F;
    a.zy_rows = (int)((((Beff + grid - 1) / grid + v.r - 1) / v.r) * v.r);
    const int ncw = v.wpg * v.rg, nsrc = v.cs * v.rg;
    const size_t recv_b = sizeof(float) * (size_t)nsrc * a.fc;
    const size_t wsm_b = sizeof(float) * 2 * (size_t)F;
    const size_t lrecv_b = sizeof(float) * 2 * (size_t)nsrc;
    const size_t red_b = sizeof(float) * 2 * (size_t)ncw * v.r;
    const size_t zy_b = sizeof(float) * 2 * 2 * (size_t)a.zy_rows;
    a.ycap = (a.zy_rows / v.r + v.rg - 1) / v.rg * v.r;  // passes per row group * R
    const size_t yb_b = sizeof(float) * 2 * (size_t)v.rg * a.ycap;
    const size_t slot_b = (size_t)v.rs * a.row_bytes;
    const size_t prev_b = 8 * (2 * (size_t)a.fc + 4);  // + the loss accumulators' values
    const size_t fixed = recv_b + wsm_b + lrecv_b + red_b + zy_b + yb_b + prev_b + 8 * (2 * 32 + 6 + 16) + 512;
    if (h->smem_optin < fixed + (size_t)v.rg * slot_b) { *why = "the ring does not fit in shared memory"; return false; }
    size_t nsg = (h->smem_optin - fixed) / ((size_t)v.rg * slot_b);
    nsg = std::min<size_t>(nsg, (size_t)(32 / v.rg));  // one producer lane per slot
    if (h->opt_ring > 0) nsg = std::max<size_t>(1, std::min<size_t>(nsg, (size_t)h->opt_ring));
    a.nsg = (int)nsg;
    size_t off = (size_t)v.rg * nsg * slot_b;
    off = align_up(off, 16); a.off_bar = (int)off;    off += 8 * (2 * (size_t)v.rg * nsg + 6 + (size_t)v.cs);
    off = align_up(off, 16); a.off_recv = (int)off;   off += recv_b;
    off = align_up(off, 16); a.off_wsm = (int)off;    off += wsm_b;
    off = align_up(off, 16); a.off_lrecv = (int)off;  off += lrecv_b;
    off = align_up(off, 16); a.off_red = (int)off;    off += red_b;
    off = align_up(off, 16); a.off_zy = (int)off;     off += zy_b;
    off = align_up(off, 16); a.off_yb = (int)off;     off += yb_b;
    off = align_up(off, 16); a.off_prev = (int)off;   off += prev_b;
    if (off > h->smem_optin) { *why = "shared memory layout exceeds the SM"; return false; }
    const int ranks = fused ? std::max(1, h->peer_n) : 1;
    a.part_limit = 16777216.0f / (float)((long long)ncl * ranks);  // 2^(63-7-32) over all adders (one per cluster and rank)
    a.eta_over_b = h->eta / ((double)Beff * (double)ranks);
    a.inv_b = 1.0 / (double)Beff;
    a.poll_delay = h->opt_poll_delay >= 0 ? (int)h->opt_poll_delay : 0;
    a.poll_gap = h->opt_poll_gap >= 0 ? (int)h->opt_poll_gap : 0;
    a.xpoll_delay = h->opt_xpoll_delay >= 0 ? (int)h->opt_xpoll_delay : 0;
    a.xpoll_gap = h->opt_xpoll_gap >= 0 ? (int)h->opt_xpoll_gap : 0;
    a.wait_all_owners = h->opt_wait_all != 0 ? 1 : 0;
    p->smem = off;
    p->variant = vi;
    p->grid = grid;
    p->threads = v.warps * 32;
    return true;
}

size_t xblock_gx_offset(int F) { return sizeof(uint4) * 2 * 8 * (size_t)F + 8 * 3 * (size_t)F; }
size_t xblock_gx_bytes(int F) { return 8 * 2 * 8 * (size_t)F; }  // K1c: [2 buffers][8 ranks][F] tagged totals
size_t xblock_bytes_for(int F) { return xblock_gx_offset(F) + xblock_gx_bytes(F); }

// One launch of the cluster kernel between ev0 and ev1 as an instantiated graph (see sl_handle::cl_graph).
// Returns false when graphs cannot carry this launch here; the caller then launches directly.
bool cluster_graph_launch(sl_handle* h, const ClusterPlan& plan, const ClVariant& v, const void* kfn, slk::ClArgs* a) {
    if (h->cl_graph_ok == 0) return false;
    void* params[] = {a};
    cudaKernelNodeParams kp;
    memset(&kp, 0, sizeof(kp));
    kp.func = const_cast<void*>(kfn);
    const int gkey = plan.variant * 2 + (kfn == v.fn_dbg ? 1 : 0);
    kp.gridDim = dim3((unsigned)plan.grid);
    kp.blockDim = dim3((unsigned)plan.threads);
    kp.sharedMemBytes = (unsigned)plan.smem;
    kp.kernelParams = params;
    const bool same = h->cl_gexec && h->cl_graph_variant == gkey && h->cl_graph_grid == plan.grid &&
                      h->cl_graph_threads == plan.threads && h->cl_graph_smem == plan.smem;
    auto give_up = [&]() {
        cudaGetLastError();
        if (h->cl_gexec) cudaGraphExecDestroy(h->cl_gexec);
        if (h->cl_graph) cudaGraphDestroy(h->cl_graph);
        h->cl_gexec = nullptr;
        h->cl_graph = nullptr;
        h->cl_graph_ok = 0;
        return false;
    };
    if (!same) {
        if (h->cl_gexec) cudaGraphExecDestroy(h->cl_gexec);
        if (h->cl_graph) cudaGraphDestroy(h->cl_graph);
        h->cl_gexec = nullptr;
        h->cl_graph = nullptr;
        cudaGraphNode_t n0 = nullptr, n2 = nullptr;
        if (cudaGraphCreate(&h->cl_graph, 0) != cudaSuccess) return give_up();
        if (cudaGraphAddEventRecordNode(&n0, h->cl_graph, nullptr, 0, h->ev0) != cudaSuccess) return give_up();
        if (cudaGraphAddKernelNode(&h->cl_gnode, h->cl_graph, &n0, 1, &kp) != cudaSuccess) return give_up();
        cudaLaunchAttributeValue val;
        memset(&val, 0, sizeof(val));
        val.clusterDim.x = (unsigned)v.cs;
        val.clusterDim.y = 1;
        val.clusterDim.z = 1;
        if (cudaGraphKernelNodeSetAttribute(h->cl_gnode, cudaLaunchAttributeClusterDimension, &val) != cudaSuccess) return give_up();
        if (h->cl_coop != 0) {
            memset(&val, 0, sizeof(val));
            val.cooperative = 1;
            if (cudaGraphKernelNodeSetAttribute(h->cl_gnode, cudaLaunchAttributeCooperative, &val) != cudaSuccess) cudaGetLastError();
        }
        if (cudaGraphAddEventRecordNode(&n2, h->cl_graph, &h->cl_gnode, 1, h->ev1) != cudaSuccess) return give_up();
        if (cudaGraphInstantiate(&h->cl_gexec, h->cl_graph, 0) != cudaSuccess) return give_up();
        h->cl_graph_variant = gkey;
        h->cl_graph_grid = plan.grid;
        h->cl_graph_threads = plan.threads;
        h->cl_graph_smem = plan.smem;
    } else if (cudaGraphExecKernelNodeSetParams(h->cl_gexec, h->cl_gnode, &kp) != cudaSuccess) {
        return give_up();
    }
    if (cudaGraphLaunch(h->cl_gexec, h->stream) != cudaSuccess) return give_up();
    h->cl_graph_ok = 1;
    return true;
}

// Enqueues the launches (events around them); the caller synchronises.  loss_dev: device-visible
// array of n_steps doubles the kernel writes, or nullptr.
int run_cluster(sl_handle* h, const ClusterPlan& plan, int n_steps, double* loss_dev, bool fused) {
    const ClVariant& v = kClVariants[plan.variant];
    const int F = h->F;
    const size_t words = 4 * (size_t)F + 8;  // gacc, gbase [2][F]; lacc [3] + pad; lbase [3] + pad
    if (!h->cl_scratch || h->cl_scratch_F != F) {
        if (h->cl_scratch) cudaFree(h->cl_scratch);
        h->cl_scratch = nullptr;
        CK(h, cudaMalloc(&h->cl_scratch, 8 * words));
        h->cl_scratch_F = F;
        h->cl_dirty = true;
    }
    if (fused && h->cl_xstep != h->cl_gstep) {
        // this replica has taken steps its peers have not (alone before it joined, or between two synchronous
        // calls): its accumulators start afresh at the group's count — buffer and tag of the cross-GPU words
        // follow xstep0, which must agree on all ranks
        h->cl_xstep = h->cl_gstep;
        h->cl_dirty = true;
    }
    if (h->cl_dirty) {  // first use, or a launch gave up half way: start from zero (rare; costs a memset)
        CK(h, cudaMemsetAsync(h->cl_scratch, 0, 8 * words, h->stream));
        h->cl_dirty = false;
    }
    const int max_per_launch = h->opt_steps_per_launch > 0 ? (int)h->opt_steps_per_launch : kMaxStepsPerLaunch;
    *h->h_err = 0;
    unsigned long long* dbg = nullptr;
    if (h->opt_dbg_step >= 0) {
        const size_t dw = (size_t)4 * plan.grid * 32;  // K1c stamps 16 points
        if (h->dbg_words < dw) {
            if (h->dbg) cudaFree(h->dbg);
            h->dbg = nullptr;
            CK(h, cudaMalloc(&h->dbg, dw * 8));
            h->dbg_words = dw;
        }
        CK(h, cudaMemsetAsync(h->dbg, 0, dw * 8, h->stream));
        dbg = h->dbg;
    }
    const void* kfn = dbg ? v.fn_dbg : v.fn;   // the timeline stamps exist only in the debug instantiation
    const bool one_graph = n_steps <= max_per_launch && !dbg && h->opt_cl_graph != 0;
    bool in_graph = false;
    if (!one_graph) CK(h, cudaEventRecord(h->ev0, h->stream));
    int done = 0;
    while (done < n_steps) {
        const int n = std::min(max_per_launch, n_steps - done);
        slk::ClArgs a = plan.args;
        a.X = reinterpret_cast<const float*>(h->shard + h->shard_off);
        a.y = a.X + (size_t)h->n_rows * F;
        a.n_rows = h->n_rows;
        a.step0 = h->step;
        a.n_steps = n;
        a.model = h->model;
        a.gacc = h->cl_scratch;
        a.gbase = h->cl_scratch + 2 * (size_t)F;
        a.lacc = h->cl_scratch + 4 * (size_t)F;
        a.lbase = a.lacc + 4;
        a.loss_out = loss_dev ? loss_dev + done : nullptr;
        a.xstep0 = h->cl_xstep;
        a.err = h->d_err;
        for (int k = 0; k < 8; ++k) { a.vX[k] = h->vX[k]; a.vy[k] = h->vy[k]; }
        if (fused) {
            a.nranks = h->peer_n;
            a.rank = h->peer_rank;
            for (int p = 0; p < h->peer_n; ++p)
                a.peer_gx[p] = h->sym_n > 0 ? h->sym_buf[p]
                                            : reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(h->peer_base[p]) + xblock_gx_offset(F));
            a.mc_gx = h->sym_n > 0 ? h->mc_gx : nullptr;
        }
        a.dbg = done == 0 ? dbg : nullptr;
        a.dbg_step = done == 0 && dbg ? (int)h->opt_dbg_step : 0;
        if (one_graph && cluster_graph_launch(h, plan, v, kfn, &a)) {
            in_graph = true;
            h->launches++;
            h->k1_launches++;
            h->step += n;
            h->cl_xstep += n;
            if (fused) h->cl_gstep += n;
            done += n;
            continue;
        }
        if (one_graph && done == 0) CK(h, cudaEventRecord(h->ev0, h->stream));
        void* params[] = {&a};
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)plan.grid);
        cfg.blockDim = dim3((unsigned)plan.threads);
        cfg.dynamicSmemBytes = plan.smem;
        cfg.stream = h->stream;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)v.cs;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeCooperative;
        at[1].val.cooperative = 1;
        cfg.attrs = at;
        // the clusters wait for one another: ask for a co-resident grid where the driver takes both attributes
        cudaError_t e = cudaErrorUnknown;
        if (h->cl_coop != 0) {
            cfg.numAttrs = 2;
            e = cudaLaunchKernelExC(&cfg, kfn, params);
            if (h->cl_coop < 0) {
                h->cl_coop = e == cudaSuccess ? 1 : 0;
                if (e != cudaSuccess) cudaGetLastError();
            }
        }
        if (h->cl_coop == 0) {
            cfg.numAttrs = 1;
            e = cudaLaunchKernelExC(&cfg, kfn, params);
        }
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(h, SL_ECUDA, "cluster kernel launch failed: %s", cudaGetErrorString(e));
        }
        h->launches++;
        h->k1_launches++;
        h->step += n;
        h->cl_xstep += n;
        if (fused) h->cl_gstep += n;
        done += n;
    }
    if (!in_graph) CK(h, cudaEventRecord(h->ev1, h->stream));
    h->path_used = fused ? 7 : 6;
    h->variant_used = 100 + plan.variant + 1;
    h->grid_used = plan.grid;
    h->ns_used = plan.args.nsg;
    h->smem_used = (long long)plan.smem;
    return SL_OK;
}

// one step of the simple path; sync != 0 adds the all-reduce between finish and apply
int launch_simple_step(sl_handle* h, double* d_loss_slot, bool sync) {
    const float* X = reinterpret_cast<const float*>(h->shard + h->shard_off);
    const float* y = X + (size_t)h->n_rows * h->F;
    const long long mb = (long long)(((__int128)h->step * h->B) % h->n_rows);
    const int wpb = 8;
    slk::lr_fwd_kernel<<<(h->B + wpb - 1) / wpb, wpb * 32, 0, h->stream>>>(
        X, y, h->n_rows, h->F, h->B, mb, h->model, h->resid, h->rowloss);
    dim3 gg((h->F + 127) / 128, h->nsplit);
    slk::lr_grad_kernel<<<gg, 128, 0, h->stream>>>(X, h->n_rows, h->F, h->B, mb, h->resid, h->gpart,
                                                   h->rows_per_split);
    const int fb = (h->F + 255) / 256;
    const double eob = h->eta / (double)h->B;
    slk::lr_finish_kernel<<<fb, 256, 0, h->stream>>>(h->gpart, h->nsplit, h->F, h->B, h->rowloss,
                                                     h->model, eob, d_loss_slot, h->gsum,
                                                     sync ? 1 : 0);
    h->launches += 3;
    if (sync) {
        if (!h->comm || !g_nccl.AllReduce) return fail(h, SL_ESTATE, "no NCCL communicator (sl_comm_init_rank)");
        int nrc = g_nccl.AllReduce(h->gsum, h->gsum, (size_t)h->F + 1, kNcclFloat64, kNcclSum,
                                   h->comm, h->stream);
        if (nrc != 0)
            return fail(h, SL_ENCCL, "ncclAllReduce failed: %s",
                        g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "?");
        slk::lr_apply_avg_kernel<<<fb, 256, 0, h->stream>>>(
            h->model, h->gsum, h->F, eob / (double)h->nranks, d_loss_slot,
            1.0 / ((double)h->B * (double)h->nranks));
        h->launches += 1;
    }
    CK(h, cudaGetLastError());
    h->step += 1;
    return SL_OK;
}

// K1m: NM <= 4 independent replicas interleaved on every SM (kernels_multi.cuh).  F <= 1024.
constexpr int kIlvR = 16, kIlvRS = 16, kIlvNCW = 8;
bool interleaved_ok(const sl_handle* h) {
    return h->npart >= 2 && h->npart <= 4 && h->F % 4 == 0 && h->F <= 4 * 32 * kIlvNCW && h->coop &&
           h->opt_multi_mode == 0 && h->opt_variant == 0;
}
int run_interleaved(sl_handle* h, int n_steps, bool want_loss) {
    const int NM = h->npart, F = h->F;
    // exchange lanes per warp: 1 = a warp per model (experiment knob "multi_lpw")
    const int lpw = h->opt_multi_lpw > 0 ? (int)std::min<long long>(h->opt_multi_lpw, NM) : (NM == 4 ? 2 : 1);
    const void* fn = nullptr;
#define SL_ILV(N, L) if (NM == N && lpw == L) fn = (const void*)slk::lr_interleaved_kernel<kIlvR, kIlvRS, kIlvNCW, N, L>;
    SL_ILV(2, 1) SL_ILV(2, 2) SL_ILV(3, 1) SL_ILV(3, 3) SL_ILV(4, 1) SL_ILV(4, 2) SL_ILV(4, 4)
#undef SL_ILV
    if (!fn) return fail(h, SL_EINVAL, "no interleaved kernel for %d replicas with %d exchange lanes per warp", NM, lpw);
    const int xwarps = (NM + lpw - 1) / lpw;
    int grid = h->num_sms;
    if (h->opt_grid > 0) grid = (int)std::min<long long>(h->opt_grid, h->num_sms);
    slk::LrArgs a;
    memset(&a, 0, sizeof(a));
    a.F = F; a.Fp = F; a.B = h->B; a.vB = h->B; a.n_rows = h->n_rows; a.npart = NM; a.krep = 1;
    a.row_bytes = 4 * F;
    a.zy_rows = (int)((((long long)h->B + grid - 1) / grid + kIlvR - 1) / kIlvR * kIlvR);
    const size_t red_b = sizeof(float) * 2 * kIlvNCW * kIlvR;
    const size_t zy_b = sizeof(float) * 2 * 2 * (size_t)a.zy_rows;
    const size_t vec_b = 8 * (size_t)F * NM;  // gout, gin, masters: one of each per model
    const size_t fixed = red_b + zy_b + vec_b + 1024;  // one s64 vector per model; the masters live in global scratch
    const size_t slot_b = (size_t)kIlvRS * a.row_bytes;
    if (h->smem_optin < fixed + (slot_b + 16) * (kIlvR / kIlvRS))
        return fail(h, SL_EINVAL, "the interleaved kernel does not fit in shared memory at F = %d", F);
    size_t ns = std::min<size_t>((h->smem_optin - fixed) / (slot_b + 16), 32);  // one producer lane per slot
    if (h->opt_ring > 0) ns = std::max<size_t>(kIlvR / kIlvRS, std::min<size_t>(ns, (size_t)h->opt_ring));
    a.ns = (int)ns;
    size_t off = ns * slot_b;
    off = align_up(off, 16); a.off_bar = (int)off;   off += 8 * (2 * ns + 2 * NM);
    off = align_up(off, 16); a.off_red = (int)off;   off += red_b;
    off = align_up(off, 16); a.off_zy = (int)off;    off += zy_b;
    off = align_up(off, 128); a.off_gout = (int)off; off += vec_b;
    const size_t smem = off;
    {   // per-CTA fp64 masters in global memory (L2)
        const size_t wb = sizeof(double) * (size_t)grid * NM * F;
        if (h->wmaster_bytes < wb) {
            if (h->wmaster) cudaFree(h->wmaster);
            h->wmaster = nullptr;
            h->wmaster_bytes = 0;
            CK(h, cudaMalloc(&h->wmaster, wb));
            h->wmaster_bytes = wb;
        }
    }
    const size_t need = 8 * (size_t)3 * F * NM;
    if (h->acc_bytes < need) {
        if (h->acc) cudaFree(h->acc);
        h->acc = nullptr;
        h->acc_bytes = 0;
        CK(h, cudaMalloc(&h->acc, need));
        h->acc_bytes = need;
    }
    if (!h->counter) CK(h, cudaMalloc(&h->counter, 256));
    CK(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    const int threads = (kIlvNCW + 1 + xwarps) * 32;
    CK(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, threads, smem));
    if (occ < 1) return fail(h, SL_ECUDA, "interleaved kernel does not fit on an SM");
    const int max_per_launch = h->opt_steps_per_launch > 0 ? (int)h->opt_steps_per_launch : kMaxStepsPerLaunch;
    *h->h_err = 0;
    if (want_loss) CK(h, cudaMemsetAsync(h->loss_fix, 0, 8 * (size_t)n_steps * NM, h->stream));
    int done = 0;
    while (done < n_steps) {
        const int n = std::min(max_per_launch, n_steps - done);
        a.step0 = h->step;
        a.n_steps = n;
        a.acc = h->acc;
        a.counter = h->counter;
        a.wmaster = h->wmaster;
        a.loss_fix = want_loss ? h->loss_fix + done : nullptr;
        a.loss_stride = n_steps;
        a.eta_over_b = h->eta / (double)h->B;
        a.fix_limit = 134217728.0f / (float)grid;
        a.err = h->d_err;
        a.tma_chunk = 1 << 20;
        for (int k = 0; k < 8; ++k) { a.vX[k] = h->vX[k]; a.vy[k] = h->vy[k]; a.pmodel[k] = h->pmodel[k]; }
        CK(h, cudaMemsetAsync(h->acc, 0, need, h->stream));
        CK(h, cudaMemsetAsync(h->counter, 0, 256, h->stream));
        void* params[] = {&a};
        CK(h, cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(threads), params, smem, h->stream));
        h->launches++;
        h->k1_launches++;
        h->step += n;
        done += n;
    }
    if (want_loss) {
        slk::loss_finish_kernel<<<(n_steps * NM + 255) / 256, 256, 0, h->stream>>>(
            h->loss_fix, h->d_loss, n_steps * NM, 1.0 / (double)h->B);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    h->path_used = 5;
    h->variant_used = 0;
    h->grid_used = grid;
    h->ns_used = a.ns;
    h->smem_used = (long long)smem;
    return SL_OK;
}

int train_impl(sl_handle* h, int n_steps, double* loss_out, bool sync) {
    if (n_steps < 0) return fail(h, SL_EINVAL, "n_steps < 0");
    if (!h->shard_sealed || h->n_rows <= 0)
        return fail(h, SL_ESTATE, "no sealed shard with at least one row (sl_shard_end first)");
    if (sync && !h->comm && h->peer_n == 0)
        return fail(h, SL_ESTATE, "neither sl_comm_init_rank nor sl_peer_attach has been called");
    if (sync && !h->comm && (h->opt_sync_path != 0 || h->opt_path == 2))
        return fail(h, SL_ESTATE, "the NCCL path needs sl_comm_init_rank");
    if (n_steps == 0) return SL_OK;
    DeviceGuard dg(h->device);
    const size_t nloss = (size_t)n_steps * (h->npart > 1 ? h->npart : 1);  // one loss per step and replica
    int rc = ensure_loss_cap(h, nloss);
    if (rc) return rc;
    // K1c: narrow rows as thread-block clusters (one replica, or one group job; fused across GPUs when
    // peers are attached).  "variant" selects among the older kernels, so it is only taken when unset.
    if ((h->opt_path == 0 || h->opt_path == 6) && h->opt_variant == 0 && h->npart <= 1 &&
        (!sync || (h->peer_n > 0 && h->opt_sync_path == 0))) {
        ClusterPlan cp;
        std::string cwhy;
        if (plan_cluster(h, &cp, &cwhy, sync)) {
            const bool nccl_loss = sync && loss_out && h->comm && h->nranks > 1 && g_nccl.AllReduce;
            double* loss_dev = nullptr;
            if (loss_out) loss_dev = (nccl_loss || !h->h_loss_dev) ? h->d_loss : h->h_loss_dev;
            rc = run_cluster(h, cp, n_steps, loss_dev, sync);
            if (rc) return rc;
            if (nccl_loss) {  // each rank evaluated the loss of its own rows: mean over ranks
                int nrc = g_nccl.AllReduce(h->d_loss, h->d_loss, (size_t)n_steps, kNcclFloat64, kNcclSum, h->comm, h->stream);
                if (nrc != 0) return fail(h, SL_ENCCL, "ncclAllReduce(loss) failed (%d)", nrc);
                slk::scale_kernel<<<(n_steps + 255) / 256, 256, 0, h->stream>>>(h->d_loss, n_steps, 1.0 / (double)h->nranks);
                h->launches++;
            }
            if (loss_out && loss_dev == h->d_loss)
                CK(h, cudaMemcpyAsync(h->h_loss, h->d_loss, sizeof(double) * nloss, cudaMemcpyDeviceToHost, h->stream));
            CK(h, cudaStreamSynchronize(h->stream));
            CK(h, cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
            if (*h->h_err != 0) {
                const int code = *h->h_err;
                *h->h_err = 0;
                if (code != slk::kErrRange) h->cl_dirty = true;
                return fail_watchdog(h, code);
            }
            if (loss_out) memcpy(loss_out, h->h_loss, sizeof(double) * nloss);
            return SL_OK;
        }
        if (h->opt_path == 6) return fail(h, SL_EINVAL, "cluster kernel requested but unavailable: %s", cwhy.c_str());
    } else if (h->opt_path == 6) {
        return fail(h, SL_EINVAL, "cluster kernel requested but this call cannot use it (variant set, replicas, or NCCL path)");
    }
    if (!sync && interleaved_ok(h)) {  // independent replicas, interleaved on every SM
        CK(h, cudaEventRecord(h->ev0, h->stream));
        rc = run_interleaved(h, n_steps, loss_out != nullptr);
        if (rc) return rc;
        CK(h, cudaEventRecord(h->ev1, h->stream));
        if (loss_out)
            CK(h, cudaMemcpyAsync(h->h_loss, h->d_loss, sizeof(double) * nloss, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
        CK(h, cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
        if (*h->h_err != 0) {
            const int code = *h->h_err;
            *h->h_err = 0;
            return fail_watchdog(h, code);
        }
        if (loss_out) memcpy(loss_out, h->h_loss, sizeof(double) * nloss);
        return SL_OK;
    }
    PersistPlan plan;
    std::string why;
    bool persist = !sync && h->opt_path != 2 && plan_persist(h, &plan, &why);
    bool fused = false;
    if (sync && h->peer_n > 0 && h->ipc_attached && h->opt_sync_path == 0 && h->opt_path != 2 && plan_persist(h, &plan, &why)) {
        const Variant& v = kVariants[plan.variant];
        fused = v.xb && (h->F + plan.args.xslice - 1) / plan.args.xslice <= plan.grid && h->peer_n <= 8;
    }
    if (!sync && (h->opt_path == 1 || h->forbid_simple) && !persist)
        return fail(h, SL_EINVAL, "persistent path requested but unavailable: %s", why.c_str());
    if (sync && !fused && (!h->comm || !g_nccl.AllReduce))
        return fail(h, SL_ESTATE, "the fused peer path cannot run this shape (%s) and no NCCL communicator "
                    "is initialised (sl_comm_init_rank)", why.empty() ? "not eligible" : why.c_str());
    CK(h, cudaEventRecord(h->ev0, h->stream));
    if (persist || fused) {
        rc = run_persist(h, plan, n_steps, loss_out != nullptr, fused);
        if (rc) return rc;
        if (fused && loss_out && h->comm && h->nranks > 1) {
            // each rank evaluated the loss of its own rows: mean over ranks
            int nrc = g_nccl.AllReduce(h->d_loss, h->d_loss, (size_t)n_steps, kNcclFloat64, kNcclSum,
                                       h->comm, h->stream);
            if (nrc != 0) return fail(h, SL_ENCCL, "ncclAllReduce(loss) failed (%d)", nrc);
            slk::scale_kernel<<<(n_steps + 255) / 256, 256, 0, h->stream>>>(h->d_loss, n_steps,
                                                                         1.0 / (double)h->nranks);
            h->launches++;
        }
    } else {
        rc = ensure_simple_scratch(h);
        if (rc) return rc;
        for (int s = 0; s < n_steps; ++s) {
            rc = launch_simple_step(h, loss_out ? h->d_loss + s : nullptr, sync);
            if (rc) return rc;
        }
        h->path_used = sync ? 3 : 2;
        h->variant_used = 0;
    }
    CK(h, cudaEventRecord(h->ev1, h->stream));
    if (loss_out)
        CK(h, cudaMemcpyAsync(h->h_loss, h->d_loss, sizeof(double) * nloss, cudaMemcpyDeviceToHost,
                              h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    CK(h, cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    if (*h->h_err != 0) {
        const int code = *h->h_err;
        *h->h_err = 0;
        return fail_watchdog(h, code);
    }
    if (loss_out) memcpy(loss_out, h->h_loss, sizeof(double) * nloss);
    return SL_OK;
}

}  // namespace

// =====================================================================================
// C ABI
// =====================================================================================
extern "C" {

int sl_create(int device, int D, int F, int B, double eta, double lambda, sl_handle** out) {
    if (!out) return fail(nullptr, SL_EINVAL, "out is NULL");
    *out = nullptr;
    if (D <= 0 || F <= 0 || B <= 0) return fail(nullptr, SL_EINVAL, "D, F and B must be positive");
    if (D != F)
        return fail(nullptr, SL_EINVAL,
                    "D (%d) must equal F (%d): the model is bias-free logistic regression", D, F);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, SL_ECUDA, "no CUDA device: %s (this library has no CPU fallback)",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (device < 0 || device >= ndev)
        return fail(nullptr, SL_EINVAL, "device %d out of range [0,%d)", device, ndev);
    sl_handle* h = new (std::nothrow) sl_handle();
    if (!h) return fail(nullptr, SL_ENOMEM, "out of host memory");
    h->device = device;
    h->D = D; h->F = F; h->B = B; h->eta = eta; h->lambda = lambda;
    // profilers replay kernels one at a time and do not take every launch form: SL_CL_COOP=0 / SL_CL_GRAPH=0 make
    // the cluster kernel launch without the cooperative attribute / outside a graph (same kernel, same arguments)
    if (const char* e = getenv("SL_CL_COOP")) h->cl_coop = atoi(e) ? -1 : 0;
    if (const char* e = getenv("SL_CL_GRAPH")) h->opt_cl_graph = atoi(e);
    DeviceGuard dg(device);
#define CKC(call)                                                                          \
    do {                                                                                   \
        cudaError_t e_ = (call);                                                           \
        if (e_ != cudaSuccess) {                                                           \
            fail(nullptr, SL_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_));       \
            sl_destroy(h);                                                                 \
            return SL_ECUDA;                                                               \
        }                                                                                  \
    } while (0)
    int v = 0;
    CKC(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device));
    h->num_sms = v;
    CKC(cudaDeviceGetAttribute(&v, cudaDevAttrCooperativeLaunch, device));
    h->coop = v;
    CKC(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    h->smem_optin = (size_t)v;
    CKC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CKC(cudaEventCreate(&h->ev0));
    CKC(cudaEventCreate(&h->ev1));
    CKC(cudaMalloc(&h->model, sizeof(double) * D));
    CKC(cudaMalloc(&h->old, sizeof(double) * D));
    CKC(cudaMemsetAsync(h->model, 0, sizeof(double) * D, h->stream));
    CKC(cudaMemsetAsync(h->old, 0, sizeof(double) * D, h->stream));
    CKC(cudaHostAlloc(&h->h_in, sizeof(double) * D, cudaHostAllocMapped));
    CKC(cudaHostAlloc(&h->h_out, sizeof(double) * D, cudaHostAllocMapped));
    CKC(cudaHostGetDevicePointer(&h->d_in, h->h_in, 0));
    CKC(cudaHostGetDevicePointer(&h->d_out, h->h_out, 0));
    CKC(cudaHostAlloc(&h->h_err, sizeof(int), cudaHostAllocMapped));
    CKC(cudaHostGetDevicePointer(&h->d_err, h->h_err, 0));
    *h->h_err = 0;
    CKC(cudaHostAlloc(&h->h_sum, 2 * sizeof(unsigned long long), cudaHostAllocMapped));
    CKC(cudaHostGetDevicePointer(&h->d_sum, h->h_sum, 0));
    for (int i = 0; i < kStageSlots; ++i) {
        CKC(cudaMallocHost(&h->stage[i], kStageBytes));
        CKC(cudaEventCreateWithFlags(&h->stage_ev[i], cudaEventDisableTiming));
    }
    CKC(cudaStreamSynchronize(h->stream));
#undef CKC
    *out = h;
    return SL_OK;
}

int sl_destroy(sl_handle* h) {
    if (!h) return SL_OK;
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    cudaFree(h->model); cudaFree(h->old); cudaFree(h->shard); cudaFree(h->in_buf);
    cudaFree(h->acc); cudaFree(h->counter); cudaFree(h->wmaster); cudaFree(h->w32buf); cudaFree(h->d_loss); cudaFree(h->loss_fix);
    cudaFree(h->dbg);
    cudaFree(h->cl_scratch);
    if (h->cl_gexec) cudaGraphExecDestroy(h->cl_gexec);
    if (h->cl_graph) cudaGraphDestroy(h->cl_graph);
    for (int p = 0; p < h->peer_n; ++p)
        if (p != h->peer_rank && h->peer_base[p]) cudaIpcCloseMemHandle(h->peer_base[p]);
    for (auto& r : h->remote)
        if (r.used) { cudaIpcCloseMemHandle(r.model); cudaIpcCloseMemHandle(r.old); }
    cudaFree(h->xblock);
    cudaFree(h->resid); cudaFree(h->rowloss); cudaFree(h->gpart); cudaFree(h->gsum);
    if (h->h_in) cudaFreeHost(h->h_in);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->h_err) cudaFreeHost(h->h_err);
    if (h->h_sum) cudaFreeHost(h->h_sum);
    if (h->h_loss) cudaFreeHost(h->h_loss);
    for (int i = 0; i < kStageSlots; ++i) {
        if (h->stage[i]) cudaFreeHost(h->stage[i]);
        if (h->stage_ev[i]) cudaEventDestroy(h->stage_ev[i]);
    }
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    delete h;
    return SL_OK;
}

const char* sl_last_error(const sl_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int sl_set_option(sl_handle* h, const char* key, long long value) {
    if (!h || !key) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    std::string k(key);
    if (k == "grid") h->opt_grid = value;
    else if (k == "variant") h->opt_variant = value;
    else if (k == "path") h->opt_path = value;  // 0 auto, 1 persistent only, 2 simple
    else if (k == "steps_per_launch") h->opt_steps_per_launch = value;
    else if (k == "step") h->step = value;
    else if (k == "debug_timeline") h->opt_dbg_step = value;
    else if (k == "ring_slots") h->opt_ring = value;
    else if (k == "krep") h->opt_krep = value;
    else if (k == "block_rows") h->opt_block_rows = value;
    else if (k == "multi_mode") h->opt_multi_mode = value;
    else if (k == "multi_lpw") h->opt_multi_lpw = value;
    else if (k == "debug_skip_compute") h->opt_dbg_skip = value;
    else if (k == "tma_chunk") h->opt_tma_chunk = value;
    else if (k == "sync_path") h->opt_sync_path = value;
    else if (k == "cl_variant") h->opt_cl_variant = value;
    else if (k == "clusters") h->opt_clusters = value;
    else if (k == "poll_delay") h->opt_poll_delay = value;
    else if (k == "poll_gap") h->opt_poll_gap = value;
    else if (k == "xpoll_delay") h->opt_xpoll_delay = value;
    else if (k == "xpoll_gap") h->opt_xpoll_gap = value;
    else if (k == "cl_graph") h->opt_cl_graph = value;
    else if (k == "wait_all_owners") h->opt_wait_all = value;
    else if (k == "cl_coop") h->cl_coop = value < 0 ? -1 : (value ? 1 : 0);
    else return fail(h, SL_EINVAL, "unknown option '%s'", key);
    return SL_OK;
}

long long sl_get_info(sl_handle* h, const char* key) {
    if (!h || !key) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    std::string k(key);
    if (k == "launches") return h->launches;
    if (k == "k1_launches") return h->k1_launches;
    if (k == "path") return h->path_used;
    if (k == "variant") return h->variant_used;
    if (k == "grid") return h->grid_used;
    if (k == "ring_slots") return h->ns_used;
    if (k == "smem") return h->smem_used;
    if (k == "num_sms") return h->num_sms;
    if (k == "step") return h->step;
    if (k == "n_rows") return h->n_rows;
    if (k == "shard_header") return h->shard_off ? 1 : 0;
    if (k == "shard_row0") return (long long)h->shard_row0;
    if (k == "nranks") return h->nranks;
    if (k == "peers") return h->peer_n;
    if (k == "multicast") return h->sym_n > 0 && h->mc_gx ? 1 : 0;
    if (k == "cl_graph") return h->cl_graph_ok;
    if (k == "num_variants") return kNumVariants;
    if (k == "num_cl_variants") return kNumClVariants;
    return SL_EINVAL;
}

// ---- shard ingest -------------------------------------------------------------------
int sl_shard_begin(sl_handle* h, size_t size_hint) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->in_mu);
    DeviceGuard dg(h->device);
    // The sealed shard stays valid and trainable; the stream lands in a second buffer that
    // sl_shard_end swaps in (the reference receives files on a gRPC thread while the training
    // thread keeps running, src/worker.cc:242-246).
    CK(h, cudaStreamSynchronize(h->copy_stream));
    h->in_len = 0;
    h->shard_open = true;
    if (size_hint) {
        int rc = ensure_buf_cap(h, &h->in_buf, &h->in_cap, 0, size_hint, true);
        if (rc) return rc;
    }
    return SL_OK;
}

int sl_shard_append(sl_handle* h, const void* bytes, size_t n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->in_mu);
    if (!h->shard_open) return fail(h, SL_ESTATE, "sl_shard_append without sl_shard_begin");
    if (n == 0) return SL_OK;
    if (!bytes) return fail(h, SL_EINVAL, "bytes is NULL");
    DeviceGuard dg(h->device);
    int rc = ensure_buf_cap(h, &h->in_buf, &h->in_cap, h->in_len, h->in_len + n);
    if (rc) return rc;
    cudaPointerAttributes at;
    bool pinned = false;
    if (cudaPointerGetAttributes(&at, bytes) == cudaSuccess) pinned = at.type == cudaMemoryTypeHost;
    else cudaGetLastError();
    if (pinned) {
        // The caller's buffer is page-locked: DMA straight from it, asynchronously.  This is the
        // one exception to "buffers are used only for the duration of the call": a page-locked
        // source must stay valid until sl_shard_end returns (it waits for the copies).
        CK(h, cudaMemcpyAsync(h->in_buf + h->in_len, bytes, n, cudaMemcpyHostToDevice,
                              h->copy_stream));
    } else {
        const unsigned char* src = static_cast<const unsigned char*>(bytes);
        size_t off = 0;
        while (off < n) {
            const size_t m = std::min(kStageBytes, n - off);
            const int s = h->stage_idx;
            CK(h, cudaEventSynchronize(h->stage_ev[s]));  // slot's previous DMA done
            memcpy(h->stage[s], src + off, m);
            CK(h, cudaMemcpyAsync(h->in_buf + h->in_len + off, h->stage[s], m,
                                  cudaMemcpyHostToDevice, h->copy_stream));
            CK(h, cudaEventRecord(h->stage_ev[s], h->copy_stream));
            h->stage_idx = (s + 1) % kStageSlots;
            off += m;
        }
    }
    h->in_len += n;
    return SL_OK;
}

int sl_shard_end(sl_handle* h, long long* n_rows) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk_in(h->in_mu);
    if (!h->shard_open) return fail(h, SL_ESTATE, "sl_shard_end without sl_shard_begin");
    DeviceGuard dg(h->device);
    CK(h, cudaStreamSynchronize(h->copy_stream));   // the stream has landed (the replica lock is not held while waiting)
    // the swap itself: with the replica lock held no training call is in flight, so no kernel reads the outgoing shard
    std::lock_guard<std::mutex> lk(h->mu);
    CK(h, cudaStreamSynchronize(h->stream));
    // a self-describing stream is checked before it replaces the sealed shard
    size_t off = 0;
    unsigned long long row0 = 0;
    long long rows = (long long)(h->in_len / (4 * ((size_t)h->F + 1)));
    if (h->in_len >= sizeof(sl_shard_header)) {
        sl_shard_header hd;
        CK(h, cudaMemcpy(&hd, h->in_buf, sizeof(hd), cudaMemcpyDeviceToHost));
        if (memcmp(hd.magic, SL_SHARD_MAGIC, 8) == 0) {
            h->shard_open = false;
            const size_t in_len = h->in_len;
            h->in_len = 0;
            if (hd.version != 1) return fail(h, SL_EINVAL, "shard header version %u not understood", hd.version);
            if ((int)hd.F != h->F) return fail(h, SL_EINVAL, "shard was written for F = %u, this replica has F = %d", hd.F, h->F);
            // bound n_rows by the stream length before multiplying: a crafted count must not wrap
            const unsigned long long row_b = 4ull * ((unsigned long long)h->F + 1);
            if (hd.n_rows > (in_len - sizeof(hd)) / row_b)
                return fail(h, SL_EINVAL, "shard header promises %llu rows, the stream has room for %llu",
                            hd.n_rows, (unsigned long long)((in_len - sizeof(hd)) / row_b));
            const unsigned long long want = sizeof(hd) + hd.n_rows * row_b;
            if (want != in_len)
                return fail(h, SL_EINVAL, "shard header promises %llu bytes, the stream had %zu", want, in_len);
            if (hd.flags & SL_SHARD_HAS_SUMS) {
                h->h_sum[0] = 0;
                h->h_sum[1] = 0;
                if (in_len > sizeof(hd)) {
                    slk::checksum_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(h->in_buf + sizeof(hd), in_len - sizeof(hd), h->d_sum);
                    h->launches++;
                    CK(h, cudaGetLastError());
                }
                CK(h, cudaStreamSynchronize(h->stream));
                if (h->h_sum[0] != hd.byte_sum || h->h_sum[1] != hd.weighted)
                    return fail(h, SL_EINVAL, "shard payload does not match the sums in its header");
            }
            off = sizeof(hd);
            row0 = hd.row0;
            rows = (long long)hd.n_rows;
            h->in_len = in_len;
        }
    }
    if (rows <= 0 && h->shard_sealed && h->n_rows > 0) {
        // a stream without one complete row never replaces a trainable shard: whatever was sealed before stays
        // (the reference acks such a stream like any other, src/worker.cc:57-60, so this is no error)
        h->in_len = 0;
        h->shard_open = false;
        if (n_rows) *n_rows = 0;
        return SL_OK;
    }
    std::swap(h->shard, h->in_buf);
    std::swap(h->shard_cap, h->in_cap);
    h->shard_len = h->in_len;
    h->shard_off = off;
    h->shard_row0 = row0;
    h->in_len = 0;
    h->shard_open = false;
    h->shard_sealed = true;
    h->n_rows = rows;
    h->step = 0;
    if (n_rows) *n_rows = h->n_rows;
    return SL_OK;
}

int sl_shard_checksum(sl_handle* h, unsigned long long* byte_sum, unsigned long long* weighted) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->shard_sealed) return fail(h, SL_ESTATE, "no sealed shard");
    DeviceGuard dg(h->device);
    h->h_sum[0] = 0;
    h->h_sum[1] = 0;
    if (h->shard_len) {
        const int grid = h->num_sms * 8;
        slk::checksum_kernel<<<grid, 256, 0, h->stream>>>(h->shard, h->shard_len, h->d_sum);
        h->launches++;
        CK(h, cudaGetLastError());
    }
    CK(h, cudaStreamSynchronize(h->stream));
    if (byte_sum) *byte_sum = h->h_sum[0];
    if (weighted) *weighted = h->h_sum[1];
    return SL_OK;
}

int sl_shard_read(sl_handle* h, size_t offset, void* out, size_t n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->shard_sealed) return fail(h, SL_ESTATE, "no sealed shard");
    if (offset + n > h->shard_len) return fail(h, SL_EINVAL, "read past the end of the shard");
    DeviceGuard dg(h->device);
    CK(h, cudaMemcpyAsync(out, h->shard + offset, n, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return SL_OK;
}

int sl_debug_read(sl_handle* h, unsigned long long* out, size_t n_words) {
    if (!h || !out) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->dbg || n_words > h->dbg_words) return fail(h, SL_ESTATE, "no debug timeline recorded");
    DeviceGuard dg(h->device);
    CK(h, cudaMemcpy(out, h->dbg, n_words * 8, cudaMemcpyDeviceToHost));
    return SL_OK;
}

long long sl_shard_bytes(sl_handle* h) { return h ? (long long)h->shard_len : SL_EINVAL; }

int sl_shard_synth(sl_handle* h, long long n_rows, unsigned long long row0,
                   unsigned long long seed) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n_rows <= 0) return fail(h, SL_EINVAL, "n_rows must be positive");
    DeviceGuard dg(h->device);
    CK(h, cudaStreamSynchronize(h->stream));
    const size_t bytes = (size_t)n_rows * 4 * ((size_t)h->F + 1);
    h->shard_len = 0;
    int rc = ensure_buf_cap(h, &h->shard, &h->shard_cap, 0, bytes, true);
    if (rc) return rc;
    float* X = reinterpret_cast<float*>(h->shard);
    float* y = X + (size_t)n_rows * h->F;
    const size_t quads = ((size_t)n_rows * h->F + 3) / 4;
    const int grid = (int)std::min<size_t>((quads + 255) / 256, (size_t)h->num_sms * 16);
    slk::synth_kernel<<<grid, 256, 0, h->stream>>>(X, y, n_rows, h->F, row0, seed);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    h->shard_len = bytes;
    h->shard_off = 0;
    h->shard_row0 = row0;
    h->n_rows = n_rows;
    h->shard_sealed = true;
    h->step = 0;
    return SL_OK;
}

// ---- training -----------------------------------------------------------------------
int sl_train_tick(sl_handle* h) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    slk::tick_kernel<<<(h->D + 255) / 256, 256, 0, h->stream>>>(h->model, h->D);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    return SL_OK;
}

int sl_train_steps(sl_handle* h, int n_steps, double* loss_out) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    return train_impl(h, n_steps, loss_out, false);
}

int sl_train_steps_sync(sl_handle* h, int n_steps, double* loss_out) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    return train_impl(h, n_steps, loss_out, true);
}

int sl_train_steps_group(sl_handle* const* handles, int W, int n_steps, double* loss_out) {
    if (!handles || W < 1 || W > 8) return fail(nullptr, SL_EINVAL, "W must be 1..8");
    for (int k = 0; k < W; ++k) {
        if (!handles[k]) return fail(nullptr, SL_EINVAL, "handle %d is NULL", k);
        for (int j = 0; j < k; ++j)
            if (handles[j] == handles[k]) return fail(handles[0], SL_EINVAL, "handle %d is listed twice", k);
    }
    sl_handle* h = handles[0];
    // take every handle's lock, in address order so that two group calls cannot deadlock
    std::vector<sl_handle*> order(handles, handles + W);
    std::sort(order.begin(), order.end());
    std::vector<std::unique_lock<std::mutex>> locks;
    for (sl_handle* o : order) locks.emplace_back(o->mu);
    for (int k = 0; k < W; ++k) {
        sl_handle* w = handles[k];
        if (w->device != h->device || w->F != h->F || w->B != h->B || w->eta != h->eta)
            return fail(h, SL_EINVAL, "worker %d differs from worker 0 in device, F, B or eta", k);
        if (!w->shard_sealed || w->n_rows <= 0) return fail(h, SL_ESTATE, "worker %d has no sealed shard", k);
        if (w->n_rows != h->n_rows)
            return fail(h, SL_EINVAL, "worker %d holds %lld rows, worker 0 holds %lld: shards must be equal",
                        k, w->n_rows, h->n_rows);
        if (w->step != h->step) return fail(h, SL_ESTATE, "worker %d is at step %lld, worker 0 at %lld", k, w->step, h->step);
        if (((uintptr_t)(w->shard + w->shard_off) & 15) != 0) return fail(h, SL_EINVAL, "worker %d: shard not 16-byte aligned", k);
    }
    if (n_steps < 0) return fail(h, SL_EINVAL, "n_steps < 0");
    if (n_steps == 0) return SL_OK;
    DeviceGuard dg(h->device);
    for (int k = 1; k < W; ++k) CK(h, cudaStreamSynchronize(handles[k]->stream));  // their shards are settled
    for (int k = 0; k < W; ++k) {
        h->vX[k] = reinterpret_cast<const float*>(handles[k]->shard + handles[k]->shard_off);
        h->vy[k] = h->vX[k] + (size_t)handles[k]->n_rows * h->F;
    }
    h->vw = W;
    h->forbid_simple = true;  // the simple path has no multi-shard mode: fail instead of falling back
    int rc = h->opt_path != 2 ? train_impl(h, n_steps, loss_out, false)
                              : fail(h, SL_EINVAL, "sl_train_steps_group runs on the persistent kernels only");
    h->forbid_simple = false;
    h->vw = 0;
    if (rc) return rc;
    for (int k = 1; k < W; ++k) {  // every replica leaves with the new model (old is untouched, as in sl_train_steps)
        CK(h, cudaMemcpyAsync(handles[k]->model, h->model, sizeof(double) * (size_t)h->D, cudaMemcpyDeviceToDevice, h->stream));
        handles[k]->step = h->step;
    }
    CK(h, cudaStreamSynchronize(h->stream));
    return SL_OK;
}

// W independent replicas on one GPU as W concurrent launches of the cluster kernel: every replica on its own
// handle's stream with 1/W of the clusters that fit on the device, each the very launch sl_train_steps makes with
// option "clusters" set to that share.  The launches do not wait for one another; while one replica's step is in
// its exchange the others keep the row stream going (config 2, local steps: 9.8 us per round of four B = 4096 steps
// against 12.5 us for the interleaved kernel).  *ran = false: this shape stays on the older kernels.
static int run_multi_cluster(sl_handle* const* hs, int W, int n_steps, double* loss_out, bool* ran) {
    *ran = false;
    sl_handle* h0 = hs[0];
    if (!(h0->opt_path == 0 || h0->opt_path == 6) || h0->opt_variant != 0) return SL_OK;
    if (h0->F > 1024) {   // wider rows: a replica's share of the clusters would take four and more passes per step
        if (h0->opt_path == 6) return fail(h0, SL_EINVAL, "cluster kernel requested but sl_train_steps_multi uses it up to F = 1024");
        return SL_OK;
    }
    std::vector<ClusterPlan> plans((size_t)W);
    std::string why;
    for (int k = 0; k < W; ++k) {
        sl_handle* h = hs[k];
        const long long keep_clusters = h->opt_clusters, keep_variant = h->opt_cl_variant;
        h->opt_cl_variant = h0->opt_cl_variant;
        h->opt_clusters = 0;
        ClusterPlan full;
        bool ok = plan_cluster(h, &full, &why, false);
        if (ok) {
            long long share = full.args.ncl / W;
            if (keep_clusters > 0) share = std::min(share, keep_clusters);
            ok = share >= 1;
            if (ok) {
                h->opt_clusters = share;
                ok = plan_cluster(h, &plans[(size_t)k], &why, false);
            } else {
                why = "fewer clusters fit on the device than there are replicas";
            }
        }
        h->opt_clusters = keep_clusters;
        h->opt_cl_variant = keep_variant;
        if (!ok) {
            if (h0->opt_path == 6) return fail(h0, SL_EINVAL, "cluster kernel requested but unavailable: %s", why.c_str());
            return SL_OK;
        }
    }
    for (int k = 0; k < W; ++k) {
        int rc = ensure_loss_cap(hs[k], (size_t)n_steps);
        if (rc) return rc;
    }
    *ran = true;
    std::vector<double*> loss_dev((size_t)W, nullptr);
    int rc = SL_OK;
    for (int k = 0; k < W && rc == SL_OK; ++k) {
        sl_handle* h = hs[k];
        if (loss_out) loss_dev[(size_t)k] = h->h_loss_dev ? h->h_loss_dev : h->d_loss;
        rc = run_cluster(h, plans[(size_t)k], n_steps, loss_dev[(size_t)k], false);
        if (rc == SL_OK && loss_out && loss_dev[(size_t)k] == h->d_loss &&
            cudaMemcpyAsync(h->h_loss, h->d_loss, sizeof(double) * (size_t)n_steps, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess)
            rc = fail(h, SL_ECUDA, "copy of the losses failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    // every launch that went out is waited for, also after an error
    float whole = 0.f;
    for (int k = 0; k < W; ++k) {
        sl_handle* h = hs[k];
        if (cudaStreamSynchronize(h->stream) != cudaSuccess && rc == SL_OK)
            rc = fail(h, SL_ECUDA, "worker %d: %s", k, cudaGetErrorString(cudaGetLastError()));
    }
    if (rc) return rc;
    for (int k = 0; k < W; ++k) {
        sl_handle* h = hs[k];
        float ms = 0.f;
        CK(h, cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
        CK(h, cudaEventElapsedTime(&ms, h0->ev0, h->ev1));   // from the first launch to this replica's end
        whole = std::max(whole, ms);
        if (*h->h_err != 0) {
            const int code = *h->h_err;
            *h->h_err = 0;
            if (code != slk::kErrRange) h->cl_dirty = true;
            rc = fail_watchdog(h, code);
            if (h != h0) fail(h0, rc, "worker %d: %s", k, sl_last_error(h));
        }
        if (rc == SL_OK && loss_out) memcpy(loss_out + (size_t)k * n_steps, h->h_loss, sizeof(double) * (size_t)n_steps);
        h->path_used = 8;
    }
    h0->last_ms = whole;   // the job: all replicas' n_steps
    return rc;
}

int sl_train_steps_multi(sl_handle* const* handles, int W, int n_steps, double* loss_out) {
    if (!handles || W < 1 || W > 8) return fail(nullptr, SL_EINVAL, "W must be 1..8");
    for (int k = 0; k < W; ++k) {
        if (!handles[k]) return fail(nullptr, SL_EINVAL, "handle %d is NULL", k);
        for (int j = 0; j < k; ++j)
            if (handles[j] == handles[k]) return fail(handles[0], SL_EINVAL, "handle %d is listed twice", k);
    }
    sl_handle* h = handles[0];
    std::vector<sl_handle*> order(handles, handles + W);
    std::sort(order.begin(), order.end());
    std::vector<std::unique_lock<std::mutex>> locks;
    for (sl_handle* o : order) locks.emplace_back(o->mu);
    for (int k = 0; k < W; ++k) {
        sl_handle* w = handles[k];
        if (w->device != h->device || w->F != h->F || w->B != h->B || w->eta != h->eta)
            return fail(h, SL_EINVAL, "worker %d differs from worker 0 in device, F, B or eta", k);
        if (!w->shard_sealed || w->n_rows <= 0) return fail(h, SL_ESTATE, "worker %d has no sealed shard", k);
        if (w->n_rows != h->n_rows)
            return fail(h, SL_EINVAL, "worker %d holds %lld rows, worker 0 holds %lld: shards must be equal",
                        k, w->n_rows, h->n_rows);
        if (w->step != h->step) return fail(h, SL_ESTATE, "worker %d is at step %lld, worker 0 at %lld", k, w->step, h->step);
        if (((uintptr_t)(w->shard + w->shard_off) & 15) != 0) return fail(h, SL_EINVAL, "worker %d: shard not 16-byte aligned", k);
    }
    if (n_steps < 0) return fail(h, SL_EINVAL, "n_steps < 0");
    if (n_steps == 0) return SL_OK;
    DeviceGuard dg(h->device);
    for (int k = 1; k < W; ++k) CK(h, cudaStreamSynchronize(handles[k]->stream));
    {
        bool ran = false;
        int mrc = run_multi_cluster(handles, W, n_steps, loss_out, &ran);
        if (mrc || ran) return mrc;
    }
    for (int k = 0; k < W; ++k) {
        h->vX[k] = reinterpret_cast<const float*>(handles[k]->shard + handles[k]->shard_off);
        h->vy[k] = h->vX[k] + (size_t)handles[k]->n_rows * h->F;
        h->pmodel[k] = handles[k]->model;
    }
    h->npart = W;
    h->forbid_simple = true;
    int rc = (h->opt_path == 0 || h->opt_path == 1) ? train_impl(h, n_steps, loss_out, false)
                                                    : fail(h, SL_EINVAL, "sl_train_steps_multi runs on the persistent kernel only");
    h->forbid_simple = false;
    h->npart = 0;
    if (rc) return rc;
    for (int k = 1; k < W; ++k) handles[k]->step = h->step;
    return SL_OK;
}

float sl_last_train_ms(sl_handle* h) { return h ? h->last_ms : -1.f; }

// ---- update exchange ------------------------------------------------------------------
int sl_merge(sl_handle* h, const double* delta_in, int n_in, double* reply_out, int n_out) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n_in != h->D || n_out != h->D)
        return fail(h, SL_EINVAL, "sl_merge: lengths (%d in, %d out) must equal D = %d", n_in, n_out,
                    h->D);
    if (!delta_in || !reply_out) return fail(h, SL_EINVAL, "NULL buffer");
    DeviceGuard dg(h->device);
    memcpy(h->h_in, delta_in, sizeof(double) * h->D);
    const int thr = 128, blocks = ((h->D + 1) / 2 + thr - 1) / thr;
    slk::merge_kernel<<<blocks, thr, 0, h->stream>>>(h->model, h->old, h->d_in, h->d_out, h->D,
                                                     h->lambda);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    memcpy(reply_out, h->h_out, sizeof(double) * h->D);
    return SL_OK;
}

int sl_delta_snapshot(sl_handle* h, double* out, int n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n != h->D) return fail(h, SL_EINVAL, "sl_delta_snapshot: n (%d) must equal D = %d", n, h->D);
    if (!out) return fail(h, SL_EINVAL, "NULL buffer");
    DeviceGuard dg(h->device);
    const int thr = 128, blocks = ((h->D + 1) / 2 + thr - 1) / thr;
    slk::delta_kernel<<<blocks, thr, 0, h->stream>>>(h->model, h->old, h->d_out, h->D);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    memcpy(out, h->h_out, sizeof(double) * h->D);
    return SL_OK;
}

int sl_apply_reply(sl_handle* h, const double* back, int n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n != h->D) return fail(h, SL_EINVAL, "sl_apply_reply: n (%d) must equal D = %d", n, h->D);
    if (!back) return fail(h, SL_EINVAL, "NULL buffer");
    DeviceGuard dg(h->device);
    memcpy(h->h_in, back, sizeof(double) * h->D);
    const int thr = 128, blocks = ((h->D + 1) / 2 + thr - 1) / thr;
    slk::apply_kernel<<<blocks, thr, 0, h->stream>>>(h->model, h->old, h->d_in, h->D, h->lambda);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    return SL_OK;
}

static int read_vec(sl_handle* h, const double* dev, double* out, int n) {
    if (n != h->D) return fail(h, SL_EINVAL, "n (%d) must equal D = %d", n, h->D);
    if (!out) return fail(h, SL_EINVAL, "NULL buffer");
    DeviceGuard dg(h->device);
    CK(h, cudaMemcpyAsync(h->h_out, dev, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    memcpy(out, h->h_out, sizeof(double) * n);
    return SL_OK;
}

int sl_get_model(sl_handle* h, double* out, int n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    return read_vec(h, h->model, out, n);
}

int sl_get_old(sl_handle* h, double* out, int n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    return read_vec(h, h->old, out, n);
}

int sl_set_state(sl_handle* h, const double* model, const double* old, int n) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n != h->D) return fail(h, SL_EINVAL, "n (%d) must equal D = %d", n, h->D);
    DeviceGuard dg(h->device);
    if (model) {
        memcpy(h->h_in, model, sizeof(double) * n);
        CK(h, cudaMemcpyAsync(h->model, h->h_in, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    if (old) {
        memcpy(h->h_in, old, sizeof(double) * n);
        CK(h, cudaMemcpyAsync(h->old, h->h_in, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    return SL_OK;
}

// ---- same-box data parallel -------------------------------------------------------------
int sl_comm_unique_id(void* id128) {
    if (!id128) return SL_EINVAL;
    const char* why = load_nccl();
    if (why) return fail(nullptr, SL_ENCCL, "%s", why);
    NcclUniqueId id;
    int rc = g_nccl.GetUniqueId(&id);
    if (rc != 0) return fail(nullptr, SL_ENCCL, "ncclGetUniqueId failed (%d)", rc);
    memcpy(id128, &id, sizeof(id));
    return SL_OK;
}

int sl_comm_init_rank(sl_handle* h, int nranks, int rank, const void* id128) {
    if (!h || !id128) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(h, SL_EINVAL, "bad rank/nranks");
    const char* why = load_nccl();
    if (why) return fail(h, SL_ENCCL, "%s", why);
    DeviceGuard dg(h->device);
    NcclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    int rc = g_nccl.CommInitRank(&h->comm, nranks, id, rank);
    if (rc != 0)
        return fail(h, SL_ENCCL, "ncclCommInitRank failed: %s",
                    g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
    h->nranks = nranks;
    h->rank = rank;
    return SL_OK;
}

int sl_gossip_p2p(sl_handle* initiator, sl_handle* callee) {
    if (!initiator || !callee) return SL_EINVAL;
    if (initiator == callee) return fail(initiator, SL_EINVAL, "sl_gossip_p2p: a replica cannot gossip with itself here");
    // lock both handles in address order (two threads may gossip in opposite directions)
    sl_handle* first = initiator < callee ? initiator : callee;
    sl_handle* second = initiator < callee ? callee : initiator;
    std::lock_guard<std::mutex> l1(first->mu);
    std::lock_guard<std::mutex> l2(second->mu);
    if (initiator->D != callee->D)
        return fail(initiator, SL_EINVAL, "sl_gossip_p2p: model lengths differ (%d vs %d)", initiator->D, callee->D);
    DeviceGuard dg(callee->device);
    if (initiator->device != callee->device) {
        int can = 0;
        CK(initiator, cudaDeviceCanAccessPeer(&can, callee->device, initiator->device));
        if (!can) return fail(initiator, SL_ECUDA, "GPU %d cannot map GPU %d's memory", callee->device, initiator->device);
        cudaError_t e = cudaDeviceEnablePeerAccess(initiator->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
            return fail(initiator, SL_ECUDA, "cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
        cudaGetLastError();
    }
    // the initiator's own stream must be idle before another GPU touches its state
    {
        DeviceGuard dgi(initiator->device);
        CK(initiator, cudaStreamSynchronize(initiator->stream));
    }
    const int n = callee->D;
    slk::gossip_p2p_kernel<<<(n + 127) / 128, 128, 0, callee->stream>>>(
        initiator->model, initiator->old, callee->model, callee->old, n, initiator->lambda, callee->lambda);
    callee->launches++;
    CK(callee, cudaGetLastError());
    CK(callee, cudaStreamSynchronize(callee->stream));
    return SL_OK;
}

// ---- the same round between worker PROCESSES (the reference's deployment: one worker per process,
// src/worker.cc:233-258).  The initiator's model / old are plain cudaMalloc allocations: their IPC handles
// travel once (any channel), the callee maps them and from then on runs the round as one kernel on its own GPU.
int sl_state_export(sl_handle* h, void* handles128) {
    if (!h || !handles128) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    cudaIpcMemHandle_t ih[2];
    CK(h, cudaIpcGetMemHandle(&ih[0], h->model));
    CK(h, cudaIpcGetMemHandle(&ih[1], h->old));
    memcpy(handles128, ih, sizeof(ih));
    return SL_OK;
}

int sl_state_attach(sl_handle* h, const void* handles128, int n, double peer_lambda, int* peer_out) {
    if (!h || !handles128 || !peer_out) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (n != h->D) return fail(h, SL_EINVAL, "sl_state_attach: model lengths differ (%d vs %d)", n, h->D);
    int slot = -1;
    for (int i = 0; i < kMaxRemoteStates; ++i)
        if (!h->remote[i].used) { slot = i; break; }
    if (slot < 0) return fail(h, SL_ESTATE, "sl_state_attach: %d peer states are already attached", kMaxRemoteStates);
    DeviceGuard dg(h->device);
    cudaIpcMemHandle_t ih[2];
    memcpy(ih, handles128, sizeof(ih));
    void* ptr[2] = {nullptr, nullptr};
    for (int k = 0; k < 2; ++k) {
        cudaError_t e = cudaIpcOpenMemHandle(&ptr[k], ih[k], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            if (k == 1) cudaIpcCloseMemHandle(ptr[0]);
            return fail(h, SL_ECUDA, "cudaIpcOpenMemHandle(%s) failed: %s (same process, or no peer access?)",
                        k ? "old" : "model", cudaGetErrorString(e));
        }
    }
    h->remote[slot].model = static_cast<double*>(ptr[0]);
    h->remote[slot].old = static_cast<double*>(ptr[1]);
    h->remote[slot].lambda = peer_lambda;
    h->remote[slot].used = true;
    *peer_out = slot;
    return SL_OK;
}

int sl_state_detach(sl_handle* h, int peer) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (peer < 0 || peer >= kMaxRemoteStates || !h->remote[peer].used)
        return fail(h, SL_EINVAL, "sl_state_detach: no such peer state");
    DeviceGuard dg(h->device);
    CK(h, cudaStreamSynchronize(h->stream));
    cudaIpcCloseMemHandle(h->remote[peer].model);
    cudaIpcCloseMemHandle(h->remote[peer].old);
    h->remote[peer] = sl_handle::RemoteState();
    return SL_OK;
}

// Initiator side: nothing of this replica runs or starts between lend and reclaim (the peer's kernel owns the
// state meanwhile).  The mutex is held across the two calls, so they come from one thread.
int sl_state_lend(sl_handle* h) {
    if (!h) return SL_EINVAL;
    h->mu.lock();
    DeviceGuard dg(h->device);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) {
        fail(h, SL_ECUDA, "sl_state_lend: %s", cudaGetErrorString(e));
        h->mu.unlock();
        return SL_ECUDA;
    }
    h->lent_by = std::this_thread::get_id();
    h->lent.store(true, std::memory_order_release);
    return SL_OK;
}

int sl_state_reclaim(sl_handle* h) {
    if (!h) return SL_EINVAL;
    // only the lending thread can get here with lent == true and be the owner of mu
    if (!h->lent.load(std::memory_order_acquire) || h->lent_by != std::this_thread::get_id()) return SL_ESTATE;
    h->lent.store(false, std::memory_order_relaxed);
    h->mu.unlock();
    return SL_OK;
}

int sl_gossip_p2p_remote(sl_handle* callee, int peer) {
    if (!callee) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(callee->mu);
    if (peer < 0 || peer >= kMaxRemoteStates || !callee->remote[peer].used)
        return fail(callee, SL_EINVAL, "sl_gossip_p2p_remote: no such peer state");
    DeviceGuard dg(callee->device);
    const sl_handle::RemoteState& r = callee->remote[peer];
    const int n = callee->D;
    slk::gossip_p2p_kernel<<<(n + 127) / 128, 128, 0, callee->stream>>>(r.model, r.old, callee->model, callee->old, n,
                                                                       r.lambda, callee->lambda);
    callee->launches++;
    CK(callee, cudaGetLastError());
    CK(callee, cudaStreamSynchronize(callee->stream));
    return SL_OK;
}

int sl_peer_export(sl_handle* h, void* handle64) {
    if (!h || !handle64) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard dg(h->device);
    if (!h->xblock) {
        const size_t bytes = xblock_bytes_for(h->F);  // inboxes and totals of K1/K1w, then K1c's accumulators
        CK(h, cudaMalloc(&h->xblock, bytes));
        CK(h, cudaMemset(h->xblock, 0, bytes));
        h->xblock_bytes = bytes;
    }
    cudaIpcMemHandle_t ih;
    CK(h, cudaIpcGetMemHandle(&ih, h->xblock));
    static_assert(sizeof(ih) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &ih, sizeof(ih));
    return SL_OK;
}

int sl_peer_attach(sl_handle* h, int nranks, int rank, const void* handles) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks)
        return fail(h, SL_EINVAL, "sl_peer_attach: 1 <= nranks <= 8 and 0 <= rank < nranks");
    if (!h->xblock) return fail(h, SL_ESTATE, "sl_peer_attach before sl_peer_export");
    if (h->peer_n) return fail(h, SL_ESTATE, "peers already attached");
    if (nranks > 1 && !handles) return fail(h, SL_EINVAL, "handles is NULL");
    DeviceGuard dg(h->device);
    for (int p = 0; p < nranks; ++p) {
        if (p == rank) { h->peer_base[p] = h->xblock; continue; }
        cudaIpcMemHandle_t ih;
        memcpy(&ih, static_cast<const unsigned char*>(handles) + 64 * (size_t)p, sizeof(ih));
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(h, SL_ECUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s", p, cudaGetErrorString(e));
        }
        h->peer_base[p] = ptr;
    }
    h->peer_n = nranks;
    h->peer_rank = rank;
    h->xflag_base = 0;
    h->cl_gstep = 0;
    h->ipc_attached = true;
    if (h->nranks == 1) { h->nranks = nranks; h->rank = rank; }
    return SL_OK;
}

int sl_peer_attach_symmetric(sl_handle* h, int nranks, int rank, const void* const* bufs, void* multicast, size_t bytes) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks)
        return fail(h, SL_EINVAL, "sl_peer_attach_symmetric: 1 <= nranks <= 8 and 0 <= rank < nranks");
    if (!bufs) return fail(h, SL_EINVAL, "bufs is NULL");
    if (h->peer_n && (h->peer_n != nranks || h->peer_rank != rank))
        return fail(h, SL_ESTATE, "peers already attached as rank %d of %d", h->peer_rank, h->peer_n);
    if (bytes < xblock_gx_bytes(h->F))
        return fail(h, SL_EINVAL, "the symmetric buffer holds %zu bytes, %zu are needed (128 * F)", bytes, xblock_gx_bytes(h->F));
    for (int p = 0; p < nranks; ++p) {
        if (!bufs[p] || ((uintptr_t)bufs[p] & 7)) return fail(h, SL_EINVAL, "buffer of rank %d is NULL or not 8-byte aligned", p);
        h->sym_buf[p] = static_cast<unsigned long long*>(const_cast<void*>(bufs[p]));
    }
    h->mc_gx = static_cast<unsigned long long*>(multicast);
    h->sym_n = nranks;
    h->peer_n = nranks;
    h->peer_rank = rank;
    h->cl_gstep = 0;
    if (h->nranks == 1) { h->nranks = nranks; h->rank = rank; }
    return SL_OK;
}

// Leave the group (every member does, before a group of another size is formed: a worker joined or died,
// src/master.cc:79-91 / :177-196).  Unmaps the peers, forgets a symmetric buffer and wipes the own inbox so that
// the next group starts from step tags and flags of zero.
int sl_peer_detach(sl_handle* h) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->peer_n) return SL_OK;
    DeviceGuard dg(h->device);
    CK(h, cudaStreamSynchronize(h->stream));
    if (h->ipc_attached)
        for (int p = 0; p < h->peer_n; ++p)
            if (p != h->peer_rank && h->peer_base[p]) cudaIpcCloseMemHandle(h->peer_base[p]);
    for (int p = 0; p < 8; ++p) { h->peer_base[p] = nullptr; h->sym_buf[p] = nullptr; }
    h->ipc_attached = false;
    h->mc_gx = nullptr;
    h->sym_n = 0;
    h->peer_n = 0;
    h->peer_rank = 0;
    h->xflag_base = 0;
    h->cl_gstep = 0;
    if (!h->comm) { h->nranks = 1; h->rank = 0; }
    if (h->xblock) CK(h, cudaMemset(h->xblock, 0, h->xblock_bytes));
    CK(h, cudaDeviceSynchronize());
    return SL_OK;
}

int sl_allreduce_avg(sl_handle* h) {
    if (!h) return SL_EINVAL;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->comm) return fail(h, SL_ESTATE, "sl_comm_init_rank has not been called");
    DeviceGuard dg(h->device);
    int rc = g_nccl.AllReduce(h->model, h->model, (size_t)h->D, kNcclFloat64, kNcclSum, h->comm,
                              h->stream);
    if (rc != 0) return fail(h, SL_ENCCL, "ncclAllReduce failed (%d)", rc);
    slk::scale_kernel<<<(h->D + 255) / 256, 256, 0, h->stream>>>(h->model, h->D,
                                                                1.0 / (double)h->nranks);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    return SL_OK;
}

}  // extern "C"

// =====================================================================================
// Allocation-free minibatch step
// =====================================================================================
// Scratch of sl_lr_step_noalloc: residuals, per-row losses and the row-split gradient
// partials of ONE step.  It lives in the module image, so every GPU that loads this
// library carries these bytes until the process exits; every call on a device shares it.
#define SL_NOALLOC_MAX_B 8192
#define SL_NOALLOC_MAX_F 4096
#define SL_NOALLOC_GPART_FLOATS (512 * 1024)
#define SL_NOALLOC_SCRATCH_BYTES \
    (sizeof(float) * (2 * SL_NOALLOC_MAX_B + SL_NOALLOC_GPART_FLOATS))
static __device__ __align__(256) unsigned char sl_noalloc_scratch[SL_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNaMaxDevices = 64;
std::mutex g_na_mu;
unsigned char* g_na_ptr[kNaMaxDevices] = {nullptr};  // device address of the scratch, per device
int g_na_sms[kNaMaxDevices] = {0};
}  // namespace

extern "C" int sl_lr_step_noalloc(int device, const float* d_X, const float* d_y, long long n_rows,
                                  int F, int B, long long t, double eta, double* d_model,
                                  double* d_loss, void* stream) {
    if (!d_X || !d_y || !d_model) return fail(nullptr, SL_EINVAL, "sl_lr_step_noalloc: NULL buffer");
    if (n_rows <= 0 || F <= 0 || B <= 0 || t < 0)
        return fail(nullptr, SL_EINVAL, "sl_lr_step_noalloc: n_rows, F, B must be positive, t >= 0");
    if (B > SL_NOALLOC_MAX_B || F > SL_NOALLOC_MAX_F)
        return fail(nullptr, SL_EINVAL,
                    "sl_lr_step_noalloc: B = %d / F = %d exceed the built-in scratch "
                    "(B <= %d, F <= %d); use sl_train_steps for larger shapes",
                    B, F, SL_NOALLOC_MAX_B, SL_NOALLOC_MAX_F);
    if (device < 0 || device >= kNaMaxDevices)
        return fail(nullptr, SL_EINVAL, "sl_lr_step_noalloc: device %d out of range", device);
    if (F % 4 == 0 && ((uintptr_t)d_X & 15) != 0)
        return fail(nullptr, SL_EINVAL, "sl_lr_step_noalloc: d_X must be 16-byte aligned when F is a multiple of 4 "
                    "(rows are read as float4)");
    DeviceGuard dg(device);
    unsigned char* base = nullptr;
    int sms = 0;
    {
        // One-time set-up per device, with that device current.  The array's name is only
        // a host-side handle; the address a kernel can use comes from the runtime.
        std::lock_guard<std::mutex> lk(g_na_mu);
        if (!g_na_ptr[device]) {
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, sl_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                cudaGetLastError();
                return fail(nullptr, SL_ECUDA, "cudaGetSymbolAddress(sl_noalloc_scratch) failed: %s",
                            cudaGetErrorString(e));
            }
            int v = 0;
            e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device);
            if (e != cudaSuccess || v <= 0) {
                cudaGetLastError();
                return fail(nullptr, SL_ECUDA, "cudaDeviceGetAttribute failed: %s",
                            cudaGetErrorString(e));
            }
            g_na_sms[device] = v;
            g_na_ptr[device] = static_cast<unsigned char*>(p);
        }
        base = g_na_ptr[device];
        sms = g_na_sms[device];
    }
    float* resid = reinterpret_cast<float*>(base);
    float* rowloss = resid + SL_NOALLOC_MAX_B;
    float* gpart = rowloss + SL_NOALLOC_MAX_B;
    // same row split as sl_train_steps' simple path, capped by the scratch
    const int col_blocks = (F + 127) / 128;
    int want = std::max(1, (4 * sms + col_blocks - 1) / col_blocks);
    want = std::min(want, B);
    want = std::min<long long>(want, SL_NOALLOC_GPART_FLOATS / F);
    const int rows_per_split = (B + want - 1) / want;
    const int nsplit = (B + rows_per_split - 1) / rows_per_split;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long mb = (long long)(((__int128)t * B) % n_rows);
    const int wpb = 8;
    // nothing in the scratch has to start at zero: the three kernels overwrite every word
    // they later read
    slk::lr_fwd_kernel<<<(B + wpb - 1) / wpb, wpb * 32, 0, st>>>(d_X, d_y, n_rows, F, B, mb, d_model,
                                                               resid, rowloss);
    slk::lr_grad_kernel<<<dim3(col_blocks, nsplit), 128, 0, st>>>(d_X, n_rows, F, B, mb, resid, gpart,
                                                                  rows_per_split);
    slk::lr_finish_kernel<<<(F + 255) / 256, 256, 0, st>>>(gpart, nsplit, F, B, rowloss, d_model,
                                                           eta / (double)B, d_loss, nullptr, 0);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        return fail(nullptr, SL_ECUDA, "sl_lr_step_noalloc: launch failed: %s", cudaGetErrorString(e));
    return SL_OK;
}
