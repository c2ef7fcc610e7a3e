// kernels.cuh — hand-written sm_100a kernels for the serverless_learn hot path.
//
//   K1  lr_persist_kernel   persistent, one CTA per SM: the whole minibatch step
//                           (forward GEMV, sigmoid/loss, transposed GEMV, cross-CTA
//                           reduction, weight update) for n_steps steps in ONE launch.
//                           Takes the slot of simulate_training (src/worker.cc:221-231).
//   K1s lr_fwd / lr_grad / lr_finish   three-launch "simple" step for shapes the
//                           persistent kernel does not cover, and for the NCCL mode.
//   K2  merge_kernel        src/master.cc:105-110 == src/worker.cc:91-96
//   K3  apply_kernel        src/worker.cc:161-162 + :215
//   K4  delta_kernel        src/worker.cc:205-208
//   K6  tick_kernel         src/worker.cc:225-227
//   H1  checksum_kernel     device-side check of the ingested byte stream
//   S1  synth_kernel        synthetic shard, bit-identical to the oracle
//
// All file:line citations are relative to the reference's root.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace slk {

// ------------------------------------------------------------------------------------
// PTX helpers (sm_90+/sm_100a: mbarrier, bulk async copy = TMA engine, volatile vectors)
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier.
// SASS: UBLKCP (the TMA unit); no tensor map needed because a row block is contiguous.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                         uint32_t bar, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "l"(pol)
        : "memory");
}
// Packet loads / stores of the exchange.  SL_LL_MODE (experiments): 0 = volatile
// (STRONG.SYS), 1 = relaxed.gpu both ways, 2 = relaxed.gpu loads + weak .cg stores.
#ifndef SL_LL_MODE
#define SL_LL_MODE 0
#endif
__device__ __forceinline__ uint4 ld_vol_v4(const uint4* p) {
    uint4 r;
#if SL_LL_MODE == 0
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
#else
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
#endif
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_vol_v4(uint4* p, uint4 v) {
#if SL_LL_MODE == 0
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};"
#elif SL_LL_MODE == 1
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};"
#else
    asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};"
#endif
                 ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ int ld_vol_s32(const int* p) {
    int r;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ float4 lds_v4(uint32_t addr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "r"(addr)
                 : "memory");
    return r;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------------------------
// Watchdog: every in-kernel wait is bounded so a protocol bug ends in SL_ETIMEOUT, not
// in a hung GPU.  kSpinLimit polls of >= ~300 cycles each is on the order of a second.
// ------------------------------------------------------------------------------------
constexpr uint32_t kSpinLimit = 1u << 22;
constexpr uint32_t kPeerSpinLimit = 1u << 26;  // another GPU's kernel may start much later
constexpr int kErrTimeoutMbar = 1;
constexpr int kErrTimeoutPart = 2;
constexpr int kErrTimeoutW = 3;

#ifndef SL_POLL_SLEEP
#define SL_POLL_SLEEP 0
#endif
constexpr int kErrTimeoutPeer = 4;
constexpr int kErrRange = 5;   // a gradient partial left the fixed-point range (reported as SL_EINVAL, not as a timeout)
// The error word lives in page-locked HOST memory (the host reads it after the launch): plain loads and stores
// only — an atomic on it would be an atomic across PCIe, which this platform does not take from a whole grid.
// A posted store, no read-back (thousands of threads may report at once; reads of host memory would queue up on
// PCIe): a later code overwriting an earlier one is harmless, any non-zero value fails the call.
__device__ __forceinline__ void set_err(int* err, int code) {
    asm volatile("st.volatile.global.s32 [%0], %1;" ::"l"(err), "r"(code) : "memory");
}
__device__ __forceinline__ bool spin_abort(uint32_t& spins, int* err, int code,
                                           uint32_t limit = kSpinLimit) {
    ++spins;
    if (SL_POLL_SLEEP > 0) __nanosleep(SL_POLL_SLEEP);
    if ((spins & 1023u) == 0) {
        if (ld_vol_s32(err) != 0) return true;
        if (spins >= limit) {
            set_err(err, code);
            return true;
        }
    }
    return false;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (spin_abort(spins, err, kErrTimeoutMbar)) return;
    }
}

// ------------------------------------------------------------------------------------
// Multi-value warp reduction: N (power of two <= 32) per-lane partials -> lane L ends
// with the warp total of value row_of_lane<N>(L).  N-1 + (5-log2 N) shuffles instead of
// 5*N.  The order of additions is fixed, so results are run-to-run deterministic.
// ------------------------------------------------------------------------------------
template <int N, int OFF>
struct MultiReduce {
    static __device__ __forceinline__ void run(float* v, int lane) {
        constexpr int H = N / 2;
        const bool upper = (lane & OFF) != 0;
#pragma unroll
        for (int i = 0; i < H; ++i) {
            const float send = upper ? v[i] : v[i + H];
            const float keep = upper ? v[i + H] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
        }
        MultiReduce<H, OFF / 2>::run(v, lane);
    }
};
template <int OFF>
struct MultiReduce<1, OFF> {
    static __device__ __forceinline__ void run(float* v, int) {
#pragma unroll
        for (int off = OFF; off >= 1; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
};
template <>
struct MultiReduce<1, 0> {
    static __device__ __forceinline__ void run(float*, int) {}
};

template <int N>
__device__ __forceinline__ int row_of_lane(int lane) {
    int idx = 0;
#pragma unroll
    for (int k = 0, h = N / 2; h >= 1; ++k, h >>= 1)
        if (lane & (16 >> k)) idx += h;
    return idx;
}
template <int N>
__device__ __forceinline__ constexpr int lane_of_row(int r) {
    int lane = 0;
    for (int k = 0, h = N / 2; h >= 1; ++k, h >>= 1)
        if (r & h) lane |= (16 >> k);
    return lane;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// loss of one row, softplus(z) - y*z, summed so that the large parts cancel exactly first
// (y in {0,1}: max(z,0) - y*z is 0 or |z| without rounding; adding it to log1p(...) afterwards
// keeps the small term that softplus(z) - y*z would round away once |z| exceeds ~16)
__device__ __forceinline__ float row_loss_f(float z, float y) {
    return (fmaxf(z, 0.0f) - y * z) + log1pf(expf(-fabsf(z)));
}
__device__ __forceinline__ float sigmoid_f(float z) { return 1.0f / (1.0f + expf(-z)); }

// ------------------------------------------------------------------------------------
// K1 — persistent minibatch-step kernel
// ------------------------------------------------------------------------------------
// Warp roles inside one CTA (one CTA per SM, persistent over all steps of the launch):
//   warps [0, NCW)   compute: RG row groups x WPG warps; a thread owns CH float4 column
//                    chunks of every row its group processes and the fp64 master copy of
//                    the weights of those columns (every CTA updates all weights itself,
//                    from the same exact integer gradient sum, so all replicas stay equal)
//   warp  NCW        producer: one lane issues the TMA bulk copies of X, NS ring slots
//                    ahead of the maths; never depends on the weights
//   warp  NCW+1      exchange: one lane adds the CTA's gradient (fixed point, s64) into
//                    a global accumulator with ONE TMA bulk reduction, bumps a counter,
//                    waits until every CTA has done so and TMA-loads the total
// Cross-CTA reduction = integer adds performed by L2 => exact, order-independent,
// deterministic.  Accumulators are triple-buffered over steps and replicated KREP times
// (CTA c uses replica c % KREP) so that a line is hit by G/KREP, not G, reductions.
//
// CH  float4 column chunks per thread per row        (F <= 4*CH*32*NCW)
// R   rows per sub-batch of one row group (power of two), held in registers at once
// RS  rows per ring slot (divides R): one bulk copy and one mbarrier per slot
// XB  exchange through shared memory + TMA (F small enough for smem staging); otherwise
//     per-element red.global.add.u64 and direct loads (wide rows, where a step is long)
// WM  where the weights live.  0: fp32 copy and fp64 master of a thread's columns in its
//     registers.  1 (wide rows): fp32 in shared memory, fp64 master in a per-CTA global scratch.
//     2 (many row groups): fp32 in registers, ONE fp64 master per CTA in shared memory, updated
//     by the first F/4 threads and broadcast through shared memory — frees 8 registers per
//     column chunk, which is what lets 7 row groups hold a CTA's 28 rows of a step at once
constexpr int kFixShift = 36;                       // gradient fixed point: 2^-36 units
constexpr double kFixScale = 68719476736.0;         // 2^36
constexpr double kFixInv = 1.0 / 68719476736.0;

struct LrArgs {
    const float* X;        // [n_rows, F] row-major, resident shard
    const float* y;        // [n_rows]
    long long n_rows;
    int F;
    int B;
    long long step0;       // global step index of the first step of this launch
    int n_steps;
    double* model;         // fp64 master weights [F] (model_state, src/master.cc:58)
    unsigned long long* acc;   // [3][krep][Fp] s64 gradient accumulators (zero at launch)
    unsigned* counter;     // arrivals; zero at launch
    double* wmaster;       // WM=1 only: [grid][F] per-CTA fp64 master copies
    float* w32buf;         // K1w: [2][F] fp32 weights published by the slice owners
    // same-box data parallel (nranks > 1 or peers attached): one kernel per GPU, the kernels
    // talk through peer-mapped memory over NVLink
    int nranks, rank;      // nranks == 0: single-GPU exchange
    uint4* peer_x[8];      // rank p's inbox [2][nranks][F] of {lo, flag, hi, flag} packets
    unsigned long long* fin;   // [3][F] all-rank gradient totals of this GPU
    unsigned* counter2;    // owner CTAs that have published their slice; zero at launch
    unsigned xflag0;       // step s of this launch uses packet flag xflag0 + s + 1
    int n_own;             // CTAs that own a slice of the vector in the cross-GPU stage
    int xslice;            // elements per slice (32 in lr_persist_kernel)
    int tma_chunk;         // bytes per bulk copy (a row block is fetched as several copies on one mbarrier)
    int dbg_skip;          // diagnostics: K1w skips the per-row maths (measures the bare TMA ring rate)
    int block_rows;        // K1w: 1 = a CTA takes a contiguous block of the minibatch instead of every G-th row
    unsigned long long* loss_fix;  // device [n_steps] s64 fixed-point loss sums (zero at launch) or nullptr
    double eta_over_b;
    float fix_limit;       // |gradient partial| at or above this could overflow the 64-bit fixed-point sums
    int* err;              // 0 = ok; set by the watchdog
    // geometry, computed on the host
    int Fp;                // accumulator vector length (= F)
    int zy_rows;           // rows of the deferred-loss buffer (>= this CTA's rows per step)
    int krep;              // accumulator replicas
    int wpg;               // warps per row group = ceil(F/4/CH/32)
    int ns;                // ring slots (RS rows each)
    int row_bytes;         // 4*F
    int off_bar, off_red, off_zy, off_gsm, off_gout, off_gin, off_wsm, off_msm;  // smem offsets
    // diagnostics (sl_set_option "debug_timeline"): [4 steps][grid][16] time stamps
    unsigned long long* dbg;
    int dbg_step;
    // several worker replicas on ONE GPU trained as one job (sl_train_steps_group): the step's
    // B = nvw*vB positions are worker 0's minibatch, then worker 1's, ...; position p belongs
    // to worker p / vB and is row (step*vB + p % vB) mod n_rows of that worker's shard
    // (vX[k], vy[k]; all shards hold n_rows rows).  nvw <= 1: one shard (X, y), vB = B.
    int nvw;
    int vB;
    const float* vX[8];
    const float* vy[8];
    // several INDEPENDENT replicas on one GPU, each taking its own local steps (sl_train_steps_multi):
    // the grid is cut into npart equal partitions of CTAs; partition k trains pmodel[k] on shard
    // (vX[k], vy[k]) with its own accumulators, counter and loss slots.  npart <= 1: one replica.
    int npart;
    int loss_stride;       // loss slots per partition (= steps of the whole call)
    double* pmodel[8];
};

// where position p of the current minibatch lives: shard row pointer and label pointer;
// `room` = how many consecutive positions from p lie in consecutive rows of the same shard
struct RowRef {
    const float* x;
    const float* y;
    long long room;
};
template <class Args>
__device__ __forceinline__ RowRef locate_row(const Args& a, long long mb, long long p, int part = 0) {
    const float* X = a.npart > 1 ? a.vX[part] : a.X;
    const float* y = a.npart > 1 ? a.vy[part] : a.y;
    long long r = p, lim = a.n_rows;
    if (a.nvw > 1) {
        const int k = (int)(p / a.vB);
        r = p - (long long)k * a.vB;
        X = a.vX[k];
        y = a.vy[k];
        lim = a.vB - r;  // the next worker's rows start there
    }
    long long row = mb + r;
    if (row >= a.n_rows) {
        row -= a.n_rows;
        if (row >= a.n_rows) row %= a.n_rows;
    }
    RowRef ref;
    ref.x = X + (size_t)row * a.F;
    ref.y = y + row;
    ref.room = a.n_rows - row;  // the shard wraps there
    if (a.nvw > 1 && lim < ref.room) ref.room = lim;
    return ref;
}

template <class Args>
__device__ __forceinline__ void dbg_stamp(const Args& a, int s, int point) {
    if (a.dbg != nullptr && s >= a.dbg_step && s < a.dbg_step + 4) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        unsigned long long* d = a.dbg + ((size_t)(s - a.dbg_step) * gridDim.x + blockIdx.x) * 16;
        d[2 * point] = (unsigned long long)clock64();
        d[2 * point + 1] = t;
    }
}

__device__ __forceinline__ void bulk_reduce_add_u64(void* gdst, uint32_t ssrc, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;" ::"l"(gdst),
                 "r"(ssrc), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s_plain(uint32_t dst, const void* src, uint32_t bytes,
                                               uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar)
        : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ ulonglong2 ld_relaxed_v2u64(const unsigned long long* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u32(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// exact: scaling by 2^36 does not round in fp32, the conversion rounds once
__device__ __forceinline__ long long to_fix(float v) { return __float2ll_rn(v * 68719476736.0f); }
// Sums of these integers must fit 64 bits: a partial beyond `limit` (2^27 over the number of partials that
// are added up; NaN too) is reported through the watchdog word instead of wrapping silently.
__device__ __forceinline__ bool fix_out_of_range(const float4& t, float limit) {
    const float m = fmaxf(fmaxf(fabsf(t.x), fabsf(t.y)), fmaxf(fabsf(t.z), fabsf(t.w)));
    return !(m < limit);
}
__device__ __forceinline__ void check_fix_range(const float4& t, float limit, int* err) {
    if (fix_out_of_range(t, limit)) set_err(err, kErrRange);
}

template <int CH, int R, int RS, int NCW, bool XB, int WM>
__global__ void __launch_bounds__((NCW + 2) * 32, 1) lr_persist_kernel(const LrArgs a) {
    static_assert(R % RS == 0, "a sub-batch is a whole number of ring slots");
    constexpr bool WS = WM == 1;  // fp32 weights read from smem in the dot, fp64 master in global
    constexpr bool MS = WM == 2;  // fp32 weights in registers, fp64 master in smem (one copy per CTA)
    constexpr int SPB = R / RS;  // ring slots per sub-batch
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // grid partition of this CTA (sl_train_steps_multi): independent replicas, nothing shared
    const int G = a.npart > 1 ? (int)gridDim.x / a.npart : (int)gridDim.x;
    const int part = a.npart > 1 ? (int)blockIdx.x / G : 0;
    const int c = (int)blockIdx.x - part * G;
    double* const model = a.npart > 1 ? a.pmodel[part] : a.model;
    unsigned long long* const acc = a.acc + (size_t)part * 3 * (size_t)a.krep * (size_t)a.Fp;
    unsigned* const counter = a.counter + part * 8;
    unsigned long long* const loss_fix = a.loss_fix ? a.loss_fix + (size_t)part * a.loss_stride : nullptr;

    unsigned char* ring = smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_bar);    // [ns]
    uint64_t* empty = full + a.ns;                                     // [ns]
    uint64_t* gfull = empty + a.ns;                                    // [1] gradient staged
    uint64_t* wfull = gfull + 1;                                       // [1] total received
    float* red = reinterpret_cast<float*>(smem + a.off_red);          // [2][NCW][R]
    float2* zy = reinterpret_cast<float2*>(smem + a.off_zy);          // [2][zy_rows] (z, y) per row
    float* gsm = reinterpret_cast<float*>(smem + a.off_gsm);          // [RG-1][F]
    long long* gout = reinterpret_cast<long long*>(smem + a.off_gout);  // XB: [Fp]
    long long* gin = reinterpret_cast<long long*>(smem + a.off_gin);    // XB: [krep][Fp]
    float* wsm = reinterpret_cast<float*>(smem + a.off_wsm);          // WM != 0: [F] fp32 weights
    double* msm = reinterpret_cast<double*>(smem + a.off_msm);        // WM == 2: [F] fp64 master

    const int WPG = a.wpg, RG = NCW / WPG, CT = WPG * 32;
    const int F = a.F, Fp = a.Fp, ns = a.ns, row_bytes = a.row_bytes, KREP = a.krep;
    const uint32_t slot_bytes = (uint32_t)RS * (uint32_t)row_bytes;
    const long long lo = ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = (int)(hi - lo);             // this CTA's rows of every minibatch
    const int nsl_c = (nrows_c + RS - 1) / RS;      // ... in ring slots
    const int nst_c = (nrows_c + R - 1) / R;        // ... in sub-batch stages of R rows
    const int nsb = (nst_c + RG - 1) / RG;          // sub-batches: one stage per row group
    const long long mb0 = (long long)(((__int128)a.step0 * a.vB) % a.n_rows);
    const size_t acc_stride = (size_t)KREP * Fp;    // one step buffer
    auto wrap = [&](long long v) {  // v mod n_rows; v < 2*n_rows in the common case
        if (v >= a.n_rows) {
            v -= a.n_rows;
            if (v >= a.n_rows) v %= a.n_rows;
        }
        return v;
    };

    if (tid == 0) {
        for (int s = 0; s < ns; ++s) {
            mbar_init(smem_u32(&full[s]), 1);
            mbar_init(smem_u32(&empty[s]), WPG);
        }
        {   // arrivals that complete "gradient staged": the warps (XB) or threads that write it
            const int granules = F / 4;
            const int comb_warps = RG > 1 ? min(NCW, (granules + 31) / 32) : WPG;
            const int comb_threads = RG > 1 ? min(granules, NCW * 32) : WPG * 32;
            mbar_init(smem_u32(gfull), XB ? comb_warps : comb_threads);
        }
        mbar_init(smem_u32(wfull), 1);
        fence_mbar_init();
    }
    __syncthreads();

    // =========================================================================== producer
    if (warp == NCW) {
        if (lane == 0 && nsl_c > 0) {
            const uint64_t pol = policy_evict_first();
            uint32_t slot = 0, phase = 0;
            long long mb = mb0;
            uint32_t pos = 0;        // position of the slot being issued in this CTA's stream
            for (int s = 0; s < a.n_steps; ++s) {
                for (int q = 0; q < nsl_c; ++q, ++pos) {
                    mbar_wait(smem_u32(&empty[slot]), phase ^ 1u, a.err);
                    const int rc = min(RS, nrows_c - q * RS);
                    const uint32_t fb = smem_u32(&full[slot]);
                    mbar_arrive_expect_tx(fb, (uint32_t)rc * (uint32_t)row_bytes);
                    long long p = lo + (long long)q * RS;  // position in the minibatch
                    uint32_t dst = smem_u32(ring + (size_t)slot * slot_bytes);
                    int left = rc;
                    while (left > 0) {  // one copy unless the rows wrap around the shard end (or change worker)
                        const RowRef ref = locate_row(a, mb, p, part);
                        const int n = ref.room < (long long)left ? (int)ref.room : left;
                        {   // several moderate copies keep more requests in flight than one large one
                            const unsigned char* srcb = reinterpret_cast<const unsigned char*>(ref.x);
                            uint32_t todo = (uint32_t)n * (uint32_t)row_bytes, o = 0;
                            while (todo > 0) {
                                const uint32_t m = todo < (uint32_t)a.tma_chunk ? todo : (uint32_t)a.tma_chunk;
                                bulk_g2s(dst + o, srcb + o, m, fb, pol);
                                o += m;
                                todo -= m;
                            }
                        }
                        dst += (uint32_t)n * (uint32_t)row_bytes;
                        left -= n;
                        p += n;
                    }
                    if (++slot == (uint32_t)ns) { slot = 0; phase ^= 1u; }
                }
                mb = wrap(mb + a.vB);
            }
        }
        return;
    }

    // =========================================================================== exchange
    if (warp == NCW + 1) {
        for (int s = 0; s < a.n_steps; ++s) {
            unsigned long long* buf = acc + (size_t)(s % 3) * acc_stride;
            // the CTA's gradient is staged (XB) or has been added by the compute threads
            mbar_wait(smem_u32(gfull), (uint32_t)s & 1u, a.err);
            if (lane == 0) {
                dbg_stamp(a, s, 6);
                if (XB) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    bulk_reduce_add_u64(buf + (size_t)(c % KREP) * Fp, smem_u32(gout), (uint32_t)Fp * 8u);
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
                    dbg_stamp(a, s, 7);
                    asm volatile("fence.proxy.async;" ::: "memory");
                }
                red_release_add_u32(counter, 1u);
                dbg_stamp(a, s, 3);
                const unsigned want = (unsigned)G * (unsigned)(s + 1);
                uint32_t spins = 0;
                while ((int)(ld_relaxed_u32(counter) - want) < 0) {
                    if (spin_abort(spins, a.err, kErrTimeoutPart)) break;
                }
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                if (a.nranks == 0) {
                    if (XB) {
                        asm volatile("fence.proxy.async;" ::: "memory");
                        const uint32_t wb = smem_u32(wfull);
                        mbar_arrive_expect_tx(wb, (uint32_t)acc_stride * 8u);
                        bulk_g2s_plain(smem_u32(gin), buf, (uint32_t)acc_stride * 8u, wb);
                    } else {
                        mbar_arrive(smem_u32(wfull));
                    }
                    dbg_stamp(a, s, 4);
                }
            }
            if (XB && a.nranks > 0) {
                // ---- cross-GPU stage, fused into the same kernel: the CTA that owns a 32-element
                // slice sums this GPU's replicas of it, stores the sum into every peer's inbox over
                // NVLink (16-byte data+flag packets), adds the peers' sums as they land, publishes
                // the all-rank total locally.  s64 adds: exact, identical on every GPU.
                __syncwarp();  // lane 0 has seen every local CTA's contribution (acquire above)
                const unsigned xflag = a.xflag0 + (unsigned)s + 1u;
                const size_t half = (size_t)(xflag & 1u) * (size_t)a.nranks * (size_t)F;
                unsigned long long* fin = a.fin + (size_t)(s % 3) * F;
                if (c < a.n_own) {
                    const int e = c * 32 + lane;
                    if (e < F) {
                        long long tot = 0;
                        for (int k = 0; k < KREP; ++k) tot += (long long)ld_relaxed_u64(buf + (size_t)k * Fp + e);
                        const uint4 pkt = make_uint4((unsigned)((unsigned long long)tot & 0xffffffffu), xflag,
                                                     (unsigned)((unsigned long long)tot >> 32), xflag);
                        for (int p = 0; p < a.nranks; ++p)
                            if (p != a.rank) st_vol_v4(a.peer_x[p] + half + (size_t)a.rank * F + e, pkt);
                        const uint4* inbox = a.peer_x[a.rank] + half;
                        for (int p0 = 0; p0 < a.nranks; p0 += 4) {  // 4 peers' packets in flight
                            uint4 q[4];
                            uint32_t spins = 0;
                            for (;;) {
                                bool ok = true;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const int p = p0 + k;
                                    if (p < a.nranks && p != a.rank) {
                                        q[k] = ld_vol_v4(inbox + (size_t)p * F + e);
                                        ok = ok && q[k].y == xflag && q[k].w == xflag;
                                    }
                                }
                                if (ok) break;
                                if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                            }
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const int p = p0 + k;
                                if (p < a.nranks && p != a.rank)
                                    tot += (long long)(((unsigned long long)q[k].z << 32) | (unsigned long long)q[k].x);
                            }
                        }
                        fin[e] = (unsigned long long)tot;
                    }
                    __syncwarp();
                    if (lane == 0) { dbg_stamp(a, s, 7); red_release_add_u32(a.counter2, 1u); }
                }
                if (lane == 0) {
                    const unsigned want2 = (unsigned)a.n_own * (unsigned)(s + 1);
                    uint32_t spins = 0;
                    while ((int)(ld_relaxed_u32(a.counter2) - want2) < 0) {
                        if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                    }
                    asm volatile("fence.acq_rel.gpu;" ::: "memory");
                    asm volatile("fence.proxy.async;" ::: "memory");
                    const uint32_t wb = smem_u32(wfull);
                    mbar_arrive_expect_tx(wb, (uint32_t)F * 8u);
                    bulk_g2s_plain(smem_u32(gin), fin, (uint32_t)F * 8u, wb);
                    dbg_stamp(a, s, 4);
                }
            }
            __syncwarp();
            // Everybody has contributed to step s, hence finished reading the buffer of
            // step s-1: clear this CTA's slice of it for step s+2.  The stores precede this
            // warp's next release on the counter.
            unsigned long long* z = acc + (size_t)((s + 2) % 3) * acc_stride;
            for (size_t i = (size_t)c * 32 + lane; i < acc_stride; i += (size_t)G * 32) z[i] = 0ull;
            __syncwarp();
        }
        return;
    }

    // =========================================================================== compute
    const int rg = warp / WPG, wig = warp - rg * WPG;
    const int ct = wig * 32 + lane;
    const bool active = rg < RG;  // warps beyond RG*WPG idle through the maths (still sync)
    const int ridx = row_of_lane<R>(lane);
    const bool canon = lane == lane_of_row<R>(ridx);
    constexpr int NCT = NCW * 32;

    bool gv[CH];
    int gq[CH];
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        gq[ch] = ct + ch * CT;
        gv[ch] = active && (gq[ch] * 4 < F);
    }
    // weights: fp64 master + fp32 working copy, per thread for its own columns
    constexpr int NW = WS ? 1 : CH;
    constexpr int NM = WM == 0 ? CH : 1;
    float4 w[NW];
    double wm[NM][4];
    const int granules = F / 4;
    const bool comb = RG > 1 ? tid < granules : rg == 0;  // threads that stage the CTA's gradient
    const uint32_t wsm_a = smem_u32(wsm);
    double* my_master = WS ? a.wmaster + (size_t)c * F : nullptr;
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        if (!gv[ch]) {
            if (!WS) w[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (WM == 0) { wm[ch][0] = wm[ch][1] = wm[ch][2] = wm[ch][3] = 0.0; }
            continue;
        }
        const double* m = model + gq[ch] * 4;
        const float4 wf = make_float4((float)m[0], (float)m[1], (float)m[2], (float)m[3]);
        if (WS) {
            if (rg == 0) {  // one group seeds the CTA's master copy; every group its own smem words
                double* mm = my_master + gq[ch] * 4;
                mm[0] = m[0]; mm[1] = m[1]; mm[2] = m[2]; mm[3] = m[3];
            }
            *reinterpret_cast<float4*>(wsm + gq[ch] * 4) = wf;
        } else {
            w[ch] = wf;
            if (WM == 0) { wm[ch][0] = m[0]; wm[ch][1] = m[1]; wm[ch][2] = m[2]; wm[ch][3] = m[3]; }
        }
    }
    if (MS && tid < granules) {
        const double* m = model + tid * 4;
        msm[tid * 4 + 0] = m[0]; msm[tid * 4 + 1] = m[1]; msm[tid * 4 + 2] = m[2]; msm[tid * 4 + 3] = m[3];
    }
    if (WS) named_bar_sync(1, NCT);

    // apply the exact gradient total of step s to this thread's weights
    auto apply_total = [&](int s) {
        const unsigned long long* buf = acc + (size_t)(s % 3) * acc_stride;
        if (MS) {
            if (tid < granules) {  // one updater per column chunk; XB is implied by the variant table
                long long t[4] = {0, 0, 0, 0};
                const int nrep = a.nranks > 0 ? 1 : KREP;
                for (int k = 0; k < nrep; ++k) {
                    const longlong2* p = reinterpret_cast<const longlong2*>(gin + (size_t)k * Fp + tid * 4);
                    const longlong2 u = p[0], v = p[1];
                    t[0] += u.x; t[1] += u.y; t[2] += v.x; t[3] += v.y;
                }
                double* mm = msm + tid * 4;
                const double v0 = mm[0] - a.eta_over_b * ((double)t[0] * kFixInv);
                const double v1 = mm[1] - a.eta_over_b * ((double)t[1] * kFixInv);
                const double v2 = mm[2] - a.eta_over_b * ((double)t[2] * kFixInv);
                const double v3 = mm[3] - a.eta_over_b * ((double)t[3] * kFixInv);
                mm[0] = v0; mm[1] = v1; mm[2] = v2; mm[3] = v3;
                *reinterpret_cast<float4*>(wsm + tid * 4) = make_float4((float)v0, (float)v1, (float)v2, (float)v3);
            }
            named_bar_sync(1, NCT);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (gv[ch]) w[WS ? 0 : ch] = lds_v4(wsm_a + (uint32_t)gq[ch] * 16u);
            return;
        }
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (!gv[ch]) continue;
            long long t[4] = {0, 0, 0, 0};
            if (XB) {
                const int nrep = a.nranks > 0 ? 1 : KREP;  // multi-GPU: gin holds the one all-rank total
                for (int k = 0; k < nrep; ++k) {
                    const longlong2* p = reinterpret_cast<const longlong2*>(gin + (size_t)k * Fp + gq[ch] * 4);
                    const longlong2 u = p[0], v = p[1];
                    t[0] += u.x; t[1] += u.y; t[2] += v.x; t[3] += v.y;
                }
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) t[e] = (long long)ld_relaxed_u64(buf + gq[ch] * 4 + e);
            }
            if (WS) {
                if (rg == 0 || RG == 1) {
                    double* mm = my_master + gq[ch] * 4;
                    float4 wf;
                    double v0 = mm[0] - a.eta_over_b * ((double)t[0] * kFixInv);
                    double v1 = mm[1] - a.eta_over_b * ((double)t[1] * kFixInv);
                    double v2 = mm[2] - a.eta_over_b * ((double)t[2] * kFixInv);
                    double v3 = mm[3] - a.eta_over_b * ((double)t[3] * kFixInv);
                    mm[0] = v0; mm[1] = v1; mm[2] = v2; mm[3] = v3;
                    wf = make_float4((float)v0, (float)v1, (float)v2, (float)v3);
                    *reinterpret_cast<float4*>(wsm + gq[ch] * 4) = wf;
                }
            } else if (WM == 0) {
#pragma unroll
                for (int e = 0; e < 4; ++e) wm[ch][e] -= a.eta_over_b * ((double)t[e] * kFixInv);
                w[ch] = make_float4((float)wm[ch][0], (float)wm[ch][1], (float)wm[ch][2], (float)wm[ch][3]);
            }
        }
        if (WS && RG > 1) named_bar_sync(1, NCT);  // other groups read rg 0's smem weights
    };

    uint32_t base_slot = 0, base_phase = 0;  // ring position of this step's slot 0
    int red_buf = 0;
    long long mb = mb0;
    // label of the row this lane finishes, fetched one stage ahead of its use
    auto load_y = [&](long long mbs, int j) -> float {
        const int i = j * R + ridx;
        return (active && j < nst_c && i < nrows_c) ? __ldg(locate_row(a, mbs, lo + i, part).y) : 0.f;
    };
    float y_next = load_y(mb, rg);

    for (int s = 0; s < a.n_steps; ++s) {
        const long long mb_next = wrap(mb + a.vB);
        float4 g[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        float2* zyb = zy + (size_t)(s & 1) * a.zy_rows;
        if (tid == 0) dbg_stamp(a, s, 0);
        // a CTA without rows still takes part in every exchange
        if (nsb == 0 && s > 0) {
            mbar_wait(smem_u32(wfull), (uint32_t)(s - 1) & 1u, a.err);
            apply_total(s - 1);
        }

        for (int sb = 0; sb < nsb; ++sb) {
            const int j = sb * RG + rg;               // this group's stage of the step
            const bool have = active && j < nst_c;    // warp-uniform
            const int rc = have ? min(R, nrows_c - j * R) : 0;
            const float yv = y_next;
            {   // prefetch the label for this group's next stage (maybe next step's first)
                if (sb + 1 < nsb) y_next = load_y(mb, j + RG);
                else y_next = (s + 1 < a.n_steps) ? load_y(mb_next, rg) : 0.f;
            }
            float4 x[R][CH];
            if (have) {
                uint32_t sl = base_slot + (uint32_t)(j * SPB);
                uint32_t k = sl / (uint32_t)ns;
                uint32_t slot = sl - k * (uint32_t)ns;
                uint32_t par = (base_phase ^ k) & 1u;
#pragma unroll
                for (int q = 0; q < SPB; ++q) {
                    if (q * RS < rc) {  // warp-uniform
                        mbar_wait(smem_u32(&full[slot]), par, a.err);
                        const uint32_t sbase = smem_u32(ring + (size_t)slot * slot_bytes);
#pragma unroll
                        for (int rr = 0; rr < RS; ++rr) {
                            const int r = q * RS + rr;
#pragma unroll
                            for (int ch = 0; ch < CH; ++ch)
                                x[r][ch] = (gv[ch] && r < rc)
                                               ? lds_v4(sbase + (uint32_t)rr * (uint32_t)row_bytes +
                                                        (uint32_t)gq[ch] * 16u)
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(&empty[slot]));  // rows are in registers
                        if (++slot == (uint32_t)ns) { slot = 0; par ^= 1u; }
                    } else {
#pragma unroll
                        for (int rr = 0; rr < RS; ++rr)
#pragma unroll
                            for (int ch = 0; ch < CH; ++ch)
                                x[q * RS + rr][ch] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r)
#pragma unroll
                    for (int ch = 0; ch < CH; ++ch) x[r][ch] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (sb == 0 && s > 0) {
                // the rows of the step's first stage are already in registers: only now
                // wait for the previous step's gradient total and update the weights
                mbar_wait(smem_u32(wfull), (uint32_t)(s - 1) & 1u, a.err);
                apply_total(s - 1);
                if (tid == 0) dbg_stamp(a, s, 5);
            }
            // partial dot products over this thread's columns
            float d[R];
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = 0.f;
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                float4 wv;
                if (WS) wv = gv[ch] ? lds_v4(wsm_a + (uint32_t)gq[ch] * 16u)
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
                else wv = w[ch];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float acc = d[r];
                    acc = fmaf(x[r][ch].x, wv.x, acc);
                    acc = fmaf(x[r][ch].y, wv.y, acc);
                    acc = fmaf(x[r][ch].z, wv.z, acc);
                    acc = fmaf(x[r][ch].w, wv.w, acc);
                    d[r] = acc;
                }
            }
            MultiReduce<R, 16>::run(d, lane);
            float z = d[0];  // warp total for row ridx
            if (WPG > 1 && active) {  // warp-uniform; WPG > 1 implies RG <= 8, ids 2..9
                float* rb = red + red_buf * (NCW * R);
                if (canon) rb[(rg * WPG + wig) * R + ridx] = z;
                named_bar_sync(2 + rg, CT);
                z = 0.f;
                for (int k = 0; k < WPG; ++k) z += rb[(rg * WPG + k) * R + ridx];
                red_buf ^= 1;
            }
            float resid = 0.f;
            if (ridx < rc) {
                resid = sigmoid_f(z) - yv;
                // the loss is evaluated after the gradient has been handed off
                if (canon && wig == 0 && loss_fix) zyb[j * R + ridx] = make_float2(z, yv);
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float rr = __shfl_sync(0xffffffffu, resid, lane_of_row<R>(r));
#pragma unroll
                for (int ch = 0; ch < CH; ++ch) {
                    g[ch].x = fmaf(rr, x[r][ch].x, g[ch].x);
                    g[ch].y = fmaf(rr, x[r][ch].y, g[ch].y);
                    g[ch].z = fmaf(rr, x[r][ch].z, g[ch].z);
                    g[ch].w = fmaf(rr, x[r][ch].w, g[ch].w);
                }
            }
        }
        {   // advance the ring cursor past this step's slots
            const uint32_t sl = base_slot + (uint32_t)nsl_c;
            const uint32_t k = sl / (uint32_t)ns;
            base_slot = sl - k * (uint32_t)ns;
            base_phase ^= (k & 1u);
        }
        if (tid == 0) dbg_stamp(a, s, 1);

        // ---- combine row groups inside the CTA, hand the CTA's gradient to the exchange ----
        if (RG > 1) {
            // every row group parks its partial; after the barrier the first F/4 threads each sum
            // one column chunk over the groups (fixed order) — the combine is spread over warps
            // instead of serialised on row group 0
            if (active) {
#pragma unroll
                for (int ch = 0; ch < CH; ++ch)
                    if (gv[ch])
                        *reinterpret_cast<float4*>(gsm + (size_t)rg * F + gq[ch] * 4) = g[ch];
            }
            named_bar_sync(1, NCT);
        }
        if (comb) {
            unsigned long long* buf = acc + (size_t)(s % 3) * acc_stride;
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                int gqc;
                float4 t;
                if (RG > 1) {
                    gqc = tid + ch * NCT;  // CH*NCT >= granules: every column chunk has a thread
                    if (gqc >= granules) break;
                    t = *reinterpret_cast<const float4*>(gsm + gqc * 4);
                    for (int k = 1; k < RG; ++k) {
                        const float4 o = *reinterpret_cast<const float4*>(gsm + (size_t)k * F + gqc * 4);
                        t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
                    }
                } else {
                    if (!gv[ch]) continue;
                    gqc = gq[ch];
                    t = g[ch];
                }
                check_fix_range(t, a.fix_limit, a.err);
                if (XB) {
                    longlong2* p = reinterpret_cast<longlong2*>(gout + gqc * 4);
                    p[0] = make_longlong2(to_fix(t.x), to_fix(t.y));
                    p[1] = make_longlong2(to_fix(t.z), to_fix(t.w));
                } else {
                    red_add_u64(buf + gqc * 4 + 0, (unsigned long long)to_fix(t.x));
                    red_add_u64(buf + gqc * 4 + 1, (unsigned long long)to_fix(t.y));
                    red_add_u64(buf + gqc * 4 + 2, (unsigned long long)to_fix(t.z));
                    red_add_u64(buf + gqc * 4 + 3, (unsigned long long)to_fix(t.w));
                }
            }
            if (!XB) {
                __threadfence();  // this thread's adds are performed before the CTA signals
                mbar_arrive(smem_u32(gfull));
            }
        }
        if (XB) {
            // one arrival per warp that holds staging threads; the warp-wide sync sits outside
            // the (possibly divergent) staging branch
            const int comb_warps = RG > 1 ? min(NCW, (granules + 31) / 32) : WPG;
            __syncwarp();
            if (lane == 0 && warp < comb_warps) mbar_arrive(smem_u32(gfull));
        }
        if (tid == 0) dbg_stamp(a, s, 2);
        // off the critical path (the exchange is in flight): this step's loss
        if (warp == 0 && loss_fix) {
            float l = 0.f;
            for (int i = lane; i < nrows_c; i += 32) {
                const float2 v = zyb[i];
                l += row_loss_f(v.x, v.y);
            }
            l = warp_sum(l);
            if (lane == 0) red_add_u64(loss_fix + s, (unsigned long long)to_fix(l));
        }
        mb = mb_next;
    }

    // the last step's total
    if (a.n_steps > 0) {
        mbar_wait(smem_u32(wfull), (uint32_t)(a.n_steps - 1) & 1u, a.err);
        apply_total(a.n_steps - 1);
    }
    // master weights back to global: all CTAs hold identical copies, CTA 0 writes
    if (MS) {
        if (c == 0 && tid < granules) {
            double* m = model + tid * 4;
            m[0] = msm[tid * 4]; m[1] = msm[tid * 4 + 1]; m[2] = msm[tid * 4 + 2]; m[3] = msm[tid * 4 + 3];
        }
    } else if (c == 0 && rg == 0) {
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (!gv[ch]) continue;
            double* m = model + gq[ch] * 4;
            if (WS) {
                const double* mm = my_master + gq[ch] * 4;
                m[0] = mm[0]; m[1] = mm[1]; m[2] = mm[2]; m[3] = mm[3];
            } else if (WM == 0) {
                m[0] = wm[ch][0]; m[1] = wm[ch][1]; m[2] = wm[ch][2]; m[3] = wm[ch][3];
            }
        }
    }
}

// ------------------------------------------------------------------------------------
// K1s — simple three-launch step (any F, any B; also the local half of the NCCL mode)
// ------------------------------------------------------------------------------------
// one warp per minibatch row: z = x.w, residual and per-row loss
__global__ void __launch_bounds__(256) lr_fwd_kernel(const float* __restrict__ X,
                                                     const float* __restrict__ y,
                                                     long long n_rows, int F, int B, long long mb,
                                                     const double* __restrict__ model,
                                                     float* __restrict__ resid,
                                                     float* __restrict__ rowloss) {
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (j >= B) return;
    const long long row = (mb + j) % n_rows;
    const float* x = X + (size_t)row * F;
    float acc = 0.f;
    if ((F & 3) == 0) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        for (int q = lane; q < F / 4; q += 32) {
            const float4 v = __ldg(x4 + q);
            const double* m = model + 4 * q;
            acc = fmaf(v.x, (float)m[0], acc);
            acc = fmaf(v.y, (float)m[1], acc);
            acc = fmaf(v.z, (float)m[2], acc);
            acc = fmaf(v.w, (float)m[3], acc);
        }
    } else {
        for (int f = lane; f < F; f += 32) acc = fmaf(__ldg(x + f), (float)model[f], acc);
    }
    const float z = warp_sum(acc);
    if (lane == 0) {
        const float yy = y[row];
        resid[j] = sigmoid_f(z) - yy;
        rowloss[j] = row_loss_f(z, yy);
    }
}

// thread per column, block (x = column block, y = row split): partial g over a row range.
// The minibatch was just read by lr_fwd_kernel, so at the configured shapes this second
// pass is served by L2 (16 MiB tile vs 126 MB), not DRAM.
__global__ void __launch_bounds__(128) lr_grad_kernel(const float* __restrict__ X, long long n_rows,
                                                      int F, int B, long long mb,
                                                      const float* __restrict__ resid,
                                                      float* __restrict__ gpart, int rows_per_split) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    const int j0 = blockIdx.y * rows_per_split;
    const int j1 = min(B, j0 + rows_per_split);
    if (col >= F) return;
    float acc = 0.f;
    long long row = (mb + j0) % n_rows;
    for (int j = j0; j < j1; ++j) {
        acc = fmaf(resid[j], __ldg(X + (size_t)row * F + col), acc);
        if (++row == n_rows) row = 0;
    }
    gpart[(size_t)blockIdx.y * F + col] = acc;
}

// sums the row-split partials in a fixed order (fp64).  mode 0: update the model in
// place.  mode 1: write gsum[F] and gsum[F] = loss sum for an all-reduce.
__global__ void __launch_bounds__(256) lr_finish_kernel(const float* __restrict__ gpart, int nsplit,
                                                        int F, int B,
                                                        const float* __restrict__ rowloss,
                                                        double* __restrict__ model,
                                                        double eta_over_b, double* loss_out,
                                                        double* gsum, int mode) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col < F) {
        double acc = 0.0;
        for (int k = 0; k < nsplit; ++k) acc += (double)gpart[(size_t)k * F + col];
        if (mode == 0) model[col] -= eta_over_b * acc;
        else gsum[col] = acc;
    }
    if (blockIdx.x == 0) {
        __shared__ double sm[256];
        double acc = 0.0;
        for (int j = threadIdx.x; j < B; j += blockDim.x) acc += (double)rowloss[j];
        sm[threadIdx.x] = acc;
        __syncthreads();
        for (int s = blockDim.x / 2; s > 0; s >>= 1) {
            if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            if (mode == 0) { if (loss_out) *loss_out = sm[0] / (double)B; }
            else gsum[F] = sm[0];
        }
    }
}

__global__ void loss_finish_kernel(const unsigned long long* __restrict__ fix, double* __restrict__ out,
                                   int n, double inv_b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ((double)(long long)fix[i] * kFixInv) * inv_b;
}

// after the all-reduce of gsum[F+1] over nranks: model -= eta/(B*nranks) * gsum
__global__ void lr_apply_avg_kernel(double* __restrict__ model, const double* __restrict__ gsum,
                                    int F, double scale, double* loss_out, double loss_scale) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col < F) model[col] -= scale * gsum[col];
    if (col == 0 && loss_out) *loss_out = gsum[F] * loss_scale;
}

__global__ void scale_kernel(double* __restrict__ v, int n, double s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] *= s;
}

// ------------------------------------------------------------------------------------
// K2/K3/K4/K6 — exchange arithmetic, fp64, bit-exact with the reference's loops:
// __dmul_rn/__dadd_rn keep the compiler from contracting `m + lambda*d` into an FMA
// (the reference is compiled without -O/-march: rounded multiply, then rounded add).
// 128-bit accesses on the aligned body, scalar tail.
// ------------------------------------------------------------------------------------
__global__ void merge_kernel(double* __restrict__ model, double* __restrict__ old,
                             const double* __restrict__ din, double* __restrict__ reply, int n,
                             double lambda) {
    const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = 2 * i2;
    if (i + 1 < n) {
        const double2 m = reinterpret_cast<const double2*>(model)[i2];
        const double2 o = reinterpret_cast<const double2*>(old)[i2];
        const double2 d = reinterpret_cast<const double2*>(din)[i2];
        double2 mm, rr;
        mm.x = __dadd_rn(m.x, __dmul_rn(lambda, d.x));   // src/master.cc:106
        mm.y = __dadd_rn(m.y, __dmul_rn(lambda, d.y));
        rr.x = __dsub_rn(mm.x, o.x);                      // src/master.cc:107
        rr.y = __dsub_rn(mm.y, o.y);
        reinterpret_cast<double2*>(model)[i2] = mm;
        reinterpret_cast<double2*>(old)[i2] = mm;         // src/master.cc:110
        reinterpret_cast<double2*>(reply)[i2] = rr;
    } else if (i < n) {
        const double mm = __dadd_rn(model[i], __dmul_rn(lambda, din[i]));
        reply[i] = __dsub_rn(mm, old[i]);
        model[i] = mm;
        old[i] = mm;
    }
}

__global__ void delta_kernel(const double* __restrict__ model, const double* __restrict__ old,
                             double* __restrict__ out, int n) {
    const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = 2 * i2;
    if (i + 1 < n) {
        const double2 m = reinterpret_cast<const double2*>(model)[i2];
        const double2 o = reinterpret_cast<const double2*>(old)[i2];
        reinterpret_cast<double2*>(out)[i2] = make_double2(__dsub_rn(m.x, o.x), __dsub_rn(m.y, o.y));
    } else if (i < n) {
        out[i] = __dsub_rn(model[i], old[i]);            // src/worker.cc:207
    }
}

__global__ void apply_kernel(double* __restrict__ model, double* __restrict__ old,
                             const double* __restrict__ back, int n, double lambda) {
    const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = 2 * i2;
    if (i + 1 < n) {
        const double2 m = reinterpret_cast<const double2*>(model)[i2];
        const double2 d = reinterpret_cast<const double2*>(back)[i2];
        const double2 mm = make_double2(__dadd_rn(m.x, __dmul_rn(lambda, d.x)),
                                        __dadd_rn(m.y, __dmul_rn(lambda, d.y)));
        reinterpret_cast<double2*>(model)[i2] = mm;       // src/worker.cc:162
        reinterpret_cast<double2*>(old)[i2] = mm;         // src/worker.cc:215
    } else if (i < n) {
        const double mm = __dadd_rn(model[i], __dmul_rn(lambda, back[i]));
        model[i] = mm;
        old[i] = mm;
    }
}

// One whole gossip round between two replicas in ONE kernel (src/worker.cc:194-219 on the
// initiator A, src/worker.cc:81-100 on the callee B): A's state is read and written through
// peer-mapped pointers over NVLink when A lives on another GPU.  Same arithmetic, same order:
//   d = A.model - A.old ; B.model += l*d ; reply = B.model - B.old ; B.old = B.model ;
//   A.model += l*reply ; A.old = A.model
__global__ void gossip_p2p_kernel(double* __restrict__ a_model, double* __restrict__ a_old,
                                  double* __restrict__ b_model, double* __restrict__ b_old, int n,
                                  double lambda_a, double lambda_b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double am = a_model[i], ao = a_old[i];
    const double d = __dsub_rn(am, ao);                          // src/worker.cc:207
    const double bm = __dadd_rn(b_model[i], __dmul_rn(lambda_b, d));   // :92
    const double reply = __dsub_rn(bm, b_old[i]);                // :93
    b_model[i] = bm;
    b_old[i] = bm;                                               // :96
    const double am2 = __dadd_rn(am, __dmul_rn(lambda_a, reply));      // :162
    a_model[i] = am2;
    a_old[i] = am2;                                              // :215
}

__global__ void tick_kernel(double* __restrict__ model, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) model[i] = __dadd_rn(model[i], 1.0);      // src/worker.cc:226
}

// ------------------------------------------------------------------------------------
// H1 — checksum of the ingested byte stream: sum b[i] and sum (i+1)*b[i]  (mod 2^64).
// Integer adds commute, so the atomics give a deterministic result.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) checksum_kernel(const unsigned char* __restrict__ p,
                                                       size_t n, unsigned long long* out) {
    unsigned long long s = 0, wsum = 0;
    const size_t nvec = n / 16;
    const uint4* v = reinterpret_cast<const uint4*>(p);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (size_t)gridDim.x * blockDim.x) {
        const uint4 q = __ldg(v + i);
        const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
        const unsigned long long base = (unsigned long long)i * 16ull + 1ull;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const unsigned long long byte = (wds[k] >> (8 * b)) & 0xFFu;
                s += byte;
                wsum += (base + (unsigned long long)(4 * k + b)) * byte;
            }
        }
    }
    if (blockIdx.x == 0) {  // tail bytes
        for (size_t i = nvec * 16 + threadIdx.x; i < n; i += blockDim.x) {
            s += p[i];
            wsum += (unsigned long long)(i + 1) * p[i];
        }
    }
    for (int off = 16; off >= 1; off >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, off);
        wsum += __shfl_xor_sync(0xffffffffu, wsum, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out, s);
        atomicAdd(out + 1, wsum);
    }
}

// ------------------------------------------------------------------------------------
// S1 — synthetic shard, counter-based: x = splitmix64(seed ^ (row*F+col))
// ------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    unsigned long long z = x + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__device__ __forceinline__ float synth_x(unsigned long long seed, unsigned long long idx) {
    const unsigned long long u = splitmix64(seed ^ idx);
    const float f = (float)(u >> 40) * (1.0f / 16777216.0f);
    return f * 2.0f - 1.0f;
}
__global__ void __launch_bounds__(256) synth_kernel(float* __restrict__ X, float* __restrict__ y,
                                                    long long n_rows, int F,
                                                    unsigned long long row0,
                                                    unsigned long long seed) {
    const size_t total = (size_t)n_rows * F;
    for (size_t e = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; e < total;
         e += (size_t)gridDim.x * blockDim.x * 4) {
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const size_t ee = e + k;
            if (ee < total) {
                const unsigned long long r = ee / F, cidx = ee - r * F;
                v[k] = synth_x(seed, (row0 + r) * (unsigned long long)F + cidx);
                if (cidx == 0) y[r] = v[k] > 0.0f ? 1.0f : 0.0f;
            }
        }
        if (e + 3 < total && (F & 3) == 0) {
            *reinterpret_cast<float4*>(X + e) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            for (int k = 0; k < 4; ++k)
                if (e + k < total) X[e + k] = v[k];
        }
    }
}

}  // namespace slk
