// kernels_cluster.cuh — K1c: the minibatch step for narrow rows (F <= 1024) as thread-block clusters.
//
// lr_persist_kernel's step is one long latency chain (compute -> L2 bulk reduction -> release fence ->
// arrival counter -> poll -> TMA load of the total: ~5 us of exchange for 2.6 us of HBM streaming at
// config 3, profiles/r1_timeline.md).  This kernel keeps the row ring and the per-thread arithmetic and
// replaces the exchange by three fence-free hops:
//
//   1. reduce-scatter inside a cluster of CS CTAs through distributed shared memory: every compute
//      thread sends its fp32 gradient chunk straight from registers to the CTA that owns those columns
//      (st.async, completion counted in bytes on the owner's mbarrier) — no staging, no barrier;
//   2. ONE hop through L2 between the clusters: the owner converts every partial to fixed point, adds
//      the integers and adds the cluster's sum into a global accumulator with a single red.add.u64 whose
//      low bits count arrivals ((sum << 7) + 1).  Readers poll the accumulator itself until the count
//      says every cluster has added: data and "flag" are one word, so there is no release fence and no
//      separate counter.  Accumulators are never cleared: a reader subtracts the value it saw complete
//      two steps earlier (two buffers, by step parity), arithmetic mod 2^64;
//   3. the owner applies the total to ITS fp64 master weights (the owner of a column is the only place
//      they live; all clusters hold identical copies) and sends the new fp32 weights to the CS CTAs of
//      its cluster (st.async again); compute threads only ever see fp32 weights in shared memory.
//
// Integer sums commute: the result does not depend on arrival order, every cluster (and every GPU)
// computes the same bits.  With peers attached (sl_peer_attach) a fourth hop sums the GPUs' totals the
// same way over NVLink: add-with-count into every rank's accumulator (one multimem.red through the
// NVSwitch when a multicast mapping was attached, else one red per peer), poll the own copy.
//
// The per-step loss travels the same road one step late, off the critical path, and is written by the
// kernel itself as a double: no memset before and no conversion kernel after a launch.
//
// Takes the slot of simulate_training (src/worker.cc:221-231) like lr_persist_kernel.
#pragma once
#include "kernels.cuh"

namespace slk {

constexpr int kClCntBits = 7;                                  // arrivals per word: up to 127 adders
constexpr unsigned long long kClCntMask = (1ull << kClCntBits) - 1ull;
constexpr double kClFixInv = 1.0 / 4294967296.0;               // fixed point: 2^-32 units

struct ClArgs {
    const float* X;        // [n_rows, F] row-major, resident shard
    const float* y;        // [n_rows]
    long long n_rows;
    int F;
    int B;                 // positions per step (nvw * vB in group mode)
    long long step0;       // global step index of the first step of this launch
    int n_steps;
    double* model;         // fp64 master weights [F] (model_state, src/master.cc:58)
    unsigned long long* gacc;    // [2][F] gradient accumulators: (sum << 7) + arrivals, never cleared
    unsigned long long* gbase;   // [2][F] their values when the previous launch ended
    unsigned long long* lacc;    // [3]    loss accumulators, same encoding, by step mod 3
    unsigned long long* lbase;   // [3]
    double* loss_out;      // device [n_steps]: mean loss of every step, written by the kernel; or nullptr
    long long xstep0;      // accumulator parity of this launch's first step (continues across launches); with peers
                           // attached also buffer and tag of the cross-GPU words: the host keeps it equal on all ranks
    // same-box data parallel: one kernel per GPU, totals summed through peer-mapped memory
    int nranks, rank;      // nranks == 0: single GPU, no cross-GPU stage (1: the stage runs against itself)
    unsigned long long* peer_gx[8];  // rank p's all-rank words [2 buffers][F][8 ranks] as mapped here (own = local)
    unsigned long long* mc_gx;       // multicast alias of the same words on all ranks, or nullptr
    double eta_over_b;     // eta / (B * ranks)
    double inv_b;          // 1 / B
    float part_limit;      // |partial| above this could overflow the accumulator word
    int poll_delay, poll_gap;  // SM cycles before the first look at an accumulator / between looks
    int xpoll_delay, xpoll_gap;  // the same for the all-rank words
    int wait_all_owners;   // diagnostics: 1 = a compute thread waits for every owner's weights, not only its columns' owner
    int* err;              // 0 = ok; set by the watchdog
    // geometry, computed on the host
    int ncl;               // clusters in the grid
    int fc;                // columns owned by one CTA of a cluster = F / CS
    int nsg;               // ring slots per row group
    int row_bytes;         // 4*F
    int zy_rows;
    int ycap;              // label slots per row group and step (>= passes per group * R)
    int off_bar, off_recv, off_wsm, off_lrecv, off_red, off_zy, off_yb, off_prev;  // smem offsets
    unsigned long long* dbg;
    int dbg_step;
    // several worker replicas on ONE GPU trained as one job (sl_train_steps_group), as in LrArgs
    int nvw;
    int vB;
    int npart;             // always 0 here (locate_row reads it)
    const float* vX[8];
    const float* vy[8];
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 16 / 4 bytes from registers into another CTA's shared memory; the bytes are counted on that CTA's mbarrier
__device__ __forceinline__ void st_async_v4(uint32_t raddr, float4 v, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(raddr),
                 "r"(__float_as_uint(v.x)), "r"(__float_as_uint(v.y)), "r"(__float_as_uint(v.z)),
                 "r"(__float_as_uint(v.w)), "r"(rbar)
                 : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t raddr, uint32_t v, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr), "r"(v),
                 "r"(rbar)
                 : "memory");
}
// Wait on a barrier whose bytes were written by other CTAs of the cluster (st.async: the bytes are counted
// on the barrier as they land in THIS CTA's shared memory).  The wait is the plain CTA-scope one: a cluster-scope
// acquire compiles to an L1 invalidation (CCTL.IVALL) in EVERY turn of the spin loop — a quarter of all stall
// samples of the kernel in the first profile — and there is nothing in L1 to invalidate: the data is in shared memory.
#ifndef SL_CL_POLL_ACQ
#define SL_CL_POLL_ACQ 0
#endif
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
#ifndef SL_CL_WAIT_MODE
#define SL_CL_WAIT_MODE 0   // 0: try_wait.acquire.cluster, 1: test_wait (pure polling, CTA scope)
#endif
__device__ __forceinline__ uint32_t mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t ok;
#if SL_CL_WAIT_MODE == 0
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
#else
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
#endif
    return ok;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity, int* err) {
    uint32_t spins = 0;
    while (!mbar_try_wait_cluster(bar, parity)) {
        if (spin_abort(spins, err, kErrTimeoutMbar)) return;
    }
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// one store, delivered by the NVSwitch to the same offset in every GPU of the multicast group
__device__ __forceinline__ void multimem_st_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ ulonglong2 ld_relaxed_sys_v2u64(const unsigned long long* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void cp_async_4(uint32_t dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// diagnostics: 16 (clock64, globaltimer) pairs per CTA and step, [4 steps][grid][32].  Points 8..15 sit inside
// the passes and perturb them (0.2 us per step at config 3): compiled in only with -DSL_CL_FINE_STAMPS=1.
#ifndef SL_CL_FINE_STAMPS
#define SL_CL_FINE_STAMPS 0
#endif
// Compiled in only in the DBG instantiation of the kernel (option debug_timeline): the stamps' pointer and step
// compares otherwise cost the compute warps registers they do not have (3.82 -> 3.68 us/step at config 3).
template <bool DBG>
__device__ __forceinline__ void cl_stamp(const ClArgs& a, int s, int point) {
    if (!DBG) return;
    if (a.dbg != nullptr && s >= a.dbg_step && s < a.dbg_step + 4) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        unsigned long long* d = a.dbg + ((size_t)(s - a.dbg_step) * gridDim.x + blockIdx.x) * 32;
        d[2 * point] = (unsigned long long)clock64();
        d[2 * point + 1] = t;
    }
}
// exact: scaling by 2^32 does not round in fp32, the conversion rounds once
__device__ __forceinline__ long long to_fix32(float v) { return __float2ll_rn(v * 4294967296.0f); }

// Poll an add-with-count word until `want` adders have arrived since `prev`; returns their sum.
template <bool SYS>
__device__ __forceinline__ long long poll_total(const unsigned long long* p, unsigned long long& prev, unsigned want,
                                                int* err, int code, uint32_t limit) {
    uint32_t spins = 0;
    unsigned long long v, d;
    for (;;) {
        v = SYS ? ld_relaxed_sys_u64(p) : ld_relaxed_u64(p);
        d = v - prev;
        if ((d & kClCntMask) == (unsigned long long)want) break;
        if (spin_abort(spins, err, code, limit)) break;
    }
    prev = v;
    return (long long)(d - (unsigned long long)want) >> kClCntBits;  // exact: the low bits are zero now
}

// Warps of a CTA: compute, exchange, one producer, padded to whole warpgroups when registers are re-split.
template <int WPG, int RG, int XW, int CREG, int PW>
constexpr int cl_warps() {
    return CREG > 0 ? (WPG * RG + XW + PW + 3) / 4 * 4 : WPG * RG + XW + PW;
}

// The same for the XC columns of one exchange thread at once: every round issues all outstanding loads
// before it looks at any of them, so the columns' L2 round trips overlap.  Polls compete with the adders for
// the accumulators' L2 lines (four staggered loads in flight per column made the hop 1.5x SLOWER), so the
// first look waits `delay` cycles and later ones `gap` cycles.
template <bool SYS, int XC>
__device__ __forceinline__ void poll_totals(unsigned long long* const (&p)[XC], const bool (&act)[XC],
                                            unsigned long long (&prev)[XC], long long (&T)[XC], unsigned want,
                                            int delay, int gap, int* err, int code, uint32_t limit) {
    bool done[XC];
#pragma unroll
    for (int i = 0; i < XC; ++i) done[i] = !act[i];
    if (delay > 0) {
        const long long t0 = clock64();
        while (clock64() - t0 < delay) {}
    }
    uint32_t spins = 0;
    for (;;) {
        unsigned long long v[XC];
#pragma unroll
        for (int i = 0; i < XC; ++i)
            if (!done[i]) v[i] = SYS ? ld_relaxed_sys_u64(p[i]) : (SL_CL_POLL_ACQ ? ld_acquire_u64(p[i]) : ld_relaxed_u64(p[i]));
        bool all = true;
#pragma unroll
        for (int i = 0; i < XC; ++i) {
            if (done[i]) continue;
            const unsigned long long d = v[i] - prev[i];
            if ((d & kClCntMask) == (unsigned long long)want) {
                done[i] = true;
                prev[i] = v[i];
                T[i] = (long long)(d - (unsigned long long)want) >> kClCntBits;  // exact: the low bits are zero now
            } else {
                all = false;
            }
        }
        if (all) break;
        if (spin_abort(spins, err, code, limit)) break;
        if (gap > 0) {
            const long long t0 = clock64();
            while (clock64() - t0 < gap) {}
        }
    }
}

// WPG warps cover a row (one float4 column chunk per thread: F <= 128*WPG); RG row groups of WPG warps
// take R rows each per pass, from their own ring of RS-row slots; CS = CTAs per cluster; XW exchange
// warps share the (at most 128) columns this CTA owns.  CREG > 0: the compute warpgroups take CREG
// registers per thread from the exchange / producer warps (setmaxnreg), which keep XREG.  PW = 1: a warp of
// its own issues the TMA row copies; PW = 0: lanes of the first exchange warp do, once per step after the
// weights have gone out (for shapes where a step is one pass: a slot is then always free when it is asked for).
template <int WPG, int RG, int R, int RS, int CS, int XW, int CREG, int XREG, int PW, bool DBG>
__global__ void __launch_bounds__(cl_warps<WPG, RG, XW, CREG, PW>() * 32, 1) lr_cluster_kernel(const ClArgs a) {
    static_assert(R % RS == 0, "a pass is a whole number of ring slots");
    static_assert(WPG > 1, "the labels become visible at the row group's barrier");
    static_assert(CREG == 0 || (WPG * RG) % 4 == 0, "register re-split needs whole compute warpgroups");
    constexpr int NCW = WPG * RG;     // compute warps
    constexpr int XT = XW * 32;       // exchange threads
    constexpr int XC = (128 * WPG / CS + XT - 1) / XT;  // owned columns per exchange thread (F <= 128 * WPG)
    constexpr int SPB = R / RS;       // ring slots per pass
    constexpr int NSRC = CS * RG;     // gradient partials an owner receives per step
    constexpr int CT = WPG * 32;      // threads of a row group
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t crank = cluster_ctarank();
    const int q = (int)cluster_id_x();
    const int G = (int)gridDim.x, c = (int)blockIdx.x;
    const int F = a.F, fc = a.fc, nsg = a.nsg, row_bytes = a.row_bytes, n_steps = a.n_steps;
    const bool want_loss = a.loss_out != nullptr;

    unsigned char* ring = smem;                                          // [RG][nsg][RS rows]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_bar);      // [RG*nsg]
    uint64_t* empty = full + RG * nsg;                                   // [RG*nsg]
    uint64_t* rfull = empty + RG * nsg;                                  // gradient partials of a step landed
    uint64_t* wfull = rfull + 1;                                         // [CS] next step's weights of owner o landed
    uint64_t* lfull = wfull + CS;                                        // [2] loss partials of a step landed
    uint64_t* xdone = lfull + 2;                                         // the exchange warps have read this step's partials
    float* recv = reinterpret_cast<float*>(smem + a.off_recv);          // [NSRC][fc]
    float* wsm = reinterpret_cast<float*>(smem + a.off_wsm);            // [2][F]
    float* lrecv = reinterpret_cast<float*>(smem + a.off_lrecv);        // [2][NSRC]
    float* red = reinterpret_cast<float*>(smem + a.off_red);            // [2][NCW][R]
    float2* zy = reinterpret_cast<float2*>(smem + a.off_zy);            // [2][zy_rows]
    float* ybuf = reinterpret_cast<float*>(smem + a.off_yb);            // [2][RG][ycap] labels of a step's rows
    unsigned long long* xprev = reinterpret_cast<unsigned long long*>(smem + a.off_prev);  // [2][fc] accumulator values seen complete, [3] for the loss

    const uint32_t slot_bytes = (uint32_t)RS * (uint32_t)row_bytes;
    const long long lo = ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = (int)(hi - lo);             // this CTA's rows of every minibatch
    const int nst_c = (nrows_c + R - 1) / R;        // ... in passes of R rows
    const int nsb = (nst_c + RG - 1) / RG;          // passes per row group
    const long long mb0 = (long long)(((__int128)a.step0 * a.vB) % a.n_rows);
    auto wrap = [&](long long v) {
        if (v >= a.n_rows) {
            v -= a.n_rows;
            if (v >= a.n_rows) v %= a.n_rows;
        }
        return v;
    };
    const uint32_t recv_bytes = (uint32_t)NSRC * (uint32_t)fc * 4u;

    if (tid == 0) {
        if (!SL_CL_FINE_STAMPS) cl_stamp<DBG>(a, a.dbg_step, 14);  // kernel entry
        for (int s = 0; s < RG * nsg; ++s) {
            mbar_init(smem_u32(&full[s]), 1);
            mbar_init(smem_u32(&empty[s]), WPG);
        }
        mbar_init(smem_u32(rfull), 1);
        for (int o = 0; o < CS; ++o) mbar_init(smem_u32(&wfull[o]), 1);
        mbar_init(smem_u32(&lfull[0]), 1);
        mbar_init(smem_u32(&lfull[1]), 1);
        mbar_init(smem_u32(xdone), XW);
        fence_mbar_init();
        // what the first phases will receive (later phases are armed by their waiters)
        if (n_steps > 0) mbar_arrive_expect_tx(smem_u32(rfull), recv_bytes);
        if (n_steps > 1)
            for (int o = 0; o < CS; ++o) mbar_arrive_expect_tx(smem_u32(&wfull[o]), (uint32_t)fc * 4u);
        if (want_loss && crank == 0) {
            if (n_steps > 0) mbar_arrive_expect_tx(smem_u32(&lfull[0]), (uint32_t)NSRC * 4u);
            if (n_steps > 1) mbar_arrive_expect_tx(smem_u32(&lfull[1]), (uint32_t)NSRC * 4u);
        }
    }
    __syncthreads();
    cluster_sync_all();  // every CTA's barriers exist before anybody stores into a neighbour
    if (!SL_CL_FINE_STAMPS && tid == 0) cl_stamp<DBG>(a, a.dbg_step, 9);

    // ---- TMA producer: lane l owns ring slot l % nsg of row group l / nsg and fetches that slot's uses over the
    // whole launch (issuing a slot costs a thread several hundred cycles; one lane would bound the step)
    const bool prod_warp = PW ? warp == NCW + XW : warp == NCW;
    int pr_g = 0, pr_sl = 0, pr_nslg = 0, pr_s = n_steps, pr_k = 0, pr_prev = -1;
    uint32_t pr_phase = 0, pr_eb = 0, pr_fb = 0, pr_dst0 = 0;
    long long pr_mb = mb0;
    uint64_t pr_pol = 0;
    if (prod_warp) {
        pr_g = lane / nsg;
        pr_sl = lane - pr_g * nsg;
        // slots of one step that belong to the group: passes g, g+RG, ... of SPB slots, the last maybe short
        if (pr_g < RG)
            for (int j = pr_g; j < nst_c; j += RG) pr_nslg += (min(R, nrows_c - j * R) + RS - 1) / RS;
        if (pr_g < RG && pr_nslg > 0) {
            pr_pol = policy_evict_first();
            pr_eb = smem_u32(&empty[pr_g * nsg + pr_sl]);
            pr_fb = smem_u32(&full[pr_g * nsg + pr_sl]);
            pr_dst0 = smem_u32(ring + ((size_t)pr_g * nsg + pr_sl) * slot_bytes);
            pr_s = pr_sl / pr_nslg;           // step and slot-of-step of this lane's next use
            pr_k = pr_sl - pr_s * pr_nslg;
            for (int i = 0; i < pr_s && i < n_steps; ++i) pr_mb = wrap(pr_mb + a.vB);
        }
    }
    // issues this lane's uses, in order, as long as the slot's previous use lies at a step <= max_prev
    auto produce = [&](int max_prev) {
        while (pr_s < n_steps && pr_prev <= max_prev) {
            const int jj = pr_k / SPB, qq = pr_k - jj * SPB;
            const int row = (pr_g + RG * jj) * R + qq * RS;      // first row of the slot within the CTA's rows
            const int rc = min(RS, nrows_c - row);
            mbar_wait(pr_eb, pr_phase ^ 1u, a.err);
            mbar_arrive_expect_tx(pr_fb, (uint32_t)rc * (uint32_t)row_bytes);
            long long p = lo + row;
            uint32_t dst = pr_dst0;
            int left = rc;
            while (left > 0) {  // one copy unless the rows wrap around the shard end (or change worker)
                const RowRef ref = locate_row(a, pr_mb, p);
                const int n = ref.room < (long long)left ? (int)ref.room : left;
                bulk_g2s(dst, ref.x, (uint32_t)n * (uint32_t)row_bytes, pr_fb, pr_pol);
                dst += (uint32_t)n * (uint32_t)row_bytes;
                left -= n;
                p += n;
            }
            pr_phase ^= 1u;
            pr_prev = pr_s;
            pr_k += nsg;
            while (pr_k >= pr_nslg) {
                pr_k -= pr_nslg;
                ++pr_s;
                pr_mb = wrap(pr_mb + a.vB);
            }
        }
    };

    // Register re-split (warpgroup-uniform: the compute warps fill whole warpgroups): every role states its
    // budget at the top of its own branch.
    if (warp >= NCW + XW + PW) {
        // padding warp of the last warpgroup: nothing to do
        if (CREG > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(XREG > 0 ? XREG : 24));
    } else if (PW && warp == NCW + XW) {
        // ======================================================================= producer
        if (CREG > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(XREG > 0 ? XREG : 24));
        produce(0x7fffffff);
    } else if (warp >= NCW) {
        // ======================================================================= exchange
        if (CREG > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(XREG > 0 ? XREG : 24));
        if (!PW && warp == NCW) {  // first use of every ring slot, before anything that waits for global memory
            produce(-1);
            __syncwarp();
        }
        const int t = (warp - NCW) * 32 + lane;
        bool xact[XC];
        int col[XC];
        double master[XC];
        // accumulator values seen complete live in shared memory, one private slot per thread, column and
        // buffer ([local b=0,1][cross-GPU b=0,1]): registers are what the compute warps need
        bool xsend[XC];   // this thread forwards the column's local total to the peers
#pragma unroll
        for (int i = 0; i < XC; ++i) {
            const int tl = t + i * XT;
            xact[i] = tl < fc;
            col[i] = (int)crank * fc + tl;
            master[i] = xact[i] ? a.model[col[i]] : 0.0;
            if (xact[i]) {
                xprev[tl] = a.gbase[col[i]];
                xprev[fc + tl] = a.gbase[F + col[i]];
            }
            xsend[i] = a.nranks > 0 && xact[i] && (tl % a.ncl) == q;
        }
        const uint32_t wfull_a = smem_u32(wfull), wsm_a = smem_u32(wsm);

        for (int s = 0; s < n_steps; ++s) {
            const int b = (int)((a.xstep0 + s) & 1);
            if (!PW && warp == NCW && s == 0) {
                // cold start: the slots of step 0 are free again as soon as its rows have gone into registers —
                // refill them now, not after this step's exchange (later steps refill after the weights went out)
                produce(0);
                __syncwarp();
            }
            mbar_wait_cluster(smem_u32(rfull), (uint32_t)s & 1u, a.err);
            if (t == 0) {
                cl_stamp<DBG>(a, s, 6);
                if (s + 1 < n_steps) mbar_arrive_expect_tx(smem_u32(rfull), recv_bytes);
            }
            // the cluster's sum of every owned column: all partials are read and converted before the first
            // red goes out (the asm statements keep their order; plain loads may not move across them)
            long long T[XC], P[XC];
            bool bad = false;
#pragma unroll
            for (int i = 0; i < XC; ++i) {
                T[i] = 0;
                P[i] = 0;
                if (!xact[i]) continue;
                // the cluster's partials of a column are summed in a fixed pairwise order (every cluster, every
                // GPU does the same), then converted once: integers from here on
                float v[NSRC];
#pragma unroll
                for (int k = 0; k < NSRC; ++k) v[k] = recv[k * fc + t + i * XT];
#pragma unroll
                for (int w = 1; w < NSRC; w *= 2)
#pragma unroll
                    for (int k = 0; k + w < NSRC; k += 2 * w) v[k] += v[k + w];
                bad = bad || !(fabsf(v[0]) < a.part_limit);
                P[i] = to_fix32(v[0]);
            }
            // this cluster's sums go into the step's accumulators; then wait until every cluster's have.  (With
            // peers attached only the forwarding thread NEEDS this GPU's total, but letting the others skip ahead
            // to the all-rank words made the step 1.6x slower: they then poll those words for a microsecond longer,
            // and readers on a line slow down the stores that are on their way to it.)
            unsigned long long* lp[XC];
            unsigned long long lpv[XC];
            bool lact[XC];
#pragma unroll
            for (int i = 0; i < XC; ++i) {
                lp[i] = a.gacc + (size_t)b * F + col[i];
                lact[i] = xact[i];
                lpv[i] = lact[i] ? xprev[b * fc + t + i * XT] : 0ull;
                if (xact[i]) red_add_u64(lp[i], ((unsigned long long)P[i] << kClCntBits) + 1ull);
            }
            // the compute warps hold back their next burst of shared-memory loads (pulling the next step's rows
            // into registers) until the partials have been read: the two would queue in the same pipe
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(xdone));
            if (bad) set_err(a.err, kErrRange);
            if (t == 0) cl_stamp<DBG>(a, s, 7);
            poll_totals<false, XC>(lp, lact, lpv, T, (unsigned)a.ncl, a.poll_delay, a.poll_gap, a.err, kErrTimeoutPart, kSpinLimit);
#pragma unroll
            for (int i = 0; i < XC; ++i)
                if (lact[i]) xprev[b * fc + t + i * XT] = lpv[i];
            if (t == 0) cl_stamp<DBG>(a, s, 3);
            if (a.nranks > 0) {
                // ---- across GPUs: one 8-byte word per (rank, column) and buffer, value in the upper 57 bits and a
                // step tag in the lower 7: the thread that forwards a column stores this GPU's total into slot
                // [rank] of EVERY rank (one multimem.st through the NVSwitch when a multicast mapping is attached,
                // else one store per peer over NVLink); every rank then reads the slots of all ranks in its own
                // memory and adds them up — integers again, the same bits everywhere.  A slot's previous content
                // is two steps older (or the zero it started with) and never carries the current tag.
                const unsigned long long tag = (unsigned long long)((a.xstep0 + s + 1) & 127);
#pragma unroll
                for (int i = 0; i < XC; ++i) {
                    if (!xsend[i]) continue;
                    const unsigned long long word = ((unsigned long long)T[i] << 7) | tag;
                    const size_t off = ((size_t)b * F + col[i]) * 8 + a.rank;  // [buffer][column][rank]: a column's eight words share 64 bytes
                    if (a.mc_gx) {
                        multimem_st_u64(a.mc_gx + off, word);
                    } else {
                        for (int p = 0; p < a.nranks; ++p) st_relaxed_sys_u64(a.peer_gx[p] + off, word);
                    }
                }
                const unsigned long long* mine = a.peer_gx[a.rank] + (size_t)b * F * 8;
                unsigned pend[XC];
                long long S[XC];
#pragma unroll
                for (int i = 0; i < XC; ++i) {
                    pend[i] = xact[i] ? ((1u << a.nranks) - 1u) : 0u;
                    S[i] = 0;
                }
                // the columns take turns (eight loads in flight at a time: the exchange warps are short of registers);
                // nothing can have crossed NVLink before `xdelay` cycles, and every look costs the writers time
                if (a.xpoll_delay > 0) {
                    const long long t0 = clock64();
                    while (clock64() - t0 < a.xpoll_delay) {}
                }
                uint32_t spins = 0;
                for (;;) {
                    bool any = false;
#pragma unroll
                    for (int i = 0; i < XC; ++i) {
                        if (pend[i] == 0u) continue;
                        unsigned long long v[8];
#pragma unroll
                        for (int p = 0; p < 8; p += 2)   // two ranks' words per 16-byte load
                            if ((pend[i] >> p) & 3u) {
                                const ulonglong2 w2 = ld_relaxed_sys_v2u64(mine + (size_t)col[i] * 8 + p);
                                v[p] = w2.x;
                                v[p + 1] = w2.y;
                            }
#pragma unroll
                        for (int p = 0; p < 8; ++p)
                            if ((pend[i] >> p) & 1u) {
                                if ((v[p] & 127ull) == tag) {
                                    S[i] += (long long)v[p] >> 7;
                                    pend[i] &= ~(1u << p);
                                } else {
                                    any = true;
                                }
                            }
                    }
                    if (!any) break;
                    if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                    if (a.xpoll_gap > 0) {
                        const long long t0 = clock64();
                        while (clock64() - t0 < a.xpoll_gap) {}
                    }
                }
#pragma unroll
                for (int i = 0; i < XC; ++i) T[i] = S[i];
                if (!SL_CL_FINE_STAMPS && t == 0) cl_stamp<DBG>(a, s, 8);
            }
            // new fp32 weights to every CTA of the cluster, 16 bytes per store: the four lanes of a quad hold
            // four consecutive columns; each gathers the quad's values and serves a quarter of the CTAs
#pragma unroll
            for (int i = 0; i < XC; ++i) {
                if (xact[i]) master[i] -= a.eta_over_b * ((double)T[i] * kClFixInv);
                const float wv = (float)master[i];
                const int qb = lane & ~3;
                float4 w4;
                w4.x = __shfl_sync(0xffffffffu, wv, qb);
                w4.y = __shfl_sync(0xffffffffu, wv, qb + 1);
                w4.z = __shfl_sync(0xffffffffu, wv, qb + 2);
                w4.w = __shfl_sync(0xffffffffu, wv, qb + 3);
                if (xact[i] && s + 1 < n_steps) {
                    const uint32_t dsta = wsm_a + (uint32_t)(((s + 1) & 1) * F + (col[i] & ~3)) * 4u;
#pragma unroll
                    for (int d0 = 0; d0 < CS; d0 += 4) {
                        const int d = d0 + (lane & 3);
                        if (d < CS) st_async_v4(mapa_u32(dsta, (uint32_t)d), w4, mapa_u32(wfull_a + 8u * crank, (uint32_t)d));
                    }
                }
            }
            if (t == 0) cl_stamp<DBG>(a, s, 4);
            if (!PW && warp == NCW) {
                // refill the slots whose previous use is consumed by now: the rows of step s+1 went into registers
                // right after this step's partials had been read
                produce(s + 1);
                __syncwarp();
            }
        }
        if (q == 0) {  // all clusters hold identical masters and saw identical accumulator values
#pragma unroll
            for (int i = 0; i < XC; ++i) {
                if (!xact[i]) continue;
                a.model[col[i]] = master[i];
                const int tl = t + i * XT;
                a.gbase[col[i]] = xprev[tl];
                a.gbase[F + col[i]] = xprev[fc + tl];
            }
        }
    } else {
        // ======================================================================= compute
        if (CREG > 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CREG > 0 ? CREG : 24));
        const int rg = warp / WPG, wig = warp - rg * WPG;
        const int gq = wig * 32 + lane;                 // float4 column chunk of this thread
        const bool gv = gq * 4 < F;
        const int ridx = row_of_lane<R>(lane);
        const bool canon = lane == lane_of_row<R>(ridx);
        unsigned char* const gring = ring + (size_t)rg * nsg * slot_bytes;
        uint64_t* const gfullb = full + rg * nsg;
        uint64_t* const gemptyb = empty + rg * nsg;
        const uint32_t wsm_a = smem_u32(wsm);

        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gv) {
            const double* m = a.model + gq * 4;
            w = make_float4((float)m[0], (float)m[1], (float)m[2], (float)m[3]);
        }
        if (!SL_CL_FINE_STAMPS && tid == 0) cl_stamp<DBG>(a, a.dbg_step, 10);
        // where this thread's gradient chunk goes: the owner CTA of its columns, one slot per (source CTA, group)
        const int owner = gv ? (gq * 4) / fc : 0;
        const uint32_t send_addr = mapa_u32(smem_u32(recv) + (uint32_t)((((int)crank * RG + rg) * fc + (gq * 4 - owner * fc)) * 4), (uint32_t)owner);
        const uint32_t send_bar = mapa_u32(smem_u32(rfull), (uint32_t)owner);
        const uint32_t lsend_bar0 = mapa_u32(smem_u32(&lfull[0]), 0u), lsend_bar1 = mapa_u32(smem_u32(&lfull[1]), 0u);
        const uint32_t lsend_addr = mapa_u32(smem_u32(lrecv) + (uint32_t)(((int)crank * RG + rg) * 4), 0u);

        const bool loss_lead = want_loss && crank == 0 && tid == 0;
        unsigned long long* lprev = xprev + 2 * fc;   // [3] loss accumulator values seen complete (cluster 0 only)
        int l3 = (int)(a.xstep0 % 3);
        if (loss_lead && q == 0) { lprev[0] = a.lbase[0]; lprev[1] = a.lbase[1]; lprev[2] = a.lbase[2]; }
        uint32_t slot = 0, par = 0;  // this group's ring cursor
        int red_buf = 0;
        long long mb = mb0;
        // Labels: a global load inside a pass would put an HBM round trip into every pass.  Instead the group
        // copies the labels of a whole step into shared memory (cp.async, no registers) one step ahead, right
        // after handing off the gradient; they become visible at the first pass's barrier.
        const int ct = wig * 32 + lane;
        auto stage_labels = [&](long long mbs, int buf) {
            float* dst = ybuf + (size_t)(buf * RG + rg) * a.ycap;
            for (int u = ct; u < nsb * R; u += CT) {
                const int sbi = u / R, i = u - sbi * R;
                const int j = sbi * RG + rg, ic = j * R + i;
                if (j < nst_c && ic < nrows_c) cp_async_4(smem_u32(dst + u), locate_row(a, mbs, lo + ic).y);
            }
            cp_async_commit();
        };
        if (n_steps > 0) stage_labels(mb, 0);

        for (int s = 0; s < n_steps; ++s) {
            const long long mb_next = wrap(mb + a.vB);
            float2 glo = make_float2(0.f, 0.f), ghi = make_float2(0.f, 0.f);  // gradient of this thread's four columns
            float2* zyb = zy + (size_t)(s & 1) * a.zy_rows;
            if (tid == 0) cl_stamp<DBG>(a, s, 0);
            auto take_weights = [&]() {
                // one barrier per owner CTA: a warp starts on its columns as soon as THEIR owner's weights are here
                // (the owners finish their L2 hop up to half a microsecond apart); the thread that holds an owner's
                // first columns re-arms that barrier
                if ((fc & 127) == 0) {
                    // no warp's 128 columns span two owners (F = 1024 in clusters of 4 or 8): every lane waits for its own
                    if (gv) {
                        if (a.wait_all_owners)   // comparison switch: behave as one barrier for all weights
                            for (int o = 0; o < CS; ++o) mbar_wait_cluster(smem_u32(&wfull[o]), (uint32_t)(s - 1) & 1u, a.err);
                        mbar_wait_cluster(smem_u32(&wfull[owner]), (uint32_t)(s - 1) & 1u, a.err);
                    }
                } else {
                    // warp-uniform: a warp whose columns span several owners waits for one after the other (lanes spinning
                    // on different barriers would take turns: F = 992 measured 5.5 us/step that way, 4.0 this way)
                    const int wo_lo = (wig * 128) / fc, wo_hi = wig * 128 < F ? (min(F, wig * 128 + 128) - 1) / fc : -1;
                    const int lo_o = a.wait_all_owners ? 0 : wo_lo, hi_o = a.wait_all_owners ? CS - 1 : wo_hi;
                    if (wo_lo <= wo_hi)
                        for (int o = lo_o; o <= hi_o; ++o) mbar_wait_cluster(smem_u32(&wfull[o]), (uint32_t)(s - 1) & 1u, a.err);
                }
                if (gv && rg == 0 && gq * 4 == owner * fc && s + 1 < n_steps)
                    mbar_arrive_expect_tx(smem_u32(&wfull[owner]), (uint32_t)fc * 4u);
                if (gv) w = lds_v4(wsm_a + (uint32_t)((s & 1) * F + gq * 4) * 4u);
            };
            if (nsb == 0 && s > 0) take_weights();  // a CTA without rows still takes part in every exchange

            for (int sb = 0; sb < nsb; ++sb) {
                const int j = sb * RG + rg;               // this group's pass of the step
                const bool have = j < nst_c;              // warp-uniform
                const int rc = have ? min(R, nrows_c - j * R) : 0;
                float4 x[R];
#pragma unroll
                for (int qq = 0; qq < SPB; ++qq) {
                    if (qq * RS < rc) {  // warp-uniform
                        mbar_wait(smem_u32(&gfullb[slot]), par, a.err);
                        if (!SL_CL_FINE_STAMPS && tid == 0 && s == 0 && sb == 0 && qq == 0) cl_stamp<DBG>(a, a.dbg_step, 11);
                        const uint32_t sbase = smem_u32(gring + (size_t)slot * slot_bytes);
#pragma unroll
                        for (int rr = 0; rr < RS; ++rr) {
                            const int r = qq * RS + rr;
                            x[r] = (gv && r < rc) ? lds_v4(sbase + (uint32_t)rr * (uint32_t)row_bytes + (uint32_t)gq * 16u)
                                                  : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(&gemptyb[slot]));  // rows are in registers
                        if (++slot == (uint32_t)nsg) { slot = 0; par ^= 1u; }
                    } else {
#pragma unroll
                        for (int rr = 0; rr < RS; ++rr) x[qq * RS + rr] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
                if (SL_CL_FINE_STAMPS && tid == 0 && sb == 1) cl_stamp<DBG>(a, s, 12);
                if (sb == 0 && s > 0) {
                    // the rows of the step's first pass are already in registers: only now wait for the weights
                    take_weights();
                    if (tid == 0) cl_stamp<DBG>(a, s, 5);
                }
                // packed fp32 pairs (FFMA2): half the issue slots of the two inner loops
                float d[R];
                const float2 wlo = make_float2(w.x, w.y), whi = make_float2(w.z, w.w);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float2 acc = __ffma2_rn(make_float2(x[r].x, x[r].y), wlo, make_float2(0.f, 0.f));
                    acc = __ffma2_rn(make_float2(x[r].z, x[r].w), whi, acc);
                    d[r] = acc.x + acc.y;
                }
                MultiReduce<R, 16>::run(d, lane);
                float z = d[0];  // warp total for row ridx
                if (SL_CL_FINE_STAMPS && tid == 0 && sb < 2) cl_stamp<DBG>(a, s, sb == 0 ? 8 : 13);
                if (WPG > 1) {
                    float* rb = red + red_buf * (NCW * R);
                    if (canon) rb[(rg * WPG + wig) * R + ridx] = z;
                    if (sb == 0) cp_async_wait_all();  // this thread's share of the step's labels has landed
                    named_bar_sync(2 + rg, CT);
                    z = 0.f;
#pragma unroll
                    for (int k = 0; k < WPG; ++k) z += rb[(rg * WPG + k) * R + ridx];
                    red_buf ^= 1;
                }
                if (SL_CL_FINE_STAMPS && tid == 0 && sb < 2) cl_stamp<DBG>(a, s, sb == 0 ? 9 : 14);
                float resid = 0.f;
                if (ridx < rc) {
                    const float yv = ybuf[(size_t)((s & 1) * RG + rg) * a.ycap + sb * R + ridx];
                    resid = __frcp_rn(1.0f + __expf(-z)) - yv;  // MUFU.EX2 / MUFU.RCP: ~2 ulp, far inside the tolerance
                    if (canon && wig == 0 && want_loss) zyb[j * R + ridx] = make_float2(z, yv);
                }
                if (SL_CL_FINE_STAMPS && tid == 0 && sb == 0) cl_stamp<DBG>(a, s, 10);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float rr = __shfl_sync(0xffffffffu, resid, lane_of_row<R>(r));
                    const float2 rr2 = make_float2(rr, rr);
                    glo = __ffma2_rn(rr2, make_float2(x[r].x, x[r].y), glo);
                    ghi = __ffma2_rn(rr2, make_float2(x[r].z, x[r].w), ghi);
                }
                if (SL_CL_FINE_STAMPS && tid == 0 && sb < 2) cl_stamp<DBG>(a, s, sb == 0 ? 11 : 15);
            }
            if (tid == 0) cl_stamp<DBG>(a, s, 1);
            // ---- hand the gradient chunk to the owner of its columns, straight from registers
            if (gv) st_async_v4(send_addr, make_float4(glo.x, glo.y, ghi.x, ghi.y), send_bar);
            if (tid == 0) cl_stamp<DBG>(a, s, 2);
            // ---- off the critical path: this group's loss of the step
            if (wig == 0 && want_loss) {
                float l = 0.f;
                __syncwarp();
                for (int sb = 0; sb < nsb; ++sb) {
                    const int j = sb * RG + rg;
                    for (int i = lane; i < R; i += 32) {
                        if (j < nst_c && j * R + i < nrows_c) {
                            const float2 v = zyb[j * R + i];
                            l += row_loss_f(v.x, v.y);
                        }
                    }
                }
                l = warp_sum(l);
                if (lane == 0)
                    st_async_b32(lsend_addr + (uint32_t)((s & 1) * NSRC * 4), __float_as_uint(l), (s & 1) ? lsend_bar1 : lsend_bar0);
            }
            // The first compute thread of a cluster's rank-0 CTA sums the cluster's loss partials and adds them to
            // the step's accumulator; in cluster 0 it also reads the PREVIOUS step's total (complete by now) and
            // writes it out.  A compute warp has the whole exchange to do this in; an exchange thread has not.
            // Three accumulators by step: adds of step s+2 need the weights of s+2, i.e. this CTA's gradient of
            // s+1, which this thread hands off after it has read s-1.
            if (loss_lead) {
                const uint32_t lb = smem_u32(&lfull[s & 1]);
                mbar_wait_cluster(lb, (uint32_t)(s >> 1) & 1u, a.err);
                float ls = 0.f;
#pragma unroll
                for (int k = 0; k < NSRC; ++k) ls += lrecv[(s & 1) * NSRC + k];
                if (s + 2 < n_steps) mbar_arrive_expect_tx(lb, (uint32_t)NSRC * 4u);
                red_add_u64(a.lacc + l3, ((unsigned long long)to_fix32(ls) << kClCntBits) + 1ull);
                if (q == 0 && s > 0) {
                    const int lp = l3 == 0 ? 2 : l3 - 1;
                    unsigned long long pv = lprev[lp];
                    const long long Lt = poll_total<false>(a.lacc + lp, pv, (unsigned)a.ncl, a.err, kErrTimeoutPart, kSpinLimit);
                    lprev[lp] = pv;
                    a.loss_out[s - 1] = ((double)Lt * kClFixInv) * a.inv_b;
                }
                l3 = l3 == 2 ? 0 : l3 + 1;
            }
            __syncwarp();
            if (s + 1 < n_steps) stage_labels(mb_next, (s + 1) & 1);
            mbar_wait(smem_u32(xdone), (uint32_t)s & 1u, a.err);
            mb = mb_next;
        }
        if (loss_lead && q == 0 && n_steps > 0) {  // the last step's loss, and where the accumulators stand
            const int lp = l3 == 0 ? 2 : l3 - 1;
            unsigned long long pv = lprev[lp];
            const long long Lt = poll_total<false>(a.lacc + lp, pv, (unsigned)a.ncl, a.err, kErrTimeoutPart, kSpinLimit);
            lprev[lp] = pv;
            a.loss_out[n_steps - 1] = ((double)Lt * kClFixInv) * a.inv_b;
            a.lbase[0] = lprev[0];
            a.lbase[1] = lprev[1];
            a.lbase[2] = lprev[2];
        }
    }
    // nobody leaves while a neighbour may still store into its shared memory
    __syncwarp();
    cluster_sync_all();
    if (!SL_CL_FINE_STAMPS && tid == 0) cl_stamp<DBG>(a, a.dbg_step, 15);  // kernel exit
}

}  // namespace slk
