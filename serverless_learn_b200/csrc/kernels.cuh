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
__device__ __forceinline__ uint4 ld_vol_v4(const uint4* p) {
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_vol_v4(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
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
constexpr int kErrTimeoutMbar = 1;
constexpr int kErrTimeoutPart = 2;
constexpr int kErrTimeoutW = 3;

__device__ __forceinline__ bool spin_abort(uint32_t& spins, int* err, int code) {
    ++spins;
    if ((spins & 1023u) == 0) {
        if (ld_vol_s32(err) != 0) return true;
        if (spins >= kSpinLimit) {
            atomicCAS(err, 0, code);
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

__device__ __forceinline__ float softplus_f(float z) {
    return fmaxf(z, 0.0f) + log1pf(expf(-fabsf(z)));
}
__device__ __forceinline__ float sigmoid_f(float z) { return 1.0f / (1.0f + expf(-z)); }

// ------------------------------------------------------------------------------------
// K1 — persistent minibatch-step kernel
// ------------------------------------------------------------------------------------
struct LrArgs {
    const float* X;        // [n_rows, F] row-major, resident shard
    const float* y;        // [n_rows]
    long long n_rows;
    int F;
    int B;
    long long step0;       // global step index of the first step of this launch
    int n_steps;
    double* model;         // fp64 master weights [F] (model_state, src/master.cc:58)
    uint4* part_ll;        // [grid][Q] gradient-partial packets  {v0, flag, v1, flag}
    uint4* w_ll;           // [Q]       new-weight packets        {w0, flag, w1, flag}
    double* loss_out;      // device [n_steps] or nullptr
    unsigned flag0;        // step s of this launch uses flag0 + s + 1
    double eta_over_b;
    int* err;              // 0 = ok; set by the watchdog
    // geometry, computed on the host
    int Q;                 // packets per vector: F/2 weights pairs + 1 loss packet
    int qpc;               // packets owned (reduced) per CTA
    int wpg;               // warps per row group = ceil(F/4/CH/32)
    int ns;                // ring slots (one row each)
    int row_bytes;         // 4*F
    int off_full, off_red, off_lsm, off_own, off_gsm;  // smem offsets
};

// CH  float4 column chunks per thread per row        (F <= 4*CH*32*NCW)
// R   rows per row group per sub-batch (power of two)
// NCW warps per CTA.  There is no dedicated producer warp: the leader thread of a row
// group re-arms a ring slot with the row NS positions ahead as soon as its group has
// pulled the slot's row into registers ("refill on release"), so the TMA stream runs
// NS rows ahead of the maths and never depends on the weights.
template <int CH, int R, int NCW>
__global__ void __launch_bounds__(NCW * 32, 1) lr_persist_kernel(const LrArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = blockIdx.x, G = gridDim.x;

    unsigned char* ring = smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_full);
    float* red = reinterpret_cast<float*>(smem + a.off_red);    // [2][NCW][R]
    float* lsm = reinterpret_cast<float*>(smem + a.off_lsm);    // [NCW]
    double* own_w = reinterpret_cast<double*>(smem + a.off_own);  // [2*qpc]
    float* gsm = reinterpret_cast<float*>(smem + a.off_gsm);    // [RG-1][F]

    const int WPG = a.wpg, RG = NCW / WPG, CT = WPG * 32;
    const int F = a.F, ns = a.ns, row_bytes = a.row_bytes;
    const long long lo = ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = (int)(hi - lo);  // this CTA's rows of every minibatch
    const int rows_per_sb = R * RG;
    const int nsb = (nrows_c + rows_per_sb - 1) / rows_per_sb;

    const long long total_rows = (long long)a.n_steps * nrows_c;  // rows this CTA streams
    const uint64_t pol = policy_evict_first();
    // issue the TMA copy of this CTA's seq-th row into its ring slot
    auto issue_row = [&](long long seq) {
        const long long sp = seq / nrows_c;
        const int ip = (int)(seq - sp * nrows_c);
        const long long mbp = (long long)(((__int128)(a.step0 + sp) * a.B) % a.n_rows);
        const long long grow = (mbp + lo + ip) % a.n_rows;
        const uint32_t slot = (uint32_t)(seq % ns);
        const uint32_t fb = smem_u32(&full[slot]);
        mbar_arrive_expect_tx(fb, (uint32_t)row_bytes);
        bulk_g2s(smem_u32(ring + (size_t)slot * row_bytes), a.X + (size_t)grow * F,
                 (uint32_t)row_bytes, fb, pol);
    };
    if (tid == 0) {
        for (int s = 0; s < ns; ++s) mbar_init(smem_u32(&full[s]), 1);
        fence_mbar_init();
        const long long pre = total_rows < ns ? total_rows : ns;
        for (long long q = 0; q < pre; ++q) issue_row(q);
    }
    __syncthreads();

    const int rg = warp / WPG, wig = warp - rg * WPG;
    const int ct = wig * 32 + lane;
    const bool active = rg < RG;  // warps beyond RG*WPG idle through the maths (still sync)
    const bool leader = active && wig == 0 && lane == 0;  // re-arms this row group's slots
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

    // weights: fp32 working copy of the fp64 master
    float4 w[CH];
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        w[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gv[ch]) {
            const double* m = a.model + gq[ch] * 4;
            w[ch] = make_float4((float)m[0], (float)m[1], (float)m[2], (float)m[3]);
        }
    }
    // packets this CTA reduces, and the fp64 master copy of the weights in them
    const int q0 = min(a.Q, c * a.qpc), q1 = min(a.Q, (c + 1) * a.qpc);
    for (int i = tid; i < 2 * (q1 - q0); i += NCT) {
        const int col = 2 * q0 + i;
        own_w[i] = col < F ? a.model[col] : 0.0;
    }
    named_bar_sync(1, NCT);

    uint32_t base_slot = 0, base_phase = 0;
    int red_buf = 0;
    long long mb = (long long)(((__int128)a.step0 * a.B) % a.n_rows);

    for (int s = 0; s < a.n_steps; ++s) {
        const unsigned flag = a.flag0 + (unsigned)s + 1u;
        float4 g[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        float loss_acc = 0.f;

        for (int sb = 0; sb < nsb; ++sb) {
            // label of the row this lane finishes (issued early: the latency hides
            // behind the shared-memory waits below)
            const int iy = sb * rows_per_sb + ridx * RG + rg;
            const bool vy = active && iy < nrows_c;
            float yv = 0.f;
            if (vy) yv = __ldg(a.y + (size_t)((mb + lo + iy) % a.n_rows));

            float4 x[R][CH];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int i = sb * rows_per_sb + r * RG + rg;
                if (active && i < nrows_c) {  // warp-uniform
                    const uint32_t sl = base_slot + (uint32_t)i;
                    const uint32_t k = sl / (uint32_t)ns;
                    const uint32_t slot = sl - k * (uint32_t)ns;
                    mbar_wait(smem_u32(&full[slot]), (base_phase ^ k) & 1u, a.err);
                    const uint32_t rb = smem_u32(ring + (size_t)slot * row_bytes);
#pragma unroll
                    for (int ch = 0; ch < CH; ++ch)
                        x[r][ch] = gv[ch] ? lds_v4(rb + (uint32_t)gq[ch] * 16u)
                                          : make_float4(0.f, 0.f, 0.f, 0.f);
                } else {
#pragma unroll
                    for (int ch = 0; ch < CH; ++ch) x[r][ch] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            // partial dot products over this thread's columns
            float d[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                float acc = 0.f;
#pragma unroll
                for (int ch = 0; ch < CH; ++ch) {
                    acc = fmaf(x[r][ch].x, w[ch].x, acc);
                    acc = fmaf(x[r][ch].y, w[ch].y, acc);
                    acc = fmaf(x[r][ch].z, w[ch].z, acc);
                    acc = fmaf(x[r][ch].w, w[ch].w, acc);
                }
                d[r] = acc;
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
            // Every warp of the group has its rows in registers (the dot products above
            // consumed them, and for WPG > 1 the barrier was passed): re-arm the slots
            // with the rows NS positions further down this CTA's stream.
            if (leader) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll 1
                for (int r = 0; r < R; ++r) {
                    const int i = sb * rows_per_sb + r * RG + rg;
                    if (i < nrows_c) {
                        const long long nxt = (long long)s * nrows_c + i + ns;
                        if (nxt < total_rows) issue_row(nxt);
                    }
                }
            }
            float resid = 0.f;
            if (vy) {
                resid = sigmoid_f(z) - yv;
                if (canon && wig == 0) loss_acc += softplus_f(z) - yv * z;
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
        {   // advance the ring cursor past this step's rows
            const uint32_t sl = base_slot + (uint32_t)nrows_c;
            const uint32_t k = sl / (uint32_t)ns;
            base_slot = sl - k * (uint32_t)ns;
            base_phase ^= (k & 1u);
        }

        // ---- combine row groups inside the CTA, publish this CTA's partial ----------
        if (active && wig == 0) {
            const float l = warp_sum(loss_acc);
            if (lane == 0) lsm[rg] = l;
        }
        if (RG > 1) {
            if (active && rg > 0) {
#pragma unroll
                for (int ch = 0; ch < CH; ++ch)
                    if (gv[ch])
                        *reinterpret_cast<float4*>(gsm + (size_t)(rg - 1) * F + gq[ch] * 4) = g[ch];
            }
            named_bar_sync(1, NCT);
        }
        uint4* mine = a.part_ll + (size_t)c * a.Q;
        if (tid == 0) {
            float l = 0.f;
            for (int k = 0; k < RG; ++k) l += lsm[k];
            st_vol_v4(mine + (a.Q - 1), make_uint4(__float_as_uint(l), flag, 0u, flag));
        }
        if (rg == 0) {
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (!gv[ch]) continue;
                float4 t = g[ch];
                for (int k = 1; k < RG; ++k) {
                    const float4 o =
                        *reinterpret_cast<const float4*>(gsm + (size_t)(k - 1) * F + gq[ch] * 4);
                    t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
                }
                st_vol_v4(mine + 2 * gq[ch],
                          make_uint4(__float_as_uint(t.x), flag, __float_as_uint(t.y), flag));
                st_vol_v4(mine + 2 * gq[ch] + 1,
                          make_uint4(__float_as_uint(t.z), flag, __float_as_uint(t.w), flag));
            }
        }

        // ---- owner: sum the packets of every CTA in a fixed order, update, publish ----
        for (int q = q0 + warp; q < q1; q += NCW) {
            double a0 = 0.0, a1 = 0.0;
            for (int sbase = 0; sbase < G; sbase += 128) {
                uint4 p[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int src = sbase + k * 32 + lane;
                    if (src < G) p[k] = ld_vol_v4(a.part_ll + (size_t)src * a.Q + q);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int src = sbase + k * 32 + lane;
                    if (src < G) {
                        uint32_t spins = 0;
                        while (p[k].y != flag || p[k].w != flag) {
                            if (spin_abort(spins, a.err, kErrTimeoutPart)) break;
                            p[k] = ld_vol_v4(a.part_ll + (size_t)src * a.Q + q);
                        }
                        a0 += (double)__uint_as_float(p[k].x);
                        a1 += (double)__uint_as_float(p[k].z);
                    }
                }
            }
            a0 = warp_sum(a0);
            a1 = warp_sum(a1);
            if (lane == 0) {
                if (q == a.Q - 1) {
                    if (a.loss_out) a.loss_out[s] = a0 / (double)a.B;
                } else {
                    const int i = 2 * (q - q0);
                    const double w0 = own_w[i] - a.eta_over_b * a0;
                    const double w1 = own_w[i + 1] - a.eta_over_b * a1;
                    own_w[i] = w0;
                    own_w[i + 1] = w1;
                    st_vol_v4(a.w_ll + q, make_uint4(__float_as_uint((float)w0), flag,
                                                     __float_as_uint((float)w1), flag));
                }
            }
        }

        // ---- everybody: pick up the new weights ------------------------------------
        if (s + 1 < a.n_steps) {
            uint4 p[CH][2];
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (gv[ch]) {
                    p[ch][0] = ld_vol_v4(a.w_ll + 2 * gq[ch]);
                    p[ch][1] = ld_vol_v4(a.w_ll + 2 * gq[ch] + 1);
                }
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (gv[ch]) {
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        uint32_t spins = 0;
                        while (p[ch][hh].y != flag || p[ch][hh].w != flag) {
                            if (spin_abort(spins, a.err, kErrTimeoutW)) break;
                            p[ch][hh] = ld_vol_v4(a.w_ll + 2 * gq[ch] + hh);
                        }
                    }
                    w[ch] = make_float4(__uint_as_float(p[ch][0].x), __uint_as_float(p[ch][0].z),
                                        __uint_as_float(p[ch][1].x), __uint_as_float(p[ch][1].z));
                }
            mb = (mb + a.B) % a.n_rows;
        }
    }

    // master weights back to global (each packet's owner lane wrote own_w itself)
    for (int q = q0 + warp; q < q1; q += NCW) {
        if (lane == 0 && q < a.Q - 1) {
            const int i = 2 * (q - q0);
            a.model[2 * q] = own_w[i];
            a.model[2 * q + 1] = own_w[i + 1];
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
        rowloss[j] = softplus_f(z) - yy * z;
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
