// kernels_wide.cuh — K1w: the persistent minibatch-step kernel for WIDE rows (one row does not
// fit a thread's registers together with its weights: F up to 4*CH*32*NCW, e.g. 16384).
//
// Same protocol as lr_persist_kernel (kernels.cuh) — TMA row ring, fixed-point gradient summed
// by L2, counter, optional cross-GPU stage — with the roles laid out for long steps:
//   * every compute warp works on the SAME row (one row group); a thread keeps the fp32 weights
//     and the gradient of its CH float4 column chunks in registers and reads the row from
//     shared memory twice (dot product, then gradient) instead of holding it in registers;
//   * that frees the shared memory a weight copy would take: three 64 KB ring slots at F=16384;
//   * the gradient leaves through a 16 KB staging buffer, one TMA bulk reduction per 2048-column
//     chunk (instead of one red per element), the total is read back straight from L2;
//   * the fp64 master weights live in a per-CTA global scratch (L2 resident), touched once a step.
#pragma once
#include "kernels.cuh"

namespace slk {

template <int CH, int NCW>
__global__ void __launch_bounds__((NCW + 2) * 32, 1) lr_wide_kernel(const LrArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = blockIdx.x, G = gridDim.x;
    constexpr int NCT = NCW * 32;
    constexpr int CHUNK = NCT * 4;  // columns staged per bulk reduction

    unsigned char* ring = smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_bar);   // [ns]
    uint64_t* empty = full + a.ns;                                    // [ns]
    uint64_t* gfull = empty + a.ns;                                   // chunk staged
    uint64_t* gfree = gfull + 1;                                      // staging buffer read
    uint64_t* wfull = gfree + 1;                                      // total available
    float* red = reinterpret_cast<float*>(smem + a.off_red);          // [2][NCW]
    float2* zy = reinterpret_cast<float2*>(smem + a.off_zy);          // [2][zy_rows]
    long long* gout = reinterpret_cast<long long*>(smem + a.off_gout);  // [CHUNK]

    const int F = a.F, ns = a.ns, row_bytes = a.row_bytes;
    const long long lo = ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = (int)(hi - lo);
    const int nchunk = (F + CHUNK - 1) / CHUNK;  // <= CH
    const long long mb0 = (long long)(((__int128)a.step0 * a.B) % a.n_rows);
    auto wrap = [&](long long v) {
        if (v >= a.n_rows) {
            v -= a.n_rows;
            if (v >= a.n_rows) v %= a.n_rows;
        }
        return v;
    };

    if (tid == 0) {
        for (int s = 0; s < ns; ++s) {
            mbar_init(smem_u32(&full[s]), 1);
            mbar_init(smem_u32(&empty[s]), NCW);
        }
        mbar_init(smem_u32(gfull), NCW);
        mbar_init(smem_u32(gfree), 1);
        mbar_init(smem_u32(wfull), 1);
        fence_mbar_init();
    }
    __syncthreads();

    // =========================================================================== producer
    if (warp == NCW) {
        if (lane == 0 && nrows_c > 0) {
            const uint64_t pol = policy_evict_first();
            uint32_t slot = 0, phase = 0;
            long long mb = mb0;
            for (int s = 0; s < a.n_steps; ++s) {
                long long grow = wrap(mb + lo);
                for (int i = 0; i < nrows_c; ++i) {
                    mbar_wait(smem_u32(&empty[slot]), phase ^ 1u, a.err);
                    const uint32_t fb = smem_u32(&full[slot]);
                    mbar_arrive_expect_tx(fb, (uint32_t)row_bytes);
                    bulk_g2s(smem_u32(ring + (size_t)slot * row_bytes), a.X + (size_t)grow * F,
                             (uint32_t)row_bytes, fb, pol);
                    if (++slot == (uint32_t)ns) { slot = 0; phase ^= 1u; }
                    if (++grow == a.n_rows) grow = 0;
                }
                mb = wrap(mb + a.B);
            }
        }
        return;
    }

    // =========================================================================== exchange
    if (warp == NCW + 1) {
        uint32_t kc = 0;  // chunks handed over so far (phase of gfull / gfree)
        for (int s = 0; s < a.n_steps; ++s) {
            unsigned long long* buf = a.acc + (size_t)(s % 3) * a.Fp;
            if (lane == 0) {
                for (int ch = 0; ch < nchunk; ++ch, ++kc) {
                    mbar_wait(smem_u32(gfull), kc & 1u, a.err);
                    const int n = min(CHUNK, F - ch * CHUNK);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    bulk_reduce_add_u64(buf + (size_t)ch * CHUNK, smem_u32(gout), (uint32_t)n * 8u);
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging reusable
                    mbar_arrive(smem_u32(gfree));
                }
                asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
                asm volatile("fence.proxy.async;" ::: "memory");
                red_release_add_u32(a.counter, 1u);
                dbg_stamp(a, s, 3);
                const unsigned want = (unsigned)G * (unsigned)(s + 1);
                uint32_t spins = 0;
                while ((int)(ld_relaxed_u32(a.counter) - want) < 0) {
                    if (spin_abort(spins, a.err, kErrTimeoutPart)) break;
                }
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
            }
            if (a.nranks > 0) {
                // cross-GPU stage (see lr_persist_kernel): slice owners trade s64 sums with the
                // peers through NVLink inboxes and publish the all-rank total in `fin`
                __syncwarp();
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                const unsigned xflag = a.xflag0 + (unsigned)s + 1u;
                const size_t half = (size_t)(xflag & 1u) * (size_t)a.nranks * (size_t)F;
                unsigned long long* fin = a.fin + (size_t)(s % 3) * F;
                if (c < a.n_own) {
                    const int e1 = min(F, (c + 1) * a.xslice);
                    for (int e = c * a.xslice + lane; e < e1; e += 32) {
                        long long tot = (long long)ld_relaxed_u64(buf + e);
                        const uint4 pkt = make_uint4((unsigned)((unsigned long long)tot & 0xffffffffu), xflag,
                                                     (unsigned)((unsigned long long)tot >> 32), xflag);
                        for (int p = 0; p < a.nranks; ++p)
                            if (p != a.rank) st_vol_v4(a.peer_x[p] + half + (size_t)a.rank * F + e, pkt);
                    }
                    const uint4* inbox = a.peer_x[a.rank] + half;
                    for (int e = c * a.xslice + lane; e < e1; e += 32) {
                        long long tot = (long long)ld_relaxed_u64(buf + e);
                        for (int p = 0; p < a.nranks; ++p) {
                            if (p == a.rank) continue;
                            uint4 q = ld_vol_v4(inbox + (size_t)p * F + e);
                            uint32_t spins = 0;
                            while (q.y != xflag || q.w != xflag) {
                                if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                                q = ld_vol_v4(inbox + (size_t)p * F + e);
                            }
                            tot += (long long)(((unsigned long long)q.z << 32) | (unsigned long long)q.x);
                        }
                        fin[e] = (unsigned long long)tot;
                    }
                    __syncwarp();
                    if (lane == 0) red_release_add_u32(a.counter2, 1u);
                }
                if (lane == 0) {
                    const unsigned want2 = (unsigned)a.n_own * (unsigned)(s + 1);
                    uint32_t spins = 0;
                    while ((int)(ld_relaxed_u32(a.counter2) - want2) < 0) {
                        if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                    }
                    asm volatile("fence.acq_rel.gpu;" ::: "memory");
                }
            }
            if (lane == 0) {
                mbar_arrive(smem_u32(wfull));  // the compute threads read the total from L2
                dbg_stamp(a, s, 4);
            }
            __syncwarp();
            // clear this CTA's slice of the buffer of step s-1 for step s+2 (everybody is past it)
            unsigned long long* z = a.acc + (size_t)((s + 2) % 3) * a.Fp;
            for (size_t i = (size_t)c * 32 + lane; i < (size_t)a.Fp; i += (size_t)G * 32) z[i] = 0ull;
            __syncwarp();
        }
        return;
    }

    // =========================================================================== compute
    const int ct = warp * 32 + lane;
    // column chunk ch of this thread: float4 index ct + ch*NCT (valid while it is inside the row)
#define GQ(ch) (ct + (ch) * NCT)
#define GV(ch) ((ct + (ch) * NCT) * 4 < F)
    double* my_master = a.wmaster + (size_t)c * F;
    float4 w[CH];
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        w[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GV(ch)) {
            const double2 m0 = *reinterpret_cast<const double2*>(a.model + GQ(ch) * 4);
            const double2 m1 = *reinterpret_cast<const double2*>(a.model + GQ(ch) * 4 + 2);
            *reinterpret_cast<double2*>(my_master + GQ(ch) * 4) = m0;
            *reinterpret_cast<double2*>(my_master + GQ(ch) * 4 + 2) = m1;
            w[ch] = make_float4((float)m0.x, (float)m0.y, (float)m1.x, (float)m1.y);
        }
    }
    // w <- w - (eta/B) * total of step s, in batches of 2 chunks (all loads of a batch in flight)
    auto apply_total = [&](int s) {
        const unsigned long long* tot = a.nranks > 0 ? a.fin + (size_t)(s % 3) * F
                                                     : a.acc + (size_t)(s % 3) * a.Fp;
#pragma unroll
        for (int c0 = 0; c0 < CH; c0 += 2) {
            ulonglong2 u[2], v[2];
            double2 m0[2], m1[2];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int ch = c0 + k;
                if (ch < CH && GV(ch)) {
                    u[k] = ld_relaxed_v2u64(tot + GQ(ch) * 4);
                    v[k] = ld_relaxed_v2u64(tot + GQ(ch) * 4 + 2);
                    m0[k] = __ldcg(reinterpret_cast<const double2*>(my_master + GQ(ch) * 4));
                    m1[k] = __ldcg(reinterpret_cast<const double2*>(my_master + GQ(ch) * 4 + 2));
                }
            }
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int ch = c0 + k;
                if (ch < CH && GV(ch)) {
                    m0[k].x -= a.eta_over_b * ((double)(long long)u[k].x * kFixInv);
                    m0[k].y -= a.eta_over_b * ((double)(long long)u[k].y * kFixInv);
                    m1[k].x -= a.eta_over_b * ((double)(long long)v[k].x * kFixInv);
                    m1[k].y -= a.eta_over_b * ((double)(long long)v[k].y * kFixInv);
                    __stcg(reinterpret_cast<double2*>(my_master + GQ(ch) * 4), m0[k]);
                    __stcg(reinterpret_cast<double2*>(my_master + GQ(ch) * 4 + 2), m1[k]);
                    w[ch] = make_float4((float)m0[k].x, (float)m0[k].y, (float)m1[k].x, (float)m1[k].y);
                }
            }
        }
    };

    uint32_t slot = 0, par = 0;
    int red_buf = 0;
    uint32_t kc = 0;
    long long mb = mb0;
    float y_next = (nrows_c > 0 && lane == 0) ? __ldg(a.y + (size_t)wrap(mb + lo)) : 0.f;

    for (int s = 0; s < a.n_steps; ++s) {
        const long long mb_next = wrap(mb + a.B);
        float4 g[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        float2* zyb = zy + (size_t)(s & 1) * a.zy_rows;
        if (tid == 0) dbg_stamp(a, s, 0);
        if (s > 0) {
            mbar_wait(smem_u32(wfull), (uint32_t)(s - 1) & 1u, a.err);
            apply_total(s - 1);
            if (tid == 0) dbg_stamp(a, s, 5);
        }
        for (int i = 0; i < nrows_c; ++i) {
            const float yv = __shfl_sync(0xffffffffu, y_next, 0);
            if (lane == 0) {  // label of the next row (maybe the next step's first)
                if (i + 1 < nrows_c) y_next = __ldg(a.y + (size_t)wrap(mb + lo + i + 1));
                else if (s + 1 < a.n_steps) y_next = __ldg(a.y + (size_t)wrap(mb_next + lo));
            }
            mbar_wait(smem_u32(&full[slot]), par, a.err);
            const uint32_t rb = smem_u32(ring + (size_t)slot * row_bytes);
            float acc = 0.f;
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (GV(ch)) {
                    const float4 x = lds_v4(rb + (uint32_t)GQ(ch) * 16u);
                    acc = fmaf(x.x, w[ch].x, acc);
                    acc = fmaf(x.y, w[ch].y, acc);
                    acc = fmaf(x.z, w[ch].z, acc);
                    acc = fmaf(x.w, w[ch].w, acc);
                }
            }
            acc = warp_sum(acc);
            float* rbuf = red + red_buf * NCW;
            if (lane == 0) rbuf[warp] = acc;
            named_bar_sync(1, NCT);
            float z = 0.f;
#pragma unroll
            for (int k = 0; k < NCW; ++k) z += rbuf[k];
            red_buf ^= 1;
            const float resid = sigmoid_f(z) - yv;
            if (tid == 0 && a.loss_fix) zyb[i] = make_float2(z, yv);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (GV(ch)) {
                    const float4 x = lds_v4(rb + (uint32_t)GQ(ch) * 16u);
                    g[ch].x = fmaf(resid, x.x, g[ch].x);
                    g[ch].y = fmaf(resid, x.y, g[ch].y);
                    g[ch].z = fmaf(resid, x.z, g[ch].z);
                    g[ch].w = fmaf(resid, x.w, g[ch].w);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&empty[slot]));
            if (++slot == (uint32_t)ns) { slot = 0; par ^= 1u; }
        }
        if (tid == 0) dbg_stamp(a, s, 1);
        // ---- hand the gradient to the exchange warp, 2048 columns at a time ---------------------
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (ch < nchunk) {  // uniform
                if (kc > 0) mbar_wait(smem_u32(gfree), (kc - 1) & 1u, a.err);  // previous chunk was read
                if (GV(ch)) {
                    longlong2* p = reinterpret_cast<longlong2*>(gout + ct * 4);
                    p[0] = make_longlong2(to_fix(g[ch].x), to_fix(g[ch].y));
                    p[1] = make_longlong2(to_fix(g[ch].z), to_fix(g[ch].w));
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(gfull));
                ++kc;
            }
        }
        if (tid == 0) dbg_stamp(a, s, 2);
        if (warp == 0 && a.loss_fix) {  // off the critical path: this step's loss
            float l = 0.f;
            for (int i = lane; i < nrows_c; i += 32) {
                const float2 v = zyb[i];
                l += softplus_f(v.x) - v.y * v.x;
            }
            l = warp_sum(l);
            if (lane == 0) red_add_u64(a.loss_fix + s, (unsigned long long)to_fix(l));
        }
        mb = mb_next;
    }
    if (a.n_steps > 0) {
        mbar_wait(smem_u32(wfull), (uint32_t)(a.n_steps - 1) & 1u, a.err);
        apply_total(a.n_steps - 1);
    }
    __threadfence();
    if (c == 0) {
#pragma unroll
        for (int ch = 0; ch < CH; ++ch)
            if (GV(ch)) {
                *reinterpret_cast<double2*>(a.model + GQ(ch) * 4) =
                    __ldcg(reinterpret_cast<const double2*>(my_master + GQ(ch) * 4));
                *reinterpret_cast<double2*>(a.model + GQ(ch) * 4 + 2) =
                    __ldcg(reinterpret_cast<const double2*>(my_master + GQ(ch) * 4 + 2));
            }
    }
}

#undef GQ
#undef GV

}  // namespace slk
