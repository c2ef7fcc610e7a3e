// kernels_wide.cuh — K1w: the persistent minibatch-step kernel for WIDE rows (one row does not
// fit a thread's registers together with its weights: F up to 4*CH*32*NCW, e.g. 16384).
//
// Same row ring and fixed-point arithmetic as lr_persist_kernel (kernels.cuh), with the roles laid
// out for long steps:
//   * the compute warps form NG row groups; all warps of a group work on the SAME R rows (R = 1
//     or 2 per pass); a thread keeps the fp32 weights and the gradient of its CH float4 column
//     chunks in registers and reads the row from shared memory twice (dot product, then
//     gradient) instead of holding it in registers;
//   * that frees the shared memory a weight copy would take: three 64 KB ring slots at F=16384;
//   * the gradient is summed in two phases THROUGH L2 WITH PLAIN STORES AND LOADS, not atomics
//     (148 CTAs x 128 KB of L2 reductions per step measured ~100 us at F=16384): every row group
//     stores its fp32 partial vector to `acc[c*NG+grp][F]`, a counter closes the phase, then the
//     CTA that owns a slice of the vector converts the partials of its slice to fixed point and
//     adds them (integer adds: any order gives the same bits);
//   * ONE fp64 master (`model`) per GPU: the slice owner updates it from the total (after trading
//     the slice with the other GPUs when peers are attached: 16-byte data+flag packets over
//     NVLink, see lr_persist_kernel) and publishes fp32 weights; a second counter closes that
//     phase and everybody loads 4F bytes of weights.
#pragma once
#include "kernels.cuh"

namespace slk {

__device__ __forceinline__ float4 ld_vol_f32x4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

template <int CH, int NCW, int NG, int R>
__global__ void __launch_bounds__((NCW + 1) * 32, 1) lr_wide_kernel(const LrArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = blockIdx.x, G = gridDim.x;
    constexpr int NCT = NCW * 32;
    constexpr int GW = NCW / NG, GT = GW * 32;  // NG row groups of GW warps: a group works on one row
    static_assert(GW * NG == NCW, "row groups must divide the compute warps");
    static_assert(R >= 1 && R <= 4, "a group takes one to four rows per pass");

    unsigned char* ring = smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_bar);   // [ns]
    uint64_t* empty = full + a.ns;                                    // [ns]
    float* red = reinterpret_cast<float*>(smem + a.off_red);          // [2][NCW][R]
    float2* zy = reinterpret_cast<float2*>(smem + a.off_zy);          // [2][zy_rows]
    long long* gsm = reinterpret_cast<long long*>(smem + a.off_gout);  // [groups][xslice] owner partial sums

    const int F = a.F, ns = a.ns, row_bytes = a.row_bytes;
    // this CTA's positions of every minibatch: every G-th position (at any moment the grid then
    // reads one contiguous window of the shard: measured 4-5 % faster at F >= 8192 than giving every
    // CTA its own contiguous block, which is what `block_rows` restores)
    const bool ilv = a.block_rows == 0;
    const long long lo = ilv ? c : ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = ilv ? (int)((a.B - c + G - 1) / G) : (int)(hi - ((long long)c * a.B) / G);
    const int pstride = ilv ? G : 1;
    const long long mb0 = (long long)(((__int128)a.step0 * a.vB) % a.n_rows);
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
            mbar_init(smem_u32(&empty[s]), GW);
        }
        fence_mbar_init();
    }
    __syncthreads();

    // =========================================================================== producer
    // Lane l of the producer warp owns ring slot l (ns <= 32): it fetches rows l, l+ns, l+2ns, ...
    // of this CTA's row sequence over the whole launch.  Issuing a row costs one thread ~600
    // cycles (wait, address, expect_tx, copy); in parallel lanes that cost no longer bounds
    // narrow rows.  A slot's fills are all issued by the same lane, in order, and (ns is a multiple
    // of NG*R) consumed by the same row group, in order, so the parity waits on its barriers never
    // alias even though fills of different slots complete out of order.
    if (warp == NCW) {
        if (lane < ns && nrows_c > 0) {
            const uint64_t pol = policy_evict_first();
            const uint32_t Q = (uint32_t)a.n_steps * (uint32_t)nrows_c;  // rows of the launch (host: < 2^31)
            const uint32_t eb = smem_u32(&empty[lane]), fb = smem_u32(&full[lane]);
            const uint32_t dstb = smem_u32(ring + (size_t)lane * row_bytes);
            uint32_t s = (uint32_t)lane / (uint32_t)nrows_c, i = (uint32_t)lane % (uint32_t)nrows_c;
            uint32_t s_cur = 0, phase = 0;
            long long mb = mb0;
            for (uint32_t q = (uint32_t)lane; q < Q; q += (uint32_t)ns) {
                while (s_cur < s) { mb = wrap(mb + a.vB); ++s_cur; }
                const RowRef ref = locate_row(a, mb, lo + (long long)i * pstride);
                mbar_wait(eb, phase ^ 1u, a.err);
                mbar_arrive_expect_tx(fb, (uint32_t)row_bytes);
                const unsigned char* srcb = reinterpret_cast<const unsigned char*>(ref.x);
                for (uint32_t o = 0; o < (uint32_t)row_bytes; o += (uint32_t)a.tma_chunk) {
                    const uint32_t m = min((uint32_t)a.tma_chunk, (uint32_t)row_bytes - o);
                    bulk_g2s(dstb + o, srcb + o, m, fb, pol);
                }
                phase ^= 1u;
                i += (uint32_t)ns;
                while (i >= (uint32_t)nrows_c) { i -= (uint32_t)nrows_c; ++s; }
            }
        }
        return;
    }

    // =========================================================================== compute
    const int grp = warp / GW;                  // my row group: rows grp, grp+NG, ... of the CTA's rows
    const int ct = (warp % GW) * 32 + lane;     // my index inside the group
    // column chunk ch of this thread: float4 index ct + ch*GT (valid while it is inside the row)
#define GQ(ch) (ct + (ch) * GT)
#define GV(ch) ((ct + (ch) * GT) * 4 < F)
    float4 w[CH];
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        w[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GV(ch)) {
            const double2 m0 = *reinterpret_cast<const double2*>(a.model + GQ(ch) * 4);
            const double2 m1 = *reinterpret_cast<const double2*>(a.model + GQ(ch) * 4 + 2);
            w[ch] = make_float4((float)m0.x, (float)m0.y, (float)m1.x, (float)m1.y);
        }
    }
    // pick up the weights the owners published for step s (all loads in flight: one L2 round trip)
    auto apply_total = [&](int s) {
        const float* wpub = a.w32buf + (size_t)(s & 1) * F;
#pragma unroll
        for (int ch = 0; ch < CH; ++ch)
            if (GV(ch)) {
                const uint4 q = ld_vol_v4(reinterpret_cast<const uint4*>(wpub + GQ(ch) * 4));
                w[ch] = make_float4(__uint_as_float(q.x), __uint_as_float(q.y), __uint_as_float(q.z),
                                    __uint_as_float(q.w));
            }
    };

    // thread 0 watches a grid-wide arrival counter, the other compute threads wait at the barrier
    auto wait_counter = [&](const unsigned* ctr, unsigned want, int code) {
        if (tid == 0) {
            uint32_t spins = 0;
            while ((int)(ld_relaxed_u32(ctr) - want) < 0) {
                if (spin_abort(spins, a.err, code, kPeerSpinLimit)) break;
            }
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
        }
        named_bar_sync(1, NCT);
    };

    // owner stage geometry: this CTA owns columns [c*L, c*L+L); a thread adds the partials
    // p = grp, grp+ngrp, ... of four columns
    const int L = a.xslice, quads = L >> 2;
    const int P = G * NG;                                // partial vectors: one per row group
    const int ngrp = max(1, min(NCT / quads, P));
    // (a slice wider than the CTA — small grids — is walked in several rounds)
    float* part = reinterpret_cast<float*>(a.acc);  // [P][Fp] fp32 partial gradients

    int red_buf = 0;
    uint32_t slot = 0, par = 0;
    long long mb = mb0;
    // lanes 0..R-1 of every warp hold the labels of the group's next R rows
    float y_next = (lane < R && grp * R + lane < nrows_c) ? __ldg(locate_row(a, mb, lo + (long long)(grp * R + lane) * pstride).y) : 0.f;

    for (int s = 0; s < a.n_steps; ++s) {
        const long long mb_next = wrap(mb + a.vB);
        float4 g[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
        float2* zyb = zy + (size_t)(s & 1) * a.zy_rows;
        if (tid == 0) dbg_stamp(a, s, 0);
        if (s > 0) {
            wait_counter(a.counter2, (unsigned)a.n_own * (unsigned)s, kErrTimeoutPeer);
            apply_total(s - 1);
            if (tid == 0) dbg_stamp(a, s, 5);
        }
        // ring position of my group's first row of this step (the producer fills slots in row
        // order; a group takes R consecutive rows per pass, the groups interleave)
        if constexpr (NG * R > 1) {
            const uint32_t q0 = (uint32_t)s * (uint32_t)nrows_c + (uint32_t)(grp * R);
            slot = q0 % (uint32_t)ns;
            par = (q0 / (uint32_t)ns) & 1u;
        }
        for (int i0 = grp * R; i0 < nrows_c; i0 += NG * R) {
            float yv[R];
#pragma unroll
            for (int r = 0; r < R; ++r) yv[r] = __shfl_sync(0xffffffffu, y_next, r);
            if (lane < R) {  // labels of my group's next pass (maybe the next step's first)
                const int in = i0 + NG * R + lane;
                if (in < nrows_c) y_next = __ldg(locate_row(a, mb, lo + (long long)in * pstride).y);
                else if (i0 + NG * R >= nrows_c && s + 1 < a.n_steps && grp * R + lane < nrows_c)
                    y_next = __ldg(locate_row(a, mb_next, lo + (long long)(grp * R + lane) * pstride).y);
            }
            uint32_t rb[R], sl[R];
            bool valid[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                valid[r] = i0 + r < nrows_c;
                uint32_t sr = slot + (uint32_t)r, pr = par;
                if (sr >= (uint32_t)ns) { sr -= (uint32_t)ns; pr ^= 1u; }
                sl[r] = sr;
                rb[r] = smem_u32(ring + (size_t)sr * row_bytes);
                if (valid[r]) mbar_wait(smem_u32(&full[sr]), pr, a.err);
            }
            if (a.dbg_skip) {  // diagnostics only: consume the slots without touching them
                __syncwarp();
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (valid[r] && lane == 0) mbar_arrive(smem_u32(&empty[sl[r]]));
            } else {
                float acc[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float a0 = 0.f, a1 = 0.f;
                    if (valid[r]) {
#pragma unroll
                        for (int ch = 0; ch < CH; ++ch) {
                            if (GV(ch)) {
                                const float4 x = lds_v4(rb[r] + (uint32_t)GQ(ch) * 16u);
                                a0 = fmaf(x.x, w[ch].x, a0);
                                a1 = fmaf(x.y, w[ch].y, a1);
                                a0 = fmaf(x.z, w[ch].z, a0);
                                a1 = fmaf(x.w, w[ch].w, a1);
                            }
                        }
                    }
                    acc[r] = a0 + a1;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                    for (int r = 0; r < R; ++r) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], o);
                }
                float* rbuf = red + (red_buf * NCW + grp * GW) * R;
                {
                    float mine = acc[0];  // lane r publishes the group's partial dot product of row r
#pragma unroll
                    for (int r = 1; r < R; ++r)
                        if (lane == r) mine = acc[r];
                    if (lane < R) rbuf[(warp % GW) * R + lane] = mine;
                }
                named_bar_sync(2 + grp, GT);
                red_buf ^= 1;
                float resid[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float z = 0.f;
#pragma unroll
                    for (int k = 0; k < GW; ++k) z += rbuf[k * R + r];
                    resid[r] = valid[r] ? sigmoid_f(z) - yv[r] : 0.f;
                    if (ct == 0 && a.loss_fix && valid[r]) zyb[i0 + r] = make_float2(z, yv[r]);
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (valid[r]) {
#pragma unroll
                        for (int ch = 0; ch < CH; ++ch) {
                            if (GV(ch)) {
                                const float4 x = lds_v4(rb[r] + (uint32_t)GQ(ch) * 16u);
                                g[ch].x = fmaf(resid[r], x.x, g[ch].x);
                                g[ch].y = fmaf(resid[r], x.y, g[ch].y);
                                g[ch].z = fmaf(resid[r], x.z, g[ch].z);
                                g[ch].w = fmaf(resid[r], x.w, g[ch].w);
                            }
                        }
                    }
                }
                __syncwarp();
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (valid[r] && lane == 0) mbar_arrive(smem_u32(&empty[sl[r]]));
            }
            slot += NG * R;
            while (slot >= (uint32_t)ns) { slot -= (uint32_t)ns; par ^= 1u; }
        }
        if (tid == 0) dbg_stamp(a, s, 1);
        // ---- phase 1: this row group's partial gradient, fixed point, straight from registers to L2 ----
        // (one buffer is enough: it is rewritten only after the weights of this step arrived,
        // i.e. after every owner finished reading it)
        {
            float* mine = part + (size_t)(c * NG + grp) * a.Fp;
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (GV(ch)) __stcg(reinterpret_cast<float4*>(mine + (size_t)GQ(ch) * 4), g[ch]);
        }
        named_bar_sync(1, NCT);
        if (tid == 0) {
            red_release_add_u32(a.counter, 1u);
            dbg_stamp(a, s, 2);
        }
        if (warp == 1 && a.loss_fix) {  // off the critical path: this step's loss
            float l = 0.f;
            for (int i = lane; i < nrows_c; i += 32) {
                const float2 v = zyb[i];
                l += row_loss_f(v.x, v.y);
            }
            l = warp_sum(l);
            if (lane == 0) red_add_u64(a.loss_fix + s, (unsigned long long)to_fix(l));
        }
        wait_counter(a.counter, (unsigned)G * (unsigned)(s + 1), kErrTimeoutPart);
        if (tid == 0) dbg_stamp(a, s, 3);
        // ---- phase 2: the owner of a slice adds the G partials of it ------------------------------
        if (c < a.n_own) {
            for (int u = tid; u < quads * ngrp; u += NCT) {
                const int oj = u % quads, ogrp = u / quads;
                const int oe = c * L + 4 * oj;  // first of my four columns
                if (oe >= F) continue;
                // each partial goes to fixed point on its own, the sum is an integer sum: the same
                // bits in any order (and the same bits as summing s64 partials would give)
                long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
                bool bad = false;   // a partial outside the fixed-point range: reported once, after the loop
                const float* src = part + oe;
                int pp = ogrp;
                for (; pp + 3 * ngrp < P; pp += 4 * ngrp) {
                    const float4 v0 = ld_vol_f32x4(src + (size_t)pp * a.Fp);
                    const float4 v1 = ld_vol_f32x4(src + (size_t)(pp + ngrp) * a.Fp);
                    const float4 v2 = ld_vol_f32x4(src + (size_t)(pp + 2 * ngrp) * a.Fp);
                    const float4 v3 = ld_vol_f32x4(src + (size_t)(pp + 3 * ngrp) * a.Fp);
                    bad = bad || fix_out_of_range(v0, a.fix_limit) || fix_out_of_range(v1, a.fix_limit) ||
                          fix_out_of_range(v2, a.fix_limit) || fix_out_of_range(v3, a.fix_limit);
                    t0 += (to_fix(v0.x) + to_fix(v1.x)) + (to_fix(v2.x) + to_fix(v3.x));
                    t1 += (to_fix(v0.y) + to_fix(v1.y)) + (to_fix(v2.y) + to_fix(v3.y));
                    t2 += (to_fix(v0.z) + to_fix(v1.z)) + (to_fix(v2.z) + to_fix(v3.z));
                    t3 += (to_fix(v0.w) + to_fix(v1.w)) + (to_fix(v2.w) + to_fix(v3.w));
                }
                for (; pp < P; pp += ngrp) {
                    const float4 v0 = ld_vol_f32x4(src + (size_t)pp * a.Fp);
                    bad = bad || fix_out_of_range(v0, a.fix_limit);
                    t0 += to_fix(v0.x); t1 += to_fix(v0.y); t2 += to_fix(v0.z); t3 += to_fix(v0.w);
                }
                if (bad) set_err(a.err, kErrRange);
                longlong2* dst = reinterpret_cast<longlong2*>(gsm + (size_t)ogrp * L + 4 * oj);
                dst[0] = make_longlong2(t0, t1);
                dst[1] = make_longlong2(t2, t3);
            }
            named_bar_sync(1, NCT);
            for (int t = tid; t < L; t += NCT) {
                const int ue = c * L + t;  // column I update as the owner
                if (ue >= F) break;
                long long tot = 0;
                for (int k = 0; k < ngrp; ++k) tot += gsm[(size_t)k * L + t];
                if (a.nranks > 1) {  // trade the slice total with the other GPUs
                    const unsigned xflag = a.xflag0 + (unsigned)s + 1u;
                    const size_t half = (size_t)(xflag & 1u) * (size_t)a.nranks * (size_t)F;
                    const uint4 pkt = make_uint4((unsigned)((unsigned long long)tot & 0xffffffffu), xflag,
                                                 (unsigned)((unsigned long long)tot >> 32), xflag);
                    for (int p = 0; p < a.nranks; ++p)
                        if (p != a.rank) st_vol_v4(a.peer_x[p] + half + (size_t)a.rank * F + ue, pkt);
                    const uint4* inbox = a.peer_x[a.rank] + half;
                    for (int p0 = 0; p0 < a.nranks; p0 += 2) {  // two peers' packets in flight at a time
                        const int p1 = p0 + 1;
                        const bool u0 = p0 != a.rank, u1 = p1 < a.nranks && p1 != a.rank;
                        uint4 q0 = make_uint4(0u, xflag, 0u, xflag), q1 = q0;
                        if (u0) q0 = ld_vol_v4(inbox + (size_t)p0 * F + ue);
                        if (u1) q1 = ld_vol_v4(inbox + (size_t)p1 * F + ue);
                        uint32_t spins = 0;
                        while (q0.y != xflag || q0.w != xflag) {
                            if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                            q0 = ld_vol_v4(inbox + (size_t)p0 * F + ue);
                        }
                        while (q1.y != xflag || q1.w != xflag) {
                            if (spin_abort(spins, a.err, kErrTimeoutPeer, kPeerSpinLimit)) break;
                            q1 = ld_vol_v4(inbox + (size_t)p1 * F + ue);
                        }
                        tot += (long long)(((unsigned long long)q0.z << 32) | (unsigned long long)q0.x);
                        tot += (long long)(((unsigned long long)q1.z << 32) | (unsigned long long)q1.x);
                    }
                }
                const double m = __ldcg(a.model + ue) - a.eta_over_b * ((double)tot * kFixInv);
                __stcg(a.model + ue, m);
                __stcg(a.w32buf + (size_t)(s & 1) * F + ue, (float)m);
            }
            named_bar_sync(1, NCT);
            if (tid == 0) {
                red_release_add_u32(a.counter2, 1u);
                dbg_stamp(a, s, 4);
            }
        }
        mb = mb_next;
    }
    if (a.n_steps > 0) {
        wait_counter(a.counter2, (unsigned)a.n_own * (unsigned)a.n_steps, kErrTimeoutPeer);
        apply_total(a.n_steps - 1);
    }
}

#undef GQ
#undef GV

}  // namespace slk
