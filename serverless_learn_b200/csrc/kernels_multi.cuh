// kernels_multi.cuh — K1m: NM independent replicas interleaved on every SM (sl_train_steps_multi).
//
// One synchronous step of ONE model is a latency chain (compute -> reduce -> fence -> counter ->
// load, ~7 us at config 3) in which the SMs and HBM idle most of the time.  Several
// independent replicas that share a GPU (the reference's workers: own model, own shard, coupled only
// by gossip) give the SMs something to do meanwhile: every CTA works through "virtual steps"
//     (step 0, model 0), (step 0, model 1), ..., (step 0, model NM-1), (step 1, model 0), ...
// and the exchange of (s, m) — its own accumulators, counter, staging and landing buffers, its own
// exchange warp — runs while the CTA computes the next NM-1 virtual steps.  The total of (s, m) is
// only needed at (s+1, m).  The row stream (one TMA producer warp, one ring) never waits.
//
// Arithmetic per model is exactly lr_persist_kernel<1,R,RS,NCW,true,*>'s on the full grid (same row
// split, same reduction orders, same fixed point), so every replica ends bit-identical to
// sl_train_steps on its own.  F <= 4*32*NCW (one float4 column chunk per thread).
#pragma once
#include "kernels.cuh"

namespace slk {

template <int R, int RS, int NCW, int NM, int LPW>
__global__ void __launch_bounds__((NCW + 1 + (NM + LPW - 1) / LPW) * 32, 1) lr_interleaved_kernel(const LrArgs a) {
    static_assert(R % RS == 0, "a pass is a whole number of ring slots");
    constexpr int SPB = R / RS;  // ring slots per pass
    constexpr int NCT = NCW * 32;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = blockIdx.x, G = gridDim.x;

    unsigned char* ring = smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.off_bar);    // [ns]
    uint64_t* empty = full + a.ns;                                     // [ns]
    uint64_t* gfull = empty + a.ns;                                    // [NM] gradient of model m staged
    uint64_t* wfull = gfull + NM;                                      // [NM] total of model m landed
    float* red = reinterpret_cast<float*>(smem + a.off_red);          // [2][NCW][R]
    float2* zy = reinterpret_cast<float2*>(smem + a.off_zy);          // [2][zy_rows]
    // one s64 vector per model: first the CTA's staged gradient (source of the bulk reduction), later the
    // landed total of the same step (the reduction has read the buffer long before the total arrives, and
    // the next gradient is staged only after the total has been applied)
    long long* gout = reinterpret_cast<long long*>(smem + a.off_gout);  // [NM][Fp]
    long long* gin = gout;
    // fp64 masters: this CTA's private copy in global memory (stays in L2), [NM][F]; only the thread that
    // owns a column chunk ever touches its four doubles.  Shared memory goes to the row ring instead.
    double* msm = a.wmaster + (size_t)blockIdx.x * NM * (size_t)a.F;

    const int F = a.F, Fp = a.Fp, ns = a.ns, row_bytes = a.row_bytes;
    const uint32_t slot_bytes = (uint32_t)RS * (uint32_t)row_bytes;
    const long long lo = ((long long)c * a.B) / G;
    const long long hi = ((long long)(c + 1) * a.B) / G;
    const int nrows_c = (int)(hi - lo);             // this CTA's rows of every minibatch
    const int nsl_c = (nrows_c + RS - 1) / RS;      // ... in ring slots
    const int nsb = (nrows_c + R - 1) / R;          // ... in passes of R rows
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
        for (int m = 0; m < NM; ++m) {
            mbar_init(smem_u32(&gfull[m]), NCW);
            mbar_init(smem_u32(&wfull[m]), 1);
        }
        fence_mbar_init();
    }
    __syncthreads();

    // =========================================================================== producer
    // Lane l of the producer warp owns ring slot l (ns <= 32) and fetches slots l, l+ns, ... of the
    // CTA's slot sequence over the whole launch: issuing a slot costs one thread ~600 cycles, and
    // with back-to-back virtual steps (7 slots every ~2.4 us at config 3) one lane would be the limit.
    if (warp == NCW) {
        if (lane < ns && nrows_c > 0) {
            const uint64_t pol = policy_evict_first();
            const uint32_t Q = (uint32_t)a.n_steps * (uint32_t)NM * (uint32_t)nsl_c;  // slots of the launch
            const uint32_t eb = smem_u32(&empty[lane]), fb = smem_u32(&full[lane]);
            const uint32_t dst0 = smem_u32(ring + (size_t)lane * slot_bytes);
            uint32_t v = (uint32_t)lane / (uint32_t)nsl_c, qq = (uint32_t)lane % (uint32_t)nsl_c;
            uint32_t s_cur = 0, phase = 0;
            long long mb = mb0;
            for (uint32_t q = (uint32_t)lane; q < Q; q += (uint32_t)ns) {
                const uint32_t s = v / NM, m = v % NM;
                while (s_cur < s) { mb = wrap(mb + a.B); ++s_cur; }
                mbar_wait(eb, phase ^ 1u, a.err);
                const int rc = min(RS, nrows_c - (int)qq * RS);
                mbar_arrive_expect_tx(fb, (uint32_t)rc * (uint32_t)row_bytes);
                long long p = lo + (long long)qq * RS;
                uint32_t dst = dst0;
                int left = rc;
                while (left > 0) {  // one copy unless the rows wrap around the shard end
                    const RowRef ref = locate_row(a, mb, p, (int)m);
                    const int n = ref.room < (long long)left ? (int)ref.room : left;
                    bulk_g2s(dst, ref.x, (uint32_t)n * (uint32_t)row_bytes, fb, pol);
                    dst += (uint32_t)n * (uint32_t)row_bytes;
                    left -= n;
                    p += n;
                }
                phase ^= 1u;
                qq += (uint32_t)ns;
                while (qq >= (uint32_t)nsl_c) { qq -= (uint32_t)nsl_c; ++v; }
            }
        }
        return;
    }

    // =========================================================================== exchange (one LANE per model)
    // Lane m of the exchange warp runs model m's exchange loop on its own (the lanes diverge and
    // stay diverged; every wait is a spin that yields): one warp instead of NM keeps the CTA at
    // NCW+2 warps, i.e. 168 registers per thread for the compute warps.
    if (warp > NCW) {  // LPW models per exchange warp
        const int m = (warp - (NCW + 1)) * LPW + lane;
        if (lane < LPW && m < NM) {
            unsigned long long* const macc = a.acc + (size_t)m * 3 * (size_t)Fp;  // [3][Fp], one replica
            unsigned* const counter = a.counter + m * 8;
            const uint32_t gsrc = smem_u32(gout + (size_t)m * Fp), gdst = smem_u32(gin + (size_t)m * Fp);
            for (int s = 0; s < a.n_steps; ++s) {
                unsigned long long* buf = macc + (size_t)(s % 3) * Fp;
                mbar_wait(smem_u32(&gfull[m]), (uint32_t)s & 1u, a.err);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                bulk_reduce_add_u64(buf, gsrc, (uint32_t)Fp * 8u);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
                asm volatile("fence.proxy.async;" ::: "memory");
                red_release_add_u32(counter, 1u);
                const unsigned want = (unsigned)G * (unsigned)(s + 1);
                uint32_t spins = 0;
                while ((int)(ld_relaxed_u32(counter) - want) < 0) {
                    if (spin_abort(spins, a.err, kErrTimeoutPart)) break;
                }
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                asm volatile("fence.proxy.async;" ::: "memory");
                const uint32_t wb = smem_u32(&wfull[m]);
                mbar_arrive_expect_tx(wb, (uint32_t)Fp * 8u);
                bulk_g2s_plain(gdst, buf, (uint32_t)Fp * 8u, wb);
                // everybody has contributed to step s of this model, hence finished reading the buffer
                // of step s-1: clear this CTA's share of it for step s+2
                unsigned long long* z = macc + (size_t)((s + 2) % 3) * Fp;
                for (int i = c; i < Fp; i += G) z[i] = 0ull;
            }
        }
        return;
    }

    // =========================================================================== compute
    const bool gv = tid * 4 < F;                 // this thread's float4 column chunk exists
    const int ridx = row_of_lane<R>(lane);       // the row whose warp total this lane ends up with
    const bool canon = lane == lane_of_row<R>(ridx);
    float4 w[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) {
        w[m] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gv) {
            const double* src = a.pmodel[m] + tid * 4;
            double* mm = msm + (size_t)m * F + tid * 4;
            __stcg(reinterpret_cast<double2*>(mm), make_double2(src[0], src[1]));
            __stcg(reinterpret_cast<double2*>(mm) + 1, make_double2(src[2], src[3]));
            w[m] = make_float4((float)src[0], (float)src[1], (float)src[2], (float)src[3]);
        }
    }

    uint32_t base_slot = 0, base_phase = 0;  // ring position of the current virtual step's slot 0
    int red_buf = 0;
    long long mb = mb0;
    auto load_y = [&](long long mbs, int m, int j) -> float {
        const int i = j * R + ridx;
        return (j < nsb && i < nrows_c) ? __ldg(locate_row(a, mbs, lo + i, m).y) : 0.f;
    };
    float y_next = load_y(mb, 0, 0);

    for (int s = 0; s < a.n_steps; ++s) {
        const long long mb_next = wrap(mb + a.B);
#pragma unroll
        for (int m = 0; m < NM; ++m) {
            // the gradient total of (s-1, m) has had NM-1 virtual steps to arrive
            auto apply_total = [&]() {
                double2 m01 = make_double2(0.0, 0.0), m23 = m01;
                double* mm = msm + (size_t)m * F + tid * 4;
                if (gv) {  // the master's L2 round trip overlaps the wait for the total
                    m01 = __ldcg(reinterpret_cast<const double2*>(mm));
                    m23 = __ldcg(reinterpret_cast<const double2*>(mm) + 1);
                }
                mbar_wait(smem_u32(&wfull[m]), (uint32_t)(s - 1) & 1u, a.err);
                if (gv) {
                    const longlong2* p = reinterpret_cast<const longlong2*>(gin + (size_t)m * Fp + tid * 4);
                    const longlong2 u = p[0], v = p[1];
                    const double v0 = m01.x - a.eta_over_b * ((double)u.x * kFixInv);
                    const double v1 = m01.y - a.eta_over_b * ((double)u.y * kFixInv);
                    const double v2 = m23.x - a.eta_over_b * ((double)v.x * kFixInv);
                    const double v3 = m23.y - a.eta_over_b * ((double)v.y * kFixInv);
                    __stcg(reinterpret_cast<double2*>(mm), make_double2(v0, v1));
                    __stcg(reinterpret_cast<double2*>(mm) + 1, make_double2(v2, v3));
                    w[m] = make_float4((float)v0, (float)v1, (float)v2, (float)v3);
                }
            };
            const int vstep = s * NM + m;
            float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
            float2* zyb = zy + (size_t)(vstep & 1) * a.zy_rows;
            if (nsb == 0 && s > 0) apply_total();  // a CTA without rows still follows every exchange
            for (int sb = 0; sb < nsb; ++sb) {
                const int rc = min(R, nrows_c - sb * R);
                const float yv = y_next;
                if (sb + 1 < nsb) y_next = load_y(mb, m, sb + 1);
                else if (m + 1 < NM) y_next = load_y(mb, m + 1, 0);
                else y_next = (s + 1 < a.n_steps) ? load_y(mb_next, 0, 0) : 0.f;
                float4 x[R];
                {
                    // wait for all slots of the pass, issue every load, and only then hand the slots back:
                    // one exposed latency per pass instead of one per slot
                    uint32_t sl = base_slot + (uint32_t)(sb * SPB);
                    uint32_t k = sl / (uint32_t)ns;
                    uint32_t slot = sl - k * (uint32_t)ns;
                    uint32_t par = (base_phase ^ k) & 1u;
                    uint32_t slot_of[SPB];
#pragma unroll
                    for (int q = 0; q < SPB; ++q) {
                        slot_of[q] = slot;
                        if (q * RS < rc) mbar_wait(smem_u32(&full[slot]), par, a.err);  // warp-uniform
                        if (++slot == (uint32_t)ns) { slot = 0; par ^= 1u; }
                    }
#pragma unroll
                    for (int q = 0; q < SPB; ++q) {
                        const uint32_t sbase = smem_u32(ring + (size_t)slot_of[q] * slot_bytes);
#pragma unroll
                        for (int rr = 0; rr < RS; ++rr) {
                            const int r = q * RS + rr;
                            x[r] = (gv && r < rc) ? lds_v4(sbase + (uint32_t)rr * (uint32_t)row_bytes + (uint32_t)tid * 16u)
                                                  : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    }
                    __syncwarp();
                    if (lane == 0) {
#pragma unroll
                        for (int q = 0; q < SPB; ++q)
                            if (q * RS < rc) mbar_arrive(smem_u32(&empty[slot_of[q]]));  // rows are in registers
                    }
                }
                if (sb == 0 && s > 0) apply_total();
                float d[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float acc = 0.f;
                    acc = fmaf(x[r].x, w[m].x, acc);
                    acc = fmaf(x[r].y, w[m].y, acc);
                    acc = fmaf(x[r].z, w[m].z, acc);
                    acc = fmaf(x[r].w, w[m].w, acc);
                    d[r] = acc;
                }
                MultiReduce<R, 16>::run(d, lane);
                float z = d[0];
                {
                    float* rb = red + red_buf * (NCW * R);
                    if (canon) rb[warp * R + ridx] = z;
                    named_bar_sync(2, NCT);
                    z = 0.f;
                    for (int k = 0; k < NCW; ++k) z += rb[k * R + ridx];
                    red_buf ^= 1;
                }
                float resid = 0.f;
                if (ridx < rc) {
                    resid = sigmoid_f(z) - yv;
                    if (canon && warp == 0 && a.loss_fix) zyb[sb * R + ridx] = make_float2(z, yv);
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float rr = __shfl_sync(0xffffffffu, resid, lane_of_row<R>(r));
                    g.x = fmaf(rr, x[r].x, g.x);
                    g.y = fmaf(rr, x[r].y, g.y);
                    g.z = fmaf(rr, x[r].z, g.z);
                    g.w = fmaf(rr, x[r].w, g.w);
                }
            }
            {   // advance the ring cursor past this virtual step's slots
                const uint32_t sl = base_slot + (uint32_t)nsl_c;
                const uint32_t k = sl / (uint32_t)ns;
                base_slot = sl - k * (uint32_t)ns;
                base_phase ^= (k & 1u);
            }
            // hand the gradient of (s, m) to model m's exchange warp
            if (gv) {
                longlong2* p = reinterpret_cast<longlong2*>(gout + (size_t)m * Fp + tid * 4);
                check_fix_range(g, a.fix_limit, a.err);
                p[0] = make_longlong2(to_fix(g.x), to_fix(g.y));
                p[1] = make_longlong2(to_fix(g.z), to_fix(g.w));
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&gfull[m]));
            if (warp == 0 && a.loss_fix) {  // this virtual step's loss, off the critical path
                float l = 0.f;
                for (int i = lane; i < nrows_c; i += 32) {
                    const float2 v = zyb[i];
                    l += row_loss_f(v.x, v.y);
                }
                l = warp_sum(l);
                if (lane == 0) red_add_u64(a.loss_fix + (size_t)m * a.loss_stride + s, (unsigned long long)to_fix(l));
            }
        }
        mb = mb_next;
    }

    // the last totals, then the masters to the replicas' models (all CTAs hold identical copies: CTA 0 writes)
    if (a.n_steps > 0) {
        const int s = a.n_steps;
#pragma unroll
        for (int m = 0; m < NM; ++m) {
            mbar_wait(smem_u32(&wfull[m]), (uint32_t)(s - 1) & 1u, a.err);
            if (gv && c == 0) {
                const longlong2* p = reinterpret_cast<const longlong2*>(gin + (size_t)m * Fp + tid * 4);
                const longlong2 u = p[0], v = p[1];
                const double* mm = msm + (size_t)m * F + tid * 4;
                const double2 m01 = __ldcg(reinterpret_cast<const double2*>(mm));
                const double2 m23 = __ldcg(reinterpret_cast<const double2*>(mm) + 1);
                double* dst = a.pmodel[m] + tid * 4;
                dst[0] = m01.x - a.eta_over_b * ((double)u.x * kFixInv);
                dst[1] = m01.y - a.eta_over_b * ((double)u.y * kFixInv);
                dst[2] = m23.x - a.eta_over_b * ((double)v.x * kFixInv);
                dst[3] = m23.y - a.eta_over_b * ((double)v.y * kFixInv);
            }
        }
    }
}

}  // namespace slk
