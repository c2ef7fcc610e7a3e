// Diagnostic micro-benchmark (not product): the cross-CTA exchange of the persistent
// kernel in isolation — every CTA publishes Q packets, owners reduce over all CTAs and
// publish the result to everyone — to find the protocol with the shortest round.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o xbench xbench.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
namespace cg = cooperative_groups;
__device__ int g_stuck = 0;
__device__ __forceinline__ void pause(int cyc) { if (cyc > 0) { long long t = clock64(); while (clock64() - t < cyc) {} } }
#define SPIN(cond, code) { unsigned sp_ = 0; while (cond) { if (++sp_ > (1u << 22) || *(volatile int*)&g_stuck) { if (!g_stuck) g_stuck = code; break; } } }

__device__ __forceinline__ uint4 ldv(const uint4* p) {
    uint4 r;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void stv(uint4* p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float wsum(float v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// MODE 0: push/push  — partial[src][q] ; owner polls; pushes w to mailbox[dst][q]; all poll own mailbox
// MODE 1: push/pull  — owner writes w once to shared wbuf[q]; all poll wbuf
// MODE 2: like 0 but only `NPOLL` threads of a CTA poll the mailbox (others wait at bar.sync)
template <int MODE>
__global__ void __launch_bounds__(512) xchg(uint4* part, uint4* box, int Q, int iters, long long* out,
                                            int nwriters, int npoll, int delay, int first_only) {
    const int c = blockIdx.x, G = gridDim.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int qpc = (Q + G - 1) / G;
    const int q0 = min(Q, c * qpc), q1 = min(Q, (c + 1) * qpc);
    float acc = 0.f;
    __shared__ float sink[1024];
    cg::this_grid().sync();
    const long long t0 = clock64();
    for (int it = 1; it <= iters; ++it) {
        const unsigned f = (unsigned)it;
        // 1. publish my Q packets (threads < nwriters, strided)
        for (int q = tid; q < Q; q += nwriters)
            if (tid < nwriters) stv(part + (size_t)c * Q + q, make_uint4(__float_as_uint(acc + q), f, __float_as_uint(1.f), f));
        // 2. owner warps: warp w reduces packet q0+w over all sources, publishes
        if (warp < 4) {
            for (int q = q0 + warp; q < q1; q += 4) {
                float a0 = 0.f, a1 = 0.f;
                uint4 p[5];
                unsigned sp_ = 0;
                for (;;) {   // every round re-issues ALL loads together: one L2 round trip per round
                    bool ok = true;
#pragma unroll
                    for (int k = 0; k < 5; ++k) {
                        const int src = k * 32 + lane;
                        if (src < G && (k == 0 || !first_only || sp_ > 0x80000000u)) p[k] = ldv(part + (size_t)src * Q + q);
                    }
                    if (first_only && sp_ <= 0x80000000u) {
                        ok = !(lane < G) || (p[0].y == f && p[0].w == f);
                        if (__all_sync(0xffffffffu, ok)) { sp_ = 0x80000001u; continue; }  // now fetch all
                        pause(delay); ++sp_; continue;
                    }
#pragma unroll
                    for (int k = 0; k < 5; ++k) {
                        const int src = k * 32 + lane;
                        if (src < G && (p[k].y != f || p[k].w != f)) ok = false;
                    }
                    if (__all_sync(0xffffffffu, ok)) break;
                    pause(delay);
                    if ((++sp_ & 0x7fffffffu) > (1u << 20)) { g_stuck = 1; break; }
                }
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const int src = k * 32 + lane;
                    if (src < G) { a0 += __uint_as_float(p[k].x); a1 += __uint_as_float(p[k].z); }
                }
                a0 = wsum(a0); a1 = wsum(a1);
                const uint4 pkt = make_uint4(__float_as_uint(a0), f, __float_as_uint(a1), f);
                if (MODE == 1) { if (lane == 0) stv(box + q, pkt); }
                else for (int d = lane; d < G; d += 32) stv(box + (size_t)d * Q + q, pkt);
            }
        }
        // 3. everybody picks up all Q results
        const uint4* mybox = MODE == 1 ? box : box + (size_t)c * Q;
        float got = 0.f;
        if (tid < npoll) {
            for (int qb = 0; qb < Q; qb += 4 * npoll) {
                uint4 v[4];
                unsigned sp_ = 0;
                for (;;) {
                    bool ok = true;
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const int q = qb + k * npoll + tid; if (q < Q && (k == 0 || !first_only || sp_ > 0x80000000u)) v[k] = ldv(mybox + q); }
                    if (first_only && sp_ <= 0x80000000u) {
                        const int q = qb + tid;
                        if (q >= Q || (v[0].y == f && v[0].w == f)) { sp_ = 0x80000001u; continue; }
                        pause(delay); ++sp_; continue;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const int q = qb + k * npoll + tid; if (q < Q && (v[k].y != f || v[k].w != f)) ok = false; }
                    if (ok) break;
                    pause(delay);
                    if ((++sp_ & 0x7fffffffu) > (1u << 20)) { g_stuck = 2; break; }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) { const int q = qb + k * npoll + tid; if (q < Q) { got += __uint_as_float(v[k].x); sink[q & 1023] = got; } }
            }
        }
        __syncthreads();
        acc = got * 1e-9f + sink[tid & 1023] * 1e-9f;
    }
    const long long t1 = clock64();
    if (tid == 0) out[c] = t1 - t0;
}

template <int MODE> void run(const char* name, int Q, int nthreads, int nwriters, int npoll, int delay = 0, int first_only = 0) {
    int G = 148;
    uint4 *part, *box; long long* out;
    cudaMalloc(&part, sizeof(uint4) * (size_t)G * Q); cudaMalloc(&box, sizeof(uint4) * (size_t)G * Q);
    cudaMalloc(&out, 8 * G);
    cudaMemset(part, 0, sizeof(uint4) * (size_t)G * Q); cudaMemset(box, 0, sizeof(uint4) * (size_t)G * Q);
    int iters = 500;
    void* args[] = {&part, &box, &Q, &iters, &out, &nwriters, &npoll, &delay, &first_only};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)xchg<MODE>, dim3(G), dim3(nthreads), args, 0, 0);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, out, 8 * G, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < G; ++i) mx = h[i] > mx ? h[i] : mx;
    int stuck = 0; cudaMemcpyFromSymbol(&stuck, g_stuck, 4); int zero = 0; cudaMemcpyToSymbol(g_stuck, &zero, 4);
    if (stuck) printf("STUCK code %d: ", stuck);
    printf("%-10s Q=%5d threads=%3d writers=%3d pollers=%3d delay=%4d first_only=%d : %8.0f cycles/round  (%s)\n", name, Q, nthreads,
           nwriters, npoll, delay, first_only, (double)mx / iters, cudaGetErrorString(e == cudaSuccess ? cudaGetLastError() : e));
    fflush(stdout);
    cudaFree(part); cudaFree(box); cudaFree(out);
}

int main() {
    for (int Q : {5, 513}) {
        for (int fo : {0, 1})
            for (int delay : {0, 100, 300, 600}) {
                run<0>("push/push", Q, 128, 128, 128, delay, fo);
                run<1>("push/pull", Q, 128, 128, 128, delay, fo);
            }
    }
    return 0;
}
