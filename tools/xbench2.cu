// Diagnostic micro-benchmark (not product): one-hop exchange through TMA bulk reductions.
// Every CTA adds its vector (n u64 fixed-point values) into a global accumulator with one
// cp.reduce.async.bulk, bumps a counter, waits for all CTAs, and TMA-loads the total.
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
namespace cg = cooperative_groups;
__device__ int g_stuck = 0;

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(512) xred(unsigned long long* acc /*[3][n]*/, unsigned* counter, int n,
                                            int iters, long long* out, unsigned long long* check, int opt, int nchunk, int K) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* sbuf = (unsigned long long*)smem;            // my contribution
    unsigned long long* gbuf = sbuf + n;                              // received total
    uint64_t* bar = (uint64_t*)(gbuf + (size_t)n * K);
    const int c = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
    const uint32_t bytes = (uint32_t)n * 8u;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cg::this_grid().sync();
    unsigned long long got = 0;
    long long ta = 0, tb = 0, tc = 0, td = 0, te = 0, tf = 0, tmark = 0;
    const long long t0 = clock64();
    for (int it = 1; it <= iters; ++it) {
        unsigned long long* dstall = acc + (size_t)(it % 3) * n * K;
        unsigned long long* dst = dstall + (size_t)(c % K) * n;
        for (int i = tid; i < n; i += blockDim.x) sbuf[i] = (unsigned long long)(c + 1) + (got & 1);
        __syncthreads();
        tmark = clock64();
        if (tid < nchunk) {
            const uint32_t cb = bytes / (uint32_t)nchunk;   // chunked issue: several TMA ops in flight
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;" ::"l"((char*)dst + (size_t)tid * cb), "r"(s32(sbuf) + tid * cb), "r"(cb) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            if (tid == 0) { long long t = clock64(); ta += t - tmark; tmark = t; }
            if (opt & 1) {
                asm volatile("fence.proxy.async;" ::: "memory");
                asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
            } else {
                asm volatile("fence.proxy.async;" ::: "memory");
                __threadfence();
                atomicAdd(counter, 1u);
            }
        }
        if (tid == 0) {
            { long long t = clock64(); tb += t - tmark; tmark = t; }
            const unsigned want = (unsigned)G * (unsigned)it * (unsigned)nchunk;
            unsigned sp = 0, v;
            do {
                if (opt & 2) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
                else asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
                if (++sp > (1u << 22)) { g_stuck = 1; break; }
            } while ((int)(v - want) < 0);
            { long long t = clock64(); tc += t - tmark; tmark = t; }
            if (opt & 2) asm volatile("fence.acq_rel.gpu;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes * K) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(gbuf)), "l"(dstall), "r"(bytes * K), "r"(s32(bar)) : "memory");
        }
        {   // everyone waits for the total to land in shared memory
            uint32_t ok = 0, sp = 0;
            const uint32_t par = (uint32_t)(it - 1) & 1u;
            while (!ok) {
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(s32(bar)), "r"(par) : "memory");
                if (++sp > (1u << 22)) { g_stuck = 2; break; }
            }
        }
        if (tid == 0) { long long t = clock64(); td += t - tmark; tmark = t; }
        got = 0; for (int k = 0; k < K; ++k) got += gbuf[(size_t)k * n + tid % n];
        // zero my slice of the buffer used two rounds from now (last read in round it-1)
        unsigned long long* z = acc + (size_t)((it + 2) % 3) * n * K;
        for (int i = c * blockDim.x + tid; i < n * K; i += G * blockDim.x) z[i] = 0ull;
        if (it == iters && c == 0 && tid < 8) check[tid] = got;
        __syncthreads();
    }
    const long long t1 = clock64();
    if (tid == 0) out[c] = t1 - t0;
    if (tid == 0 && (c == 0 || c == 77)) printf("  cta %3d: reduce+wait %lld  signal %lld  poll %lld  load %lld (cycles avg)\n", c, ta / iters, tb / iters, tc / iters, td / iters);
}

int main() {
    for (int n : {16, 1024})
      for (int opt : {1})
        for (int nchunk : {1}) for (int K : {1, 2, 4, 8}) {
        int G = 148, iters = 500, threads = 256;
        unsigned long long *acc, *check; unsigned* counter; long long* out;
        cudaMalloc(&acc, 8 * 3 * (size_t)n * K); cudaMalloc(&counter, 4); cudaMalloc(&out, 8 * G); cudaMalloc(&check, 64);
        cudaMemset(acc, 0, 8 * 3 * (size_t)n * K); cudaMemset(counter, 0, 4);
        size_t smem = 8 * (size_t)n * (K + 1) + 64;
        cudaFuncSetAttribute(xred, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        void* args[] = {&acc, &counter, &n, &iters, &out, &check, &opt, &nchunk, &K};
        cudaError_t e = cudaLaunchCooperativeKernel((void*)xred, dim3(G), dim3(threads), args, smem, 0);
        cudaDeviceSynchronize();
        long long h[148]; unsigned long long ck[8];
        cudaMemcpy(h, out, 8 * G, cudaMemcpyDeviceToHost); cudaMemcpy(ck, check, 64, cudaMemcpyDeviceToHost);
        long long mx = 0; for (int i = 0; i < G; ++i) mx = h[i] > mx ? h[i] : mx;
        int stuck = 0; cudaMemcpyFromSymbol(&stuck, g_stuck, 4); int zero = 0; cudaMemcpyToSymbol(g_stuck, &zero, 4);
        printf("bulk-reduce n=%6d u64 K=%d opt=%d nchunk=%d: %8.0f cycles/round  stuck=%d  total[0]=%llu (expect %d..%d)  (%s)\n", n,
               K, opt, nchunk, (double)mx / iters, stuck, ck[0], 148 * 149 / 2, 148 * 149 / 2 + 148,
               cudaGetErrorString(e == cudaSuccess ? cudaGetLastError() : e));
        fflush(stdout);
        cudaFree(acc); cudaFree(counter); cudaFree(out); cudaFree(check);
    }
    return 0;
}
