// Diagnostic micro-benchmark (not product): one-hop bulk-reduce exchange with a cluster-level
// pre-reduction through distributed shared memory: the CS CTAs of a cluster reduce-scatter their
// vectors in DSMEM, each then adds only its 1/CS slice into the global accumulator.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/xbench3 tools/xbench3.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
namespace cg = cooperative_groups;
__device__ int g_stuck = 0;

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ unsigned long long ld_dsmem_u64(uint32_t addr) {
    unsigned long long v;
    asm volatile("ld.shared::cluster.u64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t par) {
    uint32_t ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(bar), "r"(par) : "memory");
    return ok;
}

template <int CS>
__global__ void __launch_bounds__(256) xred(unsigned long long* acc /*[3][n]*/, unsigned* counter, int n, int iters,
                                            long long* out, unsigned long long* check) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* sbuf = (unsigned long long*)smem;   // my contribution [n]
    unsigned long long* gsum = sbuf + n;                     // my slice of the cluster sum [n/CS]
    unsigned long long* gbuf = gsum + n / CS;                // received total [n]
    uint64_t* bar = (uint64_t*)(gbuf + n);                   // [0] total loaded, [1] cluster staged
    const int c = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
    uint32_t rank = 0;
    if (CS > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int sl = n / CS;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar[1])), "r"(CS));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (CS > 1) { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
    cg::this_grid().sync();
    unsigned long long got = 0;
    const long long t0 = clock64();
    for (int it = 1; it <= iters; ++it) {
        const uint32_t par = (uint32_t)(it - 1) & 1u;
        unsigned long long* dst = acc + (size_t)(it % 3) * n;
        for (int i = tid; i < n; i += blockDim.x) sbuf[i] = (unsigned long long)(c + 1) + (got & 1);
        __syncthreads();
        const unsigned long long* src = sbuf;
        uint32_t off = 0, bytes = (uint32_t)n * 8u;
        if (CS > 1) {
            if (tid < CS) {  // tell every CTA of the cluster (me included) that my vector is staged
                const uint32_t rb = mapa(s32(&bar[1]), (uint32_t)tid);
                asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(rb) : "memory");
            }
            unsigned sp = 0;
            while (!try_wait(s32(&bar[1]), par)) { if (++sp > (1u << 22)) { g_stuck = 3; break; } }
            for (int i = tid; i < sl; i += blockDim.x) {
                unsigned long long v = 0;
#pragma unroll
                for (int r = 0; r < CS; ++r) v += ld_dsmem_u64(mapa(s32(&sbuf[rank * sl + i]), (uint32_t)r));
                gsum[i] = v;
            }
            __syncthreads();
            src = gsum; off = rank * (uint32_t)sl * 8u; bytes = (uint32_t)sl * 8u;
        }
        if (tid == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;" ::"l"((char*)dst + off), "r"(s32(src)), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
            const unsigned want = (unsigned)G * (unsigned)it;
            unsigned sp = 0, v;
            do {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
                if (++sp > (1u << 22)) { g_stuck = 1; break; }
            } while ((int)(v - want) < 0);
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[0])), "r"((uint32_t)n * 8u) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(gbuf)), "l"(dst), "r"((uint32_t)n * 8u), "r"(s32(&bar[0])) : "memory");
        }
        {
            unsigned sp = 0;
            while (!try_wait(s32(&bar[0]), par)) { if (++sp > (1u << 22)) { g_stuck = 2; break; } }
        }
        got = gbuf[tid % n];
        unsigned long long* z = acc + (size_t)((it + 2) % 3) * n;
        for (int i = c * blockDim.x + tid; i < n; i += G * blockDim.x) z[i] = 0ull;
        if (it == iters && c == 0 && tid < 8) check[tid] = got;
        __syncthreads();
    }
    const long long t1 = clock64();
    if (tid == 0) out[c] = t1 - t0;
    if (CS > 1) { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
}

template <int CS> void run(int G, int n) {
    int iters = 500, threads = 256;
    unsigned long long *acc, *check; unsigned* counter; long long* out;
    cudaMalloc(&acc, 8 * 3 * (size_t)n); cudaMalloc(&counter, 4); cudaMalloc(&out, 8 * G); cudaMalloc(&check, 64);
    cudaMemset(acc, 0, 8 * 3 * (size_t)n); cudaMemset(counter, 0, 4);
    size_t smem = 8 * ((size_t)n * 2 + n / CS) + 64;
    cudaFuncSetAttribute(xred<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(G); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = 0;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
    at[1].id = cudaLaunchAttributeClusterDimension; at[1].val.clusterDim.x = CS; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = CS > 1 ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, xred<CS>, acc, counter, n, iters, out, check);
    cudaError_t e2 = cudaDeviceSynchronize();
    long long h[148] = {0}; unsigned long long ck[8] = {0};
    cudaMemcpy(h, out, 8 * G, cudaMemcpyDeviceToHost); cudaMemcpy(ck, check, 64, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < G; ++i) mx = h[i] > mx ? h[i] : mx;
    int stuck = 0; cudaMemcpyFromSymbol(&stuck, g_stuck, 4); int zero = 0; cudaMemcpyToSymbol(g_stuck, &zero, 4);
    printf("cluster=%d grid=%3d n=%5d u64: %8.0f cycles/round  stuck=%d  total[0]=%llu (expect %d..%d)  (%s / %s)\n", CS, G, n,
           (double)mx / iters, stuck, ck[0], G * (G + 1) / 2, G * (G + 1) / 2 + G, cudaGetErrorString(e), cudaGetErrorString(e2));
    fflush(stdout);
    cudaFree(acc); cudaFree(counter); cudaFree(out); cudaFree(check);
}

int main() {
    for (int n : {16, 1024}) {
        run<1>(148, n);
        run<1>(132, n);
        run<2>(148, n);
        run<4>(132, n);
        run<4>(148, n);
        run<8>(128, n);
    }
    return 0;
}
