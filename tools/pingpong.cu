// Micro-benchmark (diagnostic, not product): flag ping-pong latency between two CTAs
// through L2 with different load/store flavours.  nvcc -arch=sm_100a -o pingpong pingpong.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

template <int MODE> __device__ __forceinline__ void st16(uint4* p, uint4 v) {
    if (MODE == 0) asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 1) asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 2) asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 3) asm volatile("st.release.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int MODE> __device__ __forceinline__ uint4 ld16(const uint4* p) {
    uint4 r;
    if (MODE == 0) asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    if (MODE == 1) asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    if (MODE == 2) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    if (MODE == 3) asm volatile("ld.acquire.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}

// block 0 and block `peer` ping-pong; nthreads threads each poll their own packet
template <int MODE>
__global__ void pingpong(uint4* a, uint4* b, int iters, int peer, long long* out) {
    if (blockIdx.x != 0 && blockIdx.x != peer) return;
    const bool first = blockIdx.x == 0;
    uint4* mine = (first ? a : b) + threadIdx.x;
    uint4* theirs = (first ? b : a) + threadIdx.x;
    long long t0 = clock64();
    for (int i = 1; i <= iters; ++i) {
        const unsigned f = (unsigned)i;
        if (first) {
            st16<MODE>(theirs, make_uint4(i, f, i, f));
            uint4 v;
            do { v = ld16<MODE>(mine); } while (v.y != f || v.w != f);
        } else {
            uint4 v;
            do { v = ld16<MODE>(mine); } while (v.y != f || v.w != f);
            st16<MODE>(theirs, make_uint4(i, f, i, f));
        }
    }
    long long t1 = clock64();
    if (first && threadIdx.x == 0) *out = (t1 - t0);
}

// one-to-all: block 0 writes a flag, all other blocks poll it and reply into their slot;
// block 0 waits for all replies (the shape of the exchange: gather + broadcast)
template <int MODE>
__global__ void star(uint4* down, uint4* up, int iters, long long* out) {
    const int G = gridDim.x;
    long long t0 = clock64();
    for (int i = 1; i <= iters; ++i) {
        const unsigned f = (unsigned)i;
        if (blockIdx.x == 0) {
            for (int d = threadIdx.x; d < G; d += blockDim.x) st16<MODE>(down + d, make_uint4(i, f, i, f));
            for (int s = threadIdx.x; s < G; s += blockDim.x) {
                if (s == 0) continue;
                uint4 v;
                do { v = ld16<MODE>(up + s); } while (v.y != f || v.w != f);
            }
            __syncthreads();
        } else {
            if (threadIdx.x == 0) {
                uint4 v;
                do { v = ld16<MODE>(down + blockIdx.x); } while (v.y != f || v.w != f);
                st16<MODE>(up + blockIdx.x, make_uint4(i, f, i, f));
            }
        }
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = (t1 - t0);
}

template <int MODE> void run(const char* name) {
    uint4 *a, *b; long long* out; long long h;
    cudaMalloc(&a, 1 << 20); cudaMalloc(&b, 1 << 20); cudaMalloc(&out, 8);
    const int iters = 2000;
    for (int nthr : {1, 32, 256}) {
        for (int peer : {1, 2, 75, 147}) {
            cudaMemset(a, 0, 1 << 20); cudaMemset(b, 0, 1 << 20);
            pingpong<MODE><<<148, nthr>>>(a, b, iters, peer, out);
            cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
            printf("%-14s pingpong threads=%3d peer=%3d  round trip %7.0f cycles\n", name, nthr, peer, (double)h / iters);
        }
    }
    cudaMemset(a, 0, 1 << 20); cudaMemset(b, 0, 1 << 20);
    star<MODE><<<148, 160>>>(a, b, iters, out);
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaGetLastError();
    printf("%-14s star 148 blocks: %7.0f cycles per gather+broadcast round (%s)\n", name, (double)h / iters, cudaGetErrorString(e));
    cudaFree(a); cudaFree(b); cudaFree(out);
}

int main() {
    run<0>("volatile");
    run<1>("relaxed.gpu");
    run<2>("cg");
    run<3>("rel/acq.gpu");
    return 0;
}
