// Latency probes used to size the single-CTA rrLU (csrc/sweep.cu): dependent FP64 chain, shared-memory load, redux.sync,
// __syncthreads with 8 / 32 warps, IEEE double division.  nvcc -arch=sm_100a -O3 --fmad=false lat.cu -o lat && ./lat
#include <cstdio>
#include <cuda_runtime.h>
__global__ void probe(long long* out, double seed, int n) {
    __shared__ double sm[1024];
    __shared__ int si[1024];
    const int tid = threadIdx.x;
    sm[tid % 1024] = seed + tid; si[tid % 1024] = (tid * 7 + 1) % 1024;
    __syncthreads();
    double a = seed, b = seed * 0.5;
    long long t0, t1;
    // 1. dependent DMUL+DADD chain
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) { a = __dmul_rn(a, b); a = __dsub_rn(a, b); }
    t1 = clock64();
    if (tid == 0) out[0] = (t1 - t0) / n;            // cycles per (DMUL + DADD)
    // 2. dependent LDS chain (pointer chase)
    int p = tid % 1024;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) p = si[p];
    t1 = clock64();
    if (tid == 0) out[1] = (t1 - t0) / n;
    // 3. redux.sync chain
    unsigned u = (unsigned)p + tid;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) u = __reduce_max_sync(0xffffffffu, u + i);
    t1 = clock64();
    if (tid == 0) out[2] = (t1 - t0) / n;
    // 4. __syncthreads
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) __syncthreads();
    t1 = clock64();
    if (tid == 0) out[3] = (t1 - t0) / n;
    // 5. IEEE division chain
    double q = seed + 3.0;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) q = (q + 1.5) / b;
    t1 = clock64();
    if (tid == 0) out[4] = (t1 - t0) / n;
    // 6. shfl chain
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) u = __shfl_xor_sync(0xffffffffu, u, 1) + 1;
    t1 = clock64();
    if (tid == 0) out[5] = (t1 - t0) / n;
    // 7. independent DMUL throughput: 8 independent chains
    double c0 = a, c1 = a + 1, c2 = a + 2, c3 = a + 3, c4 = a + 4, c5 = a + 5, c6 = a + 6, c7 = a + 7;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) { c0 = __dmul_rn(c0, b); c1 = __dmul_rn(c1, b); c2 = __dmul_rn(c2, b); c3 = __dmul_rn(c3, b); c4 = __dmul_rn(c4, b); c5 = __dmul_rn(c5, b); c6 = __dmul_rn(c6, b); c7 = __dmul_rn(c7, b); }
    t1 = clock64();
    if (tid == 0) out[6] = (t1 - t0) * 10 / n / 8;   // tenths of a cycle per DMUL per warp
    if (a + q + u + p + c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7 == 12345.678) out[7] = 1;
}
int main() {
    long long* d; cudaMalloc(&d, 64); long long h[8];
    for (int nt : {32, 256, 1024}) {
        probe<<<1, nt>>>(d, 1.000001, 2000); cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
        printf("threads %4d: dmul+dadd %lld  lds %lld  redux %lld  syncthreads %lld  ddiv %lld  shfl %lld  dmul-throughput(x0.1 cyc) %lld\n", nt, h[0], h[1], h[2], h[3], h[4], h[5], h[6]);
    }
    return 0;
}
