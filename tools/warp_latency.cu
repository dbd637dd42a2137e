// warp_latency.cu — what a lone warp pays per dependent operation on this part (sizing aid for the commit kernel, which is one
// warp per component walking a dependent chain).  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o warp_latency warp_latency.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define N 4096

template <int MODE>
__global__ void k(long long* out, int* sink, int threads_active) {
  __shared__ unsigned int sm[1024];
  const int lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (i * 7 + 1) & 1023;
  __syncthreads();
  if ((threadIdx.x >> 5) != 0) { __syncthreads(); return; }
  unsigned int x = lane + 1, m = 0xffffffffu;
  const long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < (MODE == 7 ? 0 : N); ++i) {
    if (MODE == 0) { x = x * 3 + 1; }                                                      // dependent IMAD
    if (MODE == 1) { x = __shfl_sync(0xffffffffu, x, (x + 1) & 31); }                      // dependent SHFL
    if (MODE == 2) { m = __ballot_sync(0xffffffffu, (x ^ m) & 1); x += m; }                // dependent VOTE
    if (MODE == 3) { x = sm[x & 1023]; }                                                   // dependent LDS
    if (MODE == 4) { if (lane == 0) atomicOr(&sm[x & 1023], 1u); x += 1; }                 // ATOMS by one lane, no dependence
    if (MODE == 5) { const int p = __ffs(m) - 1; x = __shfl_sync(0xffffffffu, x, p & 31); m &= ~__ballot_sync(0xffffffffu, lane == (int)(x & 31)); if (!m) m = 0xffffffffu; }  // ffs -> shfl -> ballot -> mask
    if (MODE == 6) { const int p = __ffs(m) - 1; x = __shfl_sync(0xffffffffu, x, p & 31); m &= ~__ballot_sync(0xffffffffu, lane == (int)(x & 31)); if (!m) m = 0xffffffffu;
                     if (lane == 0) { atomicOr(&sm[x & 1023], 1u); sink[i & 1023] = x; } }  // + ATOMS + STG by lane 0
  }
  if (MODE == 7) {  // the union-window step: uniform free masks, row masks from shared memory, no collective on the chain
    uint32_t free0 = 0xffffffffu, free1 = 0xffffffffu;
    int my_pos = -1;
    const uint2* rm = reinterpret_cast<const uint2*>(sm);
    for (int b = 0; b < N / 32; ++b) {
      const int my_k = (lane * 5 + b) & 7;
#pragma unroll 1
      for (int i = 0; i < 32; ++i) {
        const int kk = __shfl_sync(0xffffffffu, my_k, i);
        const uint2 r = rm[kk];
        const uint32_t a0 = (r.x | 0xf0f0f0f0u) & free0, a1 = (r.y | 0x0f0f0f0fu) & free1;
        if ((a0 | a1) == 0u) { free0 = free1 = 0xffffffffu; continue; }
        const bool lo = a0 != 0u;
        const uint32_t av = lo ? a0 : a1;
        const int p = __ffs(av) - 1;
        const uint32_t bit = 1u << p;
        free0 &= lo ? ~bit : 0xffffffffu;
        free1 &= lo ? 0xffffffffu : ~bit;
        my_pos = lane == i ? p + (lo ? 0 : 32) : my_pos;
        if (free0 == 0u) { free0 = free1; free1 = 0xffffffffu; }
      }
      x += my_pos;
    }
  }
  const long long t1 = clock64();
  if (lane == 0) out[MODE] = t1 - t0;
  sink[threadIdx.x] = x + m;
  __syncthreads();
}

int main() {
  long long* d; int* sink;
  cudaMalloc(&d, 64 * 8); cudaMalloc(&sink, 4096 * 4);
  for (int threads : {32, 1024}) {
    k<0><<<1, threads>>>(d, sink, threads); k<1><<<1, threads>>>(d, sink, threads); k<2><<<1, threads>>>(d, sink, threads);
    k<3><<<1, threads>>>(d, sink, threads); k<4><<<1, threads>>>(d, sink, threads); k<5><<<1, threads>>>(d, sink, threads);
    k<6><<<1, threads>>>(d, sink, threads); k<7><<<1, threads>>>(d, sink, threads);
    long long h[8];
    cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
    printf("CTA of %4d threads, one warp working, cycles per iteration: imad %.1f  shfl %.1f  vote %.1f  lds %.1f  atoms(lane0) %.1f  ffs+shfl+vote %.1f  +atoms+stg %.1f  union-step %.1f\n",
           threads, h[0] / (double)N, h[1] / (double)N, h[2] / (double)N, h[3] / (double)N, h[4] / (double)N, h[5] / (double)N, h[6] / (double)N, h[7] / (double)N);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
