// Dependent-chain latency of the warp-wide primitives a lone warp's replay loop can use (cycles per link, one warp on an SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/redux_latency tools/redux_latency.cu && tools/redux_latency
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(long long* out, int* sink) {
  const int lane = threadIdx.x;
  int x = lane;
  long long t0, t1;
  const int N = 4096;
  // 1. select -> redux.min -> compare -> select (the offer-side step's chain)
  t0 = clock64();
  for (int i = 0; i < N; ++i) { const int v = ((i >> (lane & 3)) & 1) ? x : 0x7fffffff; const int mn = __reduce_min_sync(0xffffffffu, v); x = (v == mn) ? x + 32 : x; }
  t1 = clock64(); if (lane == 0) out[0] = (t1 - t0) / N;
  // 2. redux.min alone
  t0 = clock64();
  for (int i = 0; i < N; ++i) x = __reduce_min_sync(0xffffffffu, x + lane);
  t1 = clock64(); if (lane == 0) out[1] = (t1 - t0) / N;
  // 3. ballot + ffs chain
  t0 = clock64();
  for (int i = 0; i < N; ++i) { const unsigned b = __ballot_sync(0xffffffffu, (x + lane) & 1); x += __ffs(b); }
  t1 = clock64(); if (lane == 0) out[2] = (t1 - t0) / N;
  // 4. shfl chain
  t0 = clock64();
  for (int i = 0; i < N; ++i) x = __shfl_sync(0xffffffffu, x, (x + 1) & 31) + 1;
  t1 = clock64(); if (lane == 0) out[3] = (t1 - t0) / N;
  // 5. plain ALU chain (and -> neg -> and -> xor)
  unsigned f = 0xffffffffu ^ lane;
  t0 = clock64();
  for (int i = 0; i < N; ++i) { const unsigned a0 = f & (0x55555555u + i); const unsigned bit = a0 & (0u - a0); f ^= bit; f += 0x9e3779b9u; }
  t1 = clock64(); if (lane == 0) out[4] = (t1 - t0) / N;
  // 6. redux.or chain
  t0 = clock64();
  for (int i = 0; i < N; ++i) x = (int)__reduce_or_sync(0xffffffffu, (unsigned)(x + lane)) >> 1;
  t1 = clock64(); if (lane == 0) out[5] = (t1 - t0) / N;
  // 7. match_any chain
  t0 = clock64();
  for (int i = 0; i < N; ++i) x += __popc(__match_any_sync(0xffffffffu, (x + lane) & 3));
  t1 = clock64(); if (lane == 0) out[6] = (t1 - t0) / N;
  sink[lane] = x + (int)f;
}
int main() {
  long long* d; int* s; cudaMalloc(&d, 64); cudaMalloc(&s, 128);
  k<<<1, 32>>>(d, s); k<<<1, 32>>>(d, s);
  long long h[8]; cudaMemcpy(h, d, 56, cudaMemcpyDeviceToHost);
  const char* names[] = {"select+redux.min+compare+select", "redux.min", "ballot+ffs", "shfl", "alu x4", "redux.or", "match_any"};
  for (int i = 0; i < 7; ++i) printf("%-34s %lld cycles per link\n", names[i], h[i]);
  return 0;
}
