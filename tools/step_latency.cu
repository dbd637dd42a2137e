// What one step of the offer-side replay costs a lone warp, piece by piece (cycles per step, 32 unrolled steps per chunk).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/step_latency tools/step_latency.cu && tools/step_latency
#include <cstdio>
#include <cuda_runtime.h>
#define NONE 0x7fffffff
__device__ __forceinline__ int lds_s32(unsigned addr) { int v; asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
template <int VARIANT>
__global__ void k(long long* out, int* sink, int n_rows, int chunks) {
  __shared__ int ring[164 * 32];
  const int lane = threadIdx.x;
  for (int k = 0; k < 164; ++k) ring[k * 32 + lane] = (k * n_rows + lane) * 3 + 7;   // lane's queue: ascending, interleaved across lanes
  __syncwarp();
  const unsigned ring_base = (unsigned)__cvta_generic_to_shared(ring) + lane * 4u;
  int h = lane < n_rows ? ring[lane] : NONE, n1 = lane < n_rows ? ring[32 + lane] : NONE;
  unsigned mybits0 = 0x9e3779b9u * (lane + 1) | 0x11111111u;
  if (lane >= n_rows) mybits0 = 0;
  int match_acc = 0;
  const long long t0 = clock64();
  for (int c = 0; c < chunks; ++c) {
    const unsigned mybits = mybits0 ^ (c * 0x01010101u);
    unsigned ra = ring_base + 2u * 128u;
    int match = NONE;
    bool won_prev = false; int nxt_prev = NONE;
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      if (VARIANT != 2) n1 = won_prev ? nxt_prev : n1;
      const int nxt = VARIANT != 2 ? lds_s32(ra) : n1 + 96;
      const bool listed = (mybits >> e) & 1u;
      const int v = listed ? h : NONE;
      const int mn = __reduce_min_sync(0xffffffffu, v);
      const bool won = VARIANT == 3 ? (listed & (v == mn)) : (listed & (v == mn) & (mn != NONE));
      if (VARIANT != 4) match = lane == e ? mn : match;
      h = won ? n1 : h;
      if (VARIANT == 2) n1 = won ? nxt : n1;
      ra += won ? 128u : 0u;
      won_prev = won; nxt_prev = nxt;
    }
    n1 = won_prev ? nxt_prev : n1;
    match_acc += match;
    h = lane < n_rows ? (h & 0xffff) + c : NONE;   // keep the values bounded
  }
  const long long t1 = clock64();
  if (lane == 0) out[VARIANT] = (t1 - t0) / (chunks * 32);
  sink[lane] = h + n1 + match_acc;
}

// The kernel's own ring layout: lane-major, 164 words per lane (lanes 8 apart share a bank: every LDS of a step is a 4-way conflict).
// DEPTH 1: the entry behind n1 is read at the top of a step and used at the top of the next one (what k_commit_offers did);
// DEPTH 2: read two steps ahead (n1, n2 in registers).
template <int DEPTH>
__global__ void k_lane_major(long long* out, int* sink, int n_rows, int chunks) {
  __shared__ __align__(16) int ring[164 * 32];
  const int lane = threadIdx.x;
  for (int k = 0; k < 164; ++k) ring[lane * 164 + k] = (k * n_rows + lane) * 3 + 7;
  __syncwarp();
  const unsigned ring_base = (unsigned)__cvta_generic_to_shared(ring) + lane * 164u * 4u;
  int h = lane < n_rows ? ring[lane * 164] : NONE, n1 = lane < n_rows ? ring[lane * 164 + 1] : NONE, n2 = lane < n_rows ? ring[lane * 164 + 2] : NONE;
  unsigned mybits0 = 0x9e3779b9u * (lane + 1) | 0x11111111u;
  if (lane >= n_rows) mybits0 = 0;
  int match_acc = 0;
  const long long t0 = clock64();
  for (int c = 0; c < chunks; ++c) {
    const unsigned mybits = mybits0 ^ (c * 0x01010101u);
    unsigned ra = ring_base + (DEPTH == 1 ? 2u : 3u) * 4u;
    int match = NONE;
    if (DEPTH == 1) {
      bool won_prev = false; int nxt_prev = NONE;
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        n1 = won_prev ? nxt_prev : n1;
        const int nxt = lds_s32(ra);
        const bool listed = (mybits >> e) & 1u;
        const int v = listed ? h : NONE;
        const int mn = __reduce_min_sync(0xffffffffu, v);
        const bool won = listed & (v == mn) & (mn != NONE);
        match = lane == e ? mn : match;
        h = won ? n1 : h;
        ra += won ? 4u : 0u;
        won_prev = won; nxt_prev = nxt;
      }
      n1 = won_prev ? nxt_prev : n1;
    } else {
      // h, n1, n2 = entries pos, pos+1, pos+2; ra = address of entry pos+3.  A win shifts the window: h<-n1, n1<-n2, n2<-entry pos+3,
      // which was read a step (or more) ago: x1 = the entry at ra as of the previous step's read.
      int x_prev = lds_s32(ra);            // entry pos+3, read before the chunk
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        const bool listed = (mybits >> e) & 1u;
        const int v = listed ? h : NONE;
        const int mn = __reduce_min_sync(0xffffffffu, v);
        const bool won = listed & (v == mn) & (mn != NONE);
        match = lane == e ? mn : match;
        h = won ? n1 : h;
        n1 = won ? n2 : n1;
        n2 = won ? x_prev : n2;
        ra += won ? 4u : 0u;
        x_prev = lds_s32(ra);              // entry (new pos)+3: used at the earliest by the NEXT step's win, i.e. a step from now... 
      }
    }
    match_acc += match;
    h = lane < n_rows ? (h & 0xffff) + c : NONE;
  }
  const long long t1 = clock64();
  if (lane == 0) out[DEPTH] = (t1 - t0) / (chunks * 32);
  sink[lane] = h + n1 + n2 + match_acc;
}
int main() {
  long long* d; int* s; cudaMalloc(&d, 64); cudaMalloc(&s, 128);
  for (int rows : {2, 8, 32}) {
    for (int rep = 0; rep < 2; ++rep) { k<1><<<1, 32>>>(d, s, rows, 200); k<2><<<1, 32>>>(d, s, rows, 200); k<3><<<1, 32>>>(d, s, rows, 200); k<4><<<1, 32>>>(d, s, rows, 200); }
    long long h[8]; cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
    for (int rep = 0; rep < 2; ++rep) { k_lane_major<1><<<1, 32>>>(d, s, rows, 200); k_lane_major<2><<<1, 32>>>(d, s, rows, 200); }
    long long g[8]; cudaMemcpy(g, d, 64, cudaMemcpyDeviceToHost);
    printf("rows %2d: lane-major ring (4-way bank conflicts), entry read one step ahead %lld | read a window deeper %lld  cycles per step\n", rows, g[1], g[2]);
    printf("rows %2d: full step %lld | queue in registers (no LDS) %lld | without the nobody-waits test %lld | without keeping the match %lld  cycles per step\n", rows, h[1], h[2], h[3], h[4]);
  }
  return 0;
}
