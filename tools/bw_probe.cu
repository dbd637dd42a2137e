// bw_probe.cu — measures what the matrix-expansion write pattern can reach on this GPU, to size K1x's tiles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bw_probe tools/bw_probe.cu && tools/bw_probe
// Patterns (all write the same [M][stride] u32 matrix, 12.5 GB at C5 size):
//   fill          grid-stride uint4 streaming stores of a constant (write-only ceiling)
//   expand W      K1x's pattern: CTA = (W-byte column slab, block of rows); row j copies class row req_row[j]
//   bulk W        class rows of one W-byte slab staged in shared memory, one cp.async.bulk (smem -> global) per row
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void k_init(uint32_t* F, size_t n, int64_t stride) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    F[i] = (uint32_t)(i / stride) * 0x9E3779B1u + (uint32_t)(i % stride);
}

__global__ void k_fill(uint4* out, size_t n4) {
  const uint4 v = make_uint4(1, 2, 3, 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) __stcs(out + i, v);
}

template <int MODE>
__device__ __forceinline__ void store_mode(uint4* p, const uint4 v) {
  if (MODE == 0) *p = v;
  else if (MODE == 1) __stcs(p, v);
  else if (MODE == 2) __stcg(p, v);
  else __stwt(p, v);
}
template <int MODE>
__global__ void k_fill_mode(uint4* out, size_t n4) {
  const uint4 v = make_uint4(1, 2, 3, 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) store_mode<MODE>(out + i, v);
}
// block-contiguous variant: each CTA owns a contiguous chunk
template <int MODE>
__global__ void k_fill_chunk(uint4* out, size_t n4) {
  const uint4 v = make_uint4(1, 2, 3, 4);
  const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
  const size_t b = (size_t)blockIdx.x * per, e = b + per < n4 ? b + per : n4;
  for (size_t i = b + threadIdx.x; i < e; i += blockDim.x) store_mode<MODE>(out + i, v);
}
template <int MODE>
__global__ void k_expand_mode(const uint32_t* __restrict__ F, const int32_t* __restrict__ req_row, uint32_t* __restrict__ out, int64_t stride,
                              int64_t M, int rows_per_block) {
  const int64_t w4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (w4 >= stride) return;
  const int64_t j0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t j1 = min(M, j0 + rows_per_block);
  for (int64_t j = j0 + threadIdx.y; j < j1; j += blockDim.y) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)__ldg(req_row + j) * stride + w4));
    store_mode<MODE>(reinterpret_cast<uint4*>(out + j * stride + w4), a);
  }
}

// slab of `lanes` uint4 per row (lanes * 16 bytes), blockDim = (lanes, rows_in_flight)
__global__ void k_expand(const uint32_t* __restrict__ F, const int32_t* __restrict__ req_row, uint32_t* __restrict__ out, int64_t stride,
                         int64_t M, int rows_per_block) {
  const int64_t w4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (w4 >= stride) return;
  const int64_t j0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t j1 = min(M, j0 + rows_per_block);
  for (int64_t j = j0 + threadIdx.y; j < j1; j += blockDim.y) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)__ldg(req_row + j) * stride + w4));
    __stcs(reinterpret_cast<uint4*>(out + j * stride + w4), a);
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// CTA = (slab of W bytes, block of rows).  All threads stage the C class rows of the slab; then every thread issues bulk
// stores for its share of the rows.
__global__ void k_bulk(const uint32_t* __restrict__ F, int C, const int32_t* __restrict__ req_row, uint32_t* __restrict__ out, int64_t stride,
                       int64_t M, int rows_per_block, int slab_words) {
  extern __shared__ __align__(128) uint32_t sm[];
  const int64_t w0 = (int64_t)blockIdx.x * slab_words;
  const int nt = blockDim.x;
  for (int q = threadIdx.x; q < C * slab_words / 4; q += nt) {
    const int c = q / (slab_words / 4), k = q % (slab_words / 4);
    reinterpret_cast<uint4*>(sm)[q] = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)c * stride + w0) + k);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const int64_t j0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t j1 = min(M, j0 + rows_per_block);
  const uint32_t bytes = slab_words * 4;
  for (int64_t j = j0 + threadIdx.x; j < j1; j += nt) {
    const int c = __ldg(req_row + j);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + j * stride + w0), "r"(smem_u32(sm + c * slab_words)),
                 "r"(bytes)
                 : "memory");
  }
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

template <class Fn>
float time_ms(Fn fn, int reps = 5) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  fn();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(a);
    fn();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main(int argc, char** argv) {
  const int64_t N = 1000000, M = argc > 1 ? atoll(argv[1]) : 100000;
  const int C = 32;
  const int64_t stride = ((N + 31) / 32 + 127) / 128 * 128;
  const size_t bytes = (size_t)M * stride * 4;
  uint32_t *F, *out; int32_t* req;
  CK(cudaMalloc(&F, (size_t)C * stride * 4)); CK(cudaMalloc(&out, bytes)); CK(cudaMalloc(&req, M * 4));
  k_init<<<1024, 256>>>(F, (size_t)C * stride, stride);
  std::vector<int32_t> h(M);
  uint32_t s = 12345;
  for (auto& x : h) { s = s * 1664525u + 1013904223u; x = (s >> 8) % C; }
  CK(cudaMemcpy(req, h.data(), M * 4, cudaMemcpyHostToDevice));
  printf("matrix %.2f GB, stride %lld words\n", bytes / 1e9, (long long)stride);
  auto report = [&](const char* name, int w, float ms) { printf("%-10s W=%5d  %8.3f ms  %8.1f GB/s\n", name, w, ms, bytes / ms / 1e6); fflush(stdout); };

  report("memset", 0, time_ms([&] { cudaMemsetAsync(out, 0, bytes); }));
  for (int mult : {4, 8, 16}) report("fill", mult, time_ms([&] { k_fill<<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));

  for (int mult : {8, 32}) {
    report("fill/st", mult, time_ms([&] { k_fill_mode<0><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
    report("fill/cs", mult, time_ms([&] { k_fill_mode<1><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
    report("fill/cg", mult, time_ms([&] { k_fill_mode<2><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
    report("fill/wt", mult, time_ms([&] { k_fill_mode<3><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
    report("chunk/st", mult, time_ms([&] { k_fill_chunk<0><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
    report("chunk/cs", mult, time_ms([&] { k_fill_chunk<1><<<148 * mult, 256>>>((uint4*)out, bytes / 16); }));
  }
  {
    const int W = 1024, lanes = 64;
    dim3 block(lanes, 4);
    const unsigned xb = (unsigned)((stride * 4 + W - 1) / W);
    for (int target : {148 * 32, 148 * 64}) {
      int64_t yb = std::max<int64_t>(1, (target + xb - 1) / xb);
      int rpb = (int)((M + yb - 1) / yb);
      yb = (M + rpb - 1) / rpb;
      report("expand/st", target / 148, time_ms([&] { k_expand_mode<0><<<dim3(xb, (unsigned)yb), block>>>(F, req, out, stride, M, rpb); }));
      report("expand/cs", target / 148, time_ms([&] { k_expand_mode<1><<<dim3(xb, (unsigned)yb), block>>>(F, req, out, stride, M, rpb); }));
      report("expand/cg", target / 148, time_ms([&] { k_expand_mode<2><<<dim3(xb, (unsigned)yb), block>>>(F, req, out, stride, M, rpb); }));
      report("expand/wt", target / 148, time_ms([&] { k_expand_mode<3><<<dim3(xb, (unsigned)yb), block>>>(F, req, out, stride, M, rpb); }));
    }
  }
  for (int W : {1024}) {
    const int lanes = W / 16;
    dim3 block(lanes, 256 / lanes);
    const unsigned xb = (unsigned)((stride * 4 + W - 1) / W);
    for (int target : {148 * 8, 148 * 16, 148 * 32}) {
      int64_t yb = std::max<int64_t>(1, (target + xb - 1) / xb);
      int rpb = (int)((M + yb - 1) / yb);
      yb = (M + rpb - 1) / rpb;
      char nm[32]; snprintf(nm, sizeof nm, "expand/%d", target / 148);
      report(nm, W, time_ms([&] { k_expand<<<dim3(xb, (unsigned)yb), block>>>(F, req, out, stride, M, rpb); }));
    }
  }
  CK(cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (int W : {512}) {
    const int slab_words = W / 4;
    if (stride % slab_words) continue;
    const unsigned xb = (unsigned)(stride / slab_words);
    const size_t smem = (size_t)C * W;
    for (int target : {148 * 2, 148 * 4, 148 * 8}) {
      int64_t yb = std::max<int64_t>(1, (target + xb - 1) / xb);
      int rpb = (int)((M + yb - 1) / yb);
      yb = (M + rpb - 1) / rpb;
      for (int nt : {32, 128}) {
        char nm[32]; snprintf(nm, sizeof nm, "bulk/%d/%d", target / 148, nt);
        report(nm, W, time_ms([&] { k_bulk<<<dim3(xb, (unsigned)yb), nt, smem>>>(F, C, req, out, stride, M, rpb, slab_words); }));
      }
    }
  }
  // verify the last bulk run against the expansion definition on a few rows
  std::vector<uint32_t> row(stride), cls(stride);
  int bad = 0;
  for (int64_t j : {(int64_t)0, M / 2, M - 1}) {
    CK(cudaMemcpy(row.data(), out + j * stride, stride * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cls.data(), F + (int64_t)h[j] * stride, stride * 4, cudaMemcpyDeviceToHost));
    for (int64_t k = 0; k < stride; ++k) bad += row[k] != cls[k];
  }
  printf("verify: %d mismatching words\n", bad);
  return bad != 0;
}
