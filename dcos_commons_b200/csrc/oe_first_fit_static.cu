// oe_first_fit_static.cu — ordered first-fit over feasibility bit rows for callers that must not allocate.
//
// Who this is for.  The engine handle (oe_engine.cu) grows its device buffers with cudaMalloc/cudaFree whenever a
// cycle is larger than the last one.  Two kinds of scheduler hosts cannot live with that:
//   * a multi-service scheduler that runs one engine per service on the same GPU (the reference's
//     MultiServiceEventClient fans one offer batch out to every service): cudaMalloc and cudaFree synchronise the whole
//     device, so one service growing a buffer stalls the cycles of all the others in mid-flight;
//   * a host whose device memory is handed out up front by some other component (a pooled allocator that owns the whole
//     card, or an operator-set memory budget): a cudaMalloc from inside this library would fail or break the budget.
// Such callers keep the class-factored bit rows, the requirement→row map and the result column in device memory that
// THEY own, and call oe_first_fit_static(): it launches one kernel and allocates nothing, not on the device and not
// through any library.  Its only working memory is the fixed array below, which every GPU that loads this library
// carries (about 544 KiB) from load until the process exits.
//
// What it computes: the commit of S/scheduler/plan/PlanScheduler.java:56-67 — requirements in order, each takes the
// lowest-index offer whose bit is set in its row and that no earlier requirement took — i.e. the `offer_row` column
// of oe_evaluate_cycle, from the same class bits.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/oe_abi.h"

namespace {

constexpr int64_t kUsedWords = OE_STATIC_MAX_OFFERS / 32;
constexpr int64_t kScratchWords = kUsedWords + OE_STATIC_MAX_ROWS;

// Working memory of k_first_fit_static: [0, kUsedWords) = one bit per offer, "taken by an earlier requirement";
// [kUsedWords, kScratchWords) = per bit row, the first word that may still hold a free offer.
// In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a device address.
static __device__ __align__(256) uint32_t g_first_fit_scratch[kScratchWords];

// One CTA.  All threads clear the part of the scratch this call uses (it still holds the previous call's state), then
// warp 0 walks the requirements in order; its 32 lanes test 32 words of the row per step.
__global__ void __launch_bounds__(1024) k_first_fit_static(uint32_t* __restrict__ scratch, const uint32_t* __restrict__ bits,
                                                           int64_t stride, int n_rows, int64_t n_offers,
                                                           const int32_t* __restrict__ req_row, int64_t n_reqs,
                                                           int32_t* __restrict__ offer_row) {
  uint32_t* used = scratch;
  uint32_t* cursor = scratch + kUsedWords;
  const int64_t words = (n_offers + 31) / 32;
  for (int64_t k = threadIdx.x; k < words; k += blockDim.x) used[k] = 0u;
  for (int k = threadIdx.x; k < n_rows; k += blockDim.x) cursor[k] = 0u;
  __syncthreads();
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  const uint32_t tail_mask = (n_offers & 31) ? ((1u << (n_offers & 31)) - 1u) : 0xffffffffu;
  for (int64_t j = 0; j < n_reqs; ++j) {
    const int row = req_row[j];
    int32_t pick = -1;
    if (row >= 0 && row < n_rows) {
      const uint32_t* r = bits + (int64_t)row * stride;
      int64_t k = cursor[row];
      while (k < words) {
        const int64_t w = k + lane;
        uint32_t free_bits = 0u;
        if (w < words) {
          free_bits = r[w] & ~used[w];
          if (w == words - 1) free_bits &= tail_mask;
        }
        const uint32_t any = __ballot_sync(0xffffffffu, free_bits != 0u);
        if (any) {
          const int first = __ffs(any) - 1;
          const uint32_t fb = __shfl_sync(0xffffffffu, free_bits, first);
          const int b = __ffs(fb) - 1;
          pick = (int32_t)((k + first) * 32 + b);
          if (lane == first) {
            used[w] |= 1u << b;
            cursor[row] = (uint32_t)w;  // earlier words of this row hold nothing free any more, and never will again
          }
          break;
        }
        k += 32;
      }
      if (pick < 0 && lane == 0) cursor[row] = (uint32_t)words;
    }
    if (lane == 0) offer_row[j] = pick;
    __syncwarp();  // the next requirement's loads must see this one's claim
  }
}

std::mutex g_setup_mutex;
constexpr int kMaxDevices = 64;
uint32_t* g_scratch_addr[kMaxDevices];  // device address of g_first_fit_scratch on each device; null = not looked up yet

int fail(char* err, int32_t cap, int code, const char* fmt, ...) {
  if (err && cap > 0) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, (size_t)cap, fmt, ap);
    va_end(ap);
  }
  return code;
}

}  // namespace

extern "C" int32_t oe_first_fit_static(int32_t device, void* stream, const uint32_t* d_row_bits, int64_t stride_words,
                                       int32_t n_rows, int64_t n_offers, const int32_t* d_req_row, int64_t n_reqs,
                                       int32_t* d_offer_row, char* err_text, int32_t err_cap) {
  if (err_text && err_cap > 0) err_text[0] = 0;
  // sizes first: these checks need no device
  if (n_offers < 0 || n_reqs < 0 || n_rows < 0 || stride_words < 0)
    return fail(err_text, err_cap, OE_ERR_INVALID, "oe_first_fit_static: negative size");
  if (n_offers > OE_STATIC_MAX_OFFERS)
    return fail(err_text, err_cap, OE_ERR_LIMIT,
                "oe_first_fit_static: %lld offers exceed OE_STATIC_MAX_OFFERS (%d): the fixed scratch array has one bit per offer; "
                "use oe_evaluate_cycle for larger cycles",
                (long long)n_offers, (int)OE_STATIC_MAX_OFFERS);
  if (n_rows > OE_STATIC_MAX_ROWS)
    return fail(err_text, err_cap, OE_ERR_LIMIT,
                "oe_first_fit_static: %d bit rows exceed OE_STATIC_MAX_ROWS (%d): the fixed scratch array has one cursor per row",
                (int)n_rows, (int)OE_STATIC_MAX_ROWS);
  if (stride_words * 32 < n_offers)
    return fail(err_text, err_cap, OE_ERR_INVALID, "oe_first_fit_static: stride of %lld words cannot hold %lld offers",
                (long long)stride_words, (long long)n_offers);
  if (n_reqs > 0 && (!d_req_row || !d_offer_row || (n_rows > 0 && n_offers > 0 && !d_row_bits)))
    return fail(err_text, err_cap, OE_ERR_INVALID, "oe_first_fit_static: null device pointer");
  if (device < 0 || device >= kMaxDevices) return fail(err_text, err_cap, OE_ERR_INVALID, "oe_first_fit_static: device ordinal %d", (int)device);
  if (n_reqs == 0) return OE_OK;

  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e != cudaSuccess) return fail(err_text, err_cap, OE_ERR_NO_DEVICE, "oe_first_fit_static: no CUDA device (%s); there is no CPU path", cudaGetErrorString(e));
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(err_text, err_cap, OE_ERR_CUDA, "oe_first_fit_static: cudaSetDevice(%d): %s", (int)device, cudaGetErrorString(e));

  // Set-up, once per device and with that device current: turn the array's host-side name into its address on this
  // device.  This is a lookup, not an allocation.
  uint32_t* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[device]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_first_fit_scratch);
      if (e != cudaSuccess || !p) {
        cudaSetDevice(prev);
        return fail(err_text, err_cap, OE_ERR_CUDA, "oe_first_fit_static: cudaGetSymbolAddress on device %d: %s", (int)device, cudaGetErrorString(e));
      }
      g_scratch_addr[device] = static_cast<uint32_t*>(p);
    }
    scratch = g_scratch_addr[device];
  }

  k_first_fit_static<<<1, 1024, 0, (cudaStream_t)stream>>>(scratch, d_row_bits, stride_words, n_rows, n_offers, d_req_row, n_reqs, d_offer_row);
  e = cudaGetLastError();
  cudaSetDevice(prev);
  if (e != cudaSuccess) return fail(err_text, err_cap, OE_ERR_CUDA, "oe_first_fit_static: launch: %s", cudaGetErrorString(e));
  return OE_OK;
}
