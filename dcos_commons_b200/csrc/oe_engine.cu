// oe_engine.cu — B200 (sm_100a) offer-evaluation engine: kernels + the C ABI of include/oe_abi.h.
//
// Kernels (all integer / fp64 compare-and-mask work, HBM- or latency-bound; no tensor cores by design):
//   K0  k_count_tasks / k_rr_summary    task table → per-key tallies that the placement rules read
//   K0c k_rule_tables                   zone/region-only rule verdicts per (row, zone, region), looked up by K1
//   K1  k_class_feasibility             offers × requirement classes → class-factored feasibility bits; one launch per landed
//                                       offer chunk while the table is arriving; the PEERS variant also stores its words into
//                                       every shard's exchange region over NVLink (k_peer_signal / k_peer_wait order the readers)
//   K1x k_expand_matrix                 class bits → full offers × requirements bit matrix (the HBM-bound pass)
//   K3  k_tile_popc / k_row_offsets / k_compact_rows (+ k_label_components, k_union_bits, k_union_masks)
//                                       class bits → one ascending offer list per component with row-membership masks
//   K3d k_row_overlap                   which bit rows share an offer (components that commit independently)
//   K2a k_sort_hist / k_sort_scatter    requirements grouped by component, plan order kept
//   K2  k_commit_union (k_commit)       ordered greedy first-fit: one warp per component replays its requirements
//   K2c k_count_unplaced / k_collect_unplaced / k_merge_unplaced   commit in two passes while the offer table is arriving
//   K2b k_materialize                   winners → chosen dynamic ports / MOUNT slots
//   K4  k_apply_launches                multi-cycle replay: placed pods → task rows
//   loads: k_validate_* (id and range checks), k_fetch_host (small tables read from pinned host memory), k_scatter_rows
//
// There is no CPU fallback: every entry point needs a CUDA device and fails with an oe_status otherwise.
#include <cuda_pipeline.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <climits>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <limits>
#include <string>
#include <unordered_map>
#include <vector>

#include "oe_device.cuh"

using namespace oe;

// ---------------------------------------------------------------------------------------------------------------------
// Kernels
// ---------------------------------------------------------------------------------------------------------------------

// Flags ids outside their dictionary so that later gathers cannot fault.  err[0] receives a non-zero code.
__global__ void k_validate_ids(const int32_t* __restrict__ col, int64_t n, int32_t lo, int32_t hi, int code, int* err) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int32_t v = col[i];
    if (v < lo || v >= hi) atomicMax(err, code);
  }
}

// Which requirements re-deploy a pod that is already running (a stored task of the same pod type and index,
// S/offer/TaskUtils.java:136-145), and which task rows are that pod's?  Usually none, sometimes thousands (a rolling restart).
// The device answers (OE_CFG_ASYNC_LOADS): the requirements go into a hash table by pod, every deployed task looks its pod up
// and emits one record per requirement found.  The host reads only those records: nothing per requirement or per task.
#define OE_POD_EMPTY 0xffffffffffffffffull
__device__ __forceinline__ uint64_t pod_key(int32_t type, int32_t index) { return ((uint64_t)(uint32_t)type << 32) | (uint32_t)index; }
__device__ __forceinline__ uint32_t pod_slot(uint64_t key, uint32_t mask) { return (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & mask; }
// The requirements whose class reads tallies, hashed by pod: keys[slot] = (type, index), head[slot] / next[j] chain the requirements
// of that pod (one, unless the caller lists a pod twice).
struct PodHash {
  unsigned long long* keys;
  int32_t* head;
  int32_t* next;
  uint32_t mask;
};
__global__ void k_pod_hash_insert(const int32_t* __restrict__ cls, const int32_t* __restrict__ type, const int32_t* __restrict__ index, int64_t M, int C,
                                  const uint8_t* __restrict__ sensitive, PodHash hs) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const int32_t c = cls[j];
  if (c < 0 || c >= C || !sensitive[c]) return;   // (an id outside the dictionary is reported by the validation kernels)
  const uint64_t key = pod_key(type[j], index[j]);
  if (key == OE_POD_EMPTY) return;
  for (uint32_t q = pod_slot(key, hs.mask);; q = (q + 1) & hs.mask) {
    const unsigned long long old = atomicCAS(hs.keys + q, OE_POD_EMPTY, (unsigned long long)key);
    if (old == OE_POD_EMPTY || old == key) { hs.next[j] = atomicExch(hs.head + q, (int32_t)j); return; }
  }
}
// One record (requirement, task row, class of the requirement, 0) per deployed task of a pod that a requirement re-deploys.
// count[0] = records found (those beyond `cap` are not stored: the host then searches by itself).
__global__ void k_match_tasks(const int32_t* __restrict__ type, const int32_t* __restrict__ index, int64_t T, PodHash hs, const int32_t* __restrict__ cls,
                              int4* __restrict__ pairs, int cap, int* __restrict__ count) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= T) return;
  const int32_t ty = type[r], ix = index[r];
  if (ty < 0 || ix == OE_NO_INDEX) return;
  const uint64_t key = pod_key(ty, ix);
  for (uint32_t q = pod_slot(key, hs.mask);; q = (q + 1) & hs.mask) {
    const unsigned long long e = hs.keys[q];
    if (e == OE_POD_EMPTY) return;
    if (e != key) continue;
    for (int32_t j = hs.head[q]; j >= 0; j = hs.next[j]) {
      const int at = atomicAdd(count, 1);
      if (at < cap) pairs[at] = make_int4(j, (int32_t)r, cls[j], 0);
    }
    return;
  }
}
// req_row[twin[v]] = first_row + v: the requirements that re-deploy a pod read a bit row of their own
__global__ void k_set_variant_rows(int32_t* __restrict__ req_row, const int32_t* __restrict__ twin, int V, int32_t first_row) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < V) req_row[twin[v]] = first_row + v;
}
// merged numbering (merge_equal_variants): rows below C stay, variant v reads bit row row_of_variant[v]
__global__ void k_merged_rows(const int32_t* __restrict__ req_row, int64_t M, int C, const int32_t* __restrict__ row_of_variant, int32_t* __restrict__ out) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const int32_t r = req_row[j];
  out[j] = r < C ? r : row_of_variant[r - C];
}
// a CSR offset column: begin[0] == 0 and ascending
__global__ void k_validate_offsets(const int32_t* __restrict__ begin, int64_t n, int code, int* err) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j > n) return;
  if (j == 0 ? begin[0] != 0 : begin[j] < begin[j - 1]) atomicMax(err, code);
}

// The `ports` ranges of one (offer, role slot) must be ascending and disjoint (the host merges them the way antlr's
// IntervalSet does, S/offer/RangeUtils.java:23-47): the folded port test counts offered ports by range length.
// oe_update_offers: dst[slot][rows[k]] = src[slot][k] for a slot-major column (4- or 8-byte elements)
template <class T>
__global__ void k_scatter_rows(T* __restrict__ dst, int64_t N, const T* __restrict__ src, int64_t n, int slots, const int64_t* __restrict__ rows) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * slots) return;
  const int64_t slot = q / n, k = q - slot * n;
  dst[slot * N + rows[k]] = src[q];
}

// oe_load_offers_packed: rows [r0, r1) of a slot-major column widened on the device from what crossed the bus.  Ids: unsigned
// 8 / 16 bit, all-ones = -1.  Presence masks: zero-extended.  Scalars: fp32 -> fp64 (exact; the encoder sent fp32 only because
// every value survives the round trip).
template <class TS, bool ALL_ONES_IS_ABSENT>
__global__ void k_widen_ids(int32_t* __restrict__ dst, const TS* __restrict__ src, int slots, int64_t N, int64_t r0, int64_t r1) {
  const int64_t n = r1 - r0, q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * slots) return;
  const int64_t slot = q / n, i = r0 + (q - slot * n);
  const TS v = src[slot * N + i];
  dst[slot * N + i] = (ALL_ONES_IS_ABSENT && v == (TS) ~(TS)0) ? -1 : (int32_t)v;
}
__global__ void k_widen_f32(double* __restrict__ dst, const float* __restrict__ src, int slots, int64_t N, int64_t r0, int64_t r1) {
  const int64_t n = r1 - r0, q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * slots) return;
  const int64_t slot = q / n, i = r0 + (q - slot * n);
  dst[slot * N + i] = (double)src[slot * N + i];
}
__global__ void k_fill_i32(int32_t* __restrict__ dst, int32_t v, int64_t r0, int64_t r1) {
  const int64_t i = r0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < r1) dst[i] = v;
}
// ... and the `ports` column rebuilt from the table's distinct range lists: port[s][r][i] = tuples[id[s][i]][r]; the id is kept
// (as int32) for the feasibility kernel's per-list lookup.  An id outside the list dictionary raises code 20.
template <class TS>
__global__ void k_expand_tuples(uint32_t* __restrict__ port, int32_t* __restrict__ id_out, const TS* __restrict__ ids, const uint32_t* __restrict__ tuples,
                                int n_tuples, int S, int R, int64_t N, int64_t r0, int64_t r1, int* err) {
  const int64_t n = r1 - r0, q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * S) return;
  const int64_t s = q / n, i = r0 + (q - s * n);
  int64_t t = (int64_t)ids[s * N + i];
  if (t >= n_tuples) { atomicMax(err, 20); t = -1; }
  id_out[s * N + i] = (int32_t)t;
  for (int r = 0; r < R; ++r) port[(s * R + r) * N + i] = t >= 0 ? __ldg(tuples + t * R + r) : OE_PORT_EMPTY;
}

struct OfferLimits { int32_t n_agents, n_hostnames, n_zones, n_regions, n_attrs, n_rr_values; };
// oe_load_offers_packed, one launch per landed chunk: everything that came narrow is widened into the device columns, the
// columns that did not travel are filled in (hostname = agent, one allocation role), the `ports` column is rebuilt from the
// distinct range lists, and the rows are checked like k_validate_offer_rows checks them (same codes; 20 = list id out of range).
struct WireCols {
  const void *scalar, *present, *mprof, *zone, *region, *attr, *rr, *exec, *tuple_id;   // staging (null: the column arrived at full width, in place)
  uint8_t present_b, mprof_b, zone_b, region_b, attr_b, rr_b, exec_b, tuple_b;          // bytes per element of each staged column
  uint8_t host_is_agent, alloc_is_const;
  int32_t alloc_const;
  const uint32_t* tuples;
  int32_t n_tuples;
};
struct OfferDst {
  double* scalar; uint32_t* present; uint32_t* port; int32_t *mprof, *alloc, *host, *zone, *region, *attr, *rr, *exec, *ptuple;
};
__device__ __forceinline__ int32_t wire_id(const void* col, int bytes, int64_t idx) {
  if (bytes == 1) { const uint8_t v = static_cast<const uint8_t*>(col)[idx]; return v == 0xffu ? -1 : (int32_t)v; }
  if (bytes == 2) { const uint16_t v = static_cast<const uint16_t*>(col)[idx]; return v == 0xffffu ? -1 : (int32_t)v; }
  return static_cast<const int32_t*>(col)[idx];
}
__global__ void __launch_bounds__(256) k_unpack_offer_rows(WireCols w, OfferDst d, oe::DevOffers o, OfferLimits lim, int64_t r0, int64_t r1, int* err) {
  const int64_t i = r0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= r1) return;
  const int64_t N = o.N;
  int code = 0;
  auto chk = [&](int32_t v, int32_t lo, int32_t hi, int c) { if (v < lo || v >= hi) code = max(code, c); };
  if (w.scalar) for (int q = 0; q < o.S * o.K; ++q) d.scalar[q * N + i] = (double)static_cast<const float*>(w.scalar)[q * N + i];
  if (w.present) d.present[i] = w.present_b == 1 ? (uint32_t)static_cast<const uint8_t*>(w.present)[i] : (uint32_t)static_cast<const uint16_t*>(w.present)[i];
  const int32_t agent = o.agent[i];
  chk(agent, 0, max(lim.n_agents, 1), 11);
  const int32_t host = w.host_is_agent ? agent : o.host[i];
  if (w.host_is_agent) d.host[i] = host;
  chk(host, 0, max(lim.n_hostnames, 1), 12);
  if (w.alloc_is_const) d.alloc[i] = w.alloc_const;
  const int32_t zone = w.zone ? wire_id(w.zone, w.zone_b, i) : o.zone[i];
  if (w.zone) d.zone[i] = zone;
  chk(zone, -1, lim.n_zones, 13);
  const int32_t region = w.region ? wire_id(w.region, w.region_b, i) : o.region[i];
  if (w.region) d.region[i] = region;
  chk(region, -1, lim.n_regions, 14);
  for (int a = 0; a < o.A; ++a) {
    const int32_t v = w.attr ? wire_id(w.attr, w.attr_b, a * N + i) : o.attr[a * N + i];
    if (w.attr) d.attr[a * N + i] = v;
    chk(v, -1, lim.n_attrs, 15);
  }
  for (int q = 0; q < o.Q; ++q) {
    const int32_t v = w.rr ? wire_id(w.rr, w.rr_b, q * N + i) : o.rr[q * N + i];
    if (w.rr) d.rr[q * N + i] = v;
    chk(v, -1, lim.n_rr_values, 16);
  }
  for (int e = 0; e < o.E; ++e) if (w.exec) d.exec[e * N + i] = wire_id(w.exec, w.exec_b, e * N + i);
  for (int q = 0; q < o.D; ++q) {
    const int32_t v = w.mprof ? wire_id(w.mprof, w.mprof_b, q * N + i) : o.mprof[q * N + i];
    if (w.mprof) d.mprof[q * N + i] = v;
    chk(v, -1, OE_MAX_PROFILES, 17);
  }
  for (int s = 0; s < o.S; ++s) {
    if (w.tuple_id) {
      int64_t t = w.tuple_b == 2 ? (int64_t)static_cast<const uint16_t*>(w.tuple_id)[s * N + i] : (int64_t)static_cast<const uint32_t*>(w.tuple_id)[s * N + i];
      if (t >= w.n_tuples) { code = max(code, 20); t = -1; }
      d.ptuple[s * N + i] = (int32_t)t;
      for (int r = 0; r < o.R; ++r) d.port[(s * o.R + r) * N + i] = t >= 0 ? __ldg(w.tuples + t * o.R + r) : OE_PORT_EMPTY;   // the lists were checked by k_validate_tuples
    } else {
      int prev_e = -1;
      for (int r = 0; r < o.R; ++r) {
        const uint32_t pr = o.port[((int64_t)s * o.R + r) * N + i];
        const int b = pr & 0xFFFF, e = pr >> 16;
        if (b > e) continue;
        if (b <= prev_e) code = max(code, 18);
        prev_e = e;
      }
    }
  }
  if (code) atomicMax(err, code);
}

// the same check for range lists stored list-major ([n][R]: oe_offer_packing.port_tuples)
__global__ void k_validate_tuples(const uint32_t* __restrict__ tuples, int n, int R, int code, int* err) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  int prev_e = -1;
  for (int r = 0; r < R; ++r) {
    const uint32_t pr = tuples[(int64_t)t * R + r];
    const int b = pr & 0xFFFF, e = pr >> 16;
    if (b > e) continue;
    if (b <= prev_e) atomicMax(err, code);
    prev_e = e;
  }
}

// The distinct `ports` range lists of the resident table (oe_load_offers_packed) stay a dictionary while rows are replaced
// (oe_update_offers sends full range lists): a hash table over the lists' contents, slot = list id + 1 (0 = empty).  A replaced
// row looks its list up; a list nobody had is appended (two rows that bring the same new list at once may append it twice: two
// ids with equal contents, both right).  count[0] runs past `cap` when the dictionary is full: the host then stops using the ids.
struct TupleHash {
  int32_t* slot;
  uint32_t mask;
};
__device__ __forceinline__ uint32_t tuple_hash_of(const uint32_t* w, int R) {
  uint32_t x = 2166136261u;
  for (int r = 0; r < R; ++r) { x ^= w[r]; x *= 16777619u; x ^= x >> 15; }
  return x;
}
__global__ void k_tuple_hash_insert(const uint32_t* __restrict__ tuples, int n, int R, TupleHash th) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  uint32_t w[OE_MAX_PORT_RANGES];
  for (int r = 0; r < R; ++r) w[r] = tuples[(int64_t)t * R + r];
  for (uint32_t q = tuple_hash_of(w, R) & th.mask;; q = (q + 1) & th.mask) {
    const int32_t old = atomicCAS(th.slot + q, 0, t + 1);
    if (old == 0) return;
    bool same = true;
    for (int r = 0; r < R; ++r) same = same && tuples[(int64_t)(old - 1) * R + r] == w[r];
    if (same) return;   // the caller's dictionary lists one range list twice: the first id answers look-ups, both stay valid
  }
}
// st: the replaced rows' `ports` column as staged by oe_update_offers ([S * R][n], slot-major); id_out: [S][N]
__global__ void k_reintern_tuples(const uint32_t* __restrict__ st, int64_t n, int S, int R, const int64_t* __restrict__ rows, int64_t N,
                                  uint32_t* tuples, int cap, TupleHash th, int* count, int32_t* __restrict__ id_out) {
  const int64_t q0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q0 >= n * S) return;
  const int64_t s = q0 / n, k = q0 - s * n;
  uint32_t w[OE_MAX_PORT_RANGES];
  for (int r = 0; r < R; ++r) w[r] = st[(s * R + r) * n + k];
  int32_t mine = -1, id = -1;
  for (uint32_t q = tuple_hash_of(w, R) & th.mask;; q = (q + 1) & th.mask) {
    int32_t v = *(volatile int32_t*)(th.slot + q);
    if (v == 0) {
      if (mine < 0) {
        mine = atomicAdd(count, 1);
        if (mine >= cap) break;   // dictionary full
        for (int r = 0; r < R; ++r) tuples[(int64_t)mine * R + r] = w[r];
        __threadfence();
      }
      v = atomicCAS(th.slot + q, 0, mine + 1);
      if (v == 0) { id = mine; break; }
    }
    bool same = true;
    for (int r = 0; r < R; ++r) same = same && *(volatile uint32_t*)(tuples + (int64_t)(v - 1) * R + r) == w[r];
    if (same) { id = v - 1; break; }
  }
  id_out[s * N + rows[k]] = id;
}

__global__ void k_validate_ranges(const uint32_t* __restrict__ port, int64_t n, int S, int R, int code, int* err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int s = 0; s < S; ++s) {
    int prev_e = -1;
    for (int r = 0; r < R; ++r) {
      const uint32_t pr = port[((int64_t)s * R + r) * n + i];
      const int b = pr & 0xFFFF, e = pr >> 16;
      if (b > e) continue;
      if (b <= prev_e) atomicMax(err, code);
      prev_e = e;
    }
  }
}

// The same checks for the rows [r0, r1) of a freshly copied offer chunk, in one launch (OE_CFG_ASYNC_LOADS: the table arrives
// in row chunks and every chunk is checked as soon as it has landed).  Same codes as the per-column checks.
__global__ void k_validate_offer_rows(oe::DevOffers o, OfferLimits lim, int64_t r0, int64_t r1, int* err) {
  const int64_t i = r0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= r1) return;
  int code = 0;
  auto chk = [&](int32_t v, int32_t lo, int32_t hi, int c) { if (v < lo || v >= hi) code = max(code, c); };
  chk(o.agent[i], 0, max(lim.n_agents, 1), 11);
  chk(o.host[i], 0, max(lim.n_hostnames, 1), 12);
  chk(o.zone[i], -1, lim.n_zones, 13);
  chk(o.region[i], -1, lim.n_regions, 14);
  for (int a = 0; a < o.A; ++a) chk(o.attr[(int64_t)a * o.N + i], -1, lim.n_attrs, 15);
  for (int q = 0; q < o.Q; ++q) chk(o.rr[(int64_t)q * o.N + i], -1, lim.n_rr_values, 16);
  for (int d = 0; d < o.D; ++d) chk(o.mprof[(int64_t)d * o.N + i], -1, OE_MAX_PROFILES, 17);
  for (int s = 0; s < o.S; ++s) {
    int prev_e = -1;
    for (int r = 0; r < o.R; ++r) {
      const uint32_t pr = o.port[((int64_t)s * o.R + r) * o.N + i];
      const int b = pr & 0xFFFF, e = pr >> 16;
      if (b > e) continue;
      if (b <= prev_e) code = max(code, 18);
      prev_e = e;
    }
  }
  if (code) atomicMax(err, code);
}

// Small tables while a large offer table occupies the host-to-device copy engine (OE_CFG_ASYNC_LOADS): copies are served in
// submission order, so a 400 KB column queued behind 100 MB of offers would arrive last.  The SMs read such a table straight
// out of the caller's pinned host memory over PCIe instead, beside the running DMA.  16 bytes per thread and step.
__global__ void __launch_bounds__(256) k_fetch_host(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16, unsigned char* __restrict__ dst_tail,
                                                    const unsigned char* __restrict__ src_tail, int tail) {
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n16; q += (size_t)gridDim.x * blockDim.x) dst[q] = src[q];
  if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

// K0: one thread per deployed task; every counter set whose filter the task passes gets the task's keys tallied.
__global__ void k_count_tasks(DevTasks t, DevProg p, int32_t* __restrict__ cnt, int32_t* __restrict__ type_total,
                              int32_t* __restrict__ type_agent) {
  int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= t.T) return;
  for (int c = 0; c < p.n_csets; ++c) {
    const oe_counter_set cs = p.csets[c];
    if (!task_filter(t, (int32_t)row, cs.filter_id)) continue;
    int32_t* base = cnt + p.cs_off[c];
    switch (cs.keyspace) {
      case OE_KS_HOSTNAME: { int32_t k = t.host[row]; if (k >= 0) atomicAdd(base + k, 1); } break;
      case OE_KS_ZONE: { int32_t k = t.zone[row]; if (k >= 0) atomicAdd(base + k, 1); } break;
      case OE_KS_REGION: { int32_t k = t.region[row]; if (k >= 0) atomicAdd(base + k, 1); } break;
      case OE_KS_ATTR:
        for (int a = 0; a < t.TA; ++a) { int32_t k = t.attr[(int64_t)a * t.cap + row]; if (k >= 0) atomicAdd(base + k, 1); }
        break;
      case OE_KS_RR_ATTR: { int32_t k = t.rr[(int64_t)cs.rr_col * t.cap + row]; if (k >= 0) atomicAdd(base + k, 1); } break;
    }
  }
  const int32_t ty = t.type[row];
  for (int s = 0; s < p.n_tsets; ++s) {
    if (p.ts_type[s] != ty) continue;
    atomicAdd(type_total + s, 1);
    atomicAdd(type_agent + (int64_t)s * p.n_agents + t.agent[row], 1);
  }
}

// K0b: per counter set (blockIdx.x) and, for blockIdx.y > 0, per variant y-1 with that variant's self rows taken out:
// {min over keys with a tally, max, number of keys with a tally}.  Only the sets a RoundRobin rule reads (rr_used) — a MaxPer
// rule looks at single tallies.  gridDim.z CTAs share a large key space (hostnames: one key per agent) and merge with atomics
// into a cell preset to {INT_MAX, 0, 0}; k_rr_finish then turns "no key" into min = 0.
__global__ void __launch_bounds__(256) k_rr_summary(DevCtx cx, const int32_t* __restrict__ cnt, const uint8_t* __restrict__ rr_used, int4* __restrict__ out_base,
                                                    int4* __restrict__ out_var) {
  const int c = blockIdx.x;
  if (!rr_used[c]) return;
  const int var = (int)blockIdx.y - 1;
  const oe_counter_set cs = cx.p.csets[c];
  const int64_t n = cx.p.cs_off[c + 1] - cx.p.cs_off[c];
  const int32_t* base = cnt + cx.p.cs_off[c];
  int mn = 0x7fffffff, mx = 0, nk = 0;
  for (int64_t k = (int64_t)blockIdx.z * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.z * blockDim.x) {
    int v = base[k];
    if (var >= 0 && v > 0) v -= self_key_count(cx, var, cs, (int32_t)k);
    if (v > 0) { mn = min(mn, v); mx = max(mx, v); ++nk; }
  }
  __shared__ int smn[256], smx[256], snk[256];
  smn[threadIdx.x] = mn; smx[threadIdx.x] = mx; snk[threadIdx.x] = nk;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      smn[threadIdx.x] = min(smn[threadIdx.x], smn[threadIdx.x + s]);
      smx[threadIdx.x] = max(smx[threadIdx.x], smx[threadIdx.x + s]);
      snk[threadIdx.x] += snk[threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    int4* cell = var < 0 ? out_base + c : out_var + (int64_t)var * cx.p.n_csets + c;
    if (gridDim.z == 1) {
      *cell = make_int4(smn[0] == 0x7fffffff ? 0 : smn[0], smx[0], snk[0], 0);
    } else {
      atomicMin(&cell->x, smn[0]); atomicMax(&cell->y, smx[0]); atomicAdd(&cell->z, snk[0]);
    }
  }
}
__global__ void k_rr_preset(int4* __restrict__ cells, int64_t n) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < n) cells[q] = make_int4(0x7fffffff, 0, 0, 0);
}
__global__ void k_rr_finish(int4* __restrict__ cells, int64_t n) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < n && cells[q].x == 0x7fffffff) cells[q].x = 0;
}

// K1: thread = one offer row, blockIdx.y = a tile of factored rows (classes first, then variants).  The CTA stages the
// tile's folded class descriptors, the rule nodes and its 128 offers' scalar columns in shared memory; a thread loads
// its offer's keys and port ranges into registers once and then walks the tile's rows.  Each warp emits one 32-offer
// word per row with __ballot_sync.  gridDim.x covers the padded stride so every word is written.
#define OE_K1_THREADS 128
#ifndef OE_K1_MIN_CTAS
#define OE_K1_MIN_CTAS 6
#endif
#define OE_K1_MAX_SMEM_NODES 1024
#define OE_PEER_MAX_WORLD 16
#define OE_PEER_HEADER_WORDS 64  // the flag words in front of an exchange buffer (256 B)
// Peer exchange (world > 1, peers attached): where K1 publishes its words.  base[p] = shard p's exchange buffer of this
// cycle's parity, already advanced to THIS rank's [rows][stride] block, so the layout is that of RowBits below.  Shard p's
// memory is written over NVLink peer access (or is plain device memory when p is this rank).
struct PeerOut {
  uint32_t* base[OE_PEER_MAX_WORLD];
  int n;
};
struct K1Args {
  uint32_t* F;
  int64_t stride;
  int rows, rows_per_tile;
  const FastRow* fast;   // [C]
  int n_nodes;           // rule nodes staged in shared memory (0 = read them from global memory)
  int sc_cols;           // scalar columns staged in shared memory (0 = read them through L1)
  RuleTab tab;
  PortTab ptab;
  const int32_t* row_map;  // bit row -> factored row (class or re-deployed requirement) it is evaluated as; null = identity
  PeerOut peers;
  int64_t xb0;           // first 128-offer block of this launch (the offer table may be evaluated in row chunks)
};

// K0c: rules whose leaves read only the offer's zone / region are evaluated once per (factored row, zone, region)
// instead of once per offer.  thread = one (zone, region) combination, blockIdx.y = factored row.
__global__ void __launch_bounds__(128) k_rule_tables(DevCtx cx, const FastRow* __restrict__ fast, uint32_t* __restrict__ tab, int words,
                                                     int zdim, int rdim) {
  const int row = blockIdx.y;
  const int var = row < cx.p.C ? -1 : row - cx.p.C;
  const int cls = var < 0 ? row : cx.v.cls[var];
  const FastRow& fr = fast[cls];
  if (!(fr.flags & OE_FR_RULE_TAB)) return;
  const int idx = blockIdx.x * 128 + threadIdx.x;
  bool v = false;
  if (idx < zdim * rdim) {
    OfferKeys ok{0, 0, idx % zdim - 1, idx / zdim - 1};
    v = eval_rule(cx, cx.p.nodes, 0, ok, var, fr.rule_begin, fr.rule_count);
  }
  const uint32_t bits = __ballot_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0 && (idx >> 5) < words) tab[(int64_t)row * words + (idx >> 5)] = bits;
}
// K0p: the port stages of every folded class over every distinct `ports` range list of the offer table (slot 0), once —
// thread = one list, blockIdx.y = class.  The stages run literally (static ports, dynamic ports with or without a range
// constraint, the pod's excluded ports: PortEvaluationStage.java:72-155,242-298) on the list instead of on an offer; K1
// (PORTS == 2, and the lean kernel) then looks the verdict up instead of testing ranges per offer.
__global__ void __launch_bounds__(128) k_port_tables(DevProg p, const FastRow* __restrict__ fast, const uint32_t* __restrict__ tuples, int n_tuples, int R,
                                                     uint32_t* __restrict__ out, int words) {
  const int cls = blockIdx.y;
  const int t = blockIdx.x * 128 + threadIdx.x;
  bool v = false;
  if (t < n_tuples && (fast[cls].flags & OE_FR_FAST)) {
    const uint32_t* rg = tuples + (int64_t)t * R;
    auto offered = [&](int port) {
      for (int r = 0; r < R; ++r) { const uint32_t pr = rg[r]; if ((int)(pr & 0xFFFF) <= port && port <= (int)(pr >> 16)) return true; }
      return false;
    };
    int taken[OE_DEV_MAX_TAKEN], n_taken = 0;
    auto is_taken = [&](int port) { for (int q = 0; q < n_taken; ++q) if (taken[q] == port) return true; return false; };
    v = true;
    for (int x = p.op_begin[cls]; x < p.op_begin[cls + 1] && v; ++x) {
      const oe_stage_op op = p.ops[x];
      if (op.opcode == OE_OP_PORT_STATIC) {
        v = offered(op.arg) && !is_taken(op.arg);
        if (v) taken[n_taken++] = op.arg;
      } else if (op.opcode == OE_OP_PORT_DYNAMIC) {
        int pick = -1;
        for (int r = 0; r < R && pick < 0; ++r) {
          const uint32_t pr = rg[r];
          const int b = pr & 0xFFFF, e = pr >> 16;
          for (int port = b; port <= e && pick < 0; ) {
            bool blocked = is_taken(port);
            for (int q = p.excl_begin[cls]; q < p.excl_begin[cls + 1] && !blocked; ++q) blocked = p.excl[q] == port;
            if (blocked) { ++port; continue; }
            if (op.arg >= 0) {   // inside the PortSpec's ranges, else on to the next of them
              bool in = false;
              int next = 0x7fffffff;
              for (int q = p.rs_begin[op.arg]; q < p.rs_begin[op.arg + 1]; ++q) {
                const int rb = p.rs[q] & 0xFFFF, re = p.rs[q] >> 16;
                if (rb <= port && port <= re) in = true;
                else if (rb > port && rb < next) next = rb;
              }
              if (!in) { port = next; continue; }
            }
            pick = port;
          }
        }
        v = pick >= 0;
        if (v) taken[n_taken++] = pick;
      }
    }
  }
  const uint32_t bits = __ballot_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0 && (t >> 5) < words) out[(int64_t)cls * words + (t >> 5)] = bits;
}

// PORTS: how a folded class's port stages are tested — 0 = the offer's ranges through L1, 1 = its (at most eight) ranges in
// registers, 2 = one lookup in the table k_port_tables built over the table's distinct range lists.
template <int PORTS, bool PEERS, int NQ>
__global__ void __launch_bounds__(OE_K1_THREADS, OE_K1_MIN_CTAS) k_class_feasibility(DevCtx cx, K1Args a) {
  constexpr bool REG_RANGES = PORTS == 1;
  extern __shared__ __align__(16) unsigned char k1_smem[];
  FastRow* s_fast = reinterpret_cast<FastRow*>(k1_smem);
  oe_rule_node* s_nodes = reinterpret_cast<oe_rule_node*>(s_fast + a.rows_per_tile);
  double* s_sc = reinterpret_cast<double*>(s_nodes + a.n_nodes);
  const int r0 = blockIdx.y * a.rows_per_tile;
  const int r1 = min(a.rows, r0 + a.rows_per_tile);
  const int64_t i = (a.xb0 + blockIdx.x) * OE_K1_THREADS + threadIdx.x;
  const bool valid = i < cx.o.N;
  {  // stage
    const uint4* src = reinterpret_cast<const uint4*>(a.fast);
    uint4* dst = reinterpret_cast<uint4*>(s_fast);
    for (int q = threadIdx.x; q < (r1 - r0) * OE_FR_U4; q += OE_K1_THREADS) {
      const int lrow = q / OE_FR_U4, part = q - lrow * OE_FR_U4;
      const int row = a.row_map ? __ldg(a.row_map + r0 + lrow) : r0 + lrow;
      const int cls = row < cx.p.C ? row : cx.v.cls[row - cx.p.C];
      dst[q] = __ldg(src + (int64_t)cls * OE_FR_U4 + part);
    }
    const uint4* nsrc = reinterpret_cast<const uint4*>(cx.p.nodes);
    uint4* ndst = reinterpret_cast<uint4*>(s_nodes);
    for (int q = threadIdx.x; q < a.n_nodes; q += OE_K1_THREADS) ndst[q] = __ldg(nsrc + q);
    if (valid)
      for (int q = 0; q < a.sc_cols; ++q) s_sc[q * OE_K1_THREADS + threadIdx.x] = __ldg(cx.o.scalar + (int64_t)q * cx.o.N + i);
  }
  __syncthreads();
  OfferRegs regs;
  regs.ok = OfferKeys{0, 0, -1, -1};
  if (valid) load_offer_regs(cx.o, i, REG_RANGES, regs);
  const int32_t ptuple = (PORTS == 2 && valid) ? __ldg(cx.o.ptuple + i) : 0;
  const oe_rule_node* nodes = a.n_nodes ? s_nodes : cx.p.nodes;
  const double* sc = a.sc_cols ? s_sc : nullptr;
  const int64_t word = i >> 5;
  __shared__ __align__(16) uint32_t s_words[PEERS ? 128 : 1][OE_K1_THREADS / 32];  // the CTA's four words of every row of the tile (peer exchange)
  // what changes from row to row is kept as running pointers: the row's descriptor, its output word, its rule-table word
  const bool writer = (threadIdx.x & 31) == 0 && word < a.stride;
  uint32_t* out = a.F + (int64_t)r0 * a.stride + word;
  const int tab_idx = (regs.ok.region + 1) * a.tab.zdim + (regs.ok.zone + 1);
  const int tab_bit = tab_idx & 31;
  const uint32_t* tab_col = a.tab.bits + (tab_idx >> 5);
  const uint32_t* tab_run = tab_col + (int64_t)r0 * a.tab.words;   // the row's table word when bit rows are factored rows (no row_map)
  const FastRow* frp = s_fast;
  for (int row = r0; row < r1; ++row, ++frp, out += a.stride, tab_run += a.tab.words) {
    const FastRow& fr = *frp;
    const int frow = a.row_map ? __ldg(a.row_map + row) : row;   // warp-uniform
    const int var = frow < cx.p.C ? -1 : frow - cx.p.C;
    const uint32_t* tab_word = a.row_map ? tab_col + (int64_t)frow * a.tab.words : tab_run;
    bool pass = false;
    if (fr.flags & OE_FR_FAST) {
      if (valid) pass = eval_fast<PORTS, NQ>(cx, nodes, fr, sc, i, regs, var, tab_word, tab_bit, a.ptab, ptuple);
    } else if (valid) {
      const int cls = var < 0 ? frow : cx.v.cls[var];
      pass = eval_pair<false>(cx, i, regs.ok, cls, var, nullptr);
    }
    const uint32_t bits = __ballot_sync(0xffffffffu, pass);
    if (writer) {
      *out = bits;
      if (PEERS) s_words[row - r0][threadIdx.x >> 5] = bits;
    }
  }
  // The all-gather of a sharded cycle, fused into the pass: the CTA stores its 16 bytes of every row straight into every
  // shard's exchange buffer (one 128-bit store per row and shard), so the rows cross NVLink while the other CTAs still compute
  // and no collective runs afterwards — k_peer_signal / k_peer_wait order the readers behind these stores.
  if (PEERS) {
    __syncthreads();
    const int64_t w0 = (a.xb0 + blockIdx.x) * (OE_K1_THREADS / 32);
    const int nrows = r1 - r0;
    if (w0 < a.stride)
      for (int q = threadIdx.x; q < nrows * a.peers.n; q += OE_K1_THREADS) {
        const int p = q / nrows, rr = q - p * nrows;
        *reinterpret_cast<uint4*>(a.peers.base[p] + (int64_t)(r0 + rr) * a.stride + w0) = *reinterpret_cast<const uint4*>(s_words[rr]);
      }
  }
}

// K1, lean form.  When every class is a folded new-pod pipeline whose scalar pools fit four registers, whose rule (if any) is
// tabulated over (zone, region) and whose port stages are tabulated over the table's distinct range lists — the headline shape —
// a (offer, bit row) pair needs no interpretation at all: four fp64 compares, one presence mask, the allocation-role test and two
// bits that the thread fetched ONCE per tile of 32 rows (k_transpose_tables lays the two tables out as [zone-region combination
// or range list][32 rows per word]).  ~25 instructions per pair instead of ~110.
struct LeanRow {
  double thr[4];
  uint32_t need;        // pool entries that must be present; all ones = the class can never fit (a port requested twice)
  int32_t role;         // pod role when the class consumes from the pool (offers allocated to another role fail), else -1
  uint32_t pad[2];
};
static_assert(sizeof(LeanRow) == 48, "LeanRow is staged as three uint4");
struct LeanArgs {
  uint32_t* F;
  int64_t stride;
  int rows, row_words;            // bit rows; 32-row words per table entry
  const LeanRow* lean;            // [rows]
  const uint32_t* ruleT;          // [zone-region combinations][row_words] or null (no row has a rule)
  const uint32_t* portT;          // [range lists][row_words] or null (no row has port stages)
  int zdim;
  PeerOut peers;
  int64_t xb0;
};
// bit rows -> the two verdict tables transposed: warp = (table entry e, 32-row word w), lane = row of the word
__global__ void __launch_bounds__(128) k_transpose_tables(DevCtx cx, const FastRow* __restrict__ fast, const int32_t* __restrict__ row_map, int rows, int row_words,
                                                          const uint32_t* __restrict__ rule_tab, int tab_words, int n_combos, uint32_t* __restrict__ ruleT,
                                                          const uint32_t* __restrict__ port_tab, int ptab_words, int n_tuples, uint32_t* __restrict__ portT,
                                                          LeanRow* __restrict__ lean) {
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), k = threadIdx.x & 31;
  const int n_entries = max(1, max(n_combos, n_tuples));   // entry 0 also writes the rows' descriptors: it always exists
  if (q >= n_entries * row_words) return;
  const int e = q / row_words, w = q - e * row_words;
  const int row = w * 32 + k;
  bool rule = false, port = false;
  if (row < rows) {
    const int frow = row_map ? row_map[row] : row;
    const int cls = frow < cx.p.C ? frow : cx.v.cls[frow - cx.p.C];
    const FastRow& fr = fast[cls];
    rule = !(fr.flags & OE_FR_RULE_TAB) || (e < n_combos && ((rule_tab[(int64_t)frow * tab_words + (e >> 5)] >> (e & 31)) & 1u));
    port = !(fr.n_ports | fr.n_dyn) || (e < n_tuples && ((port_tab[(int64_t)cls * ptab_words + (e >> 5)] >> (e & 31)) & 1u));
    if (e == 0) {   // the row's descriptor, once
      LeanRow lr;
      for (int x = 0; x < 4; ++x) lr.thr[x] = fr.thr[x];
      lr.need = (fr.flags & OE_FR_NEVER) ? 0xffffffffu : *reinterpret_cast<const uint32_t*>(fr.thr_q);
      lr.role = (fr.flags & OE_FR_NEEDS_POOL) ? fr.pod_role : -1;
      lr.pad[0] = lr.pad[1] = 0u;
      lean[row] = lr;
    }
  }
  const uint32_t rbits = __ballot_sync(0xffffffffu, rule), pbits = __ballot_sync(0xffffffffu, port);
  if (k == 0) {
    if (ruleT && e < n_combos) ruleT[(int64_t)e * row_words + w] = rbits;
    if (portT && e < n_tuples) portT[(int64_t)e * row_words + w] = pbits;
  }
}
template <bool PEERS>
__global__ void __launch_bounds__(OE_K1_THREADS, 8) k_class_feasibility_lean(DevOffers o, LeanArgs a) {
  __shared__ __align__(16) LeanRow s_rows[32];
  __shared__ __align__(16) uint32_t s_words[PEERS ? 32 : 1][OE_K1_THREADS / 32];
  const int tile = blockIdx.y, r0 = tile * 32, nrows = min(32, a.rows - r0);
  for (int q = threadIdx.x; q < nrows * 3; q += OE_K1_THREADS)
    reinterpret_cast<uint4*>(s_rows)[q] = __ldg(reinterpret_cast<const uint4*>(a.lean + r0) + q);
  const int64_t i = (a.xb0 + blockIdx.x) * OE_K1_THREADS + threadIdx.x;
  const bool valid = i < o.N;
  double sv[4] = {0.0, 0.0, 0.0, 0.0};
  uint32_t present = 0u, both = 0u;
  int32_t alloc = -1;
  if (valid) {
    const int SK = o.S * o.K;
#pragma unroll
    for (int q = 0; q < 4; ++q) if (q < SK) sv[q] = __ldg(o.scalar + (int64_t)q * o.N + i);
    present = __ldg(o.present + i);
    alloc = __ldg(o.alloc + i);
    both = 0xffffffffu;
    if (a.ruleT) both &= __ldg(a.ruleT + (int64_t)((__ldg(o.region + i) + 1) * a.zdim + (__ldg(o.zone + i) + 1)) * a.row_words + tile);
    if (a.portT) { const int32_t t = __ldg(o.ptuple + i); both &= t >= 0 ? __ldg(a.portT + (int64_t)t * a.row_words + tile) : 0u; }
  }
  __syncthreads();
  const int64_t word = i >> 5;
  const bool writer = (threadIdx.x & 31) == 0 && word < a.stride;
  uint32_t* out = a.F + (int64_t)r0 * a.stride + word;
#pragma unroll 4
  for (int k = 0; k < nrows; ++k, out += a.stride) {
    const LeanRow& r = s_rows[k];
    bool pass = ((both >> k) & 1u) & ((present & r.need) == r.need);
    pass &= !(sv[0] < r.thr[0]) & !(sv[1] < r.thr[1]) & !(sv[2] < r.thr[2]) & !(sv[3] < r.thr[3]);
    pass &= !((r.role >= 0) & (alloc >= 0) & (alloc != r.role));   // consumableResource (MesosResourcePool.java:87-97)
    const uint32_t bits = __ballot_sync(0xffffffffu, pass);
    if (writer) {
      *out = bits;
      if (PEERS) s_words[k][threadIdx.x >> 5] = bits;
    }
  }
  if (PEERS) {   // as in k_class_feasibility: the CTA's 16 bytes of every row into every shard's exchange region
    __syncthreads();
    const int64_t w0 = (a.xb0 + blockIdx.x) * (OE_K1_THREADS / 32);
    if (w0 < a.stride)
      for (int q = threadIdx.x; q < nrows * a.peers.n; q += OE_K1_THREADS) {
        const int p = q / nrows, rr = q - p * nrows;
        *reinterpret_cast<uint4*>(a.peers.base[p] + (int64_t)(r0 + rr) * a.stride + w0) = *reinterpret_cast<const uint4*>(s_words[rr]);
      }
  }
}

// Peer exchange, the ordering half.  After K1 (same stream) every rank raises flag[rank] = epoch in every shard's exchange
// header with a system-scope release; before the commit path reads its exchange buffer a rank waits until all of its own
// flags have reached the epoch.  A peer that never arrives (a crashed rank) ends the wait after timeout_ns with *err = shard + 1.
struct PeerFlags {
  uint32_t* flags[OE_PEER_MAX_WORLD];  // each shard's flag words
  int n, rank;
  uint32_t epoch;
};
__global__ void k_peer_signal(PeerFlags f) {
  const int p = threadIdx.x;
  if (p < f.n) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.flags[p] + f.rank), "r"(f.epoch) : "memory");
  }
}
__global__ void k_peer_wait(const uint32_t* my_flags, int n, uint32_t epoch, unsigned long long timeout_ns, int* err) {
  const int p = threadIdx.x;
  if (p >= n) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flags + p) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) { atomicExch(err, p + 1); break; }
    __nanosleep(100);
  }
}

// K1x: out[j][w] = F[req_row[j]][w].  CTA = (1 KiB column slab) x (a block of requirements): 64 lanes of uint4 across,
// 4 requirements in flight down.  The source rows of a slab stay L1/L2 resident; the 128-bit streaming stores are the
// only HBM traffic.  Sized by tools/bw_probe.cu: 1 KiB slabs with ~32 CTAs per SM write at the copy-bandwidth peak.
#define OE_KX_LANES 64
#define OE_KX_ROWS 4
__global__ void __launch_bounds__(OE_KX_LANES * OE_KX_ROWS) k_expand_matrix(const uint32_t* __restrict__ F, const int32_t* __restrict__ req_row,
                                                                            uint32_t* __restrict__ out, int64_t stride, int64_t M,
                                                                            int reqs_per_block, int64_t xb0, int64_t w_end) {
  const int64_t w4 = ((xb0 + blockIdx.x) * OE_KX_LANES + threadIdx.x) * 4;  // first word of this lane's uint4
  if (w4 >= w_end) return;
  const int64_t j0 = (int64_t)blockIdx.y * reqs_per_block;
  const int64_t j1 = min(M, j0 + reqs_per_block);
  int64_t j = j0 + threadIdx.y;
  for (; j + OE_KX_ROWS < j1; j += 2 * OE_KX_ROWS) {
    const int32_t ra = __ldg(req_row + j), rb = __ldg(req_row + j + OE_KX_ROWS);
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)ra * stride + w4));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)rb * stride + w4));
    __stcs(reinterpret_cast<uint4*>(out + j * stride + w4), a);
    __stcs(reinterpret_cast<uint4*>(out + (j + OE_KX_ROWS) * stride + w4), b);
  }
  for (; j < j1; j += OE_KX_ROWS) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(F + (int64_t)__ldg(req_row + j) * stride + w4));
    __stcs(reinterpret_cast<uint4*>(out + j * stride + w4), a);
  }
}

// View of the factored bit rows in GLOBAL offer order: word gw of row r lives in shard gw / shard_words.
struct RowBits {
  const uint32_t* F;
  int64_t stride;       // words per (shard, row)
  int rows;
  int64_t shard_words;  // words of one shard in global word index space (== stride when not sharded)
  int64_t total_words;  // global words that hold offers
  const int* skip_ptr;  // commit passes after the first (offer table arriving in chunks): nothing to do once *skip_ptr >= skip_at
  int skip_at;          // (every requirement placed); null: always run
  __device__ __forceinline__ bool skipped() const { return skip_ptr && *skip_ptr >= skip_at; }
  __device__ __forceinline__ uint32_t word(int row, int64_t gw) const {
    const int64_t shard = gw / shard_words, lw = gw - shard * shard_words;
    return F[(shard * rows + row) * stride + lw];
  }
};

// K3a: set bits per (row, tile of 1024 words).
#define OE_TILE_WORDS 1024
__global__ void __launch_bounds__(256) k_tile_popc(RowBits rb, int tiles, int32_t* __restrict__ tile_cnt) {
  const int row = blockIdx.y, tile = blockIdx.x;
  int c = 0;
  for (int q = threadIdx.x; q < OE_TILE_WORDS; q += 256) {
    const int64_t gw = (int64_t)tile * OE_TILE_WORDS + q;
    if (gw < rb.total_words) c += __popc(rb.word(row, gw));
  }
  __shared__ int sc[256];
  sc[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sc[threadIdx.x] += sc[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_cnt[(int64_t)row * tiles + tile] = sc[0];
}

// K3b: one CTA.  Per row: exclusive scan of its tile counts (in place) and the list length min(total, cap); then the
// exclusive scan of the lengths gives every row's base in the candidate array.  row_base[rows] = total entries.
__global__ void __launch_bounds__(1024) k_row_offsets(int rows, int tiles, int64_t cap, int32_t* __restrict__ tile_cnt,
                                                      int64_t* __restrict__ row_base, int32_t* __restrict__ row_len) {
  __shared__ int64_t carry;
  __shared__ int64_t warp_sum[32];
  // per-row tile scans: a warp per row
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int row = warp; row < rows; row += 32) {
    int32_t* c = tile_cnt + (int64_t)row * tiles;
    int64_t run = 0;
    for (int t0 = 0; t0 < tiles; t0 += 32) {
      const int t = t0 + lane;
      int v = t < tiles ? c[t] : 0;
      int incl = v;
      for (int d = 1; d < 32; d <<= 1) { int n = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += n; }
      if (t < tiles) c[t] = (int32_t)min((int64_t)0x7fffffff, run + incl - v);
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) row_len[row] = (int32_t)min(run, cap);
  }
  __syncthreads();
  // exclusive scan of row_len over rows
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int r0 = 0; r0 < rows; r0 += 1024) {
    const int r = r0 + threadIdx.x;
    int64_t v = r < rows ? row_len[r] : 0;
    int64_t incl = v;
    for (int d = 1; d < 32; d <<= 1) { int64_t n = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += n; }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int64_t w = warp_sum[lane], wi = w;
      for (int d = 1; d < 32; d <<= 1) { int64_t n = __shfl_up_sync(0xffffffffu, wi, d); if (lane >= d) wi += n; }
      warp_sum[lane] = wi - w;
    }
    __syncthreads();
    const int64_t excl = carry + warp_sum[warp] + incl - v;
    if (r < rows) row_base[r] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) row_base[rows] = carry;
}

// K3c: ascending list of feasible GLOBAL offer indices per row (first row_len[r] of them).
__global__ void __launch_bounds__(256) k_compact_rows(RowBits rb, int tiles, const int32_t* __restrict__ tile_off,
                                                      const int64_t* __restrict__ row_base, const int32_t* __restrict__ row_len,
                                                      int32_t* __restrict__ cand) {
  const int row = blockIdx.y, tile = blockIdx.x;
  const int len = row_len[row];
  const int64_t toff = tile_off[(int64_t)row * tiles + tile];
  if (toff >= len) return;
  // each thread owns 4 consecutive words of the tile
  const int64_t gw0 = (int64_t)tile * OE_TILE_WORDS + threadIdx.x * 4;
  uint32_t w[4];
  int c = 0;
  for (int q = 0; q < 4; ++q) {
    w[q] = (gw0 + q < rb.total_words) ? rb.word(row, gw0 + q) : 0u;
    c += __popc(w[q]);
  }
  __shared__ int warp_tot[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = c;
  for (int d = 1; d < 32; d <<= 1) { int n = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += n; }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  int before = 0;
  for (int q = 0; q < warp; ++q) before += warp_tot[q];
  int64_t pos = toff + before + incl - c;
  int32_t* out = cand + row_base[row];
  for (int q = 0; q < 4; ++q) {
    uint32_t x = w[q];
    while (x) {
      const int b = __ffs(x) - 1;
      x &= x - 1;
      if (pos < len) out[pos] = (int32_t)((gw0 + q) * 32 + b);
      ++pos;
    }
  }
}

// K3d: which bit rows share an offer.  adj[r] gets bit q set iff rows r and q are both feasible for some offer.  Rows that
// share nothing can be committed independently of each other.  thread = one 32-offer word; only used when rows <= 64.
#define OE_ADJ_MAX_ROWS 64
#define OE_ADJ_SPLIT 8   // threads per word: each takes every eighth row r (a single thread per word left the SMs at 4-8 warps
                         // each, walking a chain of rows^2 dependent tests: 41 us on C5)
__global__ void __launch_bounds__(128 * OE_ADJ_SPLIT) k_row_overlap(RowBits rb, unsigned long long* __restrict__ adj) {
  __shared__ uint32_t tile[OE_ADJ_MAX_ROWS][128];
  if (rb.skipped()) return;
  const int x = threadIdx.x & 127, y = threadIdx.x >> 7;
  const int64_t gw = (int64_t)blockIdx.x * 128 + x;
  for (int r = y; r < rb.rows; r += OE_ADJ_SPLIT) tile[r][x] = gw < rb.total_words ? rb.word(r, gw) : 0u;
  __syncthreads();
  for (int r = y; r < rb.rows; r += OE_ADJ_SPLIT) {
    const uint32_t w = tile[r][x];
    unsigned long long acc = 0ull;
    if (w)
      for (int q = 0; q < rb.rows; ++q) acc |= (unsigned long long)((w & tile[q][x]) != 0u) << q;
    const uint32_t lo = __reduce_or_sync(0xffffffffu, (uint32_t)acc), hi = __reduce_or_sync(0xffffffffu, (uint32_t)(acc >> 32));   // (REDUX: two instead of ten shuffles)
    acc = ((unsigned long long)hi << 32) | lo;
    if ((threadIdx.x & 31) == 0 && acc) atomicOr(adj + r, acc);
  }
}

// K2: the greedy commit of PlanScheduler.resourceOffers (S/scheduler/plan/PlanScheduler.java:56-67): requirements in
// order, each claims the lowest-index feasible offer that no earlier requirement of the cycle claimed.
//
// The order only matters between requirements that can compete for an offer.  Bit rows are grouped into connected
// components of the "share an offer" relation (K3d); requirements of different components commute, so each component
// is replayed strictly in plan order by ONE warp, up to 32 components side by side.
// Prologue (whole CTA): label the components, then a stable counting sort of the requirements by component.
//
// A lone warp retires a dependent instruction every ~6 cycles, so what a requirement costs is the NUMBER of instructions on
// its path (measured: a step of ~100 instructions — window from shared memory, claimed words, ballot, shuffle — took ~600
// cycles, and probing four requirements side by side took four times that).  replay_component therefore keeps everything a
// step needs in registers: lane l holds entry l of the current 32-candidate window of each of the component's R rows, and one
// warp-uniform mask per row says which entries are still free.  A requirement of row k takes the lowest set bit of mask k;
// the claimed offer is then struck from every row's mask with one compare + ballot per row (rows of one component may list
// the same offer).  Shared memory (claimed bitmap) and global memory (the row's list) are touched only when a row's window
// has no free entry left; the window after next is prefetched into registers, so a refill does not wait for memory either.
struct RowState { int cursor, len; long long base; };
struct CommitArgs {
  const int32_t* cand;
  const int64_t* row_base;
  const int32_t* row_len;
  int rows;
  int64_t used_words;
  const int32_t* req_row;
  int64_t M;
  oe_decision* dec;
  uint32_t* used_g;                  // global-memory fallbacks when the bitmap / row states exceed shared memory
  struct RowState* rows_g;
  int used_in_smem, rows_in_smem;
  const unsigned long long* adj;     // [rows] from K3d, or null (one component)
  int2* order;                       // [M] scratch: (requirement, row) grouped by component, plan order kept inside a group
  int* n_placed;
};
#define OE_COMMIT_THREADS 1024  // 32 warps: one per component
#define OE_COMP_ROWS 8          // rows per component that replay_component keeps in registers (more: list-walking fallback)

// One component, R <= OE_COMP_ROWS rows, replayed by one warp (see above).  wrow[k] = bit row of local row k; lrow[row] = k.
//
// Measured on this part (tools/warp_latency.cu): a taken branch ~19 cycles, SHFL / VOTE / LDS ~30 each when dependent, and an
// `if (lane == 0) { atomicOr; store; }` inside the loop ~100 — so the step below has no divergent code and no memory
// operation at all: the lane whose turn it is keeps the pick in a register, and claims (shared-memory bitmap) and decisions
// (global memory) are written 32 at a time at the end of a batch, or just before a window refill reads the bitmap.
template <int R>
__device__ __forceinline__ int replay_component(const CommitArgs& a, RowState* rs, uint32_t* used, const int* wrow, const int* lrow,
                                                int* win, const int n_rows, const int begin, const int end, const int lane) {
  volatile uint32_t* vused = reinterpret_cast<volatile uint32_t*>(used);
  int c[R], cn[R];        // lane's entry of row k's current window / of the window after it
  uint32_t mine_free = 0;  // bit k: this lane's entry of row k is still free (lane-local, so striking an offer needs no ballot)
#pragma unroll
  for (int k = 0; k < R; ++k) {
    c[k] = cn[k] = -1;
    if (k < n_rows) {
      const RowState st = rs[wrow[k]];
      if (lane < st.len) c[k] = __ldg(a.cand + st.base + lane);
      if (32 + lane < st.len) cn[k] = __ldg(a.cand + st.base + 32 + lane);
    }
    if (c[k] >= 0) mine_free |= 1u << k;   // nothing of this component is claimed yet
    if (k < n_rows) win[k * 32 + lane] = c[k];  // a copy in shared memory: "my entry of row k" for a run-time k is one load, not R selects
  }
  __syncwarp();
  // next window of local row k (a uniform value): registers <- prefetched registers, then test against the claimed bitmap
  auto refill = [&](const int k) {
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (q != k) continue;
      for (;;) {
        const RowState st = rs[wrow[q]];
        const int start = st.cursor + 32;
        if (start >= st.len) break;  // the row is exhausted: nothing free stays nothing free
        c[q] = cn[q];
        win[q * 32 + lane] = c[q];
        cn[q] = start + 32 + lane < st.len ? __ldg(a.cand + st.base + start + 32 + lane) : -1;
        if (lane == 0) rs[wrow[q]].cursor = start;
        __syncwarp();
        const uint32_t w = vused[max(c[q], 0) >> 5];
        const bool is_free = c[q] >= 0 && !((w >> (c[q] & 31)) & 1u);
        mine_free = is_free ? mine_free | (1u << q) : mine_free & ~(1u << q);
        if (__ballot_sync(0xffffffffu, is_free)) break;
      }
    }
  };
  int placed = 0;
  int2 nx = begin + lane < end ? a.order[begin + lane] : make_int2(0, 0);
  for (int b0 = begin; b0 < end; b0 += 32) {
    const int my_req = nx.x, my_k = lrow[nx.y];               // plan position b0 + lane
    if (b0 + 32 + lane < end) nx = a.order[b0 + 32 + lane];   // a batch ahead of its use
    const int cnt = min(32, end - b0);
    int my_o = -1;                                            // what the requirement at b0 + lane takes
    auto flush = [&]() {  // idempotent: claims of this batch so far into the bitmap
      if (my_o >= 0) atomicOr(&used[my_o >> 5], 1u << (my_o & 31));
      __syncwarp();
    };
    if (R == 1) {
      // one row: nothing to arbitrate, the i-th requirement of the batch takes the i-th free entry — all lanes at once
      int done = 0;
      while (done < cnt) {
        uint32_t fm = __ballot_sync(0xffffffffu, mine_free & 1u);
        if (fm == 0u) { flush(); refill(0); fm = __ballot_sync(0xffffffffu, mine_free & 1u); if (fm == 0u) break; }
        const int take = min(__popc(fm), cnt - done);
        const bool mine = lane >= done && lane < done + take;
        const int p = mine ? (int)__fns(fm, 0, lane - done + 1) : 0;
        const int o = __shfl_sync(0xffffffffu, c[0], p);
        if (mine) my_o = o;
        const uint32_t gone = __reduce_or_sync(0xffffffffu, mine ? 1u << p : 0u);
        mine_free &= ~((gone >> lane) & 1u);
        done += take;
      }
      placed += done;
    } else {
      for (int i = 0; i < cnt; ++i) {
        const int k = __shfl_sync(0xffffffffu, my_k, i);
        const int ck = win[k * 32 + lane];  // issued early: k and the windows are known long before fmk is
        uint32_t fmk = __ballot_sync(0xffffffffu, (mine_free >> k) & 1u);   // the free entries of row k's window
        bool fmk_refilled = false;
        if (fmk == 0u) {  // rare: once per 32 claimed entries of a row
          flush();
          refill(k);
          __syncwarp();
          fmk = __ballot_sync(0xffffffffu, (mine_free >> k) & 1u);
          fmk_refilled = true;
        }
        const int o = __shfl_sync(0xffffffffu, fmk_refilled ? win[k * 32 + lane] : ck, (__ffs(fmk) - 1) & 31);
        const bool got = fmk != 0u;
        // the offer is gone from every row that lists it (rows of one component may share offers): each lane strikes its own
        // entries, no ballot.  Window padding is -1 and o >= 0 whenever `got`.
#pragma unroll
        for (int q = 0; q < R; ++q) mine_free = (got && c[q] == o) ? mine_free & ~(1u << q) : mine_free;
        my_o = (lane == i && got) ? o : my_o;
        placed += got;
      }
    }
    flush();
    if (lane < cnt) a.dec[my_req].offer_row = my_o;
  }
  return placed;
}

// SMEM = row states and claimed bitmap both live in dynamic shared memory.  It is a template parameter and not a run-time choice
// of pointer because a pointer that MAY be global makes every bitmap access a generic one (LD.E / ATOM.E instead of LDS / ATOMS),
// and a generic atomic on a shared-memory address costs several hundred cycles on the step's dependent chain.
template <bool SMEM>
__global__ void __launch_bounds__(OE_COMMIT_THREADS) k_commit(CommitArgs a) {
  extern __shared__ __align__(16) unsigned char commit_smem[];
  __shared__ int comp_of[OE_ADJ_MAX_ROWS];
  __shared__ int n_comp_s, placed_s;
  __shared__ int cnt[32][33];   // [component][warp], [component][32] = total
  __shared__ int comp_begin[33];
  __shared__ int lrow[OE_ADJ_MAX_ROWS];              // row -> its index among the rows of its warp's component(s)
  __shared__ int wrow[32][OE_COMP_ROWS];             // warp -> the rows it replays (first OE_COMP_ROWS of them)
  __shared__ int wrows[32];                          // warp -> how many rows it replays
  __shared__ int win[OE_ADJ_MAX_ROWS * 32];          // the rows' current windows, warp by warp (a copy of the registers, see replay_component)
  __shared__ int woff[32];                           // warp -> where its windows start in win[]
  const bool in_regs = SMEM && a.rows <= OE_ADJ_MAX_ROWS && a.M < (1 << 28);  // (requirement, local row) packs into one register
  RowState* rs;    // [rows]
  uint32_t* used;  // [used_words]
  if (SMEM) {
    rs = reinterpret_cast<RowState*>(commit_smem);
    used = reinterpret_cast<uint32_t*>(commit_smem + (size_t)a.rows * sizeof(RowState));
  } else {
    rs = a.rows_in_smem ? reinterpret_cast<RowState*>(commit_smem) : a.rows_g;
    used = a.used_in_smem ? reinterpret_cast<uint32_t*>(commit_smem + (a.rows_in_smem ? (size_t)a.rows * sizeof(RowState) : 0)) : a.used_g;
  }
  const int tid = threadIdx.x;
  for (int64_t k = tid; k < a.used_words; k += OE_COMMIT_THREADS) used[k] = 0u;
  for (int r = tid; r < a.rows; r += OE_COMMIT_THREADS) { rs[r].cursor = 0; rs[r].len = a.row_len[r]; rs[r].base = a.row_base[r]; }
  // ---- components of the rows (labels 0..n_comp-1 in order of their lowest row; folded onto 32 warps)
  if (tid == 0) {
    placed_s = 0;
    int n = 1;
    if (a.adj) {
      n = 0;
      unsigned long long seen = 0ull;
      for (int r = 0; r < a.rows; ++r) {
        if ((seen >> r) & 1ull) continue;
        unsigned long long group = 1ull << r, frontier = group;
        while (frontier) {
          const int q = __ffsll((long long)frontier) - 1;
          frontier &= frontier - 1;
          const unsigned long long nb = a.adj[q] & ~group;
          group |= nb;
          frontier |= nb;
        }
        seen |= group;
        for (int q = 0; q < a.rows; ++q) if ((group >> q) & 1ull) comp_of[q] = n & 31;
        ++n;
      }
      n = n > 32 ? 32 : n;
    }
    n_comp_s = n;
    if (a.rows <= OE_ADJ_MAX_ROWS) {
      for (int w = 0; w < 32; ++w) wrows[w] = 0;
      for (int r = 0; r < a.rows; ++r) {
        const int w = a.adj ? comp_of[r] : 0;
        const int k = wrows[w]++;
        lrow[r] = k & (OE_COMP_ROWS - 1);
        if (k < OE_COMP_ROWS) wrow[w][k] = r;
      }
      for (int w = 0, run = 0; w < 32; ++w) { woff[w] = run * 32; run += wrows[w]; }
    }
  }
  __syncthreads();
  const int n_comp = n_comp_s;
  // ---- stable counting sort of the requirements by component.  A warp owns a contiguous chunk of the plan and walks it 32
  // requirements at a time; lanes of one component find each other with match_any, so a batch costs a handful of instructions
  // whatever the number of components (cnt[c][w] = requirements of component c in warp w's chunk, then its running offset).
  const int warp = tid >> 5, lane = tid & 31;
  const int64_t chunk = ((a.M + 31) / 32 + 31) / 32 * 32;  // per warp, a multiple of 32
  const int64_t j_lo = min(a.M, (int64_t)warp * chunk), j_hi = min(a.M, j_lo + chunk);
  cnt[lane][warp] = 0;
  __syncwarp();
  for (int64_t j0 = j_lo; j0 < j_hi; j0 += 32) {
    const int64_t j = j0 + lane;
    const int c = j < j_hi ? (a.adj ? comp_of[a.req_row[j]] : 0) : 32 + lane;  // past the end: a group of its own, not counted
    const uint32_t same = __match_any_sync(0xffffffffu, c);
    if (c < 32 && lane == __ffs(same) - 1) cnt[c][warp] += __popc(same);
    __syncwarp();
  }
  __syncthreads();
  if (tid < n_comp) {  // exclusive scan over the warps, per component
    int run = 0;
    for (int w = 0; w < 32; ++w) { const int v = cnt[tid][w]; cnt[tid][w] = run; run += v; }
    cnt[tid][32] = run;
  }
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int c = 0; c < n_comp; ++c) { comp_begin[c] = run; run += cnt[c][32]; }
    comp_begin[n_comp] = run;
  }
  __syncthreads();
  for (int64_t j0 = j_lo; j0 < j_hi; j0 += 32) {
    const int64_t j = j0 + lane;
    const int row = j < j_hi ? a.req_row[j] : 0;
    const int c = j < j_hi ? (a.adj ? comp_of[row] : 0) : 32 + lane;
    const uint32_t same = __match_any_sync(0xffffffffu, c);
    if (c < 32) a.order[comp_begin[c] + cnt[c][warp] + __popc(same & ((1u << lane) - 1u))] = make_int2((int)j, row);
    __syncwarp();
    if (c < 32 && lane == __ffs(same) - 1) cnt[c][warp] += __popc(same);
    __syncwarp();
  }
  __syncthreads();
  // ---- replay: warp = component
  int placed = 0;
  const int begin = warp < n_comp ? comp_begin[warp] : 0, end = warp < n_comp ? comp_begin[warp + 1] : 0;
  // the component's (requirement, row) pairs come in 32 at a time, one per lane, a batch ahead of their use: a load per step, even
  // issued a step early, would put an L2 round trip on every step
  int2 batch = begin + lane < end ? a.order[begin + lane] : make_int2(0, 0);
  int2 ahead = begin + 32 + lane < end ? a.order[begin + 32 + lane] : make_int2(0, 0);
  auto next_pair = [&](int pos) {
    const int q = (pos - begin) & 31;
    if (q == 0 && pos != begin) {
      batch = ahead;
      ahead = pos + 32 + lane < end ? a.order[pos + 32 + lane] : make_int2(0, 0);
    }
    return make_int2(__shfl_sync(0xffffffffu, batch.x, q), __shfl_sync(0xffffffffu, batch.y, q));
  };
  if (in_regs && begin < end && wrows[warp] <= OE_COMP_ROWS) {
    const int n_rows = wrows[warp];
    if (n_rows == 1) placed = replay_component<1>(a, rs, used, wrow[warp], lrow, win + woff[warp], n_rows, begin, end, lane);
    else if (n_rows == 2) placed = replay_component<2>(a, rs, used, wrow[warp], lrow, win + woff[warp], n_rows, begin, end, lane);
    else if (n_rows <= 4) placed = replay_component<4>(a, rs, used, wrow[warp], lrow, win + woff[warp], n_rows, begin, end, lane);
    else placed = replay_component<8>(a, rs, used, wrow[warp], lrow, win + woff[warp], n_rows, begin, end, lane);
  } else
  for (int pos = begin; pos < end; ++pos) {
    const int2 jr = next_pair(pos);
    const RowState st = rs[jr.y];
    const int32_t* list = a.cand + st.base;
    int pick = -1;
    int k = st.cursor;
    while (k < st.len) {
      const int idx = k + lane;
      const int c = idx < st.len ? __ldg(list + idx) : -1;
      // warps of other components share bitmap words, never bits: the OR must be atomic, the test need not be
      const bool is_free = c >= 0 && !(reinterpret_cast<volatile uint32_t*>(used)[c >> 5] & (1u << (c & 31)));
      const uint32_t m = __ballot_sync(0xffffffffu, is_free);
      if (m) {
        const int f = __ffs(m) - 1;
        pick = __shfl_sync(0xffffffffu, c, f);
        if (lane == f) atomicOr(&used[c >> 5], 1u << (c & 31));
        k += f + 1;
        break;
      }
      k += 32;
    }
    if (lane == 0) {
      rs[jr.y].cursor = min(k, st.len);  // everything before it is claimed
      a.dec[jr.x].offer_row = pick;
    }
    placed += pick >= 0;
    __syncwarp();
  }
  if (lane == 0 && placed) atomicAdd(&placed_s, placed);
  __syncthreads();
  if (tid == 0) *a.n_placed = placed_s;
}

// ---- K2 over union lists (rows <= OE_UNION_MAX_ROWS) --------------------------------------------------------------------------
// A component's rows mostly list the SAME offers (that is what makes them a component), so per-row windows hold the same
// offers several times over, drain together and have to be struck one by one.  Here every component gets ONE ascending list of
// the offers any of its rows can use, with a membership mask per entry (bit k = the component's k-th row lists it):
//   K3u-a k_label_components   components of the rows -> CommitMeta (one warp)
//   K3u-b k_union_bits         U[w] = OR of the bit rows of component w
//   K3a-c                      (as for rows) ascending list of U[w]
//   K3u-c k_union_masks        membership mask of every list entry
// The replay then keeps a 64-entry window of that list in registers (two halves + a prefetched third) with two warp-uniform
// "still free" words; a requirement of row k ANDs them with row k's membership words (shared memory, written when a half is
// loaded), takes the lowest set bit and clears it: a handful of ALU instructions, no collective, no memory write.
#define OE_UNION_MAX_ROWS 32
#define OE_RING 4   // list chunks of 32 entries a warp keeps in flight towards shared memory (cp.async), beyond its 64-entry window
struct CommitMeta {
  int n_comp;
  int comp_of[OE_UNION_MAX_ROWS];   // row -> warp (component, folded onto 32 warps)
  int lrow[OE_UNION_MAX_ROWS];      // row -> index among its warp's rows
  int wrows[32];                    // warp -> number of rows
  int woff[33];                     // warp -> first slot in wrow[]
  int wrow[OE_UNION_MAX_ROWS];      // the rows, warp by warp
};

__global__ void k_label_components(int rows, const unsigned long long* __restrict__ adj, CommitMeta* __restrict__ meta) {
  // one warp; the labelling itself is serial (thread 0), but on a copy in shared memory: as a chain of read-modify-writes on
  // global memory it took 16 us
  __shared__ CommitMeta m;
  __shared__ unsigned long long s_adj[OE_UNION_MAX_ROWS];
  for (int r = threadIdx.x; r < rows; r += 32) s_adj[r] = adj ? adj[r] : 0ull;
  for (int q = threadIdx.x; q < (int)(sizeof(CommitMeta) / sizeof(int)); q += 32) reinterpret_cast<int*>(&m)[q] = 0;
  __syncwarp();
  if (threadIdx.x == 0) {
    int n = 0;
    unsigned long long seen = 0ull;
    for (int r = 0; r < rows; ++r) {
      if ((seen >> r) & 1ull) continue;
      unsigned long long group = 1ull << r, frontier = group;
      while (adj && frontier) {
        const int q = __ffsll((long long)frontier) - 1;
        frontier &= frontier - 1;
        const unsigned long long nb = s_adj[q] & ~group;
        group |= nb;
        frontier |= nb;
      }
      seen |= group;
      for (int q = 0; q < rows; ++q) if ((group >> q) & 1ull) m.comp_of[q] = n & 31;
      ++n;
    }
    m.n_comp = n > 32 ? 32 : n;
    for (int w = 0; w < 32; ++w) m.wrows[w] = 0;
    for (int r = 0; r < rows; ++r) m.lrow[r] = m.wrows[m.comp_of[r]]++;
    int run = 0;
    for (int w = 0; w < 32; ++w) { m.woff[w] = run; run += m.wrows[w]; }
    m.woff[32] = run;
    int fill[32];
    for (int w = 0; w < 32; ++w) fill[w] = 0;
    for (int r = 0; r < rows; ++r) { const int w = m.comp_of[r]; m.wrow[m.woff[w] + fill[w]++] = r; }
  }
  __syncwarp();
  const int* src = reinterpret_cast<const int*>(&m);
  int* dst = reinterpret_cast<int*>(meta);
  for (int q = threadIdx.x; q < (int)(sizeof(CommitMeta) / sizeof(int)); q += 32) dst[q] = src[q];
}

// thread = (union row w, word): OR of the member rows' words
__global__ void __launch_bounds__(256) k_union_bits(RowBits rb, const CommitMeta* __restrict__ meta, uint32_t* __restrict__ U, int64_t ustride) {
  const int w = blockIdx.y;
  const int64_t gw = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (gw >= rb.total_words) return;
  uint32_t x = 0u;
  const int b = meta->woff[w], e = meta->woff[w + 1];
  for (int q = b; q < e; ++q) x |= rb.word(meta->wrow[q], gw);
  U[(int64_t)w * ustride + gw] = x;
}

// thread = one list entry: which of the component's rows list this offer
__global__ void __launch_bounds__(256) k_union_masks(RowBits rb, const CommitMeta* __restrict__ meta, const int32_t* __restrict__ ucand,
                                                     const int64_t* __restrict__ ubase, const int32_t* __restrict__ ulen,
                                                     uint32_t* __restrict__ umask) {
  const int w = blockIdx.y;
  const int len = ulen[w];
  const int b = meta->woff[w], n = meta->woff[w + 1] - b;
  for (int e = blockIdx.x * 256 + threadIdx.x; e < len; e += gridDim.x * 256) {
    const int o = ucand[ubase[w] + e];
    uint32_t m = 0u;
    for (int k = 0; k < n; ++k) m |= ((rb.word(meta->wrow[b + k], o >> 5) >> (o & 31)) & 1u) << k;
    umask[ubase[w] + e] = m;
  }
}

struct UnionArgs {
  const CommitMeta* meta;
  const int32_t* ucand;      // union lists, component by component
  const uint32_t* umask;
  const int64_t* ubase;      // [32]
  const int32_t* ulen;       // [32]
  int64_t used_words;
  const int32_t* req_row;
  int64_t M;
  oe_decision* dec;
  int2* order;
  const int* comp_begin;     // [33] from k_sort_scatter
  int* n_placed;
  long long* trace;          // optional (OE_COMMIT_TRACE=1): per warp {steps, cycles, slides, top-ups, deferred, searches beyond, prologue cycles}
};

// K2a: stable counting sort of the requirements by component, two small multi-CTA kernels: per-block histograms, then every
// block ranks its own chunk (tiles of 256 requirements, in order) behind the blocks before it.  order[] = (requirement, local
// row) grouped by component, plan order kept inside a group; comp_begin[c] = where component c starts.
#define OE_SORT_BLOCKS 64
__global__ void __launch_bounds__(256) k_sort_hist(const int32_t* __restrict__ req_row, int64_t M, const CommitMeta* __restrict__ meta,
                                                   int* __restrict__ hist) {
  __shared__ int h[32];
  __shared__ int comp_of[OE_UNION_MAX_ROWS];
  if (threadIdx.x < 32) { h[threadIdx.x] = 0; comp_of[threadIdx.x] = meta->comp_of[threadIdx.x]; }
  __syncthreads();
  const int64_t chunk = ((M + OE_SORT_BLOCKS - 1) / OE_SORT_BLOCKS + 255) / 256 * 256;
  const int64_t lo = min(M, (int64_t)blockIdx.x * chunk), hi = min(M, lo + chunk);
  const int lane = threadIdx.x & 31;
  for (int64_t j0 = lo; j0 < hi; j0 += 256) {
    const int64_t j = j0 + threadIdx.x;
    const int c = j < hi ? comp_of[req_row[j]] : 32 + lane;   // past the end: a group of its own, not counted
    const uint32_t same = __match_any_sync(0xffffffffu, c);
    if (c < 32 && lane == __ffs(same) - 1) atomicAdd(&h[c], __popc(same));
  }
  __syncthreads();
  if (threadIdx.x < 32) hist[blockIdx.x * 32 + threadIdx.x] = h[threadIdx.x];
}
__global__ void __launch_bounds__(256) k_sort_scatter(const int32_t* __restrict__ req_row, int64_t M, const CommitMeta* __restrict__ meta,
                                                      const int* __restrict__ hist, int2* __restrict__ order, int* __restrict__ comp_begin_out) {
  __shared__ int run[32];          // next free slot of every component for this block
  __shared__ int wc[8][32];        // per warp of the tile: requirements of every component
  __shared__ int comp_of[OE_UNION_MAX_ROWS], lrow[OE_UNION_MAX_ROWS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid < 32) {
    comp_of[tid] = meta->comp_of[tid]; lrow[tid] = meta->lrow[tid];
    int before = 0, total = 0;   // of component `tid`: in the blocks before this one / in all blocks
    for (int b = 0; b < OE_SORT_BLOCKS; ++b) { const int v = hist[b * 32 + tid]; total += v; if (b < (int)blockIdx.x) before += v; }
    int begin = total;           // exclusive scan of the totals over the components
    for (int d = 1; d < 32; d <<= 1) { const int n = __shfl_up_sync(0xffffffffu, begin, d); if (lane >= d) begin += n; }
    begin -= total;
    run[tid] = begin + before;
    if (blockIdx.x == 0) { comp_begin_out[tid] = begin; if (tid == 31) comp_begin_out[32] = begin + total; }
  }
  __syncthreads();
  const int64_t chunk = ((M + OE_SORT_BLOCKS - 1) / OE_SORT_BLOCKS + 255) / 256 * 256;
  const int64_t lo = min(M, (int64_t)blockIdx.x * chunk), hi = min(M, lo + chunk);
  for (int64_t j0 = lo; j0 < hi; j0 += 256) {
    const int64_t j = j0 + tid;
    const int row = j < hi ? req_row[j] : 0;
    const int c = j < hi ? comp_of[row] : 32 + lane;
    wc[warp][lane] = 0;
    __syncwarp();
    const uint32_t same = __match_any_sync(0xffffffffu, c);
    if (c < 32 && lane == __ffs(same) - 1) wc[warp][c] = __popc(same);
    __syncthreads();
    if (c < 32) {
      int at = run[c] + __popc(same & ((1u << lane) - 1u));
      for (int w = 0; w < warp; ++w) at += wc[w][c];
      order[at] = make_int2((int)j, lrow[row]);
    }
    __syncthreads();
    if (tid < 32) { int add = 0; for (int w = 0; w < 8; ++w) add += wc[w][tid]; run[tid] += add; }
    __syncthreads();
  }
}

// K2: one CTA of one warp per component, each on an SM of its own (a lone warp retires a dependent instruction every few cycles;
// 23 such warps on ONE SM compete for its four schedulers and each runs at less than half that rate).
__global__ void __launch_bounds__(32, 1) k_commit_union(UnionArgs a) {
  extern __shared__ __align__(16) unsigned char commit_smem[];
  uint32_t* used = reinterpret_cast<uint32_t*>(commit_smem);   // claims made BEYOND a window (the window's own live in registers)
  int2* rings = reinterpret_cast<int2*>(commit_smem + (a.used_words * 4 + 15) / 16 * 16);   // [OE_RING][lane] (offer, mask)
  __shared__ CommitMeta meta;
  __shared__ uint2 rmem[OE_UNION_MAX_ROWS];   // per local row: which entries of the two window halves list the row
  __shared__ int bcur[OE_UNION_MAX_ROWS];     // per local row: where a search beyond the window resumes
  const int lane = threadIdx.x, warp = blockIdx.x;   // "warp" = component
  const long long t_start = clock64();
  int n_slides = 0, n_topups = 0, n_deferred = 0, n_searches = 0;
  for (int k = lane; k < (int)(sizeof(CommitMeta) / 4); k += 32) reinterpret_cast<int*>(&meta)[k] = reinterpret_cast<const int*>(a.meta)[k];
  __syncwarp();
  const int n_comp = meta.n_comp;
  if (warp >= n_comp) return;
  for (int64_t k = lane; k < (a.used_words + 3) / 4; k += 32) reinterpret_cast<uint4*>(used)[k] = make_uint4(0u, 0u, 0u, 0u);
  bcur[lane] = 0;
  __syncwarp();
  const int* comp_begin = a.comp_begin;
  // ---- replay: warp = component
  int placed = 0;
  const long long t_replay = clock64();
  const int begin = comp_begin[warp], end = comp_begin[warp + 1];
  if (begin < end) {
    volatile uint32_t* vused = reinterpret_cast<volatile uint32_t*>(used);
    const int R = meta.wrows[warp];
    uint2* rm = rmem;
    int* bc = bcur;
    const int32_t* L = a.ucand + a.ubase[warp];
    const uint32_t* Mk = a.umask + a.ubase[warp];
    const int len = a.ulen[warp];
    int rd = 64;          // list index of the first entry that is not in the window yet
    int nf = 2 + OE_RING;  // next list chunk to send for
    int u0 = lane < len ? __ldg(L + lane) : -1, u1 = 32 + lane < len ? __ldg(L + 32 + lane) : -1;
    uint32_t m0 = lane < len ? __ldg(Mk + lane) : 0u, m1 = 32 + lane < len ? __ldg(Mk + 32 + lane) : 0u;
    // The list chunks after the window travel to shared memory asynchronously, OE_RING chunks ahead (one cp.async group per
    // chunk; a lane copies, and later reads, only its own entry): when the matrix expansion runs beside this kernel a load
    // takes several microseconds, far more than the 32 steps that a one-chunk-ahead register prefetch covers.
    int2* ring = rings;
    auto fetch_chunk = [&](int chunk) {
      int2* slot = ring + (chunk & (OE_RING - 1)) * 32 + lane;
      const int idx = chunk * 32 + lane;
      if (idx < len) {
        __pipeline_memcpy_async(&slot->x, L + idx, 4);
        __pipeline_memcpy_async(&slot->y, Mk + idx, 4);
      } else {
        *slot = make_int2(-1, 0);
      }
      __pipeline_commit();
    };
    for (int c = 2; c < 2 + OE_RING; ++c) fetch_chunk(c);
    const volatile int* ring_flat = reinterpret_cast<const volatile int*>(ring);   // entry idx (< nf * 32) sits at [(idx & (OE_RING * 32 - 1)) * 2]
    auto landed = [&](int upto) {   // entries up to list index `upto` are in shared memory, for every lane
      const int newer = nf - (upto >> 5) - 1;   // groups sent for after the one that carries `upto`
      __pipeline_wait_prior(newer < 0 ? 0 : newer > 7 ? 7 : newer);
      __syncwarp();
    };
    auto advanced = [&]() {         // rd moved: chunks that lie wholly before it give their ring slots to the chunks OE_RING later
      __syncwarp();
      while (nf - OE_RING < (rd >> 5)) fetch_chunk(nf++);
    };
    auto membership = [&]() {       // rm[k] <- which entries of the two halves list row k
      uint2 mine = make_uint2(0u, 0u);
      for (int q = 0; q < R; ++q) {
        const uint32_t x = __ballot_sync(0xffffffffu, (m0 >> q) & 1u), y = __ballot_sync(0xffffffffu, (m1 >> q) & 1u);
        if (lane == q) mine = make_uint2(x, y);
      }
      if (lane < R) rm[lane] = mine;
      __syncwarp();
    };
    uint32_t free0 = __ballot_sync(0xffffffffu, u0 >= 0), free1 = __ballot_sync(0xffffffffu, u1 >= 0);
    membership();
    int2 nx = begin + lane < end ? a.order[begin + lane] : make_int2(0, 0);
    int2 nx2 = begin + 32 + lane < end ? a.order[begin + 32 + lane] : make_int2(0, 0);
    for (int b0 = begin; b0 < end; b0 += 32) {
      const int my_req = nx.x, my_k = nx.y;                       // plan position b0 + lane
      nx = nx2;
      if (b0 + 64 + lane < end) nx2 = a.order[b0 + 64 + lane];    // two batches ahead of its use
      const int cnt_b = min(32, end - b0);
      int my_o = -1;
      int my_pos = -1;  // window position taken by the requirement at b0 + lane; -2 = nothing of its row was free in the window
      // the picks made inside the window, read out of the registers
      auto read_out = [&]() {
        const int o0 = __shfl_sync(0xffffffffu, u0, my_pos & 31), o1 = __shfl_sync(0xffffffffu, u1, my_pos & 31);
        if (my_pos >= 0) { my_o = my_pos < 32 ? o0 : o1; my_pos = -1; }
      };
      // The window is topped up to 64 FREE entries: what is still free moves to the front (order kept), the list supplies the
      // rest.  Needed when the free entries left in half 0 are ones that the rows asking now do not list: half 0 never empties,
      // the window stops moving, and row after row finds nothing in it.
      auto compact = [&]() {
        const int n0 = __popc(free0), n1 = __popc(free1), keep = n0 + n1, n_new = 64 - keep;
        ++n_topups;
        landed(rd + n_new - 1);
        int nu[2];
        uint32_t nm[2];
        bool nfree[2];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int d = hh * 32 + lane;
          const int s0 = d < n0 ? (int)__fns(free0, 0, d + 1) : 0, s1 = (d >= n0 && d < keep) ? (int)__fns(free1, 0, d - n0 + 1) : 0;
          const int ua = __shfl_sync(0xffffffffu, u0, s0), ub = __shfl_sync(0xffffffffu, u1, s1);
          const uint32_t ma = __shfl_sync(0xffffffffu, m0, s0), mb = __shfl_sync(0xffffffffu, m1, s1);
          const int e = ((rd + d - keep) & (OE_RING * 32 - 1)) * 2;
          const int un_ = d >= keep ? ring_flat[e] : -1;
          const uint32_t mn_ = d >= keep ? (uint32_t)ring_flat[e + 1] : 0u;
          nu[hh] = d < n0 ? ua : d < keep ? ub : un_;
          nm[hh] = d < n0 ? ma : d < keep ? mb : mn_;
          nfree[hh] = d < keep || (un_ >= 0 && !((vused[max(un_, 0) >> 5] >> (un_ & 31)) & 1u));
        }
        u0 = nu[0]; m0 = nm[0]; u1 = nu[1]; m1 = nm[1];
        free0 = __ballot_sync(0xffffffffu, nfree[0]);
        free1 = __ballot_sync(0xffffffffu, nfree[1]);
        rd += n_new;
        advanced();
        membership();
      };
      // Requirements whose row had nothing free inside the window, in plan order.  First the window is topped up; if the row
      // still finds nothing, the list is searched beyond the window, from where the last such search of the row ended.
      auto serve_deferred = [&]() {
        uint32_t pend = __ballot_sync(0xffffffffu, my_pos == -2);
        if (!pend) return;
        read_out();
        if (rd < len && __popc(free0) + __popc(free1) < 64) compact();
        for (; pend; pend &= pend - 1) {
          const int i = __ffs(pend) - 1;
          const int k = __shfl_sync(0xffffffffu, my_k, i);
          const uint2 r = rm[k];
          const uint32_t a0 = r.x & free0, a1 = r.y & free1;
          int o = -1;
          ++n_deferred;
          if (a0 | a1) {
            const int p = __ffs(a0 ? a0 : a1) - 1;
            const int o0 = __shfl_sync(0xffffffffu, u0, p), o1 = __shfl_sync(0xffffffffu, u1, p);
            o = a0 ? o0 : o1;
            if (a0) free0 ^= 1u << p; else free1 ^= 1u << p;
          } else {
            int start = max(bc[k], rd);
            ++n_searches;
            __pipeline_wait_prior(0);
            __syncwarp();
            while (start < len) {
              const int chunk = start >> 5, first = start & 31, idx = chunk * 32 + lane;
              int u = -1;
              uint32_t m = 0u;
              if (chunk < nf) {
                const int e = (idx & (OE_RING * 32 - 1)) * 2;
                u = ring_flat[e]; m = (uint32_t)ring_flat[e + 1];
              } else if (idx < len) {
                u = __ldg(L + idx); m = __ldg(Mk + idx);
              }
              const bool ok = lane >= first && u >= 0 && ((m >> k) & 1u) && !((vused[max(u, 0) >> 5] >> (u & 31)) & 1u);
              const uint32_t b = __ballot_sync(0xffffffffu, ok);
              if (b) {
                const int p = __ffs(b) - 1;
                o = __shfl_sync(0xffffffffu, u, p);
                if (lane == p) atomicOr(&used[u >> 5], 1u << (u & 31));
                start = chunk * 32 + p + 1;
                break;
              }
              start = (chunk + 1) * 32;
            }
            if (lane == 0) bc[k] = min(start, len);
            __syncwarp();
          }
          if (lane == i) { my_o = o; my_pos = -1; }
          placed += o >= 0;
        }
      };
      // Straight-line steps, eight at a time: no branch, no collective except the shuffle that fetches the row (off the
      // chain), no memory write.  What happens rarely is only noted and dealt with between the groups of eight: a requirement
      // whose row has nothing free inside the window (it takes an entry BEYOND the window, which no step inside the window
      // can touch, so deferring it changes nothing — as long as it is served before the window moves), and moving the window
      // on once half 0 is used up (until then the steps simply draw on half 1).
      for (int i0 = 0; i0 < cnt_b; i0 += 8) {
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const int i = i0 + t;
          const int k = __shfl_sync(0xffffffffu, my_k, i);
          const uint2 r = rm[k];
          // the dependent chain from one step to the next is and -> negate -> and -> xor on free0 (one select more on free1)
          const uint32_t act = i < cnt_b ? 0xffffffffu : 0u;
          const uint32_t a0 = r.x & free0 & act, a1 = r.y & free1 & act;
          const uint32_t bit0 = a0 & (0u - a0);                      // lowest free entry of the row in half 0, if any
          const uint32_t bit1 = a0 ? 0u : a1 & (0u - a1);            // ... else in half 1
          free0 ^= bit0;
          free1 ^= bit1;
          // off the chain: where that was, for the lane whose requirement this is (bfind of a one-bit word = its position)
          const uint32_t both = bit0 | bit1;
          const int none = -1 - (int)(act & 1u);
          asm volatile(
              "{ .reg .pred q0, qa, qme; .reg .s32 pos, hi, v;\n"
              "  bfind.u32 pos, %2;\n"
              "  setp.ne.u32 q0, %3, 0;\n"
              "  selp.s32 hi, 0, 32, q0;\n"
              "  add.s32 pos, pos, hi;\n"
              "  setp.ne.u32 qa, %2, 0;\n"
              "  selp.s32 v, pos, %4, qa;\n"
              "  setp.eq.s32 qme, %1, %5;\n"
              "  selp.s32 %0, v, %0, qme; }"
              : "+r"(my_pos)
              : "r"(lane), "r"(both), "r"(bit0), "r"(none), "r"(i));
          placed += both != 0u;
        }
        serve_deferred();
        if (free0 == 0u && rd < len + 64) {
          read_out();
          do {  // half 0 is used up: half 1 -> half 0, the next 32 list entries -> half 1
            ++n_slides;
            u0 = u1; m0 = m1; free0 = free1;
            landed(rd + 31);
            const int e = ((rd + lane) & (OE_RING * 32 - 1)) * 2;
            u1 = ring_flat[e]; m1 = (uint32_t)ring_flat[e + 1];
            rd += 32;
            advanced();
            free1 = __ballot_sync(0xffffffffu, u1 >= 0 && !((vused[max(u1, 0) >> 5] >> (u1 & 31)) & 1u));
            uint32_t y = 0u;
            for (int q = 0; q < R; ++q) {
              const uint32_t v = __ballot_sync(0xffffffffu, (m1 >> q) & 1u);
              if (lane == q) y = v;
            }
            if (lane < R) rm[lane] = make_uint2(rm[lane].y, y);
            __syncwarp();
          } while (free0 == 0u && rd < len + 64);
        }
      }
      serve_deferred();
      read_out();
      if (lane < cnt_b) a.dec[my_req].offer_row = my_o;
    }
  }
  if (a.trace && lane == 0) {
    long long* d = a.trace + warp * 8;
    d[0] = end - begin; d[1] = clock64() - t_replay; d[2] = n_slides; d[3] = n_topups; d[4] = n_deferred;
    d[5] = n_searches; d[6] = t_replay - t_start;
  }
  if (lane == 0 && placed) atomicAdd(a.n_placed, placed);
}

// ---- K2 from the offers' side (any number of bit rows, at most 32 per component) ----------------------------------------------
// The greedy order of PlanScheduler.resourceOffers (requirements in plan order, each takes the lowest free offer of its row) is a
// stable matching in which every requirement ranks offers by index and every offer ranks requirements by plan position; with one
// ranking per side that matching is unique, and it is reached from either side.  From the offers' side: offers in ascending
// index order, each goes to the EARLIEST unmatched requirement whose row lists it.  (Proof sketch: let offer i list requirement j.
// If j, at its turn, finds i free, then every lower offer of its row was taken by then, so j takes i unless an earlier requirement
// that lists i did — which is "the earliest unmatched requirement that lists i" once all lower offers have been handed out.)
// Requirements of one row are interchangeable, so the earliest unmatched one of a row is the head of the row's queue, and the
// replay state is one queue position per row — no bitmap of taken offers, no windows, no search.  Per component (rows that share
// offers, K3d) one warp walks the component's offers; lane k keeps the head of the component's k-th row; a step is
//     v = offer lists my row ? my head : none;   winner = warp-wide minimum;   the winning lane records the match and advances.
// The chain from one step to the next is select -> REDUX.MIN -> compare -> select on `head`; offers and masks arrive 32 at a
// time, a chunk nobody can use costs one ballot, and a pass over a later part of the offers (the table arriving in chunks) simply
// resumes from the rows' queue positions.
#define OE_OS_MAX_ROWS 2048       // bit rows (classes + re-deployed pods) this path takes; beyond: k_commit
#define OE_OS_SORT_BLOCK 256      // items per block of the stable counting sorts (one warp ranks a block: 8 rounds of 32)

// Rows > 64: components by union-find over "rows that list the same offer".  thread = one offer.
__device__ __forceinline__ int uf_find(int* parent, int x) {
  for (;;) {
    const int px = *reinterpret_cast<volatile int*>(parent + x);
    if (px == x) return x;
    const int ppx = *reinterpret_cast<volatile int*>(parent + px);
    if (ppx != px) parent[x] = ppx;   // path halving; a lost update only costs a step
    x = px;
  }
}
__global__ void __launch_bounds__(128) k_row_link(RowBits rb, int* __restrict__ parent) {
  extern __shared__ uint32_t link_tile[];   // [rows][4]
  const int64_t gw0 = (int64_t)blockIdx.x * 4;
  for (int q = threadIdx.x; q < rb.rows * 4; q += 128) {
    const int r = q >> 2;
    const int64_t gw = gw0 + (q & 3);
    link_tile[q] = gw < rb.total_words ? rb.word(r, gw) : 0u;
  }
  __syncthreads();
  const int wq = threadIdx.x >> 5, b = threadIdx.x & 31;
  int first = -1;
  for (int r = 0; r < rb.rows; ++r) {
    const uint32_t w = link_tile[r * 4 + wq];
    if (!((w >> b) & 1u)) continue;
    if (first < 0) { first = r; continue; }
    int a = first, c = r;
    for (;;) {
      a = uf_find(parent, a); c = uf_find(parent, c);
      if (a == c) break;
      if (a > c) { const int t = a; a = c; c = t; }
      if (atomicCAS(parent + c, c, a) == c) break;
    }
  }
}

// Components -> dense ids in order of their lowest row, every row's index inside its component, the rows grouped by component.
// info[0] = components, info[1] = rows of the widest one.  One CTA; the numbering itself is serial (rows <= OE_OS_MAX_ROWS).
struct RowGroups {
  int* comp_of;         // [rows]
  int* lrow;            // [rows]
  int* comp_rows;       // [rows] rows, component by component
  int* comp_row_begin;  // [rows + 1]
  int* info;            // [2]
};
__global__ void __launch_bounds__(256) k_group_rows(int rows, const unsigned long long* __restrict__ adj, int* __restrict__ parent, RowGroups g,
                                                    const int* __restrict__ skip_ptr, int skip_at) {
  extern __shared__ int grp_smem[];
  if (skip_ptr && *skip_ptr >= skip_at) {   // a later commit pass with nothing left to place: no components, so that every kernel behind this one idles
    if (threadIdx.x == 0) { g.info[0] = 0; g.info[1] = 0; }
    return;
  }
  int* comp = grp_smem;           // [rows]
  int* count = comp + rows;       // [rows]
  int* begin = count + rows;      // [rows + 1]
  __shared__ unsigned long long s_adj[64];
  for (int r = threadIdx.x; r < rows; r += blockDim.x) { comp[r] = -1; count[r] = 0; }
  if (adj) for (int r = threadIdx.x; r < rows && r < 64; r += blockDim.x) s_adj[r] = adj[r];
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    if (!parent) {            // rows <= 64: breadth-first over the adjacency masks (or one row: one component)
      unsigned long long seen = 0ull;
      for (int r = 0; r < rows; ++r) {
        if ((seen >> r) & 1ull) continue;
        unsigned long long group = 1ull << r, frontier = group;
        while (adj && frontier) {
          const int q = __ffsll((long long)frontier) - 1;
          frontier &= frontier - 1;
          const unsigned long long nb = s_adj[q] & ~group;
          group |= nb;
          frontier |= nb;
        }
        seen |= group;
        for (int q = 0; q < rows; ++q) if ((group >> q) & 1ull) comp[q] = n;
        ++n;
      }
    } else {
      // roots carry the lowest row of their set (unions attach the higher root under the lower)
      for (int r = 0; r < rows; ++r) {
        const int root = uf_find(parent, r);
        if (comp[root] < 0) comp[root] = n++;   // root <= r: numbered when its lowest row comes by
        comp[r] = comp[root];
      }
    }
    int widest = 0;
    for (int r = 0; r < rows; ++r) { g.lrow[r] = count[comp[r]]++; g.comp_of[r] = comp[r]; }
    int run = 0;
    for (int c = 0; c < n; ++c) { begin[c] = run; run += count[c]; widest = max(widest, count[c]); }
    begin[n] = run;
    for (int c = 0; c <= n; ++c) g.comp_row_begin[c] = begin[c];
    for (int r = 0; r < rows; ++r) g.comp_rows[begin[comp[r]] + g.lrow[r]] = r;
    g.info[0] = n; g.info[1] = widest;
  }
}

// Per offer: its component and which of the component's rows list it (bit k = the component's k-th row); -1 when no row does.
// CTA = OE_OS_SORT_BLOCK offers; it also counts its offers per component for the sort that follows.
__global__ void __launch_bounds__(256) k_offer_masks(RowBits rb, const int* __restrict__ comp_of, const int* __restrict__ lrow, const int* __restrict__ info,
                                                     int32_t* __restrict__ ocomp, uint32_t* __restrict__ omask, int* __restrict__ hist) {
  extern __shared__ int om_smem[];
  const int rows = rb.rows;
  const int n_comp = info[0];
  if (n_comp == 0 && rb.skip_ptr) return;   // (k_group_rows found nothing left to place)
  int* s_comp = om_smem;                 // [rows]
  int* s_lrow = s_comp + rows;           // [rows]
  int* s_hist = s_lrow + rows;           // [rows] (n_comp <= rows)
  uint32_t* s_words = reinterpret_cast<uint32_t*>(s_hist + rows);   // [rows][8]: the 8 words of this pass of 256 offers
  for (int r = threadIdx.x; r < rows; r += 256) { s_comp[r] = comp_of[r]; s_lrow[r] = lrow[r]; s_hist[r] = 0; }
  const int64_t words_per_block = OE_OS_SORT_BLOCK / 32;
  for (int pass = 0; pass < OE_OS_SORT_BLOCK / 256; ++pass) {
    const int64_t gw0 = (int64_t)blockIdx.x * words_per_block + pass * 8;
    __syncthreads();
    for (int q = threadIdx.x; q < rows * 8; q += 256) {
      const int64_t gw = gw0 + (q & 7);
      s_words[q] = gw < rb.total_words ? rb.word(q >> 3, gw) : 0u;
    }
    __syncthreads();
    const int wq = threadIdx.x >> 5, b = threadIdx.x & 31;
    int c = -1;
    uint32_t m = 0u;
    for (int r = 0; r < rows; ++r) {
      const uint32_t w = s_words[r * 8 + wq];
      if ((w >> b) & 1u) { c = s_comp[r]; m |= 1u << (s_lrow[r] & 31); }
    }
    const int64_t gi = (gw0 + wq) * 32 + b;
    if (gw0 + wq < rb.total_words) { ocomp[gi] = c; omask[gi] = m; }
    if (c >= 0) atomicAdd(&s_hist[c], 1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < n_comp; c += 256) hist[(int64_t)c * gridDim.x + blockIdx.x] = s_hist[c];   // key-major: a key's counts over the blocks are contiguous
}

// Stable counting sort, three kernels.  keys[i] in [0, *n_keys) or negative (dropped); blocks of OE_OS_SORT_BLOCK items.
//   k_key_hist     per block: items per key                     (the offers' histogram comes from k_offer_masks instead)
//   k_key_scan     one CTA: where every (block, key) starts; key_begin[k] = start of key k, key_begin[n_keys] = items kept
//   k_key_scatter  one warp per block walks its items in order: out_idx[pos] = item, out_aux[pos] = aux[item]
__global__ void __launch_bounds__(256) k_key_hist(const int32_t* __restrict__ keys, int64_t n, int key_stride, int* __restrict__ hist) {
  extern __shared__ int kh_smem[];
  for (int k = threadIdx.x; k < key_stride; k += 256) kh_smem[k] = 0;
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * OE_OS_SORT_BLOCK;
  for (int q = threadIdx.x; q < OE_OS_SORT_BLOCK; q += 256) {
    const int64_t i = lo + q;
    if (i < n) { const int k = keys[i]; if (k >= 0) atomicAdd(&kh_smem[k], 1); }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < key_stride; k += 256) hist[(int64_t)k * gridDim.x + blockIdx.x] = kh_smem[k];
}
// CTA = one key: the key's counts over the blocks (contiguous: the histogram is key-major) become exclusive prefixes in place,
// totals[key] = its items; then one CTA scans the totals.
__global__ void __launch_bounds__(1024) k_key_scan_cols(int* __restrict__ hist, int n_blocks, const int* __restrict__ n_keys_ptr, int n_keys_fixed,
                                                        int* __restrict__ totals) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  const int n_keys = n_keys_ptr ? *n_keys_ptr : n_keys_fixed;
  const int k = blockIdx.x;
  if (k >= n_keys) return;
  int* col = hist + (int64_t)k * n_blocks;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n_blocks; b0 += 1024) {
    const int b = b0 + threadIdx.x;
    const int v = b < n_blocks ? col[b] : 0;
    int incl = v;
    for (int d = 1; d < 32; d <<= 1) { const int x = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += x; }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sum[lane], wi = w;
      for (int d = 1; d < 32; d <<= 1) { const int x = __shfl_up_sync(0xffffffffu, wi, d); if (lane >= d) wi += x; }
      warp_sum[lane] = wi - w;
    }
    __syncthreads();
    const int excl = carry + warp_sum[warp] + incl - v;
    if (b < n_blocks) col[b] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[k] = carry;
}
__global__ void __launch_bounds__(1024) k_key_scan(const int* __restrict__ totals, const int* __restrict__ n_keys_ptr, int n_keys_fixed, int* __restrict__ key_begin) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  const int n_keys = n_keys_ptr ? *n_keys_ptr : n_keys_fixed;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int k0 = 0; k0 < n_keys; k0 += 1024) {
    const int k = k0 + threadIdx.x;
    const int total = k < n_keys ? totals[k] : 0;
    int incl = total;
    for (int d = 1; d < 32; d <<= 1) { const int x = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += x; }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sum[lane], wi = w;
      for (int d = 1; d < 32; d <<= 1) { const int x = __shfl_up_sync(0xffffffffu, wi, d); if (lane >= d) wi += x; }
      warp_sum[lane] = wi - w;
    }
    __syncthreads();
    const int excl = carry + warp_sum[warp] + incl - total;
    if (k < n_keys) key_begin[k] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) key_begin[n_keys] = carry;
}
__global__ void __launch_bounds__(32) k_key_scatter(const int32_t* __restrict__ keys, int64_t n, int key_stride, const int* __restrict__ hist,
                                                    const int* __restrict__ key_begin, const uint32_t* __restrict__ aux, int32_t* __restrict__ out_idx,
                                                    uint32_t* __restrict__ out_aux, const int* __restrict__ n_keys_gate = nullptr) {
  extern __shared__ int ks_run[];   // [key_stride]: next slot of every key for this block
  if (n_keys_gate && *n_keys_gate == 0) return;   // no keys at all (a skipped commit pass): `keys` was not written either
  const int lane = threadIdx.x;
  const int64_t lo = (int64_t)blockIdx.x * OE_OS_SORT_BLOCK;
  // only the keys this block holds are looked at, so only those are initialised: first touch below
  for (int q = 0; q < OE_OS_SORT_BLOCK; q += 32) {
    const int64_t i = lo + q + lane;
    const int k = i < n ? keys[i] : -1;
    if (k >= 0) ks_run[k] = -1;
  }
  __syncwarp();
  for (int q = 0; q < OE_OS_SORT_BLOCK; q += 32) {
    const int64_t i = lo + q + lane;
    const int k = i < n ? keys[i] : -1;
    const uint32_t same = __match_any_sync(0xffffffffu, k);
    if (k >= 0) {
      const int leader = __ffs(same) - 1;
      int base = 0;
      if (lane == leader) {
        base = ks_run[k];
        if (base < 0) base = key_begin[k] + hist[(int64_t)k * gridDim.x + blockIdx.x];
        ks_run[k] = base + __popc(same);
      }
      base = __shfl_sync(same, base, leader);
      const int pos = base + __popc(same & ((1u << lane) - 1u));
      out_idx[pos] = (int32_t)i;
      if (out_aux) out_aux[pos] = aux[i];
    }
    __syncwarp();
  }
}

// Decisions before the replay: no offer.
__global__ void k_clear_decisions(oe_decision* __restrict__ dec, int64_t M) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  uint4* p = reinterpret_cast<uint4*>(dec + j);
  p[0] = make_uint4(0xffffffffu, 0u, 0u, 0u);   // offer_row = -1
  p[1] = make_uint4(0u, 0u, 0u, 0u);
}

struct OfferSideArgs {
  const int* info;             // [0] = components
  const int* comp_begin;       // [components + 1] into list / mask (offers sorted by component, ascending inside)
  const int32_t* list;         // offer index relative to the pass's first offer
  const uint32_t* mask;
  const int* comp_rows;
  const int* comp_row_begin;
  const int* row_begin;        // [rows + 1] into queue (requirements sorted by row, plan order inside)
  const int32_t* queue;
  int* row_cursor;             // [rows] requirements of the row matched so far (carried from pass to pass)
  int32_t offer_offset;        // global index of the pass's first offer
  oe_decision* dec;
  int* n_placed;
  long long* trace;            // optional: per component {steps, cycles, chunks, chunks skipped}
};
#define OE_OS_NONE 0x7fffffff
#define OE_OS_RING 128    // queue entries a lane keeps staged in shared memory (topped up to 96 ahead at every chunk of 32 offers)
#define OE_OS_MIRROR 36   // slots mirrored behind the ring's end (a chunk reads at most 34 consecutive entries)
#define OE_OS_LANE_SLOTS (OE_OS_RING + OE_OS_MIRROR)   // 164 words per lane: 16-byte aligned rows; lanes 8 apart share banks
// Shared memory through explicit 32-bit shared-window addresses: written as array accesses beside cp.async the compiler re-derived
// the window base (S2R CgaCtaId + LEA) inside the replay loop.
__device__ __forceinline__ int lds_s32(uint32_t addr) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void cp_async_4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__global__ void __launch_bounds__(32, 1) k_commit_offers(OfferSideArgs a) {
  // the lanes' queues, staged: ring[(k & (OE_OS_RING - 1)) * 32 + lane] = entry k of the lane's row.  A lane matches at most 32
  // requirements per chunk of 32 offers, so topping the ring up once per chunk (cp.async, a chunk ahead of its use) keeps every
  // read inside a chunk in shared memory; what a step needs from memory is then one LDS whose result is used a step later.
  __shared__ __align__(16) int ring[OE_OS_LANE_SLOTS * 32];
  const int comp = blockIdx.x, lane = threadIdx.x;
  if (comp >= a.info[0]) return;
  const long long t0 = clock64();
  const int rb0 = a.comp_row_begin[comp], n_rows = a.comp_row_begin[comp + 1] - rb0;
  const int row = lane < n_rows ? a.comp_rows[rb0 + lane] : -1;
  int pos = 0, end = 0;   // queue index of the lane's head / end of its queue
  if (row >= 0) { pos = a.row_begin[row] + a.row_cursor[row]; end = a.row_begin[row + 1]; }
  const int pos0 = pos;
  const int32_t* __restrict__ Q = a.queue;
  const int e0 = a.comp_begin[comp], e1 = a.comp_begin[comp + 1];
  int* const dec_row = &a.dec[0].offer_row;
  constexpr size_t DEC_INTS = sizeof(oe_decision) / sizeof(int);
  int steps = 0, chunks = 0, skipped = 0;
  if (n_rows == 1) {
    // One row: nothing to arbitrate — the k-th offer of the component goes to the k-th waiting requirement (every offer of the
    // list is one the row lists).  No chain at all: 32 matches per iteration, loads of several iterations in flight.
    const int p0 = __shfl_sync(0xffffffffu, pos, 0), pend = __shfl_sync(0xffffffffu, end, 0);   // lane 0 holds the row
    const int n = min(e1 - e0, pend - p0);
#pragma unroll 4
    for (int k = lane; k < n; k += 32) dec_row[(size_t)__ldg(Q + p0 + k) * DEC_INTS] = __ldg(a.list + e0 + k) + a.offer_offset;
    if (lane == 0) pos += max(n, 0);
    steps = max(n, 0);
  } else {
    // ring: lane-major, entry k of the lane's queue (k = its index in the queue array) at slot k & 127 of the lane's ring; slots
    // 0..35 are mirrored behind slot 127, so the 34 entries a chunk can touch are contiguous wherever the chunk starts and a
    // step advances its read address by one slot, no wrap.  Topped up four entries (16 bytes) per cp.async.
    const uint32_t ring_base = (uint32_t)__cvta_generic_to_shared(ring) + lane * (uint32_t)(OE_OS_LANE_SLOTS * 4);
    int filled = pos & ~3;   // entries below `filled` are staged (or on their way); groups of four, aligned in the queue array
    // One group of four entries (16 bytes, aligned in the queue array) per call, branch-free (a taken branch costs a lone warp ~19
    // cycles: the former loop was 260 cycles per chunk): the calls are spread over the steps of a chunk.  Entries behind the
    // queue's end are never staged (the array is padded, so the last group may read up to three entries of whatever follows);
    // the step replaces what it reads there by "nobody waiting" (ra_end below).
    int upto = pos + 96;
    auto stage_group = [&]() {
      const int go = filled < upto && filled < end;
      const uint32_t slot = (uint32_t)filled & (OE_OS_RING - 1);
      const uint32_t dst = ring_base + slot * 4u;
      const int go_mirror = go && slot < OE_OS_MIRROR;
      asm volatile("{ .reg .pred p; setp.ne.b32 p, %2, 0; @p cp.async.cg.shared.global [%0], [%1], 16; }" ::"r"(dst), "l"(Q + filled), "r"(go) : "memory");
      asm volatile("{ .reg .pred p; setp.ne.b32 p, %2, 0; @p cp.async.cg.shared.global [%0], [%1], 16; }" ::"r"(dst + OE_OS_RING * 4u), "l"(Q + filled), "r"(go_mirror) : "memory");
      filled += go ? 4 : 0;
    };
    auto top_up = [&]() {   // (before the first chunk: everything at once)
      upto = pos + 96;
      for (int g = 0; g < 26; ++g) stage_group();
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    top_up();
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    int h = lds_s32(ring_base + ((uint32_t)pos & (OE_OS_RING - 1)) * 4u);
    int n1 = lds_s32(ring_base + ((uint32_t)(pos + 1) & (OE_OS_RING - 1)) * 4u);
    if (pos >= end) h = OE_OS_NONE;        // (also the lanes without a row: pos = end = 0)
    if (pos + 1 >= end) n1 = OE_OS_NONE;
    int o_next = e0 + lane < e1 ? __ldg(a.list + e0 + lane) : -1;
    uint32_t m_next = e0 + lane < e1 ? __ldg(a.mask + e0 + lane) : 0u;
    // Which offers of a chunk list MY row: the 32 x 32 bit matrix (lane = offer, bit = row) transposed with five butterfly
    // exchanges, so that a step needs no shuffle to learn it (bit e of lane r <- bit r of lane e).  The exchanges for the NEXT
    // chunk are spread over the steps of this one — the step's chain leaves most issue slots idle — and so is the ring's top-up.
    auto butterfly = [&](uint32_t x, int j) -> uint32_t {
      const uint32_t keep = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u : 0x55555555u;   // bit positions with bit j clear
      const uint32_t other = __shfl_xor_sync(0xffffffffu, x, j);
      return (lane & j) ? ((x & ~keep) | ((other >> j) & keep)) : ((x & keep) | ((other << j) & ~keep));
    };
    auto transposed = [&](uint32_t x) -> uint32_t {
#pragma unroll
      for (int j = 16; j >= 1; j >>= 1) x = butterfly(x, j);
      return x;
    };
    uint32_t mybits_next = transposed(m_next);
    for (int base = e0; base < e1; base += 32) {
      const int o_mine = o_next + a.offer_offset;
      const uint32_t m_mine = m_next;
      const uint32_t mybits = mybits_next;
      {  // the chunk after this one, a chunk ahead of its use
        const int e = base + 32 + lane;
        o_next = e < e1 ? __ldg(a.list + e) : -1;
        m_next = e < e1 ? __ldg(a.mask + e) : 0u;
      }
      const uint32_t active = __ballot_sync(0xffffffffu, h != OE_OS_NONE);   // rows that still have a requirement waiting
      if (!active) break;
      const uint32_t want = __ballot_sync(0xffffffffu, (m_mine & active) != 0u);
      ++chunks;
      skipped += want == 0u;
      if (!want) { mybits_next = transposed(m_next); continue; }
      asm volatile("cp.async.wait_group 0;" ::: "memory");   // the entries asked for a chunk ago (a lane reads only what it staged itself)
      // 32 steps, one per offer of the chunk, straight line (a lone warp pays ~4 cycles per instruction issued, so the step is
      // counted in instructions).  Chain from step to step: select (my row listed ? head : none) -> REDUX.MIN -> compare ->
      // select (head <- n1).  The entry behind n1 is read at the top of a step and moves into n1 at the top of the NEXT one,
      // so its load never stalls the chain; the match is kept by the lane that holds the offer and stored after the chunk.
      const uint32_t ra0 = ring_base + ((uint32_t)(pos + 2) & (OE_OS_RING - 1)) * 4u;
      uint32_t ra = ra0;          // shared address of entry pos + 2
      const int ra_end = (int)ra0 + (end - (pos + 2)) * 4;   // address behind the queue's last entry (at or below ra0: nothing left)
      int match = OE_OS_NONE;     // requirement that gets MY offer of this chunk
      bool won_prev = false;
      int nxt_prev = OE_OS_NONE;
      uint32_t tb = 0u;           // the next chunk's masks on their way through the butterfly
      upto = pos + 96;
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        if (e < 24 && e % 3 == 1) stage_group();    // eight groups: what the chunk before consumed (at most 32 entries)
        if (e == 24) asm volatile("cp.async.commit_group;" ::: "memory");
        if (e == 17) tb = butterfly(m_next, 16);    // (m_next was asked for ~700 cycles ago)
        if (e == 20) tb = butterfly(tb, 8);
        if (e == 23) tb = butterfly(tb, 4);
        if (e == 26) tb = butterfly(tb, 2);
        if (e == 29) tb = butterfly(tb, 1);
        n1 = won_prev ? nxt_prev : n1;
        int nxt = lds_s32(ra);
        nxt = (int)ra < ra_end ? nxt : OE_OS_NONE;   // behind the queue's end: nobody waiting
        const bool listed = (mybits >> e) & 1u;
        const int v = listed ? h : OE_OS_NONE;
        const int mn = __reduce_min_sync(0xffffffffu, v);   // the requirement that gets offer e, known to every lane
        const bool won = listed & (v == mn);   // (nobody waiting at all: the listed rows "win" nothing — h stays none, pos is clamped below)
        match = lane == e ? mn : match;
        h = won ? n1 : h;
        ra += won ? 4u : 0u;
        won_prev = won;
        nxt_prev = nxt;
      }
      n1 = won_prev ? nxt_prev : n1;
      mybits_next = tb;
      pos = min(pos + (int)((ra - ra0) >> 2), end);
      steps += 32;
      if (match != OE_OS_NONE) dec_row[(size_t)match * DEC_INTS] = o_mine;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  int placed = pos - pos0;
  if (row >= 0) a.row_cursor[row] = pos - a.row_begin[row];
  for (int d = 16; d > 0; d >>= 1) placed += __shfl_xor_sync(0xffffffffu, placed, d);
  if (lane == 0 && placed) atomicAdd(a.n_placed, placed);
  if (a.trace && lane == 0) {
    long long* d = a.trace + (long long)comp * 4;
    d[0] = steps; d[1] = clock64() - t0; d[2] = chunks; d[3] = skipped;
  }
}

// Commit in two passes (the offer table still arriving, see oe_evaluate_cycle): the requirements that the pass over the prefix
// left without an offer, in plan order.  sub_idx[k] = requirement, sub_row[k] = its bit row, *count = how many.
__global__ void __launch_bounds__(1024) k_count_unplaced(const oe_decision* __restrict__ dec, int64_t M, int* __restrict__ tile_open) {
  const int64_t j = (int64_t)blockIdx.x * 1024 + threadIdx.x;
  const int n = __syncthreads_count(j < M && dec[j].offer_row < 0);
  if (threadIdx.x == 0) tile_open[blockIdx.x] = n;
}
__global__ void __launch_bounds__(1024) k_collect_unplaced(const oe_decision* __restrict__ dec, const int32_t* __restrict__ req_row, int64_t M,
                                                           const int* __restrict__ tile_open, int32_t* __restrict__ sub_idx, int32_t* __restrict__ sub_row,
                                                           int* __restrict__ count) {
  // tile = 1024 requirements; its base = the open ones of the tiles before it (every CTA adds those up itself)
  __shared__ int warp_tot[32];
  __shared__ int base_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int part = 0;
  for (int t = threadIdx.x; t < (int)blockIdx.x; t += 1024) part += tile_open[t];
  for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
  if (lane == 0) warp_tot[warp] = part;
  __syncthreads();
  if (threadIdx.x == 0) { int b = 0; for (int w = 0; w < 32; ++w) b += warp_tot[w]; base_s = b; }
  __syncthreads();
  const int64_t j = (int64_t)blockIdx.x * 1024 + threadIdx.x;
  const bool open = j < M && dec[j].offer_row < 0;
  const uint32_t b = __ballot_sync(0xffffffffu, open);
  __syncthreads();
  if (lane == 0) warp_tot[warp] = __popc(b);
  __syncthreads();
  int before = base_s;
  for (int w = 0; w < warp; ++w) before += warp_tot[w];
  if (open) {
    const int k = before + __popc(b & ((1u << lane) - 1u));
    sub_idx[k] = (int32_t)j;
    sub_row[k] = req_row[j];
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 1023) *count = before + __popc(b);
}
// ... and what the pass over the rest of the offers found for them (its offer indices start at `offset`).
__global__ void k_merge_unplaced(const oe_decision* __restrict__ dec_sub, const int32_t* __restrict__ sub_idx, int n, int32_t offset, oe_decision* __restrict__ dec) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int32_t o = dec_sub[k].offer_row;
  if (o >= 0) dec[sub_idx[k]].offer_row = o + offset;
}

// K2b: one thread per requirement re-runs the winner's pipeline with recording on (ports, MOUNT slots, new ids).
// `mismatch` receives requirement + 1 when the recorded re-run of a winner fails although the feasibility pass (folded form
// or rule table) set its bit: the two formulations disagree, and the decision's ports / MOUNT slots cannot be trusted —
// oe_evaluate_cycle turns that into OE_ERR_STATE instead of returning them.
// Sharded engines with peers attached (push.n > 0): only the shard that holds the winning offer can fill the decision in; it
// also stores the decision and the offer's placement columns (what a launch records, LaunchEvaluationStage.java:58-68) into
// every shard's exchange region, record j of the pass's parity: [8 words oe_decision][agent, host, zone, region, A attrs, Q rr].
struct DecisionPush {
  uint32_t* area[OE_PEER_MAX_WORLD];   // every shard's decision records of this pass (null when not sharded)
  int n, rec_words;
};
__global__ void k_materialize(DevCtx cx, const int32_t* __restrict__ req_row, int64_t M, oe_decision* __restrict__ dec, int* __restrict__ mismatch,
                              DecisionPush push) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  oe_decision d = dec[j];
  d.n_new_reservations = 0;
  for (int q = 0; q < OE_MAX_DYN_PORTS; ++q) d.dyn_port[q] = 0;
  for (int q = 0; q < OE_MAX_MOUNT_OPS; ++q) d.mount_slot[q] = -1;
  d.reserved0 = 0;
  const int64_t i = (int64_t)d.offer_row - cx.o.row_offset;
  const bool local = d.offer_row >= 0 && i >= 0 && i < cx.o.N;
  if (local) {
    const int row = req_row[j];
    const int var = row < cx.p.C ? -1 : row - cx.p.C;
    const int cls = var < 0 ? row : cx.v.cls[var];
    OfferKeys ok{cx.o.agent[i], cx.o.host[i], cx.o.zone[i], cx.o.region[i]};
    const bool pass = eval_pair<true>(cx, i, ok, cls, var, &d);
    if (!pass) atomicCAS(mismatch, 0, (int)min(j + 1, (int64_t)INT_MAX));  // the commit only picks offers whose bit is set
  }
  dec[j] = d;
  if (push.n > 0 && local) {
    uint32_t w[8];
    memcpy(w, &d, sizeof d);
    for (int p = 0; p < push.n; ++p) {
      uint32_t* rec = push.area[p] + (int64_t)j * push.rec_words;
      reinterpret_cast<uint4*>(rec)[0] = make_uint4(w[0], w[1], w[2], w[3]);
      reinterpret_cast<uint4*>(rec)[1] = make_uint4(w[4], w[5], w[6], w[7]);
      rec[8] = (uint32_t)cx.o.agent[i]; rec[9] = (uint32_t)cx.o.host[i]; rec[10] = (uint32_t)cx.o.zone[i]; rec[11] = (uint32_t)cx.o.region[i];
      for (int q = 0; q < cx.o.A; ++q) rec[12 + q] = (uint32_t)cx.o.attr[(int64_t)q * cx.o.N + i];
      for (int q = 0; q < cx.o.Q; ++q) rec[12 + cx.o.A + q] = (uint32_t)cx.o.rr[(int64_t)q * cx.o.N + i];
    }
  }
}
// ... and the other half: the decisions whose winner lives on another shard, read from this shard's own region once every
// shard's records of the pass have landed (k_peer_wait on the decision flags).
__global__ void k_gather_decisions(oe_decision* __restrict__ dec, int64_t M, const uint32_t* __restrict__ area, int rec_words, int64_t row_offset, int64_t N) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const int32_t o = dec[j].offer_row;
  const int64_t i = (int64_t)o - row_offset;
  if (o < 0 || (i >= 0 && i < N)) return;
  const uint4* rec = reinterpret_cast<const uint4*>(area + j * rec_words);
  uint4* out = reinterpret_cast<uint4*>(dec + j);
  out[0] = rec[0]; out[1] = rec[1];
}

// Kdbg: failed-stage mask of every (requirement in [j0, j1), offer) pair — what OfferEvaluator.java:173-214 feeds into the
// offer-outcome trackers.  thread = offer, blockIdx.y = requirement.
__global__ void __launch_bounds__(128) k_stage_failures(DevCtx cx, const int32_t* __restrict__ req_row, int64_t j0,
                                                        unsigned long long* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= cx.o.N) return;
  const int64_t j = j0 + blockIdx.y;
  const int row = req_row[j];
  const int var = row < cx.p.C ? -1 : row - cx.p.C;
  const int cls = var < 0 ? row : cx.v.cls[var];
  OfferKeys ok{cx.o.agent[i], cx.o.host[i], cx.o.zone[i], cx.o.region[i]};
  uint64_t failed = 0;
  eval_pair<false>(cx, i, ok, cls, var, nullptr, &failed);
  out[(int64_t)blockIdx.y * cx.o.N + i] = failed;
}

// Kgc: reservation garbage scan.  k_mark_taken flags the offers the last cycle accepted; k_unexpected tests every reserved
// side-table entry of every offer row against the keep set.
__global__ void k_mark_taken(const oe_decision* __restrict__ dec, int64_t M, int64_t row_offset, int64_t N, uint32_t* __restrict__ taken) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const int64_t i = (int64_t)dec[j].offer_row - row_offset;
  if (dec[j].offer_row >= 0 && i >= 0 && i < N) atomicOr(taken + (i >> 5), 1u << (i & 31));
}
__global__ void k_unexpected(DevOffers o, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ taken,
                             uint8_t* __restrict__ out, unsigned long long* __restrict__ n_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int n = 0;
  if (i < o.N && !(taken && ((taken[i >> 5] >> (i & 31)) & 1u))) {
    for (int64_t q = o.rbegin[i]; q < o.rbegin[i + 1]; ++q) {
      const int32_t id = o.rid[q];
      const bool unexpected = !((keep[id >> 5] >> (id & 31)) & 1u);
      out[q] = unexpected;
      n += unexpected;
    }
  }
  for (int d = 16; d > 0; d >>= 1) n += __shfl_xor_sync(0xffffffffu, n, d);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(n_out, (unsigned long long)n);
}

// K4: multi-cycle replay.  One thread per stored task: copies the winning offer's placement columns into a task row.
struct LaunchArgs {
  int64_t n;
  const int32_t* dst_row;   // task row to (over)write
  const int32_t* req;       // requirement that placed it
  const int32_t* name_id;
  const uint32_t* filter;   // [n][FW]
  const int32_t* req_type;
  const int32_t* req_index;
};
// the same on a sharded engine: the winning offer may live on another shard, its placement columns come from the decision
// record that shard pushed into this shard's exchange region (k_materialize)
__global__ void k_apply_launches_records(LaunchArgs a, DevTasks t, const uint32_t* __restrict__ area, int rec_words, int A, int Q) {
  const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= a.n) return;
  const int32_t j = a.req[x];
  const uint32_t* rec = area + (int64_t)j * rec_words + 8;
  const int64_t r = a.dst_row[x];
  t.name[r] = a.name_id[x];
  t.type[r] = a.req_type[j];
  t.index[r] = a.req_index[j];
  t.agent[r] = (int32_t)rec[0]; t.host[r] = (int32_t)rec[1]; t.zone[r] = (int32_t)rec[2]; t.region[r] = (int32_t)rec[3];
  for (int q = 0; q < t.TA; ++q) t.attr[(int64_t)q * t.cap + r] = q < A ? (int32_t)rec[4 + q] : -1;
  for (int q = 0; q < t.Q; ++q) t.rr[(int64_t)q * t.cap + r] = q < Q ? (int32_t)rec[4 + A + q] : -1;
  for (int q = 0; q < t.FW; ++q) t.filter[(int64_t)q * t.cap + r] = a.filter[x * t.FW + q];
}
__global__ void k_apply_launches(LaunchArgs a, DevOffers o, DevTasks t, const oe_decision* __restrict__ dec) {
  const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= a.n) return;
  const int32_t j = a.req[x];
  const int64_t i = (int64_t)dec[j].offer_row - o.row_offset;
  const int64_t r = a.dst_row[x];
  t.name[r] = a.name_id[x];
  t.type[r] = a.req_type[j];
  t.index[r] = a.req_index[j];
  t.agent[r] = o.agent[i];
  t.host[r] = o.host[i];
  t.zone[r] = o.zone[i];
  t.region[r] = o.region[i];
  for (int q = 0; q < t.TA; ++q) t.attr[(int64_t)q * t.cap + r] = q < o.A ? o.attr[(int64_t)q * o.N + i] : -1;
  for (int q = 0; q < t.Q; ++q) t.rr[(int64_t)q * t.cap + r] = q < o.Q ? o.rr[(int64_t)q * o.N + i] : -1;
  for (int q = 0; q < t.FW; ++q) t.filter[(int64_t)q * t.cap + r] = a.filter[x * t.FW + q];
}

// ---------------------------------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------------------------------
namespace {

thread_local std::string g_create_error;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return static_cast<T*>(p); }
};

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace

struct oe_engine {
  int device = 0, rank = 0, world = 1;
  uint32_t flags = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t aux_stream = nullptr;  // lowest priority: K1x runs here beside the commit kernel (neither needs the other's output)
  cudaStream_t hi_stream = nullptr;   // highest priority: the commit kernel, so that its one CTA is placed before K1x floods the SMs
  cudaEvent_t ev_commit = nullptr;
  bool fork_expand = false;           // evaluate_cycle -> run_commit: start K1x at the point where the commit kernel is launched
  cudaEvent_t ev_fork = nullptr, ev_x0 = nullptr, ev_x1 = nullptr;
  bool own_stream = false;
  std::string err;
  int smem_optin = 0, commit_static_smem = 0, union_static_smem = 0;

  // ---- programs (host copies kept for validation, variant detection and launch bookkeeping)
  bool hasP = false, hasT = false, hasO = false, hasR = false;
  oe_dict_sizes dict{};
  int C = 0, n_csets = 0, n_tsets = 0;
  std::vector<int32_t> h_op_begin;
  std::vector<oe_stage_op> h_ops;
  std::vector<oe_rule_node> h_nodes;
  std::vector<oe_counter_set> h_csets;
  std::vector<int64_t> h_cs_off;
  std::vector<uint8_t> class_self_sensitive, h_rr_used;   // h_rr_used[c]: a RoundRobin rule reads counter set c's {min, max, #keys}
  bool any_rr_used = false;
  int64_t max_rr_keys = 0;
  std::vector<int32_t> h_pod_role, h_excl_begin;
  std::vector<uint16_t> h_excl;
  int n_rule_nodes = 0;
  int fast_K = -1, fast_S = -1;  // offer-table shape the folded class descriptors were built for (-1 = stale)
  bool fast_port_lists = false;  // ... and whether the table's distinct `ports` range lists were usable then
  int tab_words = 0, tab_zdim = 1, tab_rdim = 1;
  bool any_rule_tab = false, all_lean = false, any_ports = false, any_rule = false;
  DevBuf d_fast, d_rule_tab, d_rr_used, d_lean, d_ruleT, d_portT;
  DevBuf d_op_begin, d_ops, d_nodes, d_pod_role, d_excl_begin, d_excl, d_rs_begin, d_rs, d_csets, d_cs_off, d_ts_type, d_mbits;
  DevProg dp{};

  // ---- tasks
  DevTasks dt{};
  DevBuf d_tcols[7], d_tattr, d_trr, d_tfilter;
  std::vector<int32_t> h_tname, h_ttype, h_tindex;  // host mirror of what apply_launches / self detection need
  std::unordered_map<int32_t, int32_t> name_to_row;
  bool name_map_built = false;

  // ---- offers
  DevOffers dof{};
  int64_t N = 0, N_global = 0, row_offset = 0, shard_rows = 0, n_reserved = 0;
  DevBuf d_scalar, d_present, d_port, d_msize, d_mprof, d_ocols[5], d_oattr, d_orr, d_oexec, d_rbegin, d_rid, d_rkind, d_rval;
  // oe_load_offers_packed: the narrow columns as they came over the bus (widened into the buffers above on the device), the
  // table's distinct `ports` range lists, every (slot, offer)'s list id, and the classes' port verdicts per list
  DevBuf d_wire[12], d_ptuples, d_ptuple_id, d_port_tab;
  int32_t n_port_tuples = 0;
  bool tuples_valid = false;
  DevBuf d_tuple_hash, d_ntuples;   // the range-list dictionary kept up while rows are replaced (k_reintern_tuples)
  int32_t ptuple_cap = 0;
  uint32_t tuple_mask = 0;
  bool tuple_count_pending = false; // h_os_info[2] will hold the dictionary's size after the last oe_update_offers

  // ---- requirements
  int64_t M = 0;
  int V = 0;
  std::vector<int32_t> h_class, h_rtype, h_rindex, h_req_row;
  std::vector<int32_t> h_launch_begin, h_launch_name, h_self_begin, h_self_rows, h_vcls, h_vself_begin, h_vself_rows;
  bool validation_pending = false;
  bool self_provided = false;
  std::vector<uint32_t> h_launch_filter;
  // OE_CFG_ASYNC_LOADS: host mirrors that nobody waits for are filled from the caller's tables inside the next evaluation call
  // (the tables stay valid until it returns); re-deployed pods are looked for on the device (k_find_twins) and the host-side search
  // only runs when there are any
  const int32_t *pending_t_name = nullptr, *pending_t_type = nullptr, *pending_t_index = nullptr;
  int64_t pending_T = -1;                     // >= 0: h_tname / h_ttype / h_tindex not yet copied
  // OE_CFG_ASYNC_LOADS: the host copies above are filled only when something reads them (oe_apply_launches, oe_requirement_rows,
  // the search for re-deployed pods): from the caller's tables while those are promised to be valid (until the evaluation call
  // after the load returns), from the device columns afterwards.  A plain cycle never touches them.
  bool tasks_caller_valid = false, reqs_caller_valid = false;
  DevBuf d_lr_begin, d_lr_name, d_lr_filter;  // launch rows of the requirement table, kept on the device for the same reason
  int64_t pending_launch_rows = -1;           // >= 0: h_launch_* not yet filled, this many rows in d_lr_name / d_lr_filter
  const int32_t *pending_r_class = nullptr, *pending_r_type = nullptr, *pending_r_index = nullptr;
  bool reqs_pending = false, twins_unresolved = false;
  DevBuf d_pod_hash, d_pod_head, d_pod_next, d_sensitive, d_twin, d_pairs, d_rclass, d_twin_j, d_row_of_variant;
  std::vector<int32_t> h_twin_j;              // requirement of variant v (device-found variants; empty otherwise)
  std::vector<int4> h_pairs;
  int pairs_cap = 0;
  bool req_row_eff_on_host = false;
  DevBuf d_req_row, d_rtype, d_rindex, d_vcls, d_vself_begin, d_vself_rows, d_vrrsum;
  // Re-deployed requirements whose rule verdicts over (zone, region) are identical share one bit row: V_eff rows behind the
  // classes instead of V (-1 = not merged: every such requirement has a row of its own).  row_map[bit row] = factored row it is
  // evaluated as; req_row_eff[j] = bit row of requirement j.
  int V_eff = -1;
  std::vector<int32_t> h_row_map, h_req_row_eff;
  std::vector<uint32_t> h_rule_tab;
  DevBuf d_row_map, d_req_row_eff;
  const int32_t* bit_row_of_req() const { return V_eff >= 0 ? d_req_row_eff.as<int32_t>() : d_req_row.as<int32_t>(); }

  // ---- derived / outputs
  DevBuf d_cnt, d_rrsum, d_type_total, d_type_agent, d_F, d_Fglobal, d_matrix, d_dec, d_used, d_cursor, d_err, d_nplaced, d_order, d_adj, d_dbg, d_keep, d_gc_out, d_taken;
  DevBuf d_l_dst, d_l_req, d_l_name, d_l_filter, d_tile_cnt, d_row_base, d_row_len, d_cand;
  DevBuf d_meta, d_union, d_umask;   // commit over union lists (rows <= OE_UNION_MAX_ROWS)
  DevBuf d_sort_hist, d_comp_begin;
  // commit from the offers' side: row groups, per-offer (component, mask), offers sorted by component, requirements sorted by row
  DevBuf d_os_totals, d_os_parent, d_os_groups, d_os_ocomp, d_os_omask, d_os_ohist, d_os_comp_begin, d_os_list, d_os_mask, d_os_rhist, d_os_row_begin, d_os_queue, d_os_cursor;
  int* h_os_info = nullptr;          // pinned: {components, rows of the widest component} of the last grouping
  bool want_queues = false;          // set by oe_evaluate_cycle around its feasibility pass
  int queues_rows = -1;              // rows the requirement queues on the side stream were built for (-1: not built for this pass)
  cudaEvent_t ev_queues = nullptr;
  std::vector<uint64_t> h_vfilter;   // rebuild_variants scratch, kept between calls (no allocation per cycle)
  std::vector<int32_t> h_vcand, h_vspan;
  bool task_spans_valid = false, task_odd_type = false, req_spans_valid = false;
  DevBuf d_stage, d_stage_rows;      // oe_update_offers: the changed rows on their way into the resident table
  bool counters_fresh = false, class_bits_fresh = false, cycle_done = false;
  std::vector<oe_decision> h_last_dec;
  cudaEvent_t ev[8]{};
  cudaEvent_t ev_variants = nullptr, ev_side = nullptr;
  cudaStream_t side_stream = nullptr;

  // ---- OE_CFG_ASYNC_LOADS: the offer table arrives on a copy stream of its own, in row chunks; a chunk is checked as soon as
  // it has landed and the feasibility pass evaluates it while the next ones are still on the bus
  static constexpr int kMaxChunks = 8;
  struct OfferChunk { int64_t r0 = 0, r1 = 0; cudaEvent_t landed = nullptr; };
  OfferChunk chunks[kMaxChunks];
  int n_chunks = 1;                  // 1 = the whole table at once
  bool offers_in_flight = false;     // chunk events are pending on copy_stream and nobody has waited for them yet
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_main = nullptr, ev_k1[kMaxChunks]{};
  DevBuf d_offer_err;                // one flag per chunk, read back into h_offer_err behind the chunk's checks
  int* h_offer_err = nullptr;        // pinned
  DevBuf d_sub_idx, d_sub_row, d_dec_sub, d_sub_count, d_tile_open;   // commit in two passes: the requirements left open by the prefix pass
  cudaEvent_t ev_prefix = nullptr, ev_phase_a = nullptr, ev_k1_last = nullptr;
  char* pin_arena = nullptr;         // pinned staging for the engine's own (pageable) host tables while offers are in flight
  size_t pin_cap = 0, pin_used = 0;
  int64_t expand_from_word = 0;      // evaluate_cycle: words below this were expanded chunk by chunk beside the pass
  bool expand_started = false;

  // ---- peer exchange (world > 1): [header: one flag word per shard][2 parities][world][exch_max_rows][exch_stride] words
  DevBuf d_exch, d_peer_err;
  int exch_max_rows = 0;
  int64_t exch_stride = 0;
  uint32_t* peer_base[OE_PEER_MAX_WORLD]{};   // every shard's exchange region, as this process addresses it
  bool peer_opened[OE_PEER_MAX_WORLD]{};      // mapped with cudaIpcOpenMemHandle (to be closed), as opposed to a same-process pointer
  bool peers_attached = false;
  uint32_t peer_epoch = 0;                    // passes published so far; parity = peer_epoch & 1
  unsigned long long peer_timeout_ns = 10ull * 1000 * 1000 * 1000;
  // ... followed by the decision records [2 parities][exch_max_reqs][exch_rec_words] (k_materialize pushes, oe_gather_decisions
  // and oe_apply_launches read); flag words 16.. of the header carry the pass number of every shard's records
  int64_t exch_max_reqs = 0;
  int exch_rec_words = 0;
  bool decisions_gathered = false;            // this pass's records of the other shards were merged into d_dec already
  size_t exch_parity_words() const { return (size_t)world * exch_max_rows * exch_stride; }
  uint32_t* exch_records(int shard, uint32_t epoch) const {
    return peer_base[shard] + OE_PEER_HEADER_WORDS + 2 * exch_parity_words() + (size_t)(epoch & 1u) * (size_t)exch_max_reqs * exch_rec_words;
  }
  uint32_t* exch_rows(int shard, uint32_t epoch) const {  // shard's buffer of that epoch's parity
    return peer_base[shard] + OE_PEER_HEADER_WORDS + (size_t)(epoch & 1u) * exch_parity_words();
  }
  void detach_peers() {
    for (int p = 0; p < OE_PEER_MAX_WORLD; ++p) {
      if (peer_opened[p] && peer_base[p]) cudaIpcCloseMemHandle(peer_base[p]);
      peer_base[p] = nullptr; peer_opened[p] = false;
    }
    peers_attached = false;
  }

  int fail(int code, const char* fmt, ...) {
    char buf[768];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err = buf;
    return code;
  }
  int cuda_fail(cudaError_t e, const char* what) {
    return fail(OE_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  }
  int64_t stride() const {
    const int64_t rows = (world > 1 && shard_rows > 0) ? shard_rows : N;
    return std::max<int64_t>(128, round_up((rows + 31) / 32, 128));
  }
  int rows_factored() const { return C + (V_eff >= 0 ? V_eff : V); }
  DevCtx ctx() const {
    DevCtx cx;
    cx.o = dof; cx.t = dt; cx.p = dp;
    cx.c.cnt = d_cnt.as<int32_t>(); cx.c.rrsum = d_rrsum.as<int4>();
    cx.c.type_total = d_type_total.as<int32_t>(); cx.c.type_agent = d_type_agent.as<int32_t>();
    cx.v.V = V; cx.v.cls = d_vcls.as<int32_t>(); cx.v.self_begin = d_vself_begin.as<int32_t>();
    cx.v.self_rows = d_vself_rows.as<int32_t>(); cx.v.rrsum = d_vrrsum.as<int4>();
    return cx;
  }
};

#define CU(call)                                                   \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return h->cuda_fail(e__, #call);       \
  } while (0)

namespace {

// Where k_materialize stores the decision records of this pass (sharded engines with peers attached and a decision area).
DecisionPush decision_push(const oe_engine* h) {
  DecisionPush p{};
  if (h->world > 1 && h->peers_attached && h->exch_max_reqs >= h->M && h->exch_rec_words > 0) {
    p.n = h->world; p.rec_words = h->exch_rec_words;
    for (int r = 0; r < h->world; ++r) p.area[r] = h->exch_records(r, h->peer_epoch);
  }
  return p;
}

// While the offer table is on the bus: a table of up to 16 MB that lives in pinned (device-visible) host memory is fetched by
// a kernel, not queued behind the offers on the copy engine (see k_fetch_host).  Pageable sources keep the plain copy.
bool fetch_instead_of_copy(oe_engine* h, void* dst, const void* src, size_t bytes, cudaStream_t on) {
  if (!h->offers_in_flight || bytes < 4096 || bytes > ((size_t)16 << 20) || (reinterpret_cast<uintptr_t>(dst) & 15)) return false;
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, src) != cudaSuccess) { cudaGetLastError(); return false; }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer || (reinterpret_cast<uintptr_t>(at.devicePointer) & 15)) {
    // pageable (or not 16-byte aligned): through the engine's pinned arena (reset by the next oe_load_offers), when it still has room
    const size_t need = (bytes + 15) / 16 * 16;
    if (!h->pin_arena || h->pin_used + need > h->pin_cap) return false;
    char* q = h->pin_arena + h->pin_used;
    std::memcpy(q, src, bytes);
    h->pin_used += need;
    at.devicePointer = q;   // cudaMallocHost memory has one address on both sides (unified addressing)
  }
  const size_t n16 = bytes / 16;
  const int tail = (int)(bytes - n16 * 16);
  const unsigned blocks = (unsigned)std::min<size_t>(296, (n16 + 255) / 256);
  k_fetch_host<<<blocks, 256, 0, on>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(at.devicePointer), n16,
                                       static_cast<unsigned char*>(dst) + n16 * 16, static_cast<const unsigned char*>(at.devicePointer) + n16 * 16, tail);
  return cudaGetLastError() == cudaSuccess;
}

template <class T>
int upload(oe_engine* h, DevBuf& b, const T* src, size_t n, cudaStream_t on = nullptr) {
  CU(b.ensure(std::max<size_t>(n, 1) * sizeof(T)));
  if (!on) on = h->stream;
  if (n && !fetch_instead_of_copy(h, b.p, src, n * sizeof(T), on)) CU(cudaMemcpyAsync(b.p, src, n * sizeof(T), cudaMemcpyHostToDevice, on));
  return 0;
}
// The requirement-side uploads are small and come from pageable memory; on the main stream they would queue behind a large
// offer-table copy that is still in flight (OE_CFG_ASYNC_LOADS), so they go through a side stream that the main stream joins.
int join_side_uploads(oe_engine* h) {
  CU(cudaEventRecord(h->ev_side, h->side_stream));
  CU(cudaStreamWaitEvent(h->stream, h->ev_side, 0));
  return 0;
}

int check_ids(oe_engine* h, const int32_t* dcol, int64_t n, int32_t lo, int32_t hi, int code) {
  if (n <= 0) return 0;
  k_validate_ids<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(dcol, n, lo, hi, code, h->d_err.as<int>());
  return 0;
}

int fetch_err_flag(oe_engine* h, const char* what) {
  int flag = 0;
  CU(cudaMemcpyAsync(&flag, h->d_err.p, sizeof flag, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (flag) {
    CU(cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream));
    if (flag == 23) return h->fail(OE_ERR_INVALID, "%s: launch_begin not monotone (code 23)", what);
    return h->fail(OE_ERR_INVALID, flag == 18 ? "%s: a ports column holds ranges that are not ascending and disjoint (code %d)"
                                              : "%s: id column %d holds a value outside its dictionary", what, flag);
  }
  return 0;
}

int offer_flag_error(oe_engine* h, int flag, const char* what) {
  h->hasO = false;
  if (flag == 20) return h->fail(OE_ERR_INVALID, "%s: port_tuple_id names a range list outside port_tuples (code 20)", what);
  return h->fail(OE_ERR_INVALID, flag == 18 ? "%s: a ports column holds ranges that are not ascending and disjoint (code %d)"
                                            : "%s: id column %d holds a value outside its dictionary", what, flag);
}
// OE_CFG_ASYNC_LOADS: host-side wait for offer chunk c (copies + checks), and its verdict.
int wait_offer_chunk(oe_engine* h, int c, const char* what) {
  CU(cudaEventSynchronize(h->chunks[c].landed));
  if (int flag = h->h_offer_err[c]) { h->offers_in_flight = false; return offer_flag_error(h, flag, what); }
  return 0;
}
int wait_offers(oe_engine* h, const char* what) {
  if (!h->offers_in_flight) return 0;
  for (int c = 0; c < h->n_chunks; ++c) if (int rc = wait_offer_chunk(h, c, what)) return rc;
  h->offers_in_flight = false;
  return 0;
}
// rows [r0, r1) of a slot-major column (dst and src are [slots][N]) onto the copy stream
int copy_rows(oe_engine* h, void* dst, const void* src, size_t elem, int slots, int64_t N, int64_t r0, int64_t r1, cudaStream_t on = nullptr) {
  if (slots <= 0 || r1 <= r0) return 0;
  if (!on) on = h->copy_stream;
  char* d = static_cast<char*>(dst) + (size_t)r0 * elem;
  const char* sp = static_cast<const char*>(src) + (size_t)r0 * elem;
  const size_t pitch = (size_t)N * elem, width = (size_t)(r1 - r0) * elem;
  if (slots == 1 || pitch > (size_t)0x7fffffff) {
    for (int q = 0; q < slots; ++q) CU(cudaMemcpyAsync(d + q * pitch, sp + q * pitch, width, cudaMemcpyHostToDevice, on));
  } else {
    CU(cudaMemcpy2DAsync(d, pitch, sp, pitch, width, (size_t)slots, cudaMemcpyHostToDevice, on));
  }
  return 0;
}

// Per pod type, the index range of the deployed tasks (h_vspan rows 0 and 1).  It only changes with the task table
// (oe_load_tasks, oe_apply_launches), so it is kept across requirement loads.
void ensure_task_spans(oe_engine* h) {
  const size_t n_types = (size_t)std::max(h->dict.n_pod_types, 1);
  if (h->h_vspan.size() != 4 * n_types) { h->h_vspan.assign(4 * n_types, 0); h->task_spans_valid = h->req_spans_valid = false; }
  if (h->task_spans_valid) return;
  int32_t *tmin = h->h_vspan.data(), *tmax = tmin + n_types;
  std::fill(tmin, tmin + n_types, INT32_MAX); std::fill(tmax, tmax + n_types, INT32_MIN);
  h->task_odd_type = false;   // a task type outside the dictionary (the device-side check will report it): no shortcut
  const int64_t T = (int64_t)h->h_ttype.size();
  for (int64_t r = 0; r < T; ++r) {
    const int32_t ty = h->h_ttype[r], ix = h->h_tindex[r];
    if (ty < 0 || ix == OE_NO_INDEX) continue;
    if ((size_t)ty >= n_types) { h->task_odd_type = true; continue; }
    tmin[ty] = std::min(tmin[ty], ix); tmax[ty] = std::max(tmax[ty], ix);
  }
  h->task_spans_valid = true;
}

int keyspace_size(const oe_dict_sizes& d, int ks) {
  switch (ks) {
    case OE_KS_HOSTNAME: return d.n_hostnames;
    case OE_KS_ZONE: return d.n_zones;
    case OE_KS_REGION: return d.n_regions;
    case OE_KS_ATTR: return d.n_attrs;
    case OE_KS_RR_ATTR: return d.n_rr_values;
  }
  return -1;
}

}  // namespace

extern "C" {

int32_t oe_abi_version(void) { return OE_ABI_VERSION; }

const char* oe_last_error(const oe_engine* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int32_t oe_create(const oe_config* cfg, oe_engine** out) {
  if (!cfg || !out) { g_create_error = "oe_create: null argument"; return OE_ERR_INVALID; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("oe_create: no CUDA device (") + cudaGetErrorString(e) + "); this engine has no CPU path";
    return OE_ERR_NO_DEVICE;
  }
  if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "oe_create: device ordinal out of range"; return OE_ERR_INVALID; }
  if (cfg->world < 1 || cfg->rank < 0 || cfg->rank >= cfg->world) { g_create_error = "oe_create: bad rank/world"; return OE_ERR_INVALID; }
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, cfg->device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return OE_ERR_CUDA; }
  if (prop.major != 10) {
    char b[160];
    snprintf(b, sizeof b, "oe_create: device %d is sm_%d%d; this build carries sm_100a code only", cfg->device, prop.major, prop.minor);
    g_create_error = b;
    return OE_ERR_NO_DEVICE;
  }
  if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return OE_ERR_CUDA; }
  oe_engine* h = new oe_engine();
  h->device = cfg->device; h->rank = cfg->rank; h->world = cfg->world; h->flags = cfg->flags;
  h->smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (cfg->stream) {
    h->stream = (cudaStream_t)cfg->stream;
  } else {
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) {
      g_create_error = cudaGetErrorString(e); delete h; return OE_ERR_CUDA;
    }
    h->own_stream = true;
  }
  for (auto& ev : h->ev) cudaEventCreate(&ev);
  cudaEventCreate(&h->ev_variants);
  cudaEventCreateWithFlags(&h->ev_queues, cudaEventDisableTiming);
  cudaEventRecord(h->ev_variants, h->stream);
  if (cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&h->ev_side) != cudaSuccess) {
    g_create_error = "oe_create: side stream"; delete h; return OE_ERR_CUDA;
  }
  {
    int least = 0, greatest = 0;
    cudaDeviceGetStreamPriorityRange(&least, &greatest);
    if (cudaStreamCreateWithPriority(&h->aux_stream, cudaStreamNonBlocking, least) != cudaSuccess ||
        cudaStreamCreateWithPriority(&h->hi_stream, cudaStreamNonBlocking, greatest) != cudaSuccess || cudaEventCreate(&h->ev_fork) != cudaSuccess ||
        cudaEventCreate(&h->ev_x0) != cudaSuccess || cudaEventCreate(&h->ev_x1) != cudaSuccess || cudaEventCreate(&h->ev_commit) != cudaSuccess) {
      g_create_error = "oe_create: auxiliary streams"; delete h; return OE_ERR_CUDA;
    }
  }
  {
    bool ok = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) == cudaSuccess && cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming) == cudaSuccess &&
              cudaMallocHost(reinterpret_cast<void**>(&h->h_offer_err), (oe_engine::kMaxChunks + 1) * sizeof(int)) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_prefix, cudaEventDisableTiming) == cudaSuccess && cudaEventCreateWithFlags(&h->ev_phase_a, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_k1_last, cudaEventDisableTiming) == cudaSuccess && h->d_sub_count.ensure(sizeof(int)) == cudaSuccess &&
              h->d_offer_err.ensure(oe_engine::kMaxChunks * sizeof(int)) == cudaSuccess;
    for (int c = 0; ok && c < oe_engine::kMaxChunks; ++c)
      ok = cudaEventCreateWithFlags(&h->chunks[c].landed, cudaEventDisableTiming) == cudaSuccess && cudaEventCreateWithFlags(&h->ev_k1[c], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { g_create_error = "oe_create: copy stream"; delete h; return OE_ERR_CUDA; }
    std::memset(h->h_offer_err, 0, (oe_engine::kMaxChunks + 1) * sizeof(int));
  }
  if (h->d_err.ensure(sizeof(int)) != cudaSuccess || h->d_nplaced.ensure(2 * sizeof(int)) != cudaSuccess) {
    g_create_error = "oe_create: cudaMalloc failed"; delete h; return OE_ERR_CUDA;
  }
  cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream);
  cudaMemsetAsync(h->d_nplaced.p, 0, 2 * sizeof(int), h->stream);   // [0] requirements placed, [1] winner whose recorded re-run disagreed
  {  // the commit kernel may use all the shared memory left beside its static arrays
    cudaFuncAttributes fa{};
    if ((e = cudaFuncGetAttributes(&fa, k_commit<true>)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_commit<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_optin - (int)fa.sharedSizeBytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_commit<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_optin - (int)fa.sharedSizeBytes)) != cudaSuccess) {
      g_create_error = std::string("oe_create: shared memory opt-in for the commit kernel: ") + cudaGetErrorString(e);
      delete h;
      return OE_ERR_CUDA;
    }
    h->commit_static_smem = (int)fa.sharedSizeBytes;
    cudaFuncAttributes fu{};
    if ((e = cudaFuncGetAttributes(&fu, k_commit_union)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_commit_union, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_optin - (int)fu.sharedSizeBytes)) != cudaSuccess) {
      g_create_error = std::string("oe_create: shared memory opt-in for the union commit kernel: ") + cudaGetErrorString(e);
      delete h;
      return OE_ERR_CUDA;
    }
    h->union_static_smem = (int)fu.sharedSizeBytes;
    if ((e = cudaFuncSetAttribute(k_offer_masks, cudaFuncAttributeMaxDynamicSharedMemorySize, OE_OS_MAX_ROWS * 44)) != cudaSuccess ||
        (e = cudaMallocHost(reinterpret_cast<void**>(&h->h_os_info), 4 * sizeof(int))) != cudaSuccess) {
      g_create_error = std::string("oe_create: set-up of the offer-side commit: ") + cudaGetErrorString(e);
      delete h;
      return OE_ERR_CUDA;
    }
    // K1x beside other work requests shared memory only to cap its residency (run_expand); per device, hence here
    if ((e = cudaFuncSetAttribute(k_expand_matrix, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)) != cudaSuccess) {
      g_create_error = std::string("oe_create: shared memory opt-in for the expansion kernel: ") + cudaGetErrorString(e);
      delete h;
      return OE_ERR_CUDA;
    }
  }
  *out = h;
  return OE_OK;
}

void oe_destroy(oe_engine* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  DevBuf* all[] = {&h->d_op_begin, &h->d_ops, &h->d_nodes, &h->d_pod_role, &h->d_excl_begin, &h->d_excl, &h->d_rs_begin, &h->d_rs,
                   &h->d_csets, &h->d_cs_off, &h->d_ts_type, &h->d_mbits, &h->d_tattr, &h->d_trr, &h->d_tfilter, &h->d_scalar,
                   &h->d_present, &h->d_port, &h->d_msize, &h->d_mprof, &h->d_oattr, &h->d_orr, &h->d_oexec, &h->d_rbegin, &h->d_rid,
                   &h->d_rkind, &h->d_rval, &h->d_req_row, &h->d_rtype, &h->d_rindex, &h->d_vcls, &h->d_vself_begin, &h->d_vself_rows,
                   &h->d_vrrsum, &h->d_cnt, &h->d_rrsum, &h->d_type_total, &h->d_type_agent, &h->d_F, &h->d_Fglobal, &h->d_matrix,
                   &h->d_dec, &h->d_used, &h->d_cursor, &h->d_err, &h->d_nplaced, &h->d_l_dst, &h->d_l_req, &h->d_l_name, &h->d_l_filter,
                   &h->d_tile_cnt, &h->d_row_base, &h->d_row_len, &h->d_cand, &h->d_fast, &h->d_rule_tab, &h->d_order, &h->d_adj, &h->d_dbg, &h->d_keep, &h->d_gc_out, &h->d_taken,
                   &h->d_meta, &h->d_union, &h->d_umask, &h->d_stage, &h->d_stage_rows, &h->d_sort_hist, &h->d_comp_begin, &h->d_ptuples, &h->d_ptuple_id, &h->d_port_tab, &h->d_tuple_hash, &h->d_ntuples, &h->d_os_parent, &h->d_os_groups, &h->d_os_ocomp, &h->d_os_omask, &h->d_os_ohist,
                   &h->d_os_comp_begin, &h->d_os_list, &h->d_os_mask, &h->d_os_rhist, &h->d_os_row_begin, &h->d_os_queue, &h->d_os_cursor, &h->d_row_map, &h->d_req_row_eff, &h->d_os_totals, &h->d_rr_used, &h->d_lean, &h->d_ruleT, &h->d_portT, &h->d_pod_hash, &h->d_pod_head, &h->d_pod_next, &h->d_sensitive, &h->d_twin, &h->d_pairs, &h->d_rclass, &h->d_twin_j, &h->d_row_of_variant, &h->d_lr_begin, &h->d_lr_name, &h->d_lr_filter};
  if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
  for (int c = 0; c < oe_engine::kMaxChunks; ++c) {
    if (h->chunks[c].landed) cudaEventDestroy(h->chunks[c].landed);
    if (h->ev_k1[c]) cudaEventDestroy(h->ev_k1[c]);
  }
  if (h->ev_main) cudaEventDestroy(h->ev_main);
  for (cudaEvent_t e : {h->ev_prefix, h->ev_phase_a, h->ev_k1_last}) if (e) cudaEventDestroy(e);
  h->d_sub_idx.release(); h->d_sub_row.release(); h->d_dec_sub.release(); h->d_sub_count.release(); h->d_tile_open.release();
  if (h->h_offer_err) cudaFreeHost(h->h_offer_err);
  if (h->h_os_info) cudaFreeHost(h->h_os_info);
  if (h->pin_arena) cudaFreeHost(h->pin_arena);
  h->d_offer_err.release();
  h->detach_peers();
  h->d_exch.release(); h->d_peer_err.release();
  for (auto* b : all) b->release();
  for (auto& b : h->d_tcols) b.release();
  for (auto& b : h->d_ocols) b.release();
  for (auto& b : h->d_wire) b.release();
  for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
  if (h->ev_variants) cudaEventDestroy(h->ev_variants);
  if (h->ev_queues) cudaEventDestroy(h->ev_queues);
  if (h->side_stream) { cudaStreamSynchronize(h->side_stream); cudaStreamDestroy(h->side_stream); }
  if (h->ev_side) cudaEventDestroy(h->ev_side);
  if (h->aux_stream) { cudaStreamSynchronize(h->aux_stream); cudaStreamDestroy(h->aux_stream); }
  if (h->hi_stream) { cudaStreamSynchronize(h->hi_stream); cudaStreamDestroy(h->hi_stream); }
  for (cudaEvent_t e : {h->ev_fork, h->ev_x0, h->ev_x1, h->ev_commit}) if (e) cudaEventDestroy(e);
  if (h->own_stream) cudaStreamDestroy(h->stream);
  delete h;
}

void* oe_stream(oe_engine* h) { return h ? (void*)h->stream : nullptr; }

int64_t oe_matrix_stride_words(const oe_engine* h) { return h ? h->stride() : 0; }



// ---------------------------------------------------------------------------------------------------------------------
// Loading
// ---------------------------------------------------------------------------------------------------------------------
int32_t oe_load_programs(oe_engine* h, const oe_program_table* t) {
  if (!h || !t) return OE_ERR_INVALID;
  CU(cudaSetDevice(h->device));
  // A resident offer table survives a re-load of the programs when its ids stay valid (no dictionary it indexes shrank) and the
  // new stage ops fit its shape (checked below): between two cycles the classes change with the plan, the offers mostly do not
  // (oe_update_offers).  Tasks and requirements carry filter bits / class ids of the old programs and are always re-loaded.
  const bool had_offers = h->hasO;
  const oe_dict_sizes old_dict = h->dict;
  h->hasP = h->hasT = h->hasO = h->hasR = false;
  h->pending_T = -1; h->reqs_pending = h->twins_unresolved = false;   // tables of the old programs: nothing of them is kept
  h->counters_fresh = h->class_bits_fresh = h->cycle_done = false;
  const int C = t->n_classes;
  if (C < 0 || C > OE_MAX_CLASSES) return h->fail(OE_ERR_LIMIT, "n_classes %d exceeds OE_MAX_CLASSES", C);
  if (!t->op_begin || (t->n_ops && !t->ops) || (C > 0 && !t->pod_role) || !t->port_excl_begin)
    return h->fail(OE_ERR_INVALID, "oe_load_programs: missing column");
  h->task_spans_valid = h->req_spans_valid = false;
  const oe_dict_sizes& d = t->dict;
  if (d.n_agents < 0 || d.n_hostnames < 0 || d.n_zones < 0 || d.n_regions < 0 || d.n_attrs < 0 || d.n_rr_values < 0 ||
      d.n_pod_types < 0 || d.n_filters < 0)
    return h->fail(OE_ERR_INVALID, "oe_load_programs: negative dictionary size");
  if (t->op_begin[0] != 0 || t->op_begin[C] != t->n_ops) return h->fail(OE_ERR_INVALID, "op_begin does not span ops");
  if (t->n_counter_sets < 0 || t->n_type_sets < 0 || t->n_range_sets < 0 || t->n_rule_nodes < 0 || t->n_matcher_words < 0)
    return h->fail(OE_ERR_INVALID, "oe_load_programs: negative count");
  auto words_for = [](int n) { return (int64_t)((n + 31) / 32); };
  auto matcher_ok = [&](int32_t off, int n_values) { return off >= 0 && off + words_for(n_values) <= t->n_matcher_words; };

  // counter sets → offsets into one tally array
  h->h_csets.assign(t->counter_sets, t->counter_sets + t->n_counter_sets);
  h->h_cs_off.assign((size_t)t->n_counter_sets + 1, 0);
  for (int c = 0; c < t->n_counter_sets; ++c) {
    const oe_counter_set& cs = t->counter_sets[c];
    int n = keyspace_size(d, cs.keyspace);
    if (n < 0) return h->fail(OE_ERR_INVALID, "counter set %d: unknown key space %d", c, cs.keyspace);
    if (cs.filter_id < 0 || cs.filter_id >= std::max(d.n_filters, 1)) return h->fail(OE_ERR_INVALID, "counter set %d: filter id out of range", c);
    if (cs.keyspace == OE_KS_RR_ATTR && (cs.rr_col < 0 || cs.rr_col >= OE_MAX_RR_COLS)) return h->fail(OE_ERR_INVALID, "counter set %d: rr column", c);
    h->h_cs_off[c + 1] = h->h_cs_off[c] + n;
  }
  for (int s = 0; s < t->n_type_sets; ++s)
    if (t->type_set_pod_type[s] < 0 || t->type_set_pod_type[s] >= d.n_pod_types) return h->fail(OE_ERR_INVALID, "type set %d: pod type out of range", s);

  // rule programs: node fields + postfix stack discipline
  h->class_self_sensitive.assign((size_t)C, 0);
  h->h_rr_used.assign((size_t)std::max(t->n_counter_sets, 1), 0);
  for (int c = 0; c < C; ++c) {
    const int ob = t->op_begin[c], oe_ = t->op_begin[c + 1];
    if (oe_ < ob || oe_ - ob > OE_MAX_STAGE_OPS) return h->fail(OE_ERR_LIMIT, "class %d: %d stage ops (max %d)", c, oe_ - ob, OE_MAX_STAGE_OPS);
    if (t->pod_role[c] < -1) return h->fail(OE_ERR_INVALID, "class %d: pod_role", c);
    if (t->port_excl_begin[c + 1] < t->port_excl_begin[c]) return h->fail(OE_ERR_INVALID, "class %d: port_excl_begin not monotone", c);
    int n_port_ops = 0, n_free_ports = 0, n_free_mounts = 0;
    int32_t res_seen[OE_MAX_STAGE_OPS];
    int n_res_seen = 0;
    for (int x = ob; x < oe_; ++x) {
      const oe_stage_op& op = t->ops[x];
      if (op.opcode >= OE_OP__COUNT) return h->fail(OE_ERR_INVALID, "class %d op %d: unknown opcode %d", c, x - ob, op.opcode);
      if (op.opcode == OE_OP_SCALAR_RESERVED || op.opcode == OE_OP_PORT_RESERVED || op.opcode == OE_OP_MOUNT_RESERVED || op.opcode == OE_OP_FREE) {
        // A reservation is consumed (or handed back) by the one stage that names it: the mappers pair every previously reserved
        // resource with at most one spec (S/offer/evaluate/TaskResourceMapper.java:80-194, ExecutorResourceMapper.java:66-129),
        // and consumeReserved / free take the id out of the pool (S/offer/MesosResourcePool.java:259-271,374-430).  The device
        // pipeline looks the id up in the offer without tracking that, so a second stage on the same id is refused here.
        if (op.arg < 0 || op.arg >= d.n_res_ids) return h->fail(OE_ERR_INVALID, "class %d op %d: resource id %d outside the dictionary (%d ids)", c, x - ob, op.arg, d.n_res_ids);
        for (int q = 0; q < n_res_seen; ++q)
          if (res_seen[q] == op.arg) return h->fail(OE_ERR_INVALID, "class %d op %d: resource id %d is named by two stages of one pipeline", c, x - ob, op.arg);
        res_seen[n_res_seen++] = op.arg;
      }
      const bool scalar_like = op.opcode == OE_OP_SCALAR_NEW || op.opcode == OE_OP_SCALAR_RESERVED || op.opcode == OE_OP_FREE;
      const bool slotted = scalar_like || op.opcode == OE_OP_PORT_STATIC || op.opcode == OE_OP_PORT_DYNAMIC || op.opcode == OE_OP_PORT_RESERVED;
      if (slotted && op.slot >= OE_MAX_ROLE_SLOTS) return h->fail(OE_ERR_INVALID, "class %d op %d: role slot %d", c, x - ob, op.slot);
      const bool free_port = op.opcode == OE_OP_FREE && op.kind == OE_KIND_PORTS, free_mount = op.opcode == OE_OP_FREE && op.kind == OE_KIND_MOUNT;
      if (free_port && op.aux > 65535) return h->fail(OE_ERR_INVALID, "class %d op %d: freed port", c, x - ob);
      if (free_mount && op.aux > OE_MAX_PROFILES) return h->fail(OE_ERR_INVALID, "class %d op %d: freed MOUNT profile", c, x - ob);
      if (free_port) ++n_free_ports;
      if (free_mount) ++n_free_mounts;
      if (scalar_like && !free_port && !free_mount && op.kind >= OE_MAX_SCALAR_KINDS) return h->fail(OE_ERR_INVALID, "class %d op %d: scalar kind %d", c, x - ob, op.kind);
      if (op.opcode == OE_OP_PORT_STATIC && (op.arg < 0 || op.arg > 65535)) return h->fail(OE_ERR_INVALID, "class %d op %d: port %d", c, x - ob, op.arg);
      if (op.opcode == OE_OP_PORT_RESERVED && op.aux > 65535) return h->fail(OE_ERR_INVALID, "class %d op %d: port", c, x - ob);
      if (op.opcode == OE_OP_PORT_DYNAMIC && (op.arg < -1 || op.arg >= t->n_range_sets)) return h->fail(OE_ERR_INVALID, "class %d op %d: range set %d", c, x - ob, op.arg);
      if (op.opcode == OE_OP_PORT_STATIC || op.opcode == OE_OP_PORT_DYNAMIC || op.opcode == OE_OP_PORT_RESERVED) ++n_port_ops;
      if (op.opcode == OE_OP_RULE) {
        const int32_t nb = op.arg, nc = (int32_t)op.aux;
        if (nb < 0 || nc < 1 || nc > OE_MAX_RULE_NODES || (int64_t)nb + nc > t->n_rule_nodes) return h->fail(OE_ERR_INVALID, "class %d: rule program range", c);
        int depth = 0;
        for (int n = nb; n < nb + nc; ++n) {
          const oe_rule_node& nd = t->rule_nodes[n];
          switch (nd.op) {
            case OE_R_TRUE: case OE_R_FALSE: break;
            case OE_R_AND: case OE_R_OR:
              if (nd.a < 0 || nd.a > depth) return h->fail(OE_ERR_INVALID, "class %d rule node %d: child count %d exceeds stack %d", c, n - nb, nd.a, depth);
              depth -= nd.a; break;
            case OE_R_NOT: if (depth < 1) return h->fail(OE_ERR_INVALID, "class %d rule node %d: NOT on empty stack", c, n - nb); depth -= 1; break;
            case OE_R_AGENT_EQ: break;
            case OE_R_ATTR_ANY: if (!matcher_ok(nd.a, d.n_attrs)) return h->fail(OE_ERR_INVALID, "class %d rule node %d: matcher offset", c, n - nb); break;
            case OE_R_HOST_IN: if (!matcher_ok(nd.a, d.n_hostnames)) return h->fail(OE_ERR_INVALID, "class %d rule node %d: matcher offset", c, n - nb); break;
            case OE_R_ZONE_IN: if (!matcher_ok(nd.a, d.n_zones)) return h->fail(OE_ERR_INVALID, "class %d rule node %d: matcher offset", c, n - nb); break;
            case OE_R_REGION_IN: if (!matcher_ok(nd.a, d.n_regions)) return h->fail(OE_ERR_INVALID, "class %d rule node %d: matcher offset", c, n - nb); break;
            case OE_R_LOCAL_REGION: break;
            case OE_R_MAXPER_HOST: case OE_R_MAXPER_ZONE: case OE_R_MAXPER_REGION: case OE_R_MAXPER_ATTR: case OE_R_RR: {
              if (nd.a < 0 || nd.a >= t->n_counter_sets) return h->fail(OE_ERR_INVALID, "class %d rule node %d: counter set %d", c, n - nb, nd.a);
              const int ks = t->counter_sets[nd.a].keyspace;
              const int want = nd.op == OE_R_MAXPER_HOST ? OE_KS_HOSTNAME : nd.op == OE_R_MAXPER_ZONE ? OE_KS_ZONE
                             : nd.op == OE_R_MAXPER_REGION ? OE_KS_REGION : nd.op == OE_R_MAXPER_ATTR ? OE_KS_ATTR : -1;
              if (want >= 0 && ks != want) return h->fail(OE_ERR_INVALID, "class %d rule node %d: counter set key space %d does not fit the rule", c, n - nb, ks);
              if (nd.op == OE_R_RR && ks == OE_KS_ATTR) return h->fail(OE_ERR_INVALID, "class %d rule node %d: round robin over the multi-key space", c, n - nb);
              if (nd.op == OE_R_MAXPER_ATTR && !matcher_ok(nd.c, d.n_attrs)) return h->fail(OE_ERR_INVALID, "class %d rule node %d: matcher offset", c, n - nb);
              if (nd.op == OE_R_RR) h->h_rr_used[(size_t)nd.a] = 1;
              h->class_self_sensitive[c] = 1;
            } break;
            case OE_R_TYPE_AVOID: case OE_R_TYPE_COLOCATE:
              if (nd.a < 0 || nd.a >= t->n_type_sets) return h->fail(OE_ERR_INVALID, "class %d rule node %d: type set %d", c, n - nb, nd.a);
              h->class_self_sensitive[c] = 1;
              break;
            default: return h->fail(OE_ERR_INVALID, "class %d rule node %d: unknown op %d", c, n - nb, nd.op);
          }
          if (++depth > 64) return h->fail(OE_ERR_LIMIT, "class %d: rule stack deeper than 64", c);
        }
        if (depth < 1) return h->fail(OE_ERR_INVALID, "class %d: rule program leaves an empty stack", c);
      }
    }
    if (n_port_ops > OE_DEV_MAX_TAKEN) return h->fail(OE_ERR_LIMIT, "class %d: %d host-port stages (max %d)", c, n_port_ops, OE_DEV_MAX_TAKEN);
    if (n_free_ports > OE_DEV_MAX_FREED || n_free_mounts > OE_DEV_MAX_FREED)
      return h->fail(OE_ERR_LIMIT, "class %d: more than %d orphaned ports / MOUNT disks returned to the pool", c, OE_DEV_MAX_FREED);
  }
  for (int r = 0; r < t->n_range_sets; ++r)
    if (t->range_set_begin[r + 1] < t->range_set_begin[r]) return h->fail(OE_ERR_INVALID, "range_set_begin not monotone");

  h->dict = d;
  h->C = C; h->n_csets = t->n_counter_sets; h->n_tsets = t->n_type_sets;
  h->h_op_begin.assign(t->op_begin, t->op_begin + C + 1);
  h->h_ops.assign(t->ops, t->ops + t->n_ops);
  h->h_nodes.assign(t->rule_nodes, t->rule_nodes + t->n_rule_nodes);
  h->h_pod_role.assign(t->pod_role, t->pod_role + C);
  h->h_excl_begin.assign(t->port_excl_begin, t->port_excl_begin + C + 1);
  h->h_excl.assign(t->port_excl, t->port_excl + t->port_excl_begin[C]);
  h->n_rule_nodes = t->n_rule_nodes;
  h->fast_K = -1;

  if (int rc = upload(h, h->d_op_begin, t->op_begin, (size_t)C + 1)) return rc;
  if (int rc = upload(h, h->d_ops, t->ops, (size_t)t->n_ops)) return rc;
  if (int rc = upload(h, h->d_nodes, t->rule_nodes, (size_t)t->n_rule_nodes)) return rc;
  if (int rc = upload(h, h->d_pod_role, t->pod_role, (size_t)C)) return rc;
  if (int rc = upload(h, h->d_excl_begin, t->port_excl_begin, (size_t)C + 1)) return rc;
  if (int rc = upload(h, h->d_excl, t->port_excl, (size_t)t->port_excl_begin[C])) return rc;
  static const int32_t zero_begin[1] = {0};
  if (int rc = upload(h, h->d_rs_begin, t->n_range_sets ? t->range_set_begin : zero_begin, (size_t)t->n_range_sets + 1)) return rc;
  if (int rc = upload(h, h->d_rs, t->range_set, t->n_range_sets ? (size_t)t->range_set_begin[t->n_range_sets] : 0)) return rc;
  if (int rc = upload(h, h->d_csets, t->counter_sets, (size_t)t->n_counter_sets)) return rc;
  if (int rc = upload(h, h->d_cs_off, h->h_cs_off.data(), h->h_cs_off.size())) return rc;
  if (int rc = upload(h, h->d_ts_type, t->type_set_pod_type, (size_t)t->n_type_sets)) return rc;
  if (int rc = upload(h, h->d_mbits, t->matcher_bits, (size_t)t->n_matcher_words)) return rc;
  if (int rc = upload(h, h->d_rr_used, h->h_rr_used.data(), h->h_rr_used.size())) return rc;
  if (int rc = upload(h, h->d_sensitive, h->class_self_sensitive.data(), h->class_self_sensitive.size())) return rc;
  CU(h->d_sensitive.ensure(1));
  h->any_rr_used = false; h->max_rr_keys = 0;
  for (int c = 0; c < t->n_counter_sets; ++c)
    if (h->h_rr_used[(size_t)c]) { h->any_rr_used = true; h->max_rr_keys = std::max(h->max_rr_keys, h->h_cs_off[(size_t)c + 1] - h->h_cs_off[(size_t)c]); }

  CU(h->d_cnt.ensure(std::max<size_t>((size_t)h->h_cs_off.back(), 1) * 4));
  CU(h->d_rrsum.ensure(std::max<size_t>((size_t)h->n_csets, 1) * sizeof(int4)));
  CU(h->d_type_total.ensure(std::max<size_t>((size_t)h->n_tsets, 1) * 4));
  CU(h->d_type_agent.ensure(std::max<size_t>((size_t)h->n_tsets * std::max(d.n_agents, 1), 1) * 4));

  DevProg& p = h->dp;
  p.C = C; p.n_csets = h->n_csets; p.n_tsets = h->n_tsets; p.n_agents = std::max(d.n_agents, 1);
  p.op_begin = h->d_op_begin.as<int32_t>(); p.ops = h->d_ops.as<oe_stage_op>(); p.nodes = h->d_nodes.as<oe_rule_node>();
  p.pod_role = h->d_pod_role.as<int32_t>(); p.excl_begin = h->d_excl_begin.as<int32_t>(); p.excl = h->d_excl.as<uint16_t>();
  p.rs_begin = h->d_rs_begin.as<int32_t>(); p.rs = h->d_rs.as<uint32_t>(); p.csets = h->d_csets.as<oe_counter_set>();
  p.cs_off = h->d_cs_off.as<int64_t>(); p.ts_type = h->d_ts_type.as<int32_t>(); p.mbits = h->d_mbits.as<uint32_t>();
  if (!(h->flags & OE_CFG_ASYNC_LOADS)) CU(cudaStreamSynchronize(h->stream));  // the caller may free its buffers after return
  h->hasP = true;
  if (had_offers && d.n_agents >= old_dict.n_agents && d.n_hostnames >= old_dict.n_hostnames && d.n_zones >= old_dict.n_zones && d.n_regions >= old_dict.n_regions &&
      d.n_attrs >= old_dict.n_attrs && d.n_rr_values >= old_dict.n_rr_values && d.n_exec_ids >= old_dict.n_exec_ids && d.n_res_ids >= old_dict.n_res_ids) {
    bool fits = true;
    const DevOffers& o = h->dof;
    for (auto& op : h->h_ops) {
      const bool scalar_like = op.opcode == OE_OP_SCALAR_NEW || op.opcode == OE_OP_SCALAR_RESERVED || op.opcode == OE_OP_FREE;
      const bool slotted = scalar_like || op.opcode == OE_OP_PORT_STATIC || op.opcode == OE_OP_PORT_DYNAMIC || op.opcode == OE_OP_PORT_RESERVED;
      const bool free_other = op.opcode == OE_OP_FREE && (op.kind == OE_KIND_PORTS || op.kind == OE_KIND_MOUNT);
      fits = fits && !(slotted && op.slot >= o.S) && !(scalar_like && !free_other && op.kind >= o.K);
    }
    for (auto& cs : h->h_csets) fits = fits && !(cs.keyspace == OE_KS_RR_ATTR && cs.rr_col >= o.Q);
    h->hasO = fits;   // otherwise: oe_load_offers again, as after any program load before
  }
  return OE_OK;
}

namespace {
// Grows the device task columns to `cap` rows, keeping the first `keep` rows.
int reserve_tasks(oe_engine* h, int64_t cap, int64_t keep) {
  DevTasks& t = h->dt;
  if (cap <= t.cap) return 0;
  cap = (cap + cap / 4 + 64 + 3) / 4 * 4;   // slots of a column stay 16-byte aligned
  auto grow = [&](DevBuf& b, int slots, void** field) -> int {
    DevBuf nb;
    CU(nb.ensure((size_t)cap * slots * 4));
    for (int s = 0; s < slots && keep > 0; ++s)
      CU(cudaMemcpyAsync((char*)nb.p + (size_t)s * cap * 4, (char*)b.p + (size_t)s * t.cap * 4, (size_t)keep * 4, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    b.release();
    b = nb;
    *field = b.p;
    return 0;
  };
  void** fields[7] = {(void**)&t.name, (void**)&t.type, (void**)&t.index, (void**)&t.agent, (void**)&t.host, (void**)&t.zone, (void**)&t.region};
  for (int c = 0; c < 7; ++c) if (int rc = grow(h->d_tcols[c], 1, fields[c])) return rc;
  if (int rc = grow(h->d_tattr, std::max(t.TA, 1), (void**)&t.attr)) return rc;
  if (int rc = grow(h->d_trr, std::max(t.Q, 1), (void**)&t.rr)) return rc;
  if (int rc = grow(h->d_tfilter, std::max(t.FW, 1), (void**)&t.filter)) return rc;
  t.cap = cap;
  return 0;
}
}  // namespace

namespace {
// host mirror of the task columns that variant detection and oe_apply_launches read
int absorb_tasks(oe_engine* h) {
  if (h->pending_T < 0) return 0;
  const int64_t T = h->pending_T;
  if (h->tasks_caller_valid) {
    h->h_tname.assign(h->pending_t_name, h->pending_t_name + T);
    h->h_ttype.assign(h->pending_t_type, h->pending_t_type + T);
    h->h_tindex.assign(h->pending_t_index, h->pending_t_index + T);
  } else {
    h->h_tname.resize((size_t)T); h->h_ttype.resize((size_t)T); h->h_tindex.resize((size_t)T);
    if (T > 0) {
      CU(cudaMemcpyAsync(h->h_tname.data(), h->dt.name, (size_t)T * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaMemcpyAsync(h->h_ttype.data(), h->dt.type, (size_t)T * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaMemcpyAsync(h->h_tindex.data(), h->dt.index, (size_t)T * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaStreamSynchronize(h->stream));
    }
  }
  h->task_spans_valid = false;
  h->pending_T = -1;
  return 0;
}
}  // namespace

int32_t oe_load_tasks(oe_engine* h, const oe_task_table* t) {
  if (!h || !t) return OE_ERR_INVALID;
  if (!h->hasP) return h->fail(OE_ERR_STATE, "oe_load_tasks: load programs first");
  CU(cudaSetDevice(h->device));
  h->hasT = h->hasR = false;
  h->reqs_pending = h->twins_unresolved = false;
  h->counters_fresh = h->class_bits_fresh = h->cycle_done = false;
  const int64_t T = t->n_tasks;
  if (T < 0 || T > 0x7ffffff0) return h->fail(OE_ERR_INVALID, "n_tasks");
  if (t->n_attr_slots < 0 || t->n_attr_slots > OE_MAX_ATTR_SLOTS) return h->fail(OE_ERR_LIMIT, "task attr slots");
  if (t->n_rr_cols < 0 || t->n_rr_cols > OE_MAX_RR_COLS) return h->fail(OE_ERR_LIMIT, "task rr cols");
  if (t->n_filter_words < (h->dict.n_filters + 31) / 32) return h->fail(OE_ERR_INVALID, "n_filter_words too small for n_filters");
  for (auto& cs : h->h_csets)
    if (cs.keyspace == OE_KS_RR_ATTR && cs.rr_col >= t->n_rr_cols) return h->fail(OE_ERR_INVALID, "a counter set names rr column %d, task table has %d", cs.rr_col, t->n_rr_cols);
  if (T > 0 && (!t->name_id || !t->pod_type_id || !t->pod_index || !t->agent_id || !t->hostname_id || !t->zone_id || !t->region_id ||
                (t->n_attr_slots && !t->attr_id) || (t->n_rr_cols && !t->rr_key) || (t->n_filter_words && !t->filter_bits)))
    return h->fail(OE_ERR_INVALID, "oe_load_tasks: missing column");
  DevTasks& d = h->dt;
  // a change of column widths invalidates the old layout
  if (d.TA != t->n_attr_slots || d.Q != t->n_rr_cols || d.FW != t->n_filter_words) {
    d.cap = 0; d.TA = t->n_attr_slots; d.Q = t->n_rr_cols; d.FW = t->n_filter_words;
  }
  d.T = 0;
  if (int rc = reserve_tasks(h, std::max<int64_t>(T, 1), 0)) return rc;
  const int32_t* cols[7] = {t->name_id, t->pod_type_id, t->pod_index, t->agent_id, t->hostname_id, t->zone_id, t->region_id};
  int32_t* dst[7] = {d.name, d.type, d.index, d.agent, d.host, d.zone, d.region};
  if (T > 0) {
    auto put = [&](void* to, const void* from) -> cudaError_t {
      if (fetch_instead_of_copy(h, to, from, (size_t)T * 4, h->stream)) return cudaSuccess;
      return cudaMemcpyAsync(to, from, (size_t)T * 4, cudaMemcpyHostToDevice, h->stream);
    };
    for (int c = 0; c < 7; ++c) CU(put(dst[c], cols[c]));
    for (int a = 0; a < d.TA; ++a) CU(put(d.attr + (size_t)a * d.cap, t->attr_id + (size_t)a * T));
    for (int q = 0; q < d.Q; ++q) CU(put(d.rr + (size_t)q * d.cap, t->rr_key + (size_t)q * T));
    for (int w = 0; w < d.FW; ++w) CU(put(d.filter + (size_t)w * d.cap, t->filter_bits + (size_t)w * T));
  }
  d.T = T;
  const oe_dict_sizes& dc = h->dict;
  check_ids(h, d.type, T, -1, dc.n_pod_types, 1);
  check_ids(h, d.agent, T, 0, std::max(dc.n_agents, 1), 2);
  check_ids(h, d.host, T, -1, dc.n_hostnames, 3);
  check_ids(h, d.zone, T, -1, dc.n_zones, 4);
  check_ids(h, d.region, T, -1, dc.n_regions, 5);
  for (int a = 0; a < d.TA; ++a) check_ids(h, d.attr + (size_t)a * d.cap, T, -1, dc.n_attrs, 6);
  for (int q = 0; q < d.Q; ++q) check_ids(h, d.rr + (size_t)q * d.cap, T, -1, dc.n_rr_values, 7);
  h->task_spans_valid = false;
  h->name_to_row.clear();
  h->name_map_built = false;  // only oe_apply_launches needs it: built on first use
  h->pending_t_name = t->name_id; h->pending_t_type = t->pod_type_id; h->pending_t_index = t->pod_index; h->pending_T = T;
  h->tasks_caller_valid = true;
  if (!(h->flags & OE_CFG_ASYNC_LOADS)) if (int rc = absorb_tasks(h)) return rc;   // else: when somebody needs them
  if (h->flags & OE_CFG_ASYNC_LOADS) h->validation_pending = true;
  else if (int rc = fetch_err_flag(h, "oe_load_tasks")) return rc;
  h->hasT = true;
  return OE_OK;
}

namespace {
// Shape checks shared by oe_load_offers and oe_load_offers_packed (pk != null: the columns the packing replaces may be NULL).
int check_offer_table(oe_engine* h, const oe_offer_table* t, const oe_offer_packing* pk);
}  // namespace

int32_t oe_load_offers(oe_engine* h, const oe_offer_table* t) {
  if (!h || !t) return OE_ERR_INVALID;
  if (!h->hasP) return h->fail(OE_ERR_STATE, "oe_load_offers: load programs first");
  CU(cudaSetDevice(h->device));
  h->hasO = false;
  h->class_bits_fresh = h->cycle_done = false;
  h->tuples_valid = false; h->n_port_tuples = 0; h->tuple_count_pending = false;
  h->dof.ptuple = nullptr; h->dof.n_ptuples = 0;
  if (int rc = check_offer_table(h, t, nullptr)) return rc;
  const int64_t N = t->n_offers;
  DevOffers& o = h->dof;
  o.N = N; o.row_offset = t->row_offset;
  o.S = t->n_role_slots; o.K = t->n_scalar_kinds; o.R = t->n_port_ranges; o.D = t->n_mount_slots;
  o.A = t->n_attr_slots; o.Q = t->n_rr_cols; o.E = t->n_exec_slots;
  h->N = N; h->row_offset = t->row_offset; h->shard_rows = t->shard_rows; h->n_reserved = t->n_reserved;
  h->N_global = t->n_offers_global > 0 ? t->n_offers_global : N;
  const size_t n = (size_t)N;
  if (h->offers_in_flight) {  // a table that was never evaluated is being replaced: its verdict no longer matters
    CU(cudaStreamSynchronize(h->copy_stream));
    h->offers_in_flight = false;
  }
  h->n_chunks = 1;
  h->chunks[0].r0 = 0; h->chunks[0].r1 = N;
  const bool chunked = (h->flags & OE_CFG_ASYNC_LOADS) && N > 0;
  if (chunked) {
    // The table goes over the bus in row chunks on a stream of its own; behind each chunk its checks, the flag's way back to
    // the host and an event.  The feasibility pass waits chunk by chunk (run_class_feasibility), so the first chunks are being
    // evaluated (and expanded) while the later ones are still arriving.
    CU(h->d_scalar.ensure(std::max<size_t>(n * o.S * o.K, 1) * 8)); CU(h->d_present.ensure(n * 4)); CU(h->d_port.ensure(std::max<size_t>(n * o.S * o.R, 1) * 4));
    CU(h->d_msize.ensure(std::max<size_t>(n * o.D, 1) * 8)); CU(h->d_mprof.ensure(std::max<size_t>(n * o.D, 1) * 4));
    for (int c = 0; c < 5; ++c) CU(h->d_ocols[c].ensure(n * 4));
    CU(h->d_oattr.ensure(std::max<size_t>(n * o.A, 1) * 4)); CU(h->d_orr.ensure(std::max<size_t>(n * o.Q, 1) * 4)); CU(h->d_oexec.ensure(std::max<size_t>(n * o.E, 1) * 4));
  } else {
    if (int rc = upload(h, h->d_scalar, t->scalar, n * o.S * o.K)) return rc;
    if (int rc = upload(h, h->d_present, t->scalar_present, n)) return rc;
    if (int rc = upload(h, h->d_port, t->port_range, n * o.S * o.R)) return rc;
    if (int rc = upload(h, h->d_msize, t->mount_size, n * o.D)) return rc;
    if (int rc = upload(h, h->d_mprof, t->mount_profile, n * o.D)) return rc;
    const int32_t* cols[5] = {t->alloc_role, t->agent_id, t->hostname_id, t->zone_id, t->region_id};
    for (int c = 0; c < 5; ++c) if (int rc = upload(h, h->d_ocols[c], cols[c], n)) return rc;
    if (int rc = upload(h, h->d_oattr, t->attr_id, n * o.A)) return rc;
    if (int rc = upload(h, h->d_orr, t->rr_key, n * o.Q)) return rc;
    if (int rc = upload(h, h->d_oexec, t->exec_id, n * o.E)) return rc;
  }
  o.scalar = h->d_scalar.as<double>(); o.present = h->d_present.as<uint32_t>(); o.port = h->d_port.as<uint32_t>();
  o.msize = h->d_msize.as<double>(); o.mprof = h->d_mprof.as<int32_t>();
  o.alloc = h->d_ocols[0].as<int32_t>(); o.agent = h->d_ocols[1].as<int32_t>(); o.host = h->d_ocols[2].as<int32_t>();
  o.zone = h->d_ocols[3].as<int32_t>(); o.region = h->d_ocols[4].as<int32_t>();
  o.attr = h->d_oattr.as<int32_t>(); o.rr = h->d_orr.as<int32_t>(); o.exec = h->d_oexec.as<int32_t>();
  cudaStream_t ls = chunked ? h->copy_stream : h->stream;
  if (chunked) {
    if (h->pin_used) {  // staged tables of an earlier load may still be on their way
      CU(cudaStreamSynchronize(h->side_stream));
      CU(cudaStreamSynchronize(h->stream));
      h->pin_used = 0;
    }
    if (!h->pin_arena && cudaMallocHost(reinterpret_cast<void**>(&h->pin_arena), (size_t)8 << 20) == cudaSuccess) h->pin_cap = (size_t)8 << 20;
    cudaGetLastError();
  }
  if (chunked) {  // the copy stream starts behind whatever the main stream still does with the old table
    CU(cudaEventRecord(h->ev_main, h->stream));
    CU(cudaStreamWaitEvent(h->copy_stream, h->ev_main, 0));
    CU(cudaMemsetAsync(h->d_offer_err.p, 0, oe_engine::kMaxChunks * sizeof(int), h->copy_stream));
  }
  if (t->n_reserved > 0) {
    if (int rc = upload(h, h->d_rbegin, t->reserved_begin, n + 1, ls)) return rc;
    if (int rc = upload(h, h->d_rid, t->reserved_id, (size_t)t->n_reserved, ls)) return rc;
    if (int rc = upload(h, h->d_rkind, t->reserved_kind, (size_t)t->n_reserved, ls)) return rc;
    if (int rc = upload(h, h->d_rval, t->reserved_value, (size_t)t->n_reserved, ls)) return rc;
    o.rbegin = h->d_rbegin.as<int64_t>(); o.rid = h->d_rid.as<int32_t>(); o.rkind = h->d_rkind.as<int32_t>(); o.rval = h->d_rval.as<double>();
  } else {
    o.rbegin = nullptr; o.rid = nullptr; o.rkind = nullptr; o.rval = nullptr;
  }
  const oe_dict_sizes& dc = h->dict;
  if (chunked) {
    int nch = (int)std::max<int64_t>(1, std::min<int64_t>(4, N / 131072));  // >= 128 Ki rows (~0.25 ms of PCIe) per chunk; more than four only add copy calls
    if (const char* ev = getenv("OE_OFFER_CHUNKS")) nch = std::max(1, std::min(oe_engine::kMaxChunks, atoi(ev)));
    const int64_t chunk_rows = round_up((N + nch - 1) / nch, 8192);  // whole K1x slabs (256 words)
    nch = (int)((N + chunk_rows - 1) / chunk_rows);
    h->n_chunks = nch;
    if (t->n_reserved > 0)
      k_validate_ids<<<(unsigned)((t->n_reserved + 255) / 256), 256, 0, h->copy_stream>>>(o.rid, t->n_reserved, 0, std::max(dc.n_res_ids, 1), 19, h->d_offer_err.as<int>());
    const OfferLimits lim{dc.n_agents, dc.n_hostnames, dc.n_zones, dc.n_regions, dc.n_attrs, dc.n_rr_values};
    for (int c = 0; c < nch; ++c) {
      const int64_t r0 = c * chunk_rows, r1 = std::min<int64_t>(N, r0 + chunk_rows);
      h->chunks[c].r0 = r0; h->chunks[c].r1 = r1;
      if (int rc = copy_rows(h, h->d_scalar.p, t->scalar, 8, o.S * o.K, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_present.p, t->scalar_present, 4, 1, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_port.p, t->port_range, 4, o.S * o.R, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_msize.p, t->mount_size, 8, o.D, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_mprof.p, t->mount_profile, 4, o.D, N, r0, r1)) return rc;
      const int32_t* cols[5] = {t->alloc_role, t->agent_id, t->hostname_id, t->zone_id, t->region_id};
      for (int q = 0; q < 5; ++q) if (int rc = copy_rows(h, h->d_ocols[q].p, cols[q], 4, 1, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_oattr.p, t->attr_id, 4, o.A, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_orr.p, t->rr_key, 4, o.Q, N, r0, r1)) return rc;
      if (int rc = copy_rows(h, h->d_oexec.p, t->exec_id, 4, o.E, N, r0, r1)) return rc;
      k_validate_offer_rows<<<(unsigned)((r1 - r0 + 255) / 256), 256, 0, h->copy_stream>>>(o, lim, r0, r1, h->d_offer_err.as<int>() + c);
      CU(cudaGetLastError());
      CU(cudaMemcpyAsync(h->h_offer_err + c, h->d_offer_err.as<int>() + c, sizeof(int), cudaMemcpyDeviceToHost, h->copy_stream));
      CU(cudaEventRecord(h->chunks[c].landed, h->copy_stream));
    }
    h->offers_in_flight = true;
    h->validation_pending = true;
  } else {
    check_ids(h, o.agent, N, 0, std::max(dc.n_agents, 1), 11);
    check_ids(h, o.host, N, 0, std::max(dc.n_hostnames, 1), 12);
    check_ids(h, o.zone, N, -1, dc.n_zones, 13);
    check_ids(h, o.region, N, -1, dc.n_regions, 14);
    check_ids(h, o.attr, N * o.A, -1, dc.n_attrs, 15);
    check_ids(h, o.rr, N * o.Q, -1, dc.n_rr_values, 16);
    check_ids(h, o.mprof, N * o.D, -1, OE_MAX_PROFILES, 17);
    if (t->n_reserved > 0) check_ids(h, o.rid, t->n_reserved, 0, std::max(dc.n_res_ids, 1), 19);
    if (N > 0 && o.R > 0)
      k_validate_ranges<<<(unsigned)((N + 255) / 256), 256, 0, h->stream>>>(o.port, N, o.S, o.R, 18, h->d_err.as<int>());
    if (h->flags & OE_CFG_ASYNC_LOADS) h->validation_pending = true;
    else if (int rc = fetch_err_flag(h, "oe_load_offers")) return rc;  // also makes the caller's buffers reusable
  }
  h->hasO = true;
  return OE_OK;
}

namespace {
int check_offer_table(oe_engine* h, const oe_offer_table* t, const oe_offer_packing* pk) {
  const int64_t N = t->n_offers;
  if (N < 0 || N > 0x7ffffff0) return h->fail(OE_ERR_INVALID, "n_offers");
  if (t->n_role_slots < 1 || t->n_role_slots > OE_MAX_ROLE_SLOTS) return h->fail(OE_ERR_LIMIT, "n_role_slots %d", t->n_role_slots);
  if (t->n_scalar_kinds < 1 || t->n_scalar_kinds > OE_MAX_SCALAR_KINDS) return h->fail(OE_ERR_LIMIT, "n_scalar_kinds %d", t->n_scalar_kinds);
  if (t->n_role_slots * t->n_scalar_kinds > 32) return h->fail(OE_ERR_LIMIT, "role slots x scalar kinds exceeds the 32-bit presence mask");
  if (t->n_port_ranges < 0 || t->n_port_ranges > OE_MAX_PORT_RANGES) return h->fail(OE_ERR_LIMIT, "n_port_ranges %d", t->n_port_ranges);
  if (t->n_mount_slots < 0 || t->n_mount_slots > OE_MAX_MOUNT_SLOTS) return h->fail(OE_ERR_LIMIT, "n_mount_slots %d", t->n_mount_slots);
  if (t->n_attr_slots < 0 || t->n_attr_slots > OE_MAX_ATTR_SLOTS) return h->fail(OE_ERR_LIMIT, "n_attr_slots %d", t->n_attr_slots);
  if (t->n_rr_cols < 0 || t->n_rr_cols > OE_MAX_RR_COLS) return h->fail(OE_ERR_LIMIT, "n_rr_cols %d", t->n_rr_cols);
  if (t->n_exec_slots < 0 || t->n_exec_slots > OE_MAX_EXEC_SLOTS) return h->fail(OE_ERR_LIMIT, "n_exec_slots %d", t->n_exec_slots);
  for (auto& op : h->h_ops) {
    const bool scalar_like = op.opcode == OE_OP_SCALAR_NEW || op.opcode == OE_OP_SCALAR_RESERVED || op.opcode == OE_OP_FREE;
    const bool slotted = scalar_like || op.opcode == OE_OP_PORT_STATIC || op.opcode == OE_OP_PORT_DYNAMIC || op.opcode == OE_OP_PORT_RESERVED;
    if (slotted && op.slot >= t->n_role_slots) return h->fail(OE_ERR_INVALID, "a stage op names role slot %d, offer table has %d", op.slot, t->n_role_slots);
    const bool free_other = op.opcode == OE_OP_FREE && (op.kind == OE_KIND_PORTS || op.kind == OE_KIND_MOUNT);
    if (scalar_like && !free_other && op.kind >= t->n_scalar_kinds) return h->fail(OE_ERR_INVALID, "a stage op names scalar kind %d, offer table has %d", op.kind, t->n_scalar_kinds);
  }
  for (auto& cs : h->h_csets)
    if (cs.keyspace == OE_KS_RR_ATTR && cs.rr_col >= t->n_rr_cols) return h->fail(OE_ERR_INVALID, "a counter set names rr column %d, offer table has %d", cs.rr_col, t->n_rr_cols);
  if (h->world > 1) {
    if (t->shard_rows <= 0 || t->shard_rows % 1024 != 0) return h->fail(OE_ERR_INVALID, "sharded offer table needs shard_rows, a multiple of 1024");
    if (t->row_offset != (int64_t)h->rank * t->shard_rows) return h->fail(OE_ERR_INVALID, "row_offset must be rank * shard_rows");
    if (N > t->shard_rows) return h->fail(OE_ERR_INVALID, "shard holds more than shard_rows offers");
    const int64_t Ng = t->n_offers_global;
    if (Ng < 0 || Ng > (int64_t)h->world * t->shard_rows) return h->fail(OE_ERR_INVALID, "n_offers_global %lld does not fit %d shards of %lld rows", (long long)Ng, h->world, (long long)t->shard_rows);
    const int64_t want = std::max<int64_t>(0, std::min<int64_t>(t->shard_rows, Ng - t->row_offset));
    if (N != want) return h->fail(OE_ERR_INVALID, "shard %d must hold %lld of the %lld offers (every shard but the last is full), got %lld", h->rank, (long long)want, (long long)Ng, (long long)N);
  } else if (t->row_offset != 0) {
    return h->fail(OE_ERR_INVALID, "row_offset must be 0 when not sharded");
  }
  const bool no_alloc = pk && pk->alloc_role_is_const, no_host = pk && pk->hostname_is_agent, no_port = pk && pk->port_tuple_bytes;
  if (N > 0 && (!t->scalar || !t->scalar_present || (!no_alloc && !t->alloc_role) || !t->agent_id || (!no_host && !t->hostname_id) || !t->zone_id || !t->region_id ||
                (t->n_port_ranges && !no_port && !t->port_range) || (t->n_mount_slots && (!t->mount_size || !t->mount_profile)) ||
                (t->n_attr_slots && !t->attr_id) || (t->n_rr_cols && !t->rr_key) || (t->n_exec_slots && !t->exec_id)))
    return h->fail(OE_ERR_INVALID, "oe_load_offers: missing column");
  if (t->n_reserved < 0 || (t->n_reserved > 0 && (!t->reserved_begin || !t->reserved_id || !t->reserved_kind || !t->reserved_value)))
    return h->fail(OE_ERR_INVALID, "oe_load_offers: reserved side table");
  if (t->n_reserved > 0) {
    if (t->reserved_begin[0] != 0 || t->reserved_begin[N] != t->n_reserved) return h->fail(OE_ERR_INVALID, "reserved_begin does not span the side table");
    for (int64_t i = 0; i < N; ++i)
      if (t->reserved_begin[i + 1] < t->reserved_begin[i]) return h->fail(OE_ERR_INVALID, "reserved_begin not monotone");
  }

  return 0;
}
}  // namespace

int32_t oe_load_offers_packed(oe_engine* h, const oe_offer_table* t, const oe_offer_packing* pk) {
  if (!h || !t || !pk) return OE_ERR_INVALID;
  if (!h->hasP) return h->fail(OE_ERR_STATE, "oe_load_offers_packed: load programs first");
  CU(cudaSetDevice(h->device));
  h->hasO = false;
  h->class_bits_fresh = h->cycle_done = false;
  h->tuples_valid = false; h->n_port_tuples = 0; h->tuple_count_pending = false;
  h->dof.ptuple = nullptr; h->dof.n_ptuples = 0;
  if (int rc = check_offer_table(h, t, pk)) return rc;
  auto width_ok = [](uint8_t b) { return b == 1 || b == 2 || b == 4; };
  if (!width_ok(pk->present_bytes) || !width_ok(pk->zone_bytes) || !width_ok(pk->region_bytes) || !width_ok(pk->attr_bytes) || !width_ok(pk->rr_bytes) ||
      !width_ok(pk->mount_profile_bytes) || !width_ok(pk->exec_bytes))
    return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: a column width is not 1, 2 or 4 bytes");
  if (pk->port_tuple_bytes != 0 && pk->port_tuple_bytes != 2 && pk->port_tuple_bytes != 4) return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: port_tuple_bytes must be 0, 2 or 4");
  const int64_t N = t->n_offers;
  const bool tuples = pk->port_tuple_bytes != 0 && t->n_port_ranges > 0;
  if (tuples && (pk->n_port_tuples < 1 || !pk->port_tuples || (N > 0 && !pk->port_tuple_id))) return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: port_tuples / port_tuple_id missing");
  if (tuples && pk->port_tuple_bytes == 2 && pk->n_port_tuples > 65536) return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: %d range lists do not fit 16-bit ids", pk->n_port_tuples);
  const int present_bits = t->n_role_slots * t->n_scalar_kinds;
  if (pk->present_bytes * 8 < present_bits) return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: %d presence bits do not fit %d byte(s)", present_bits, pk->present_bytes);
  auto dict_fits = [](int n, uint8_t bytes) { return bytes == 4 || (int64_t)n < (1ll << (8 * bytes)) - 1; };
  const oe_dict_sizes& dc = h->dict;
  if (!dict_fits(dc.n_zones, pk->zone_bytes) || !dict_fits(dc.n_regions, pk->region_bytes) || !dict_fits(dc.n_attrs, pk->attr_bytes) ||
      !dict_fits(dc.n_rr_values, pk->rr_bytes) || !dict_fits(OE_MAX_PROFILES, pk->mount_profile_bytes) || !dict_fits(dc.n_exec_ids, pk->exec_bytes))
    return h->fail(OE_ERR_INVALID, "oe_load_offers_packed: a dictionary is too large for its column's width (the all-ones value stands for -1)");

  DevOffers& o = h->dof;
  o.N = N; o.row_offset = t->row_offset;
  o.S = t->n_role_slots; o.K = t->n_scalar_kinds; o.R = t->n_port_ranges; o.D = t->n_mount_slots;
  o.A = t->n_attr_slots; o.Q = t->n_rr_cols; o.E = t->n_exec_slots;
  h->N = N; h->row_offset = t->row_offset; h->shard_rows = t->shard_rows; h->n_reserved = t->n_reserved;
  h->N_global = t->n_offers_global > 0 ? t->n_offers_global : N;
  const size_t n = (size_t)N;
  if (h->offers_in_flight) {
    CU(cudaStreamSynchronize(h->copy_stream));
    h->offers_in_flight = false;
  }
  CU(h->d_scalar.ensure(std::max<size_t>(n * o.S * o.K, 1) * 8)); CU(h->d_present.ensure(std::max<size_t>(n, 1) * 4)); CU(h->d_port.ensure(std::max<size_t>(n * o.S * o.R, 1) * 4));
  CU(h->d_msize.ensure(std::max<size_t>(n * o.D, 1) * 8)); CU(h->d_mprof.ensure(std::max<size_t>(n * o.D, 1) * 4));
  for (int c = 0; c < 5; ++c) CU(h->d_ocols[c].ensure(std::max<size_t>(n, 1) * 4));
  CU(h->d_oattr.ensure(std::max<size_t>(n * o.A, 1) * 4)); CU(h->d_orr.ensure(std::max<size_t>(n * o.Q, 1) * 4)); CU(h->d_oexec.ensure(std::max<size_t>(n * o.E, 1) * 4));
  o.scalar = h->d_scalar.as<double>(); o.present = h->d_present.as<uint32_t>(); o.port = h->d_port.as<uint32_t>();
  o.msize = h->d_msize.as<double>(); o.mprof = h->d_mprof.as<int32_t>();
  o.alloc = h->d_ocols[0].as<int32_t>(); o.agent = h->d_ocols[1].as<int32_t>(); o.host = h->d_ocols[2].as<int32_t>();
  o.zone = h->d_ocols[3].as<int32_t>(); o.region = h->d_ocols[4].as<int32_t>();
  o.attr = h->d_oattr.as<int32_t>(); o.rr = h->d_orr.as<int32_t>(); o.exec = h->d_oexec.as<int32_t>();

  // one wire column: where it lands (staging when narrower than the device column), and how it is widened
  enum { W_SCALAR, W_PRESENT, W_MPROF, W_ZONE, W_REGION, W_ATTR, W_RR, W_EXEC, W_TUPLE, W_COUNT };
  struct Wire { const void* src; void* dev; int slots; int wire_bytes, dev_bytes; void* stage; } w[W_COUNT] = {
      {t->scalar, o.S * o.K ? h->d_scalar.p : nullptr, o.S * o.K, pk->scalar_is_f32 ? 4 : 8, 8, nullptr},
      {t->scalar_present, h->d_present.p, 1, pk->present_bytes, 4, nullptr},
      {t->mount_profile, h->d_mprof.p, o.D, pk->mount_profile_bytes, 4, nullptr},
      {t->zone_id, h->d_ocols[3].p, 1, pk->zone_bytes, 4, nullptr},
      {t->region_id, h->d_ocols[4].p, 1, pk->region_bytes, 4, nullptr},
      {t->attr_id, h->d_oattr.p, o.A, pk->attr_bytes, 4, nullptr},
      {t->rr_key, h->d_orr.p, o.Q, pk->rr_bytes, 4, nullptr},
      {t->exec_id, h->d_oexec.p, o.E, pk->exec_bytes, 4, nullptr},
      {tuples ? pk->port_tuple_id : nullptr, nullptr, tuples ? o.S : 0, pk->port_tuple_bytes, 4, nullptr},
  };
  static_assert(W_COUNT <= 12, "d_wire");
  for (int c = 0; c < W_COUNT; ++c) {
    if (w[c].slots > 0 && w[c].wire_bytes != w[c].dev_bytes) {
      CU(h->d_wire[c].ensure(std::max<size_t>(n * w[c].slots, 1) * w[c].wire_bytes));
      w[c].stage = h->d_wire[c].p;
    }
  }
  if (tuples) {
    CU(h->d_ptuple_id.ensure(std::max<size_t>(n * o.S, 1) * 4));
    if (pk->port_tuple_bytes == 4) { CU(h->d_wire[W_TUPLE].ensure(std::max<size_t>(n * o.S, 1) * 4)); w[W_TUPLE].stage = h->d_wire[W_TUPLE].p; }
  }

  const bool async = (h->flags & OE_CFG_ASYNC_LOADS) != 0 && N > 0;
  cudaStream_t ls = async ? h->copy_stream : h->stream;
  if (async) {
    if (h->pin_used) {
      CU(cudaStreamSynchronize(h->side_stream));
      CU(cudaStreamSynchronize(h->stream));
      h->pin_used = 0;
    }
    if (!h->pin_arena && cudaMallocHost(reinterpret_cast<void**>(&h->pin_arena), (size_t)8 << 20) == cudaSuccess) h->pin_cap = (size_t)8 << 20;
    cudaGetLastError();
    CU(cudaEventRecord(h->ev_main, h->stream));
    CU(cudaStreamWaitEvent(h->copy_stream, h->ev_main, 0));
  }
  CU(cudaMemsetAsync(h->d_offer_err.p, 0, oe_engine::kMaxChunks * sizeof(int), ls));
  if (t->n_reserved > 0) {
    if (int rc = upload(h, h->d_rbegin, t->reserved_begin, n + 1, ls)) return rc;
    if (int rc = upload(h, h->d_rid, t->reserved_id, (size_t)t->n_reserved, ls)) return rc;
    if (int rc = upload(h, h->d_rkind, t->reserved_kind, (size_t)t->n_reserved, ls)) return rc;
    if (int rc = upload(h, h->d_rval, t->reserved_value, (size_t)t->n_reserved, ls)) return rc;
    o.rbegin = h->d_rbegin.as<int64_t>(); o.rid = h->d_rid.as<int32_t>(); o.rkind = h->d_rkind.as<int32_t>(); o.rval = h->d_rval.as<double>();
    k_validate_ids<<<(unsigned)((t->n_reserved + 255) / 256), 256, 0, ls>>>(o.rid, t->n_reserved, 0, std::max(dc.n_res_ids, 1), 19, h->d_offer_err.as<int>());
  } else {
    o.rbegin = nullptr; o.rid = nullptr; o.rkind = nullptr; o.rval = nullptr;
  }
  if (tuples) {   // the distinct range lists: small, first over the bus, checked like a ports column of n_port_tuples rows
    const size_t words = (size_t)pk->n_port_tuples * o.R;
    h->ptuple_cap = pk->n_port_tuples + std::max(4096, pk->n_port_tuples / 4);   // room for the lists that oe_update_offers brings
    CU(h->d_ptuples.ensure(std::max<size_t>((size_t)h->ptuple_cap * o.R, 1) * 4));
    CU(cudaMemcpyAsync(h->d_ptuples.p, pk->port_tuples, words * 4, cudaMemcpyHostToDevice, ls));
    k_validate_tuples<<<(unsigned)((pk->n_port_tuples + 255) / 256), 256, 0, ls>>>(h->d_ptuples.as<uint32_t>(), pk->n_port_tuples, o.R, 18, h->d_offer_err.as<int>());
    uint32_t hsize = 1024;
    while (hsize < (uint32_t)h->ptuple_cap * 2u) hsize <<= 1;
    h->tuple_mask = hsize - 1;
    CU(h->d_tuple_hash.ensure((size_t)hsize * 4)); CU(h->d_ntuples.ensure(sizeof(int)));
    CU(cudaMemsetAsync(h->d_tuple_hash.p, 0, (size_t)hsize * 4, ls));
    k_tuple_hash_insert<<<(unsigned)((pk->n_port_tuples + 127) / 128), 128, 0, ls>>>(h->d_ptuples.as<uint32_t>(), pk->n_port_tuples, o.R, TupleHash{h->d_tuple_hash.as<int32_t>(), h->tuple_mask});
    k_fill_i32<<<1, 32, 0, ls>>>(h->d_ntuples.as<int32_t>(), pk->n_port_tuples, 0, 1);
  }

  int nch = 1;
  if (async) {
    nch = (int)std::max<int64_t>(1, std::min<int64_t>(4, N / 131072));
    if (const char* ev = getenv("OE_OFFER_CHUNKS")) nch = std::max(1, std::min(oe_engine::kMaxChunks, atoi(ev)));
  }
  const int64_t chunk_rows = std::max<int64_t>(8192, round_up((N + nch - 1) / nch, 8192));
  nch = (int)std::max<int64_t>(1, (N + chunk_rows - 1) / chunk_rows);
  h->n_chunks = nch;
  const OfferLimits lim{dc.n_agents, dc.n_hostnames, dc.n_zones, dc.n_regions, dc.n_attrs, dc.n_rr_values};
  auto blocks = [](int64_t elems) { return (unsigned)std::max<int64_t>(1, (elems + 255) / 256); };
  for (int c = 0; c < nch; ++c) {
    const int64_t r0 = c * chunk_rows, r1 = std::min<int64_t>(N, r0 + chunk_rows);
    h->chunks[c].r0 = r0; h->chunks[c].r1 = r1;
    if (r1 > r0) {
      for (int q = 0; q < W_COUNT; ++q) {
        if (w[q].slots <= 0) continue;
        void* dst = w[q].stage ? w[q].stage : w[q].dev;
        if (int rc = copy_rows(h, dst, w[q].src, (size_t)w[q].wire_bytes, w[q].slots, N, r0, r1, ls)) return rc;
      }
      if (int rc = copy_rows(h, h->d_msize.p, t->mount_size, 8, o.D, N, r0, r1, ls)) return rc;
      if (int rc = copy_rows(h, h->d_ocols[1].p, t->agent_id, 4, 1, N, r0, r1, ls)) return rc;
      if (!tuples) if (int rc = copy_rows(h, h->d_port.p, t->port_range, 4, o.S * o.R, N, r0, r1, ls)) return rc;
      if (!pk->hostname_is_agent) if (int rc = copy_rows(h, h->d_ocols[2].p, t->hostname_id, 4, 1, N, r0, r1, ls)) return rc;
      if (!pk->alloc_role_is_const) if (int rc = copy_rows(h, h->d_ocols[0].p, t->alloc_role, 4, 1, N, r0, r1, ls)) return rc;
      WireCols wc{};
      wc.scalar = w[W_SCALAR].stage; wc.present = w[W_PRESENT].stage; wc.mprof = w[W_MPROF].slots > 0 ? w[W_MPROF].stage : nullptr;
      wc.zone = w[W_ZONE].stage; wc.region = w[W_REGION].stage; wc.attr = w[W_ATTR].slots > 0 ? w[W_ATTR].stage : nullptr;
      wc.rr = w[W_RR].slots > 0 ? w[W_RR].stage : nullptr; wc.exec = w[W_EXEC].slots > 0 ? w[W_EXEC].stage : nullptr;
      wc.tuple_id = tuples ? w[W_TUPLE].stage : nullptr;
      wc.present_b = pk->present_bytes; wc.mprof_b = pk->mount_profile_bytes; wc.zone_b = pk->zone_bytes; wc.region_b = pk->region_bytes;
      wc.attr_b = pk->attr_bytes; wc.rr_b = pk->rr_bytes; wc.exec_b = pk->exec_bytes; wc.tuple_b = pk->port_tuple_bytes;
      wc.host_is_agent = pk->hostname_is_agent; wc.alloc_is_const = pk->alloc_role_is_const; wc.alloc_const = pk->alloc_role_const;
      wc.tuples = h->d_ptuples.as<uint32_t>(); wc.n_tuples = pk->n_port_tuples;
      OfferDst od{h->d_scalar.as<double>(), h->d_present.as<uint32_t>(), h->d_port.as<uint32_t>(), h->d_mprof.as<int32_t>(), h->d_ocols[0].as<int32_t>(),
                  h->d_ocols[2].as<int32_t>(), h->d_ocols[3].as<int32_t>(), h->d_ocols[4].as<int32_t>(), h->d_oattr.as<int32_t>(), h->d_orr.as<int32_t>(),
                  h->d_oexec.as<int32_t>(), h->d_ptuple_id.as<int32_t>()};
      k_unpack_offer_rows<<<blocks(r1 - r0), 256, 0, ls>>>(wc, od, o, lim, r0, r1, h->d_offer_err.as<int>() + c);
      CU(cudaGetLastError());
    }
    CU(cudaMemcpyAsync(h->h_offer_err + c, h->d_offer_err.as<int>() + c, sizeof(int), cudaMemcpyDeviceToHost, ls));
    CU(cudaEventRecord(h->chunks[c].landed, ls));
  }
  if (tuples) {
    h->n_port_tuples = pk->n_port_tuples;
    h->tuples_valid = true;
    o.ptuple = h->d_ptuple_id.as<int32_t>();   // slot 0's ids come first
    o.n_ptuples = pk->n_port_tuples;
  }
  h->hasO = true;
  if (async) {
    h->offers_in_flight = true;
    h->validation_pending = true;
  } else {
    h->offers_in_flight = true;   // wait_offers reads every chunk's verdict
    if (int rc = wait_offers(h, "oe_load_offers_packed")) return rc;
    CU(cudaStreamSynchronize(ls));
  }
  return OE_OK;
}

int32_t oe_update_offers(oe_engine* h, const int64_t* rows, int64_t n_rows, const oe_offer_table* t) {
  if (!h || !t || (n_rows > 0 && !rows)) return OE_ERR_INVALID;
  if (!h->hasO) return h->fail(OE_ERR_STATE, "oe_update_offers: no resident offer table (oe_load_offers first)");
  if (int rc = wait_offers(h, "oe_update_offers")) return rc;
  CU(cudaSetDevice(h->device));
  const DevOffers& o = h->dof;
  if (h->n_reserved > 0 || t->n_reserved > 0)
    return h->fail(OE_ERR_STATE, "oe_update_offers: a table with a reserved side table cannot be patched in place; reload it with oe_load_offers");
  if (n_rows < 0 || n_rows > o.N || t->n_offers != n_rows) return h->fail(OE_ERR_INVALID, "oe_update_offers: n_rows");
  if (t->n_role_slots != o.S || t->n_scalar_kinds != o.K || t->n_port_ranges != o.R || t->n_mount_slots != o.D || t->n_attr_slots != o.A ||
      t->n_rr_cols != o.Q || t->n_exec_slots != o.E)
    return h->fail(OE_ERR_INVALID, "oe_update_offers: slot counts differ from the resident table");
  if (n_rows == 0) return OE_OK;
  if (!t->scalar || !t->scalar_present || !t->alloc_role || !t->agent_id || !t->hostname_id || !t->zone_id || !t->region_id ||
      (o.R && !t->port_range) || (o.D && (!t->mount_size || !t->mount_profile)) || (o.A && !t->attr_id) || (o.Q && !t->rr_key) || (o.E && !t->exec_id))
    return h->fail(OE_ERR_INVALID, "oe_update_offers: missing column");
  for (int64_t k = 0; k < n_rows; ++k)
    if (rows[k] < 0 || rows[k] >= o.N) return h->fail(OE_ERR_INVALID, "oe_update_offers: rows[%lld] = %lld is not a row of the resident table", (long long)k, (long long)rows[k]);
  h->class_bits_fresh = h->cycle_done = false;
  const size_t n = (size_t)n_rows;
  // one staging buffer, column after column, each 8-byte aligned
  const size_t slots8 = (size_t)o.S * o.K + o.D, slots4 = 1 + (size_t)o.S * o.R + o.D + 5 + o.A + o.Q + o.E;
  CU(h->d_stage.ensure(n * (slots8 * 8 + slots4 * 4) + 64 * 16));
  CU(h->d_stage_rows.ensure(n * 8));
  CU(cudaMemcpyAsync(h->d_stage_rows.p, rows, n * 8, cudaMemcpyHostToDevice, h->stream));
  const int64_t* d_rows = h->d_stage_rows.as<int64_t>();
  unsigned char* cursor = h->d_stage.as<unsigned char>();
  const oe_dict_sizes& dc = h->dict;
  auto column = [&](auto* dst, const auto* src, int slots, int32_t lo, int32_t hi, int code) -> int {
    using T = std::remove_pointer_t<decltype(dst)>;
    if (slots <= 0) return 0;
    T* st = reinterpret_cast<T*>(cursor);
    const size_t bytes = n * (size_t)slots * sizeof(T);
    cursor += (bytes + 15) / 16 * 16;
    CU(cudaMemcpyAsync(st, src, bytes, cudaMemcpyHostToDevice, h->stream));
    if constexpr (std::is_same_v<T, int32_t>) { if (code) check_ids(h, st, (int64_t)n * slots, lo, hi, code); }
    const int64_t total = (int64_t)n * slots;
    k_scatter_rows<T><<<(unsigned)((total + 255) / 256), 256, 0, h->stream>>>(dst, o.N, st, (int64_t)n, slots, d_rows);
    return 0;
  };
  if (int rc = column(h->d_scalar.as<double>(), t->scalar, o.S * o.K, 0, 0, 0)) return rc;
  if (int rc = column(h->d_present.as<uint32_t>(), t->scalar_present, 1, 0, 0, 0)) return rc;
  if (o.R > 0) {
    uint32_t* st = reinterpret_cast<uint32_t*>(cursor);
    if (int rc = column(h->d_port.as<uint32_t>(), t->port_range, o.S * o.R, 0, 0, 0)) return rc;
    k_validate_ranges<<<(unsigned)((n_rows + 255) / 256), 256, 0, h->stream>>>(st, n_rows, o.S, o.R, 18, h->d_err.as<int>());
    if (h->tuples_valid) {   // the replaced rows' list ids: looked up (or added) in the resident dictionary
      k_reintern_tuples<<<(unsigned)((n_rows * o.S + 127) / 128), 128, 0, h->stream>>>(st, n_rows, o.S, o.R, d_rows, o.N, h->d_ptuples.as<uint32_t>(), h->ptuple_cap,
                                                                                     TupleHash{h->d_tuple_hash.as<int32_t>(), h->tuple_mask}, h->d_ntuples.as<int>(),
                                                                                     h->d_ptuple_id.as<int32_t>());
      CU(cudaMemcpyAsync(h->h_os_info + 2, h->d_ntuples.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
      h->tuple_count_pending = true;
    }
  }
  if (int rc = column(h->d_msize.as<double>(), t->mount_size, o.D, 0, 0, 0)) return rc;
  if (int rc = column(h->d_mprof.as<int32_t>(), t->mount_profile, o.D, -1, OE_MAX_PROFILES, 17)) return rc;
  if (int rc = column(h->d_ocols[0].as<int32_t>(), t->alloc_role, 1, 0, 0, 0)) return rc;
  if (int rc = column(h->d_ocols[1].as<int32_t>(), t->agent_id, 1, 0, std::max(dc.n_agents, 1), 11)) return rc;
  if (int rc = column(h->d_ocols[2].as<int32_t>(), t->hostname_id, 1, 0, std::max(dc.n_hostnames, 1), 12)) return rc;
  if (int rc = column(h->d_ocols[3].as<int32_t>(), t->zone_id, 1, -1, dc.n_zones, 13)) return rc;
  if (int rc = column(h->d_ocols[4].as<int32_t>(), t->region_id, 1, -1, dc.n_regions, 14)) return rc;
  if (int rc = column(h->d_oattr.as<int32_t>(), t->attr_id, o.A, -1, dc.n_attrs, 15)) return rc;
  if (int rc = column(h->d_orr.as<int32_t>(), t->rr_key, o.Q, -1, dc.n_rr_values, 16)) return rc;
  if (int rc = column(h->d_oexec.as<int32_t>(), t->exec_id, o.E, 0, 0, 0)) return rc;
  CU(cudaGetLastError());
  if (h->flags & OE_CFG_ASYNC_LOADS) h->validation_pending = true;
  else if (int rc = fetch_err_flag(h, "oe_update_offers")) { h->hasO = false; return rc; }  // the resident table may hold bad ids now
  return OE_OK;
}

namespace {
// Variants: requirements that have deployed tasks of the same pod type + index (S/offer/TaskUtils.java:136-145) AND
// whose class reads tallies.  Each gets a factored row of its own (row C + v) evaluated with those tasks taken out.
// OE_CYCLE_TRACE=1: host-side timestamps of one oe_evaluate_cycle on stderr (where the call's wall time goes)
struct CycleTrace {
  bool on;
  std::chrono::steady_clock::time_point t0;
  std::string line;
  CycleTrace() : on(getenv("OE_CYCLE_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {}
  void mark(const char* what) {
    if (!on) return;
    char b[64];
    snprintf(b, sizeof b, " %s=%.3f", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    line += b;
  }
  const char* title = "oe cycle";
  ~CycleTrace() { if (on) fprintf(stderr, "[%s, host ms]%s\n", title, line.c_str()); }
};

thread_local CycleTrace* g_trace = nullptr;

int rebuild_variants(oe_engine* h) {
  if (int rc = absorb_tasks(h)) return rc;
  const int64_t M = h->M;
  const int64_t T = (int64_t)h->h_ttype.size();
  // members, not locals: their uploads may still be in flight when this returns (OE_CFG_ASYNC_LOADS)
  std::vector<int32_t>& vcls = h->h_vcls; std::vector<int32_t>& vself_begin = h->h_vself_begin; std::vector<int32_t>& vself_rows = h->h_vself_rows;
  CU(cudaEventSynchronize(h->ev_variants));  // ... and their previous upload must be finished before they are rewritten
  if (g_trace) g_trace->mark("prev_upload_done");
  vcls.clear(); vself_begin.assign(1, 0); vself_rows.clear(); h->h_twin_j.clear();
  h->h_req_row.assign(h->h_class.begin(), h->h_class.end());
  auto add_variant = [&](int64_t j, const int32_t* rows, int n) -> int {
    h->h_req_row[j] = h->C + (int32_t)vcls.size();
    vcls.push_back(h->h_class[j]);
    vself_rows.insert(vself_rows.end(), rows, rows + n);
    vself_begin.push_back((int32_t)vself_rows.size());
    return 0;
  };
  if (h->self_provided) {
    for (int64_t j = 0; j < M; ++j) {
      const int b = h->h_self_begin[j], e = h->h_self_begin[j + 1];
      if (e == b || !h->class_self_sensitive[h->h_class[j]]) continue;
      if (int rc = add_variant(j, h->h_self_rows.data() + b, e - b)) return rc;
    }
  } else if (T > 0 && std::any_of(h->class_self_sensitive.begin(), h->class_self_sensitive.end(), [](uint8_t x) { return x != 0; })) {
    // Which deployed tasks share (pod type, index) with a requirement of this cycle?  Usually very few (new instances) or all
    // of them (a rolling restart).  A bit filter over the requirements' keys first discards the tasks that match nothing —
    // it is small enough to stay in cache, whereas a table over all tasks costs a cache miss per task and per requirement —
    // then an open-addressing table pod key -> chain of task rows is built over the tasks that are left.
    auto key_of = [](int32_t type, int32_t index) { return ((uint64_t)(uint32_t)type << 32) | (uint32_t)index; };
    auto mix = [](uint64_t key) { return (key * 0x9E3779B97F4A7C15ull) >> 17; };
    // Before any hashing: per pod type, the index ranges of the deployed tasks and of the requirements.  Scaling a service out
    // asks for indices above everything deployed, so the ranges of most (usually all) types are disjoint and nothing can match;
    // two streaming passes over tables of n_pod_types entries settle that.
    const size_t n_types = (size_t)std::max(h->dict.n_pod_types, 1);
    std::vector<int32_t>& span = h->h_vspan;   // [4][n_types]: task min, task max, requirement min, requirement max
    if (span.size() != 4 * n_types) { span.assign(4 * n_types, 0); h->task_spans_valid = h->req_spans_valid = false; }
    int32_t *tmin = span.data(), *tmax = tmin + n_types, *rmin = tmax + n_types, *rmax = rmin + n_types;
    if (!h->req_spans_valid) { std::fill(rmin, rmin + n_types, INT32_MAX); std::fill(rmax, rmax + n_types, INT32_MIN); }
    ensure_task_spans(h);
    const bool odd_type = h->task_odd_type;
    bool any_overlap = odd_type;
    if (g_trace) g_trace->mark("task_spans");
    if (!h->req_spans_valid) {
      for (int64_t j = 0; j < M; ++j) {
        if (!h->class_self_sensitive[h->h_class[j]]) continue;
        const int32_t ty = h->h_rtype[j], ix = h->h_rindex[j];
        rmin[ty] = std::min(rmin[ty], ix); rmax[ty] = std::max(rmax[ty], ix);
      }
      h->req_spans_valid = true;
    }
    for (size_t ty = 0; ty < n_types && !any_overlap; ++ty) any_overlap = rmin[ty] <= tmax[ty] && tmin[ty] <= rmax[ty];
    auto may_match = [&](int32_t ty) { return odd_type || (size_t)ty >= n_types || (rmin[ty] <= tmax[ty] && tmin[ty] <= rmax[ty]); };
    if (any_overlap) {
    size_t fbits = 1u << 16;
    while (fbits < (size_t)M * 32 && fbits < (1u << 24)) fbits <<= 1;
    std::vector<uint64_t>& filter = h->h_vfilter;
    filter.assign(fbits / 64, 0ull);
    for (int64_t j = 0; j < M; ++j) {
      if (!h->class_self_sensitive[h->h_class[j]] || !may_match(h->h_rtype[j])) continue;
      const uint64_t q = mix(key_of(h->h_rtype[j], h->h_rindex[j])) & (fbits - 1);
      filter[q >> 6] |= 1ull << (q & 63);
    }
    std::vector<int32_t>& cand = h->h_vcand;  // task rows that may match a requirement, ascending
    cand.clear();
    for (int64_t r = 0; r < T; ++r) {
      if (h->h_ttype[r] < 0 || h->h_tindex[r] == OE_NO_INDEX || !may_match(h->h_ttype[r])) continue;
      const uint64_t q = mix(key_of(h->h_ttype[r], h->h_tindex[r])) & (fbits - 1);
      if ((filter[q >> 6] >> (q & 63)) & 1ull) cand.push_back((int32_t)r);
    }
    if (!cand.empty()) {
      size_t cap = 16;
      while (cap < cand.size() * 2) cap <<= 1;
      const uint64_t EMPTY = ~0ull;
      std::vector<uint64_t> keys(cap, EMPTY);
      std::vector<int32_t> head(cap, -1), next(cand.size(), -1);   // chains hold positions in cand[]
      auto slot_of = [&](uint64_t key) {
        size_t q = (size_t)mix(key) & (cap - 1);
        while (keys[q] != EMPTY && keys[q] != key) q = (q + 1) & (cap - 1);
        return q;
      };
      for (int64_t c = (int64_t)cand.size() - 1; c >= 0; --c) {  // descending, so that every chain lists its rows in ascending order
        const int32_t r = cand[(size_t)c];
        const uint64_t key = key_of(h->h_ttype[r], h->h_tindex[r]);
        const size_t q = slot_of(key);
        keys[q] = key;
        next[(size_t)c] = head[q];
        head[q] = (int32_t)c;
      }
      std::vector<int32_t> chain;
      for (int64_t j = 0; j < M; ++j) {
        if (!h->class_self_sensitive[h->h_class[j]] || !may_match(h->h_rtype[j])) continue;
        const size_t q = slot_of(key_of(h->h_rtype[j], h->h_rindex[j]));
        if (keys[q] == EMPTY) continue;
        chain.clear();
        for (int32_t c = head[q]; c >= 0; c = next[(size_t)c]) chain.push_back(cand[(size_t)c]);
        if (int rc = add_variant(j, chain.data(), (int)chain.size())) return rc;
      }
    }
    }  // any_overlap
  }
  h->V = (int)vcls.size();
  h->V_eff = -1;
  if (g_trace) g_trace->mark("variants");
  // the previous contents of these device buffers may still be read by kernels of the main stream: the side stream starts behind it
  CU(cudaEventRecord(h->ev_side, h->stream));
  if (!(h->flags & OE_CFG_ASYNC_LOADS)) CU(cudaStreamWaitEvent(h->side_stream, h->ev_side, 0));
  if (int rc = upload(h, h->d_req_row, h->h_req_row.data(), (size_t)M, h->side_stream)) return rc;
  if (int rc = upload(h, h->d_vcls, vcls.data(), vcls.size(), h->side_stream)) return rc;
  if (int rc = upload(h, h->d_vself_begin, vself_begin.data(), vself_begin.size(), h->side_stream)) return rc;
  if (int rc = upload(h, h->d_vself_rows, vself_rows.data(), vself_rows.size(), h->side_stream)) return rc;
  CU(h->d_vrrsum.ensure(std::max<size_t>((size_t)h->V * std::max(h->n_csets, 1), 1) * sizeof(int4)));
  CU(cudaEventRecord(h->ev_variants, h->side_stream));
  if (int rc = join_side_uploads(h)) return rc;
  return 0;
}
}  // namespace

namespace {
// The tasks a launch of each requirement's pod stores (oe_req_table.launch_*), kept on the host for oe_apply_launches.
int keep_launch_rows(oe_engine* h, const int32_t* begin, const int32_t* name, const uint32_t* filter) {
  const int64_t M = h->M;
  if (begin[0] != 0) return h->fail(OE_ERR_INVALID, "launch_begin[0] != 0");
  bool monotone = true;
  for (int64_t j = 0; j < M; ++j) monotone &= begin[j + 1] >= begin[j];
  if (!monotone) return h->fail(OE_ERR_INVALID, "launch_begin not monotone");
  const size_t nl = (size_t)begin[M];
  if (nl && (!name || (h->dt.FW && !filter))) return h->fail(OE_ERR_INVALID, "launch rows missing");
  h->h_launch_begin.assign(begin, begin + M + 1);
  h->h_launch_name.assign(name, name + nl);
  h->h_launch_filter.assign(filter, filter + nl * h->dt.FW);
  return 0;
}
}  // namespace

namespace {
// The host side of a requirement load: id checks, host mirrors, and the search for requirements that re-deploy a running pod
// (rebuild_variants).  Synchronous loads run it at once; asynchronous ones only when the device found such a requirement.
int load_requirements_host(oe_engine* h, const int32_t* class_id, const int32_t* pod_type_id, const int32_t* pod_index, int64_t M, CycleTrace& tr) {
  // one pass: the id checks, the host copies, and per pod type the index range of the requirements whose class reads tallies
  // (rebuild_variants compares it with the deployed tasks' range before it hashes anything)
  const size_t n_types = (size_t)std::max(h->dict.n_pod_types, 1);
  if (h->h_vspan.size() != 4 * n_types) { h->h_vspan.assign(4 * n_types, 0); h->task_spans_valid = false; }
  int32_t *rmin = h->h_vspan.data() + 2 * n_types, *rmax = rmin + n_types;
  std::fill(rmin, rmin + n_types, INT32_MAX); std::fill(rmax, rmax + n_types, INT32_MIN);
  h->req_spans_valid = false;
  h->h_class.resize((size_t)M); h->h_rtype.resize((size_t)M); h->h_rindex.resize((size_t)M);
  const bool sensitive_known = h->class_self_sensitive.size() == (size_t)h->C;
  for (int64_t j = 0; j < M; ++j) {
    const int32_t c = class_id[j], ty = pod_type_id[j], ix = pod_index[j];
    if (c < 0 || c >= h->C) return h->fail(OE_ERR_INVALID, "requirement %lld: class id %d", (long long)j, c);
    if (ty < 0 || (size_t)ty >= n_types) return h->fail(OE_ERR_INVALID, "requirement %lld: pod type", (long long)j);
    h->h_class[j] = c; h->h_rtype[j] = ty; h->h_rindex[j] = ix;
    if (sensitive_known && h->class_self_sensitive[c]) { rmin[ty] = std::min(rmin[ty], ix); rmax[ty] = std::max(rmax[ty], ix); }
  }
  h->req_spans_valid = sensitive_known;
  h->reqs_pending = false;
  tr.mark("checked_and_kept");
  if (int rc = rebuild_variants(h)) return rc;
  tr.mark("variants_uploaded");
  return 0;
}
// host mirror of the requirement columns when no requirement re-deploys a pod (every requirement reads its class's bit row)
int absorb_reqs(oe_engine* h) {
  if (!h->reqs_pending) return 0;
  const size_t M = (size_t)h->M;
  if (h->reqs_caller_valid) {
    h->h_class.assign(h->pending_r_class, h->pending_r_class + M);
    h->h_rtype.assign(h->pending_r_type, h->pending_r_type + M);
    h->h_rindex.assign(h->pending_r_index, h->pending_r_index + M);
  } else {
    h->h_class.resize(M); h->h_rtype.resize(M); h->h_rindex.resize(M);
    if (M > 0) {
      CU(cudaMemcpyAsync(h->h_class.data(), h->d_rclass.p, M * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaMemcpyAsync(h->h_rtype.data(), h->d_rtype.p, M * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaMemcpyAsync(h->h_rindex.data(), h->d_rindex.p, M * 4, cudaMemcpyDeviceToHost, h->stream));
      CU(cudaStreamSynchronize(h->stream));
    }
  }
  h->h_req_row = h->h_class;
  for (size_t v = 0; v < h->h_twin_j.size(); ++v) h->h_req_row[(size_t)h->h_twin_j[v]] = h->C + (int32_t)v;   // variants found on the device
  h->req_spans_valid = false;
  h->reqs_pending = false;
  return 0;
}
int absorb_launch_rows(oe_engine* h) {
  if (h->pending_launch_rows < 0) return 0;
  const size_t M = (size_t)h->M, nl = (size_t)h->pending_launch_rows, FW = (size_t)h->dt.FW;
  h->h_launch_begin.resize(M + 1); h->h_launch_name.resize(nl); h->h_launch_filter.resize(nl * FW);
  CU(cudaMemcpyAsync(h->h_launch_begin.data(), h->d_lr_begin.p, (M + 1) * 4, cudaMemcpyDeviceToHost, h->stream));
  if (nl) CU(cudaMemcpyAsync(h->h_launch_name.data(), h->d_lr_name.p, nl * 4, cudaMemcpyDeviceToHost, h->stream));
  if (nl * FW) CU(cudaMemcpyAsync(h->h_launch_filter.data(), h->d_lr_filter.p, nl * FW * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->pending_launch_rows = -1;
  return 0;
}
int absorb_host_mirrors(oe_engine* h) {
  if (int rc = absorb_tasks(h)) return rc;
  if (int rc = absorb_reqs(h)) return rc;
  return absorb_launch_rows(h);
}
// Requirements that re-deploy running pods exist (OE_CFG_ASYNC_LOADS): their variants from the records k_match_tasks left, sorted
// by (requirement, task row) so that variants are numbered by ascending requirement and list their rows ascending, exactly as
// rebuild_variants does.  More records than the buffer holds (a pod listed by many requirements): the host-side search.
int resolve_twins(oe_engine* h) {
  const int n = h->h_os_info[3];
  CycleTrace tr;
  tr.title = "re-deployed pods";
  if (n > h->pairs_cap) {
    const int32_t *rc_ = h->pending_r_class, *rt_ = h->pending_r_type, *ri_ = h->pending_r_index;
    if (!h->reqs_caller_valid) {   // an earlier evaluation call failed before it got here: the caller's tables may be gone
      if (int rc = absorb_reqs(h)) return rc;
      rc_ = h->h_class.data(); rt_ = h->h_rtype.data(); ri_ = h->h_rindex.data();
    }
    return load_requirements_host(h, rc_, rt_, ri_, h->M, tr);
  }
  std::vector<int4>& pairs = h->h_pairs;
  pairs.resize((size_t)n);
  CU(cudaMemcpyAsync(pairs.data(), h->d_pairs.p, (size_t)n * sizeof(int4), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  tr.mark("records");
  std::sort(pairs.begin(), pairs.end(), [](const int4& a, const int4& b) { return a.x != b.x ? a.x < b.x : a.y < b.y; });
  CU(cudaEventSynchronize(h->ev_variants));   // the previous upload of these vectors
  std::vector<int32_t>& vcls = h->h_vcls; std::vector<int32_t>& vself_begin = h->h_vself_begin; std::vector<int32_t>& vself_rows = h->h_vself_rows;
  vcls.clear(); vself_begin.assign(1, 0); vself_rows.clear(); h->h_twin_j.clear();
  vself_rows.reserve((size_t)n);
  for (int k = 0; k < n; ++k) {
    if (k == 0 || pairs[(size_t)k].x != pairs[(size_t)k - 1].x) {
      if (k) vself_begin.push_back((int32_t)vself_rows.size());
      vcls.push_back(pairs[(size_t)k].z);
      h->h_twin_j.push_back(pairs[(size_t)k].x);
    }
    vself_rows.push_back(pairs[(size_t)k].y);
  }
  vself_begin.push_back((int32_t)vself_rows.size());
  h->V = (int)vcls.size();
  h->V_eff = -1;
  tr.mark("variants");
  if (int rc = upload(h, h->d_vcls, vcls.data(), vcls.size())) return rc;
  if (int rc = upload(h, h->d_vself_begin, vself_begin.data(), vself_begin.size())) return rc;
  if (int rc = upload(h, h->d_vself_rows, vself_rows.data(), vself_rows.size())) return rc;
  if (int rc = upload(h, h->d_twin_j, h->h_twin_j.data(), h->h_twin_j.size())) return rc;
  CU(h->d_vrrsum.ensure(std::max<size_t>((size_t)h->V * std::max(h->n_csets, 1), 1) * sizeof(int4)));
  k_set_variant_rows<<<(unsigned)((h->V + 255) / 256), 256, 0, h->stream>>>(h->d_req_row.as<int32_t>(), h->d_twin_j.as<int32_t>(), h->V, h->C);
  CU(cudaGetLastError());
  CU(cudaEventRecord(h->ev_variants, h->stream));
  tr.mark("uploaded");
  return 0;
}
// When an evaluation entry point returns, the caller may reuse the tables it loaded (OE_CFG_ASYNC_LOADS, include/oe_abi.h).
struct CallerTablesExpire {
  oe_engine* h;
  ~CallerTablesExpire() { if (h) h->tasks_caller_valid = h->reqs_caller_valid = false; }
};
}  // namespace

int32_t oe_load_requirements(oe_engine* h, const oe_req_table* t) {
  if (!h || !t) return OE_ERR_INVALID;
  if (!h->hasP || !h->hasT) return h->fail(OE_ERR_STATE, "oe_load_requirements: load programs and tasks first");
  CU(cudaSetDevice(h->device));
  h->hasR = false;
  h->counters_fresh = h->class_bits_fresh = h->cycle_done = false;
  h->reqs_pending = h->twins_unresolved = false;
  CycleTrace tr;
  tr.title = "oe_load_requirements";
  struct TraceScope { TraceScope(CycleTrace* t) { g_trace = t; } ~TraceScope() { g_trace = nullptr; } } trace_scope(&tr);
  const int64_t M = t->n_reqs;
  if (M < 0 || M > 0x7ffffff0) return h->fail(OE_ERR_INVALID, "n_reqs");
  if (M > 0 && (!t->class_id || !t->pod_type_id || !t->pod_index)) return h->fail(OE_ERR_INVALID, "oe_load_requirements: missing column");
  h->M = M;
  h->self_provided = t->self_begin != nullptr;
  h->h_self_begin.clear(); h->h_self_rows.clear();
  static const bool host_twins = getenv("OE_HOST_TWIN_SEARCH") != nullptr;   // measurement aid: always the host-side search
  const bool lazy = (h->flags & OE_CFG_ASYNC_LOADS) && !h->self_provided && M > 0 && !host_twins && h->class_self_sensitive.size() == (size_t)h->C;
  h->h_twin_j.clear();
  if (lazy) {
    // Nothing per requirement on the host: the three columns go to the device as they are (bit row = class), the ids are checked
    // there, and k_match_tasks lists the deployed tasks of the pods that requirements re-deploy.  The next evaluation call reads
    // the number of records together with the validation flags; only when it is not zero is there more to do (resolve_twins).
    const int64_t T = h->dt.T;
    h->V = 0; h->V_eff = -1;
    h->h_vcls.clear(); h->h_vself_begin.assign(1, 0); h->h_vself_rows.clear();
    if (int rc = upload(h, h->d_rclass, t->class_id, (size_t)M, h->side_stream)) return rc;
    if (int rc = upload(h, h->d_rtype, t->pod_type_id, (size_t)M, h->side_stream)) return rc;
    if (int rc = upload(h, h->d_rindex, t->pod_index, (size_t)M, h->side_stream)) return rc;
    CU(h->d_req_row.ensure((size_t)M * 4));
    CU(h->d_vcls.ensure(4)); CU(h->d_vself_begin.ensure(8)); CU(h->d_vself_rows.ensure(4)); CU(h->d_vrrsum.ensure(sizeof(int4)));
    uint64_t size = 64;
    while (size < (uint64_t)M * 2) size <<= 1;
    h->pairs_cap = (int)std::min<int64_t>(T + M / 4 + 64, 0x3ffffff0);
    CU(h->d_pod_hash.ensure(size * 8)); CU(h->d_pod_head.ensure(size * 4)); CU(h->d_pod_next.ensure((size_t)M * 4));
    CU(h->d_pairs.ensure((size_t)h->pairs_cap * sizeof(int4))); CU(h->d_twin.ensure(sizeof(int)));
    if (int rc = join_side_uploads(h)) return rc;
    CU(cudaMemcpyAsync(h->d_req_row.p, h->d_rclass.p, (size_t)M * 4, cudaMemcpyDeviceToDevice, h->stream));
    check_ids(h, h->d_rclass.as<int32_t>(), M, 0, std::max(h->C, 1), 21);
    check_ids(h, h->d_rtype.as<int32_t>(), M, 0, std::max(h->dict.n_pod_types, 1), 22);
    CU(cudaMemsetAsync(h->d_twin.p, 0, sizeof(int), h->stream));
    h->h_os_info[3] = 0;
    if (T > 0) {
      const PodHash hs{h->d_pod_hash.as<unsigned long long>(), h->d_pod_head.as<int32_t>(), h->d_pod_next.as<int32_t>(), (uint32_t)(size - 1)};
      CU(cudaMemsetAsync(hs.keys, 0xff, size * 8, h->stream));
      CU(cudaMemsetAsync(hs.head, 0xff, size * 4, h->stream));
      k_pod_hash_insert<<<(unsigned)((M + 255) / 256), 256, 0, h->stream>>>(h->d_rclass.as<int32_t>(), h->d_rtype.as<int32_t>(), h->d_rindex.as<int32_t>(), M, h->C,
                                                                            h->d_sensitive.as<uint8_t>(), hs);
      k_match_tasks<<<(unsigned)((T + 255) / 256), 256, 0, h->stream>>>(h->dt.type, h->dt.index, T, hs, h->d_rclass.as<int32_t>(), h->d_pairs.as<int4>(), h->pairs_cap,
                                                                        h->d_twin.as<int>());
      CU(cudaGetLastError());
      CU(cudaMemcpyAsync(h->h_os_info + 3, h->d_twin.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    }
    h->pending_r_class = t->class_id; h->pending_r_type = t->pod_type_id; h->pending_r_index = t->pod_index;
    h->reqs_pending = h->twins_unresolved = true;
    h->reqs_caller_valid = true;
    h->validation_pending = true;
    tr.mark("enqueued");
  } else {
    if (h->self_provided) {
      if (int rc = absorb_tasks(h)) return rc;
      const int64_t T = (int64_t)h->h_ttype.size();
      for (int64_t j = 0; j < M; ++j) if (t->self_begin[j + 1] < t->self_begin[j]) return h->fail(OE_ERR_INVALID, "self_begin not monotone");
      h->h_self_begin.assign(t->self_begin, t->self_begin + M + 1);
      h->h_self_rows.assign(t->self_rows, t->self_rows + t->self_begin[M]);
      for (int32_t r : h->h_self_rows) if (r < 0 || r >= T) return h->fail(OE_ERR_INVALID, "self_rows out of range");
    }
    if (int rc = load_requirements_host(h, t->class_id, t->pod_type_id, t->pod_index, M, tr)) return rc;
    if (int rc = upload(h, h->d_rtype, t->pod_type_id, (size_t)M, h->side_stream)) return rc;
    if (int rc = upload(h, h->d_rindex, t->pod_index, (size_t)M, h->side_stream)) return rc;
    if (int rc = join_side_uploads(h)) return rc;
  }
  CU(h->d_dec.ensure(std::max<size_t>((size_t)M, 1) * sizeof(oe_decision)));

  h->h_launch_begin.clear(); h->h_launch_name.clear(); h->h_launch_filter.clear();
  h->pending_launch_rows = -1;
  if (t->launch_begin) {
    if (h->flags & OE_CFG_ASYNC_LOADS) {
      // only oe_apply_launches reads the launch rows: they go to the device as they are (checked there, code 23) and come back
      // when that entry is called
      const int64_t nl = t->launch_begin[M];
      if (nl < 0 || nl > 0x7ffffff0) return h->fail(OE_ERR_INVALID, "launch_begin not monotone");
      if (nl && (!t->launch_name_id || (h->dt.FW && !t->launch_filter_bits))) return h->fail(OE_ERR_INVALID, "launch rows missing");
      if (int rc = upload(h, h->d_lr_begin, t->launch_begin, (size_t)M + 1, h->side_stream)) return rc;
      if (int rc = upload(h, h->d_lr_name, t->launch_name_id, (size_t)nl, h->side_stream)) return rc;
      if (int rc = upload(h, h->d_lr_filter, t->launch_filter_bits, (size_t)nl * h->dt.FW, h->side_stream)) return rc;
      if (int rc = join_side_uploads(h)) return rc;
      k_validate_offsets<<<(unsigned)((M + 256) / 256), 256, 0, h->stream>>>(h->d_lr_begin.as<int32_t>(), M, 23, h->d_err.as<int>());
      CU(cudaGetLastError());
      h->pending_launch_rows = nl;
      h->validation_pending = true;
    } else if (int rc = keep_launch_rows(h, t->launch_begin, t->launch_name_id, t->launch_filter_bits)) {
      return rc;
    }
  }
  tr.mark("launch_rows_kept");
  if (!(h->flags & OE_CFG_ASYNC_LOADS)) CU(cudaStreamSynchronize(h->stream));
  h->hasR = true;
  return OE_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Evaluation
// ---------------------------------------------------------------------------------------------------------------------
namespace {

struct Phase { cudaEvent_t a, b; };

int run_counters(oe_engine* h, int64_t* launches) {
  const size_t ncnt = (size_t)h->h_cs_off.back();
  if (ncnt) CU(cudaMemsetAsync(h->d_cnt.p, 0, ncnt * 4, h->stream));
  if (h->n_tsets) {
    CU(cudaMemsetAsync(h->d_type_total.p, 0, (size_t)h->n_tsets * 4, h->stream));
    CU(cudaMemsetAsync(h->d_type_agent.p, 0, (size_t)h->n_tsets * h->dp.n_agents * 4, h->stream));
  }
  if (h->dt.T > 0 && (h->n_csets || h->n_tsets)) {
    k_count_tasks<<<(unsigned)((h->dt.T + 255) / 256), 256, 0, h->stream>>>(h->dt, h->dp, h->d_cnt.as<int32_t>(), h->d_type_total.as<int32_t>(),
                                                                         h->d_type_agent.as<int32_t>());
    ++*launches;
  }
  if (h->n_csets && h->any_rr_used) {
    // CTAs per counter set: one for a few thousand keys (zones, regions, attributes), up to 64 for a key per agent
    const int splits = (int)std::max<int64_t>(1, std::min<int64_t>(64, h->max_rr_keys / 16384));
    if ((1 + (int64_t)h->V) > 65535) return h->fail(OE_ERR_LIMIT, "oe: %d re-deployed requirements exceed what one tally-summary launch covers", h->V);
    dim3 grid((unsigned)h->n_csets, (unsigned)(1 + h->V), (unsigned)splits);
    if (splits > 1) {
      k_rr_preset<<<(unsigned)((h->n_csets + 255) / 256), 256, 0, h->stream>>>(h->d_rrsum.as<int4>(), h->n_csets);
      if (h->V > 0) k_rr_preset<<<(unsigned)(((int64_t)h->V * h->n_csets + 255) / 256), 256, 0, h->stream>>>(h->d_vrrsum.as<int4>(), (int64_t)h->V * h->n_csets);
    }
    k_rr_summary<<<grid, 256, 0, h->stream>>>(h->ctx(), h->d_cnt.as<int32_t>(), h->d_rr_used.as<uint8_t>(), h->d_rrsum.as<int4>(), h->d_vrrsum.as<int4>());
    ++*launches;
    if (splits > 1) {
      k_rr_finish<<<(unsigned)((h->n_csets + 255) / 256), 256, 0, h->stream>>>(h->d_rrsum.as<int4>(), h->n_csets);
      if (h->V > 0) k_rr_finish<<<(unsigned)(((int64_t)h->V * h->n_csets + 255) / 256), 256, 0, h->stream>>>(h->d_vrrsum.as<int4>(), (int64_t)h->V * h->n_csets);
    }
  }
  CU(cudaGetLastError());
  h->counters_fresh = true;
  return 0;
}

// ---- folding of new-pod pipelines into FastRow descriptors (see oe_device.cuh)
inline uint64_t double_order_key(double d) {  // monotone map of the non-NaN doubles onto unsigned integers
  uint64_t u;
  std::memcpy(&u, &d, 8);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
inline double double_from_key(uint64_t k) {
  const uint64_t u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  double d;
  std::memcpy(&d, &u, 8);
  return d;
}
// The literal stage chain on one pool entry (MesosResourcePool.java:330-372 over ValueUtils.java:103-106,144-155).
inline bool scalar_chain_passes(double v, const std::vector<double>& amounts) {
  for (double a : amounts) {
    volatile double d = a - v;
    if (d > 0.0) return false;
    volatile double nv = v - a;
    v = nv;
  }
  return true;
}
// Smallest starting value for which the chain passes; false when the chain is not of the foldable kind.
bool fold_scalar_chain(const std::vector<double>& amounts, double* thr) {
  for (double a : amounts) if (a != a) return false;
  const double inf = std::numeric_limits<double>::infinity();
  if (!scalar_chain_passes(inf, amounts)) return false;
  uint64_t lo = double_order_key(-inf), hi = double_order_key(inf);  // invariant: passes(hi)
  if (scalar_chain_passes(-inf, amounts)) { *thr = -inf; return true; }
  while (hi - lo > 1) {  // fails(lo), passes(hi)
    const uint64_t mid = lo + (hi - lo) / 2;
    if (scalar_chain_passes(double_from_key(mid), amounts)) hi = mid; else lo = mid;
  }
  *thr = double_from_key(hi);
  return true;
}

int build_fast_rows(oe_engine* h) {
  const int K = h->dof.K;
  std::vector<FastRow> rows((size_t)std::max(h->C, 1));
  bool any_tab = false;
  // (zone, region) combinations incl. "absent": tabulate rules only while a row's table stays small
  h->tab_zdim = h->dict.n_zones + 1;
  h->tab_rdim = h->dict.n_regions + 1;
  const int64_t combos = (int64_t)h->tab_zdim * h->tab_rdim;
  h->tab_words = combos <= 8192 ? (int)((combos + 31) / 32) : 0;
  for (int c = 0; c < h->C; ++c) {
    FastRow fr;
    std::memset(&fr, 0, sizeof fr);
    fr.pod_role = h->h_pod_role[c];
    bool ok = true, dyn_seen = false;
    int n_rules = 0;
    std::vector<int> qs;                       // distinct pool entries in first-use order
    std::vector<std::vector<double>> chains;
    std::vector<int> statics;
    const int eb = h->h_excl_begin[c], ee = h->h_excl_begin[c + 1];
    auto excluded = [&](int p) { for (int q = eb; q < ee; ++q) if (h->h_excl[q] == p) return true; return false; };
    for (int x = h->h_op_begin[c]; x < h->h_op_begin[c + 1] && ok; ++x) {
      const oe_stage_op& op = h->h_ops[x];
      if (op.flags) { ok = false; break; }
      switch (op.opcode) {
        case OE_OP_NOP: break;
        case OE_OP_EXEC_ID: ok = op.arg < 0; break;
        case OE_OP_RULE:
          if (++n_rules > 1) { ok = false; break; }
          fr.rule_begin = op.arg; fr.rule_count = (uint16_t)op.aux;
          break;
        case OE_OP_SCALAR_NEW: {
          const int q = op.slot * K + op.kind;
          size_t k = 0;
          while (k < qs.size() && qs[k] != q) ++k;
          if (k == qs.size()) { qs.push_back(q); chains.emplace_back(); }
          chains[k].push_back(op.amount);
          fr.flags |= OE_FR_NEEDS_POOL;
        } break;
        case OE_OP_PORT_STATIC:
          if (op.slot != 0 || (dyn_seen && !excluded(op.arg) && !h->fast_port_lists)) { ok = false; break; }
          if (dyn_seen && !excluded(op.arg)) fr.flags |= OE_FR_PORTS_BY_TABLE;   // a static port behind a dynamic one that may have taken it
          if (std::find(statics.begin(), statics.end(), op.arg) != statics.end()) fr.flags |= OE_FR_NEVER;
          else statics.push_back(op.arg);
          fr.flags |= OE_FR_NEEDS_POOL;
          break;
        case OE_OP_PORT_DYNAMIC:
          if (op.slot != 0 || fr.n_dyn == 255) { ok = false; break; }
          if (op.arg >= 0) {   // range-constrained: no closed form per offer, but a verdict per distinct range list
            if (!h->fast_port_lists) { ok = false; break; }
            fr.flags |= OE_FR_PORTS_BY_TABLE;
          }
          dyn_seen = true;
          ++fr.n_dyn;
          fr.flags |= OE_FR_NEEDS_POOL;
          break;
        case OE_OP_VOL_MOUNT:
          if (fr.has_mount) { ok = false; break; }   // two MOUNT volumes compete for the disks: the interpreter keeps track
          fr.has_mount = 1; fr.mount_amount = op.amount; fr.mount_mask = op.mask;
          fr.flags |= OE_FR_NEEDS_POOL;
          break;
        default: ok = false;
      }
    }
    if (ok && qs.size() > 8) ok = false;
    const bool dense = h->dof.S * K <= 8;
    if (dense) {
      uint32_t need = 0;
      for (int q = 0; q < 8; ++q) fr.thr[q] = -std::numeric_limits<double>::infinity();
      for (size_t k = 0; ok && k < qs.size(); ++k) {
        ok = qs[k] < 8 && fold_scalar_chain(chains[k], &fr.thr[qs[k]]);
        need |= 1u << qs[k];
      }
      std::memcpy(fr.thr_q, &need, 4);
      fr.flags |= OE_FR_DENSE;
    } else {
      for (size_t k = 0; ok && k < qs.size(); ++k) {
        if (qs[k] >= 32) { ok = false; break; }
        fr.thr_q[k] = (uint8_t)qs[k];
        ok = fold_scalar_chain(chains[k], &fr.thr[k]);
      }
    }
    fr.n_thr = (uint8_t)qs.size();
    if (ok) {
      std::vector<int> listed(statics);
      for (int q = eb; q < ee; ++q) listed.push_back(h->h_excl[q]);
      std::sort(listed.begin(), listed.end());
      listed.erase(std::unique(listed.begin(), listed.end()), listed.end());
      if (listed.size() > 16) ok = false;
      for (size_t u = 0; ok && u < listed.size(); ++u) {
        fr.port[u] = (uint16_t)listed[u];
        if (std::find(statics.begin(), statics.end(), listed[u]) != statics.end()) fr.static_mask |= 1u << u;
      }
      // ports that are only excluded matter only when something is picked dynamically
      fr.n_ports = (uint8_t)listed.size();
    }
    if (ok && fr.rule_count > 0 && h->tab_words > 0) {
      bool zone_region_only = true;
      for (int n = fr.rule_begin; n < fr.rule_begin + fr.rule_count; ++n) {
        const oe_rule_node& nd = h->h_nodes[n];
        switch (nd.op) {
          case OE_R_TRUE: case OE_R_FALSE: case OE_R_AND: case OE_R_OR: case OE_R_NOT:
          case OE_R_ZONE_IN: case OE_R_REGION_IN: case OE_R_LOCAL_REGION: case OE_R_MAXPER_ZONE: case OE_R_MAXPER_REGION: break;
          case OE_R_RR: {
            const int ks = h->h_csets[nd.a].keyspace;
            if (ks != OE_KS_ZONE && ks != OE_KS_REGION) zone_region_only = false;
          } break;
          default: zone_region_only = false;
        }
      }
      if (zone_region_only) fr.flags |= OE_FR_RULE_TAB;
    }
    if (ok) fr.flags |= OE_FR_FAST; else fr.flags = 0;
    fr.cls = (uint16_t)c;
    any_tab = any_tab || (fr.flags & OE_FR_RULE_TAB);
    rows[c] = fr;
  }
  if (int rc = upload(h, h->d_fast, rows.data(), rows.size())) return rc;
  CU(cudaStreamSynchronize(h->stream));  // `rows` is a local
  h->fast_K = K;
  h->fast_S = h->dof.S;
  h->any_rule_tab = any_tab;
  // the lean kernel takes a table whose classes are all folded, dense, and read nothing per offer but tabulated verdicts
  h->all_lean = h->C > 0; h->any_ports = false; h->any_rule = false;
  for (int c = 0; c < h->C; ++c) {
    const FastRow& fr = rows[(size_t)c];
    const bool rule_ok = fr.rule_count == 0 || (fr.flags & OE_FR_RULE_TAB);
    h->all_lean = h->all_lean && (fr.flags & OE_FR_FAST) && (fr.flags & OE_FR_DENSE) && rule_ok && !fr.has_mount;
    h->any_ports = h->any_ports || (fr.n_ports | fr.n_dyn);
    h->any_rule = h->any_rule || (fr.flags & OE_FR_RULE_TAB);
  }
  return 0;
}


// A re-deployed requirement (a pod that is already running) gets a bit row of its own because its rules must not count the pod's
// own tasks.  For classes whose rule reads only the offer's zone and region that difference is the content of the rule table
// k_rule_tables has just written: requirements of one class with the same table have the same bit row (and one with the class's
// own table needs no row at all).  Thousands of re-deployed pods of a few classes over a few zones collapse into a handful of
// rows.  The tables are small ((C + V) x tab_words words); they are compared on the host, which costs one synchronisation per
// pass — only when re-deployed requirements exist.
int merge_equal_variants(oe_engine* h) {
  const int C = h->C, V = h->V, words = h->tab_words;
  if (words <= 0) return 0;
  h->h_rule_tab.resize((size_t)(C + V) * words);
  std::vector<FastRow> fast((size_t)std::max(C, 1));
  CU(cudaMemcpyAsync(h->h_rule_tab.data(), h->d_rule_tab.p, h->h_rule_tab.size() * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(fast.data(), h->d_fast.p, (size_t)C * sizeof(FastRow), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  // key = class + table words; value = the bit row that evaluates it
  struct KeyHash {
    size_t operator()(const std::vector<uint32_t>& k) const {
      uint64_t x = 1469598103934665603ull;
      for (uint32_t w : k) { x ^= w; x *= 1099511628211ull; }
      return (size_t)x;
    }
  };
  std::unordered_map<std::vector<uint32_t>, int32_t, KeyHash> seen;
  seen.reserve((size_t)(C + V) * 2);
  std::vector<uint32_t> key((size_t)words + 1);
  auto key_of = [&](int cls, int frow) {
    key[0] = (uint32_t)cls;
    std::memcpy(key.data() + 1, h->h_rule_tab.data() + (size_t)frow * words, (size_t)words * 4);
  };
  h->h_row_map.resize((size_t)C);
  for (int c = 0; c < C; ++c) {
    h->h_row_map[(size_t)c] = c;
    if (fast[(size_t)c].flags & OE_FR_RULE_TAB) { key_of(c, c); seen.emplace(key, c); }
  }
  std::vector<int32_t> bit_row_of_variant((size_t)V);
  for (int v = 0; v < V; ++v) {
    const int cls = h->h_vcls[(size_t)v];
    int32_t row;
    if (fast[(size_t)cls].flags & OE_FR_RULE_TAB) {
      key_of(cls, C + v);
      auto it = seen.find(key);
      if (it == seen.end()) { row = (int32_t)h->h_row_map.size(); h->h_row_map.push_back(C + v); seen.emplace(key, row); }
      else row = it->second;
    } else {
      row = (int32_t)h->h_row_map.size();
      h->h_row_map.push_back(C + v);
    }
    bit_row_of_variant[(size_t)v] = row;
  }
  const int V_eff = (int)h->h_row_map.size() - C;
  if (V_eff == V) return 0;   // nothing merged: keep the plain numbering
  if (int rc = upload(h, h->d_row_map, h->h_row_map.data(), h->h_row_map.size())) return rc;
  if (int rc = upload(h, h->d_row_of_variant, bit_row_of_variant.data(), bit_row_of_variant.size())) return rc;
  CU(h->d_req_row_eff.ensure(std::max<size_t>((size_t)h->M, 1) * 4));
  k_merged_rows<<<(unsigned)((h->M + 255) / 256), 256, 0, h->stream>>>(h->d_req_row.as<int32_t>(), h->M, C, h->d_row_of_variant.as<int32_t>(), h->d_req_row_eff.as<int32_t>());
  CU(cudaGetLastError());
  h->req_row_eff_on_host = false;
  CU(cudaStreamSynchronize(h->stream));   // the vectors are rewritten by the next pass
  h->V_eff = V_eff;
  return 0;
}

int run_expand(oe_engine* h, int64_t* launches, cudaStream_t on = nullptr, int64_t w0 = 0, int64_t w1 = -1);
bool offer_side_ok(int rows);
int prepare_queues(oe_engine* h, int64_t* launches, const int32_t* req_row, int64_t M, int rows, cudaStream_t st);

// K1 over the whole offer table, or — while an OE_CFG_ASYNC_LOADS table is still arriving — chunk by chunk as the chunks land.
// expand_beside (oe_evaluate_cycle with OE_CFG_MATERIALIZE_MATRIX): the matrix columns of every chunk but the last are expanded on
// the auxiliary stream as soon as the chunk's class bits exist; the last chunk's are left to the fork at the commit launch.
int run_class_feasibility(oe_engine* h, int64_t* launches, bool expand_beside = false, bool expand_last = false,
                          const std::function<int(int)>* after_chunk = nullptr) {
  const int64_t stride = h->stride();
  h->V_eff = -1;
  const int rows_all = h->C + h->V;   // every class and every re-deployed requirement
  h->expand_from_word = 0;
  h->expand_started = false;
  K1Args a{};
  if (rows_all > 0) {
    // classes whose port stages exist only as a verdict per range list are folded exactly while the table has its lists
    const bool port_lists = h->tuples_valid && h->dof.ptuple && h->dof.R <= OE_FAST_RANGES && h->n_port_tuples > 0 &&
                            (int64_t)h->n_port_tuples * 4 <= std::max<int64_t>(h->N, 4096);
    if (h->fast_K != h->dof.K || h->fast_S != h->dof.S || h->fast_port_lists != port_lists) {
      h->fast_port_lists = port_lists;
      if (int rc = build_fast_rows(h)) return rc;   // also sizes the rule tables
    }
    a.tab.bits = nullptr; a.tab.words = h->tab_words; a.tab.zdim = h->tab_zdim;
    if (h->any_rule_tab) {
      CU(h->d_rule_tab.ensure((size_t)rows_all * h->tab_words * 4));
      a.tab.bits = h->d_rule_tab.as<uint32_t>();
      dim3 tgrid((unsigned)((h->tab_zdim * h->tab_rdim + 127) / 128), (unsigned)rows_all);
      if (rows_all > 65535) return h->fail(OE_ERR_LIMIT, "oe: %d factored rows exceed what one rule-table launch covers", rows_all);
      k_rule_tables<<<tgrid, 128, 0, h->stream>>>(h->ctx(), h->d_fast.as<FastRow>(), h->d_rule_tab.as<uint32_t>(), h->tab_words,
                                                  h->tab_zdim, h->tab_rdim);
      ++*launches;
      static const bool no_merge = getenv("OE_NO_ROW_MERGE") != nullptr;   // debugging aid: one bit row per re-deployed requirement
      // (sharded engines merge too: tasks, programs and requirements are replicated, so every rank arrives at the same rows)
      if (h->V > 0 && !no_merge) if (int rc = merge_equal_variants(h)) return rc;
    }
  }
  const int rows = h->rows_factored();
  a.row_map = h->V_eff >= 0 ? h->d_row_map.as<int32_t>() : nullptr;
  h->queues_rows = -1;
  if (h->want_queues && h->M > 0 && rows > 0 && offer_side_ok(rows)) {
    // the commit will want every bit row's requirements in plan order: that needs the requirement table only, so it is sorted
    // on the side stream while this pass runs
    CU(cudaEventRecord(h->ev_side, h->stream));
    CU(cudaStreamWaitEvent(h->side_stream, h->ev_side, 0));
    if (int rc = prepare_queues(h, launches, h->bit_row_of_req(), h->M, rows, h->side_stream)) return rc;
    CU(cudaEventRecord(h->ev_queues, h->side_stream));
    h->queues_rows = rows;
  }
  CU(h->d_F.ensure(std::max<size_t>((size_t)rows * stride, 1) * 4));
  if (rows > 0) {
    const int64_t xblocks_all = stride * 32 / OE_K1_THREADS;
    const int n_chunks = h->offers_in_flight ? h->n_chunks : 1;
    a.ptab.bits = nullptr; a.ptab.words = 0;
    {
      // few distinct `ports` range lists (stock agents offer the same ranges): a class's port stages once per list.  With
      // nearly as many lists as offers the table would cost what it saves, and the per-offer test stays.
      if (h->fast_port_lists && h->C > 0) {
        a.ptab.words = (h->n_port_tuples + 31) / 32;
        CU(h->d_port_tab.ensure((size_t)h->C * a.ptab.words * 4));
        k_port_tables<<<dim3((unsigned)((h->n_port_tuples + 127) / 128), (unsigned)h->C), 128, 0, h->stream>>>(
            h->dp, h->d_fast.as<FastRow>(), h->d_ptuples.as<uint32_t>(), h->n_port_tuples, h->dof.R, h->d_port_tab.as<uint32_t>(), a.ptab.words);
        ++*launches;
        a.ptab.bits = h->d_port_tab.as<uint32_t>();
      }
    }
    a.F = h->d_F.as<uint32_t>(); a.stride = stride; a.rows = rows;
    a.fast = h->d_fast.as<FastRow>();
    a.n_nodes = h->n_rule_nodes <= OE_K1_MAX_SMEM_NODES ? h->n_rule_nodes : 0;
    a.sc_cols = 0;  // S*K <= 8: the pool entries live in registers (OE_FR_DENSE); otherwise they are read through L1
    if (h->peers_attached) {
      if (rows > h->exch_max_rows || stride != h->exch_stride)
        return h->fail(OE_ERR_LIMIT, "oe: the peer exchange was exported for %d bit rows of %lld words, this pass has %d rows of %lld words: "
                       "export and attach again", h->exch_max_rows, (long long)h->exch_stride, rows, (long long)stride);
      ++h->peer_epoch;
      a.peers.n = h->world;
      for (int p = 0; p < h->world; ++p) a.peers.base[p] = h->exch_rows(p, h->peer_epoch) + (size_t)h->rank * rows * stride;
    }
    // the lean form (k_class_feasibility_lean): every class folded, four pool entries at most, rules tabulated, ports tabulated
    static const bool no_lean = getenv("OE_K1_GENERAL") != nullptr;   // measurement aid: always the general kernel
    bool lean_ok = !no_lean && h->all_lean && h->dof.S * h->dof.K <= 4 && (!h->any_ports || a.ptab.bits) && (!h->any_rule || a.tab.bits);
    LeanArgs la{};
    if (lean_ok) {
      la.F = a.F; la.stride = stride; la.rows = rows; la.row_words = (rows + 31) / 32; la.zdim = h->tab_zdim; la.peers = a.peers;
      const int n_combos = h->any_rule ? h->tab_zdim * h->tab_rdim : 0, n_tuples = h->any_ports ? h->n_port_tuples : 0;
      CU(h->d_lean.ensure((size_t)la.row_words * 32 * sizeof(LeanRow)));
      CU(h->d_ruleT.ensure(std::max<size_t>((size_t)n_combos * la.row_words, 1) * 4));
      CU(h->d_portT.ensure(std::max<size_t>((size_t)n_tuples * la.row_words, 1) * 4));
      la.lean = h->d_lean.as<LeanRow>();
      la.ruleT = n_combos ? h->d_ruleT.as<uint32_t>() : nullptr;
      la.portT = n_tuples ? h->d_portT.as<uint32_t>() : nullptr;
      const int n_entries = std::max(1, std::max(n_combos, n_tuples));
      k_transpose_tables<<<(unsigned)((n_entries * la.row_words + 3) / 4), 128, 0, h->stream>>>(
          h->ctx(), h->d_fast.as<FastRow>(), a.row_map, rows, la.row_words, a.tab.bits, h->tab_words, n_combos, h->d_ruleT.as<uint32_t>(), a.ptab.bits,
          a.ptab.words, n_tuples, h->d_portT.as<uint32_t>(), h->d_lean.as<LeanRow>());
      ++*launches;
    }
    if (expand_beside && n_chunks > 1) CU(h->d_matrix.ensure(std::max<size_t>((size_t)h->M * stride, 1) * 4));
    for (int c = 0; c < n_chunks; ++c) {
      // x-blocks of this launch: the chunk's rows; the last launch also covers the padding up to the stride
      const int64_t xb0 = n_chunks > 1 ? h->chunks[c].r0 / OE_K1_THREADS : 0;
      const int64_t xb1 = (n_chunks > 1 && c + 1 < n_chunks) ? h->chunks[c].r1 / OE_K1_THREADS : xblocks_all;
      const int64_t xblocks = xb1 - xb0;
      // enough row tiles to fill the machine when there are few offers, few enough to amortise the offer-row loads
      int tiles = (int)std::min<int64_t>(rows, std::max<int64_t>(1, (148 * 16 + xblocks - 1) / xblocks));
      if (const char* ev = getenv("OE_K1_TILES")) tiles = std::max(1, std::min(rows, atoi(ev)));   // tuning probe (tools/k1_tiles_probe.py)
      int rows_per_tile = std::min(128, (rows + tiles - 1) / tiles);  // 128 descriptors = 16 KiB of shared memory
      tiles = (rows + rows_per_tile - 1) / rows_per_tile;
      if (tiles > 65535) return h->fail(OE_ERR_LIMIT, "oe: %d factored rows exceed what one feasibility launch covers", rows);
      a.rows_per_tile = rows_per_tile;
      a.xb0 = xb0;
      const size_t smem = (size_t)rows_per_tile * sizeof(FastRow) + (size_t)a.n_nodes * sizeof(oe_rule_node) +
                          (size_t)a.sc_cols * OE_K1_THREADS * sizeof(double);
      if (h->offers_in_flight) {
        // the host reads the chunk's checks before any kernel gathers through its ids; the launch goes out the moment it landed
        if (int rc = wait_offer_chunk(h, c, "oe_load_offers")) return rc;
        if (g_trace) g_trace->mark("chunk");
      }
      if (lean_ok) {
        la.xb0 = xb0;
        dim3 lgrid((unsigned)xblocks, (unsigned)la.row_words);
        if (h->peers_attached) k_class_feasibility_lean<true><<<lgrid, OE_K1_THREADS, 0, h->stream>>>(h->dof, la);
        else k_class_feasibility_lean<false><<<lgrid, OE_K1_THREADS, 0, h->stream>>>(h->dof, la);
      } else {
      dim3 grid((unsigned)xblocks, (unsigned)tiles);
      const bool fast_ranges = h->dof.R <= OE_FAST_RANGES;
      const bool few = h->dof.S * h->dof.K <= 4;   // the dense form then tests four pool entries, not eight
      const int ports = a.ptab.bits ? 2 : fast_ranges ? 1 : 0;
      const int variant = ports * 4 + (h->peers_attached ? 2 : 0) + (few ? 1 : 0);
      switch (variant) {
#define OE_K1_CASE(V, P, X, Q) case V: k_class_feasibility<P, X, Q><<<grid, OE_K1_THREADS, smem, h->stream>>>(h->ctx(), a); break;
        OE_K1_CASE(0, 0, false, 8) OE_K1_CASE(1, 0, false, 4) OE_K1_CASE(2, 0, true, 8) OE_K1_CASE(3, 0, true, 4)
        OE_K1_CASE(4, 1, false, 8) OE_K1_CASE(5, 1, false, 4) OE_K1_CASE(6, 1, true, 8) OE_K1_CASE(7, 1, true, 4)
        OE_K1_CASE(8, 2, false, 8) OE_K1_CASE(9, 2, false, 4) OE_K1_CASE(10, 2, true, 8) OE_K1_CASE(11, 2, true, 4)
#undef OE_K1_CASE
      }
      }
      ++*launches;
      CU(cudaGetLastError());
      if (expand_beside && (c + 1 < n_chunks || expand_last)) {
        CU(cudaEventRecord(h->ev_k1[c], h->stream));
        CU(cudaStreamWaitEvent(h->aux_stream, h->ev_k1[c], 0));
        if (!h->expand_started) { CU(cudaEventRecord(h->ev_x0, h->aux_stream)); h->expand_started = true; }
        const int64_t w0 = h->chunks[c].r0 / 32, w1 = c + 1 < n_chunks ? h->chunks[c].r1 / 32 : stride;
        if (int rc = run_expand(h, launches, h->aux_stream, w0, w1)) return rc;
        h->expand_from_word = w1;
      }
      if (after_chunk && c + 1 < n_chunks) if (int rc = (*after_chunk)(c)) return rc;
    }
    h->offers_in_flight = false;
    if (h->peers_attached) {
      PeerFlags f{};
      f.n = h->world; f.rank = h->rank; f.epoch = h->peer_epoch;
      for (int p = 0; p < h->world; ++p) f.flags[p] = h->peer_base[p];
      k_peer_signal<<<1, 32, 0, h->stream>>>(f);
      ++*launches;
      CU(cudaGetLastError());
    }
  } else if (h->peers_attached) {
    return h->fail(OE_ERR_STATE, "oe: a peer-attached engine needs at least one requirement class");
  }
  h->class_bits_fresh = true;
  return 0;
}

int run_expand(oe_engine* h, int64_t* launches, cudaStream_t on, int64_t w0, int64_t w1) {
  if (!on) on = h->stream;
  const int64_t stride = h->stride();
  if (w1 < 0) w1 = stride;
  CU(h->d_matrix.ensure(std::max<size_t>((size_t)h->M * stride, 1) * 4));
  if (h->M > 0 && w1 > w0) {
    const unsigned xblocks = (unsigned)(((w1 - w0) / 4 + OE_KX_LANES - 1) / OE_KX_LANES);
    // ~64 CTAs per SM overall (profiles/r01e_bw_probe.txt: 6.75 TB/s vs 6.50 at 32); each CTA streams a
    // [reqs_per_block x 1 KiB] slab
    int64_t yblocks = std::max<int64_t>(1, std::min<int64_t>((h->M + 15) / 16, (148 * 64 + xblocks - 1) / xblocks));
    const int reqs_per_block = (int)((h->M + yblocks - 1) / yblocks);
    yblocks = (h->M + reqs_per_block - 1) / reqs_per_block;
    dim3 grid(xblocks, (unsigned)yblocks);
    // An expansion that runs beside other work (one chunk beside the bus and the commit over the landed chunks, or the whole
    // matrix on the auxiliary stream beside the commit) asks for 56 KiB of shared memory it
    // never touches: that caps it at 4 CTAs (1024 threads) per SM, which still writes at the HBM rate but leaves the SMs' other
    // half to the commit path's kernels — measured on C5: cycle 3.41 ms vs 3.65 ms without the cap (36 KiB 3.45, 75 KiB 3.44,
    // 110 KiB 3.68; OE_KX_SMEM_KB overrides).
    static const int pad_kb = getenv("OE_KX_SMEM_KB") ? atoi(getenv("OE_KX_SMEM_KB")) : 56;
    size_t pad = 0;
    if (pad_kb > 0 && (w0 > 0 || w1 < stride || on == h->aux_stream)) pad = (size_t)std::min(pad_kb, 200) * 1024;   // opt-in: oe_create
    k_expand_matrix<<<grid, dim3(OE_KX_LANES, OE_KX_ROWS), pad, on>>>(h->d_F.as<uint32_t>(), h->bit_row_of_req(),
                                                                         h->d_matrix.as<uint32_t>(), stride, h->M, reqs_per_block,
                                                                         w0 / (4 * OE_KX_LANES), w1);
    ++*launches;
    CU(cudaGetLastError());
  }
  return 0;
}

// The union-list commit (see k_commit_union) of requirements req_row[0..M) over the offers of `rb`, on stream `st`: K3d, the
// components, one union list per component, the sort by component, the replay.  Decisions (offer indices relative to rb) go to
// dec[0..M).  Also used on a PREFIX of the offers and then on the rest with the requirements the prefix left unplaced
// (oe_evaluate_cycle while the offer table is still arriving): first fit takes the lowest free index, so what a requirement gets
// inside a prefix does not depend on the offers behind it, and a requirement that found nothing took nothing.
bool union_commit_fits(const oe_engine* h, int rows, int64_t total_words) {
  const size_t union_smem = ((size_t)total_words * 4 + 15) / 16 * 16 + (size_t)OE_RING * 32 * sizeof(int2);
  return rows <= OE_UNION_MAX_ROWS && union_smem + 1024 <= (size_t)h->smem_optin - h->union_static_smem;
}
int commit_union_pass(oe_engine* h, int64_t* launches, const RowBits& rb, const int32_t* req_row, int64_t M, oe_decision* dec, cudaStream_t st,
                      bool zero_placed, bool fork_ok, bool materialize) {
  const int rows = rb.rows;
  const int tiles = (int)std::max<int64_t>(1, (rb.total_words + OE_TILE_WORDS - 1) / OE_TILE_WORDS);
  const size_t union_smem = ((size_t)rb.total_words * 4 + 15) / 16 * 16 + (size_t)OE_RING * 32 * sizeof(int2);
  {
    const unsigned long long* adj = nullptr;
    if (rows > 1) {
      CU(h->d_adj.ensure(OE_ADJ_MAX_ROWS * 8));
      CU(cudaMemsetAsync(h->d_adj.p, 0, OE_ADJ_MAX_ROWS * 8, st));
      k_row_overlap<<<(unsigned)std::max<int64_t>(1, (rb.total_words + 127) / 128), 128 * OE_ADJ_SPLIT, 0, st>>>(rb, h->d_adj.as<unsigned long long>());
      ++*launches;
      adj = h->d_adj.as<unsigned long long>();
    }
    CU(h->d_meta.ensure(sizeof(CommitMeta)));
    CommitMeta* meta = h->d_meta.as<CommitMeta>();
    k_label_components<<<1, 32, 0, st>>>(rows, adj, meta);
    const int64_t ustride = rb.total_words;
    CU(h->d_union.ensure(std::max<size_t>((size_t)32 * ustride, 1) * 4));
    k_union_bits<<<dim3((unsigned)std::max<int64_t>(1, (ustride + 255) / 256), 32), 256, 0, st>>>(rb, meta, h->d_union.as<uint32_t>(), ustride);
    RowBits ub{};
    ub.F = h->d_union.as<uint32_t>(); ub.stride = ustride; ub.rows = 32; ub.shard_words = ustride; ub.total_words = ustride;
    CU(h->d_tile_cnt.ensure((size_t)32 * tiles * 4));
    CU(h->d_row_base.ensure((size_t)33 * 8));
    CU(h->d_row_len.ensure((size_t)32 * 4));
    // components share no offer, so all union lists together hold at most one entry per offer: no size read-back needed
    CU(h->d_cand.ensure(std::max<size_t>((size_t)ustride * 32, 1) * 4));
    CU(h->d_umask.ensure(std::max<size_t>((size_t)ustride * 32, 1) * 4));
    dim3 ugrid((unsigned)tiles, 32);
    k_tile_popc<<<ugrid, 256, 0, st>>>(ub, tiles, h->d_tile_cnt.as<int32_t>());
    k_row_offsets<<<1, 1024, 0, st>>>(32, tiles, (int64_t)0x7fffffff, h->d_tile_cnt.as<int32_t>(), h->d_row_base.as<int64_t>(), h->d_row_len.as<int32_t>());
    k_compact_rows<<<ugrid, 256, 0, st>>>(ub, tiles, h->d_tile_cnt.as<int32_t>(), h->d_row_base.as<int64_t>(), h->d_row_len.as<int32_t>(),
                                                 h->d_cand.as<int32_t>());
    k_union_masks<<<dim3((unsigned)std::max<int64_t>(1, std::min<int64_t>(1024, (ustride * 32 + 255) / 256)), 32), 256, 0, st>>>(
        rb, meta, h->d_cand.as<int32_t>(), h->d_row_base.as<int64_t>(), h->d_row_len.as<int32_t>(), h->d_umask.as<uint32_t>());
    *launches += 6;
    CU(cudaGetLastError());
    UnionArgs ua{};
    ua.meta = meta; ua.ucand = h->d_cand.as<int32_t>(); ua.umask = h->d_umask.as<uint32_t>(); ua.ubase = h->d_row_base.as<int64_t>();
    ua.ulen = h->d_row_len.as<int32_t>(); ua.used_words = rb.total_words; ua.req_row = req_row; ua.M = M;
    ua.dec = dec; ua.n_placed = h->d_nplaced.as<int>();
    CU(h->d_order.ensure(std::max<size_t>((size_t)M, 1) * sizeof(int2)));
    ua.order = h->d_order.as<int2>();
    CU(h->d_sort_hist.ensure((size_t)OE_SORT_BLOCKS * 32 * sizeof(int)));
    CU(h->d_comp_begin.ensure(33 * sizeof(int)));
    k_sort_hist<<<OE_SORT_BLOCKS, 256, 0, st>>>(ua.req_row, ua.M, meta, h->d_sort_hist.as<int>());
    k_sort_scatter<<<OE_SORT_BLOCKS, 256, 0, st>>>(ua.req_row, ua.M, meta, h->d_sort_hist.as<int>(), ua.order, h->d_comp_begin.as<int>());
    *launches += 2;
    ua.comp_begin = h->d_comp_begin.as<int>();
    if (zero_placed) CU(cudaMemsetAsync(ua.n_placed, 0, sizeof(int), st));
    static const bool trace_on = getenv("OE_COMMIT_TRACE") != nullptr;   // debugging aid (tools/commit_probe.py): per-warp counters on stderr
    if (trace_on) {
      CU(h->d_dbg.ensure(32 * 8 * sizeof(long long)));
      CU(cudaMemsetAsync(h->d_dbg.p, 0, 32 * 8 * sizeof(long long), st));
      ua.trace = h->d_dbg.as<long long>();
    }
    bool materialized = false;
    if (h->fork_expand && fork_ok) {
      // K1x needs only the class bits, the commit does not need the matrix.  Both become runnable at this event: the commit
      // kernel on the highest-priority stream, so that its CTAs (one per component, each with the claimed bitmap in shared
      // memory) are placed while the SMs are still empty, K1x on the lowest-priority one, filling the rest.  (Started any
      // earlier, K1x occupies every SM and the commit waits for the expansion to drain.)
      h->fork_expand = false;
      CU(cudaEventRecord(h->ev_fork, st));
      CU(cudaStreamWaitEvent(h->hi_stream, h->ev_fork, 0));
      k_commit_union<<<32, 32, union_smem, h->hi_stream>>>(ua);
      // the winners' pipelines re-run right behind it, on the same stream: at the main stream's priority they would queue
      // behind the expansion's grid
      if (materialize) k_materialize<<<(unsigned)((M + 127) / 128), 128, 0, h->hi_stream>>>(h->ctx(), h->d_req_row.as<int32_t>(), M, dec, h->d_nplaced.as<int>() + 1, decision_push(h));
      materialized = true;
      CU(cudaEventRecord(h->ev_commit, h->hi_stream));
      CU(cudaStreamWaitEvent(h->aux_stream, h->ev_fork, 0));
      if (!h->expand_started) CU(cudaEventRecord(h->ev_x0, h->aux_stream));
      if (int rc = run_expand(h, launches, h->aux_stream, h->expand_from_word)) return rc;
      CU(cudaEventRecord(h->ev_x1, h->aux_stream));
      CU(cudaStreamWaitEvent(st, h->ev_commit, 0));
    } else {
      k_commit_union<<<32, 32, union_smem, st>>>(ua);
    }
    ++*launches;
    CU(cudaGetLastError());
    if (!materialized && materialize) k_materialize<<<(unsigned)((M + 127) / 128), 128, 0, st>>>(h->ctx(), h->d_req_row.as<int32_t>(), M, dec, h->d_nplaced.as<int>() + 1, decision_push(h));
    ++*launches;
    CU(cudaGetLastError());
    if (trace_on) {
      long long d[32 * 8];
      CU(cudaMemcpyAsync(d, ua.trace, sizeof d, cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      for (int w = 0; w < 32; ++w)
        if (d[w * 8])
          fprintf(stderr, "commit trace: warp %2d  %6lld requirements  %8lld cycles  %4lld slides  %4lld top-ups  %5lld deferred  %5lld searches beyond  (prologue %lld cycles)\n",
                  w, d[w * 8], d[w * 8 + 1], d[w * 8 + 2], d[w * 8 + 3], d[w * 8 + 4], d[w * 8 + 5], d[w * 8 + 6]);
    }
    return 0;
  }
}


// ---- the commit from the offers' side (k_commit_offers) -------------------------------------------------------------------------
// `OE_COMMIT=union` in the environment keeps the requirement-side union-list commit (k_commit_union) for comparison.
bool offer_side_ok(int rows) {
  static const bool forced_union = getenv("OE_COMMIT") && std::string(getenv("OE_COMMIT")) == "union";
  return !forced_union && rows >= 1 && rows <= OE_OS_MAX_ROWS;
}
// Once per cycle: the requirements of every bit row in plan order (a stable counting sort by row) and the rows' queue
// positions at zero.  Needs only the requirement table, so it is enqueued before the class bits exist.
int prepare_queues(oe_engine* h, int64_t* launches, const int32_t* req_row, int64_t M, int rows, cudaStream_t st) {
  const int nb = (int)std::max<int64_t>(1, (M + OE_OS_SORT_BLOCK - 1) / OE_OS_SORT_BLOCK);
  CU(h->d_os_rhist.ensure((size_t)nb * rows * sizeof(int)));
  CU(h->d_os_row_begin.ensure(((size_t)rows + 1) * sizeof(int)));
  CU(h->d_os_queue.ensure(((size_t)M + 8) * sizeof(int32_t)));
  CU(h->d_os_cursor.ensure((size_t)rows * sizeof(int)));
  CU(cudaMemsetAsync(h->d_os_cursor.p, 0, (size_t)rows * sizeof(int), st));
  k_key_hist<<<nb, 256, (size_t)rows * sizeof(int), st>>>(req_row, M, rows, h->d_os_rhist.as<int>());
  CU(h->d_os_totals.ensure((size_t)rows * sizeof(int)));
  k_key_scan_cols<<<rows, 1024, 0, st>>>(h->d_os_rhist.as<int>(), nb, nullptr, rows, h->d_os_totals.as<int>());
  k_key_scan<<<1, 1024, 0, st>>>(h->d_os_totals.as<int>(), nullptr, rows, h->d_os_row_begin.as<int>());
  k_key_scatter<<<nb, 32, (size_t)rows * sizeof(int), st>>>(req_row, M, rows, h->d_os_rhist.as<int>(), h->d_os_row_begin.as<int>(), nullptr,
                                                          h->d_os_queue.as<int32_t>(), nullptr);
  *launches += 3;
  CU(cudaGetLastError());
  return 0;
}
// The same, enqueued early (beside the feasibility pass) on the side stream; `st` waits for it.
int queues_ready(oe_engine* h, int64_t* launches, int rows, cudaStream_t st) {
  if (h->queues_rows == rows) { h->queues_rows = -1; CU(cudaStreamWaitEvent(st, h->ev_queues, 0)); return 0; }
  return prepare_queues(h, launches, h->bit_row_of_req(), h->M, rows, st);
}
// One pass over the offers of `rb` (global index of its first offer: offer_offset).  Returns 1 (nothing enqueued beyond the
// grouping) when a component has more than 32 rows: the caller falls back to k_commit.
int commit_offers_pass(oe_engine* h, int64_t* launches, const RowBits& rb, int32_t offer_offset, int64_t M, oe_decision* dec, cudaStream_t st,
                       bool fork_ok, bool materialize, const int32_t* req_row) {
  const int rows = rb.rows;
  const int64_t n_offers = rb.total_words * 32;
  const unsigned long long* adj = nullptr;
  int* parent = nullptr;
  if (rows > OE_ADJ_MAX_ROWS) {
    CU(h->d_os_parent.ensure((size_t)rows * sizeof(int)));
    parent = h->d_os_parent.as<int>();
    std::vector<int> iota((size_t)rows);
    for (int r = 0; r < rows; ++r) iota[(size_t)r] = r;
    CU(cudaMemcpyAsync(parent, iota.data(), (size_t)rows * sizeof(int), cudaMemcpyHostToDevice, st));   // pageable: staged before the call returns
    k_row_link<<<(unsigned)std::max<int64_t>(1, (rb.total_words + 3) / 4), 128, (size_t)rows * 16, st>>>(rb, parent);
    ++*launches;
  } else if (rows > 1) {
    CU(h->d_adj.ensure(OE_ADJ_MAX_ROWS * 8));
    CU(cudaMemsetAsync(h->d_adj.p, 0, OE_ADJ_MAX_ROWS * 8, st));
    k_row_overlap<<<(unsigned)std::max<int64_t>(1, (rb.total_words + 127) / 128), 128 * OE_ADJ_SPLIT, 0, st>>>(rb, h->d_adj.as<unsigned long long>());
    ++*launches;
    adj = h->d_adj.as<unsigned long long>();
  }
  CU(h->d_os_groups.ensure(((size_t)rows * 4 + 4) * sizeof(int)));
  RowGroups g{};
  g.comp_of = h->d_os_groups.as<int>(); g.lrow = g.comp_of + rows; g.comp_rows = g.lrow + rows; g.comp_row_begin = g.comp_rows + rows;
  CU(h->d_os_comp_begin.ensure(((size_t)rows + 4) * sizeof(int)));
  g.info = h->d_os_comp_begin.as<int>() + rows + 1;
  k_group_rows<<<1, 256, ((size_t)rows * 3 + 1) * sizeof(int), st>>>(rows, adj, parent, g, rb.skip_ptr, rb.skip_at);
  ++*launches;
  CU(cudaGetLastError());
  int n_ctas = rows;
  if (rows > 32) {   // a component of more than 32 rows does not fit the lanes of its warp: the caller must know before it goes on
    CU(cudaMemcpyAsync(h->h_os_info, g.info, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (h->h_os_info[1] > 32) return 1;
    n_ctas = std::max(1, h->h_os_info[0]);
  }
  const int nb = (int)std::max<int64_t>(1, (n_offers + OE_OS_SORT_BLOCK - 1) / OE_OS_SORT_BLOCK);
  CU(h->d_os_ocomp.ensure((size_t)nb * OE_OS_SORT_BLOCK * 4)); CU(h->d_os_omask.ensure((size_t)nb * OE_OS_SORT_BLOCK * 4));
  CU(h->d_os_ohist.ensure((size_t)nb * rows * sizeof(int)));
  CU(h->d_os_list.ensure((size_t)nb * OE_OS_SORT_BLOCK * 4)); CU(h->d_os_mask.ensure((size_t)nb * OE_OS_SORT_BLOCK * 4));
  k_offer_masks<<<nb, 256, (size_t)rows * 44, st>>>(rb, g.comp_of, g.lrow, g.info, h->d_os_ocomp.as<int32_t>(), h->d_os_omask.as<uint32_t>(), h->d_os_ohist.as<int>());
  CU(h->d_os_totals.ensure((size_t)rows * sizeof(int)));
  k_key_scan_cols<<<n_ctas, 1024, 0, st>>>(h->d_os_ohist.as<int>(), nb, g.info, 0, h->d_os_totals.as<int>());
  k_key_scan<<<1, 1024, 0, st>>>(h->d_os_totals.as<int>(), g.info, 0, h->d_os_comp_begin.as<int>());
  k_key_scatter<<<nb, 32, (size_t)rows * sizeof(int), st>>>(h->d_os_ocomp.as<int32_t>(), n_offers, rows, h->d_os_ohist.as<int>(), h->d_os_comp_begin.as<int>(),
                                                          h->d_os_omask.as<uint32_t>(), h->d_os_list.as<int32_t>(), h->d_os_mask.as<uint32_t>(), rb.skip_ptr ? g.info : nullptr);
  *launches += 3;
  CU(cudaGetLastError());
  OfferSideArgs a{};
  a.info = g.info; a.comp_begin = h->d_os_comp_begin.as<int>(); a.list = h->d_os_list.as<int32_t>(); a.mask = h->d_os_mask.as<uint32_t>();
  a.comp_rows = g.comp_rows; a.comp_row_begin = g.comp_row_begin; a.row_begin = h->d_os_row_begin.as<int>(); a.queue = h->d_os_queue.as<int32_t>();
  a.row_cursor = h->d_os_cursor.as<int>(); a.offer_offset = offer_offset; a.dec = dec; a.n_placed = h->d_nplaced.as<int>();
  static const bool trace_on = getenv("OE_COMMIT_TRACE") != nullptr;
  if (trace_on) {
    CU(h->d_dbg.ensure((size_t)n_ctas * 4 * sizeof(long long)));
    CU(cudaMemsetAsync(h->d_dbg.p, 0, (size_t)n_ctas * 4 * sizeof(long long), st));
    a.trace = h->d_dbg.as<long long>();
  }
  bool materialized = false;
  if (h->fork_expand && fork_ok) {
    // as in commit_union_pass: the replay on the highest-priority stream, the matrix expansion beside it on the lowest
    h->fork_expand = false;
    CU(cudaEventRecord(h->ev_fork, st));
    CU(cudaStreamWaitEvent(h->hi_stream, h->ev_fork, 0));
    k_commit_offers<<<n_ctas, 32, 0, h->hi_stream>>>(a);
    if (materialize) k_materialize<<<(unsigned)((M + 127) / 128), 128, 0, h->hi_stream>>>(h->ctx(), req_row, M, dec, h->d_nplaced.as<int>() + 1, decision_push(h));
    materialized = true;
    CU(cudaEventRecord(h->ev_commit, h->hi_stream));
    CU(cudaStreamWaitEvent(h->aux_stream, h->ev_fork, 0));
    if (!h->expand_started) CU(cudaEventRecord(h->ev_x0, h->aux_stream));
    if (int rc = run_expand(h, launches, h->aux_stream, h->expand_from_word)) return rc;
    CU(cudaEventRecord(h->ev_x1, h->aux_stream));
    CU(cudaStreamWaitEvent(st, h->ev_commit, 0));
  } else {
    k_commit_offers<<<n_ctas, 32, 0, st>>>(a);
  }
  ++*launches;
  CU(cudaGetLastError());
  if (!materialized && materialize) {
    k_materialize<<<(unsigned)((M + 127) / 128), 128, 0, st>>>(h->ctx(), req_row, M, dec, h->d_nplaced.as<int>() + 1, decision_push(h));
    CU(cudaGetLastError());
  }
  if (materialize) ++*launches;
  if (trace_on) {
    std::vector<long long> d((size_t)n_ctas * 4);
    CU(cudaMemcpyAsync(d.data(), a.trace, d.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (int c = 0; c < n_ctas; ++c)
      if (d[(size_t)c * 4 + 2])
        fprintf(stderr, "commit trace: component %3d  %7lld steps  %9lld cycles  %6lld chunks of 32 offers, %lld skipped\n", c, d[(size_t)c * 4], d[(size_t)c * 4 + 1],
                d[(size_t)c * 4 + 2], d[(size_t)c * 4 + 3]);
  }
  return 0;
}

int run_commit(oe_engine* h, int64_t* launches) {
  const int64_t stride = h->stride();
  const int rows = h->rows_factored();
  RowBits rb{};
  rb.stride = stride; rb.rows = rows;
  if (h->world > 1) {
    rb.F = h->peers_attached ? h->exch_rows(h->rank, h->peer_epoch) : h->d_Fglobal.as<uint32_t>();
    rb.shard_words = h->shard_rows / 32;
    rb.total_words = std::min<int64_t>((h->N_global + 31) / 32, (int64_t)h->world * rb.shard_words);
  } else {
    rb.F = h->d_F.as<uint32_t>();
    rb.shard_words = stride;
    rb.total_words = (h->N + 31) / 32;
  }
  CU(cudaMemsetAsync(h->d_dec.p, 0, std::max<size_t>((size_t)h->M, 1) * sizeof(oe_decision), h->stream));
  if (rows == 0 || h->M == 0) {
    CU(cudaMemsetAsync(h->d_nplaced.p, 0, sizeof(int), h->stream));
    return 0;
  }
  if (offer_side_ok(rows)) {
    k_clear_decisions<<<(unsigned)((h->M + 255) / 256), 256, 0, h->stream>>>(h->d_dec.as<oe_decision>(), h->M);
    CU(cudaMemsetAsync(h->d_nplaced.p, 0, sizeof(int), h->stream));
    if (int rc = queues_ready(h, launches, rows, h->stream)) return rc;
    const int rc = commit_offers_pass(h, launches, rb, 0, h->M, h->d_dec.as<oe_decision>(), h->stream, true, true, h->d_req_row.as<int32_t>());
    if (rc <= 0) return rc;
    // a component wider than a warp: the per-row lists below
    CU(cudaMemsetAsync(h->d_dec.p, 0, std::max<size_t>((size_t)h->M, 1) * sizeof(oe_decision), h->stream));
  }
  const int tiles = (int)std::max<int64_t>(1, (rb.total_words + OE_TILE_WORDS - 1) / OE_TILE_WORDS);
  if (union_commit_fits(h, rows, rb.total_words))
    return commit_union_pass(h, launches, rb, h->bit_row_of_req(), h->M, h->d_dec.as<oe_decision>(), h->stream, true, true, true);
  // K3: per-row candidate lists, capped at what one cycle can possibly consume from a row
  const int64_t cap = h->M + 16;
  CU(h->d_tile_cnt.ensure((size_t)rows * tiles * 4));
  CU(h->d_row_base.ensure(((size_t)rows + 1) * 8));
  CU(h->d_row_len.ensure((size_t)rows * 4));
  dim3 tgrid((unsigned)tiles, (unsigned)rows);
  k_tile_popc<<<tgrid, 256, 0, h->stream>>>(rb, tiles, h->d_tile_cnt.as<int32_t>());
  k_row_offsets<<<1, 1024, 0, h->stream>>>(rows, tiles, cap, h->d_tile_cnt.as<int32_t>(), h->d_row_base.as<int64_t>(), h->d_row_len.as<int32_t>());
  *launches += 2;
  CU(cudaGetLastError());
  int64_t n_cand = 0;
  CU(cudaMemcpyAsync(&n_cand, h->d_row_base.as<int64_t>() + rows, 8, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  CU(h->d_cand.ensure(std::max<size_t>((size_t)n_cand, 1) * 4));
  k_compact_rows<<<tgrid, 256, 0, h->stream>>>(rb, tiles, h->d_tile_cnt.as<int32_t>(), h->d_row_base.as<int64_t>(), h->d_row_len.as<int32_t>(),
                                               h->d_cand.as<int32_t>());
  ++*launches;
  CU(cudaGetLastError());

  CommitArgs a{};
  a.cand = h->d_cand.as<int32_t>(); a.row_base = h->d_row_base.as<int64_t>(); a.row_len = h->d_row_len.as<int32_t>();
  a.rows = rows; a.used_words = rb.total_words; a.req_row = h->bit_row_of_req(); a.M = h->M; a.dec = h->d_dec.as<oe_decision>();
  a.n_placed = h->d_nplaced.as<int>();
  CU(h->d_order.ensure(std::max<size_t>((size_t)h->M, 1) * sizeof(int2)));
  a.order = h->d_order.as<int2>();
  if (rows <= OE_ADJ_MAX_ROWS && rows > 1) {
    CU(h->d_adj.ensure(OE_ADJ_MAX_ROWS * 8));
    CU(cudaMemsetAsync(h->d_adj.p, 0, OE_ADJ_MAX_ROWS * 8, h->stream));
    k_row_overlap<<<(unsigned)std::max<int64_t>(1, (rb.total_words + 127) / 128), 128 * OE_ADJ_SPLIT, 0, h->stream>>>(rb, h->d_adj.as<unsigned long long>());
    ++*launches;
    a.adj = h->d_adj.as<unsigned long long>();
  }
  const size_t budget = (size_t)h->smem_optin - h->commit_static_smem - 1024;  // static arrays (sort counters) share the pool
  // everything in shared memory (row states, claimed bitmap): the fast variant
  const size_t all_in = (size_t)rows * sizeof(RowState) + (size_t)a.used_words * 4;
  if (all_in <= budget) {
    a.rows_in_smem = a.used_in_smem = 1;
    k_commit<true><<<1, OE_COMMIT_THREADS, all_in, h->stream>>>(a);
  } else {
    size_t smem = 0;
    a.rows_in_smem = (size_t)rows * sizeof(RowState) <= budget / 4;
    if (a.rows_in_smem) smem += (size_t)rows * sizeof(RowState);
    else { CU(h->d_cursor.ensure((size_t)rows * sizeof(RowState))); a.rows_g = h->d_cursor.as<RowState>(); }
    a.used_in_smem = smem + (size_t)a.used_words * 4 <= budget;
    if (a.used_in_smem) smem += (size_t)a.used_words * 4;
    else { CU(h->d_used.ensure(std::max<size_t>(a.used_words, 1) * 4)); a.used_g = h->d_used.as<uint32_t>(); }
    k_commit<false><<<1, OE_COMMIT_THREADS, smem, h->stream>>>(a);
  }
  ++*launches;
  CU(cudaGetLastError());
  k_materialize<<<(unsigned)((h->M + 127) / 128), 128, 0, h->stream>>>(h->ctx(), h->d_req_row.as<int32_t>(), h->M, h->d_dec.as<oe_decision>(), h->d_nplaced.as<int>() + 1, decision_push(h));
  ++*launches;
  CU(cudaGetLastError());
  return 0;
}

float ms_between(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0.f;
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

// offers_deferred: the caller is a feasibility pass, which waits for an arriving offer table chunk by chunk itself
int ready(oe_engine* h, const char* who, bool offers_deferred = false) {
  if (!h) return OE_ERR_INVALID;
  if (!(h->hasP && h->hasT && h->hasO && h->hasR)) return h->fail(OE_ERR_STATE, "%s: tables not loaded (programs, tasks, offers, requirements)", who);
  for (auto& cs : h->h_csets)
    if (cs.keyspace == OE_KS_RR_ATTR && (cs.rr_col >= h->dof.Q || cs.rr_col >= h->dt.Q)) return h->fail(OE_ERR_INVALID, "%s: rr column mismatch", who);
  if (cudaSetDevice(h->device) != cudaSuccess) return h->fail(OE_ERR_CUDA, "cudaSetDevice");
  if (h->validation_pending) {  // OE_CFG_ASYNC_LOADS: the id checks of the loads are read back here, before any kernel gathers through them
    h->validation_pending = false;
    if (int rc = fetch_err_flag(h, who)) { h->hasT = h->hasO = h->hasR = h->twins_unresolved = h->reqs_pending = false; h->pending_T = -1; return rc; }
  }
  if (h->tuple_count_pending) {   // ... and the size of the range-list dictionary after the rows oe_update_offers replaced
    h->tuple_count_pending = false;
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) return h->fail(OE_ERR_CUDA, "cudaStreamSynchronize");
    if (h->h_os_info[2] > h->ptuple_cap) h->tuples_valid = false;   // dictionary full: per-offer port tests until the table is loaded whole again
    else h->n_port_tuples = h->dof.n_ptuples = h->h_os_info[2];
  }
  if (h->twins_unresolved) {
    // ... and with them the number of (requirement, deployed task of its pod) records k_match_tasks wrote.  None: every
    // requirement reads its class's bit row, which is what was loaded.
    h->twins_unresolved = false;
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) return h->fail(OE_ERR_CUDA, "cudaStreamSynchronize");
    if (h->h_os_info[3] > 0) if (int rc = resolve_twins(h)) { h->hasR = false; return rc; }
  }
  if (!offers_deferred) if (int rc = wait_offers(h, who)) return rc;
  return 0;
}

}  // namespace

int32_t oe_class_feasibility(oe_engine* h, uint32_t* bits, oe_stats* stats) {
  CallerTablesExpire expire_guard{h};
  if (int rc = ready(h, "oe_class_feasibility", true)) return rc;
  int64_t launches = 0;
  CU(cudaEventRecord(h->ev[0], h->stream));
  if (int rc = run_counters(h, &launches)) return rc;
  CU(cudaEventRecord(h->ev[1], h->stream));
  if (int rc = run_class_feasibility(h, &launches)) return rc;
  CU(cudaEventRecord(h->ev[2], h->stream));
  if (bits) CU(cudaMemcpyAsync(bits, h->d_F.p, (size_t)h->rows_factored() * h->stride() * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaEventRecord(h->ev[3], h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (stats) {
    std::memset(stats, 0, sizeof *stats);
    stats->ms_counters = ms_between(h->ev[0], h->ev[1]);
    stats->ms_feasibility = ms_between(h->ev[1], h->ev[2]);
    stats->ms_d2h = ms_between(h->ev[2], h->ev[3]);
    stats->ms_total = ms_between(h->ev[0], h->ev[3]);
    stats->n_pairs = h->N * h->M;
    stats->kernel_launches = launches;
  }
  return OE_OK;
}

int32_t oe_requirement_rows(oe_engine* h, int32_t* rows, int32_t* n_rows) {
  if (!h) return OE_ERR_INVALID;
  if (!h->hasR || !h->class_bits_fresh) return h->fail(OE_ERR_STATE, "oe_requirement_rows: no feasibility pass since the last load");
  CU(cudaSetDevice(h->device));
  if (int rc = absorb_reqs(h)) return rc;
  if (h->V_eff >= 0 && !h->req_row_eff_on_host && h->M > 0) {
    h->h_req_row_eff.resize((size_t)h->M);
    CU(cudaMemcpyAsync(h->h_req_row_eff.data(), h->d_req_row_eff.p, (size_t)h->M * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->req_row_eff_on_host = true;
  }
  if (n_rows) *n_rows = h->rows_factored();
  if (rows && h->M > 0) std::memcpy(rows, (h->V_eff >= 0 ? h->h_req_row_eff : h->h_req_row).data(), (size_t)h->M * sizeof(int32_t));
  return OE_OK;
}

int32_t oe_feasibility(oe_engine* h, uint32_t* bits, oe_stats* stats) {
  CallerTablesExpire expire_guard{h};
  if (int rc = ready(h, "oe_feasibility", true)) return rc;
  int64_t launches = 0;
  CU(cudaEventRecord(h->ev[0], h->stream));
  if (int rc = run_counters(h, &launches)) return rc;
  CU(cudaEventRecord(h->ev[1], h->stream));
  if (int rc = run_class_feasibility(h, &launches)) return rc;
  CU(cudaEventRecord(h->ev[2], h->stream));
  if (int rc = run_expand(h, &launches)) return rc;
  CU(cudaEventRecord(h->ev[3], h->stream));
  if (bits) CU(cudaMemcpyAsync(bits, h->d_matrix.p, (size_t)h->M * h->stride() * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaEventRecord(h->ev[4], h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (stats) {
    std::memset(stats, 0, sizeof *stats);
    stats->ms_counters = ms_between(h->ev[0], h->ev[1]);
    stats->ms_feasibility = ms_between(h->ev[1], h->ev[2]);
    stats->ms_expand = ms_between(h->ev[2], h->ev[3]);
    stats->ms_d2h = ms_between(h->ev[3], h->ev[4]);
    stats->ms_total = ms_between(h->ev[0], h->ev[4]);
    stats->n_pairs = h->N * h->M;
    stats->kernel_launches = launches;
  }
  return OE_OK;
}

int32_t oe_evaluate_cycle(oe_engine* h, oe_decision* out, oe_stats* stats) {
  CycleTrace tr;
  struct TraceScope { TraceScope(CycleTrace* t) { g_trace = t; } ~TraceScope() { g_trace = nullptr; } } trace_scope(&tr);
  CallerTablesExpire expire_guard{h};
  if (int rc = ready(h, "oe_evaluate_cycle", true)) return rc;
  tr.mark("ready");
  struct WantQueues { oe_engine* e; WantQueues(oe_engine* x) : e(x) { e->want_queues = true; } ~WantQueues() { e->want_queues = false; e->queues_rows = -1; } } want_queues(h);
  if (!out && h->M > 0 && !(h->world > 1 && h->peers_attached)) return h->fail(OE_ERR_INVALID, "oe_evaluate_cycle: out is NULL");
  int64_t launches = 0;
  bool expand_overlapped = false, two_pass = false;
  CU(cudaEventRecord(h->ev[0], h->stream));
  if (h->world > 1 && h->peers_attached) {
    // sharded, peers attached: this shard's pass publishes its rows into every shard's exchange buffer (K1 itself stores
    // them); the commit path then waits on the device until every shard's rows of this epoch have landed here.
    if (!h->class_bits_fresh) {
      if (int rc = run_counters(h, &launches)) return rc;
      CU(cudaEventRecord(h->ev[1], h->stream));
      if (int rc = run_class_feasibility(h, &launches, (h->flags & OE_CFG_MATERIALIZE_MATRIX) != 0)) return rc;
      if (h->flags & OE_CFG_MATERIALIZE_MATRIX) {
        expand_overlapped = true;
        CU(h->d_matrix.ensure(std::max<size_t>((size_t)h->M * h->stride(), 1) * 4));
        h->fork_expand = true;
      }
    } else {
      CU(cudaEventRecord(h->ev[1], h->stream));
    }
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaEventRecord(h->ev[3], h->stream));
    k_peer_wait<<<1, 32, 0, h->stream>>>(h->peer_base[h->rank], h->world, h->peer_epoch, h->peer_timeout_ns, h->d_peer_err.as<int>());
    ++launches;
    CU(cudaGetLastError());
  } else if (h->world > 1) {
    // sharded: the caller ran oe_class_feasibility on every shard and gathered the rows into OE_BUF_CLASS_BITS_GLOBAL
    if (!h->class_bits_fresh) return h->fail(OE_ERR_STATE, "oe_evaluate_cycle (sharded): run oe_class_feasibility and gather first");
    if (h->d_Fglobal.cap < (size_t)h->world * h->rows_factored() * h->stride() * 4) return h->fail(OE_ERR_STATE, "gathered class bits buffer was not requested");
    CU(cudaEventRecord(h->ev[1], h->stream));
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaEventRecord(h->ev[3], h->stream));
  } else {
    if (int rc = run_counters(h, &launches)) return rc;
    CU(cudaEventRecord(h->ev[1], h->stream));
    const bool want_matrix = (h->flags & OE_CFG_MATERIALIZE_MATRIX) != 0;
    const int rows = h->rows_factored();
    static const bool no_prefix_commit = getenv("OE_NO_PREFIX_COMMIT") != nullptr;
    const bool os_two_pass = offer_side_ok(rows) && rows <= 32;   // no host round trip inside a pass (rows <= 32: no component can be too wide)
    if (h->offers_in_flight && h->n_chunks > 1 && h->M > 0 && rows > 0 && !no_prefix_commit && (os_two_pass || union_commit_fits(h, rows, (h->N + 31) / 32))) {
      // The offer table is still arriving.  First fit takes the lowest free index, so the commit over the chunks that have
      // landed already is final for every requirement it places; it runs (highest priority) while the last chunk is on the bus,
      // and only the requirements it left open are committed over the last chunk afterwards.
      two_pass = true;
      const int64_t stride = h->stride();
      // Which prefixes.  From the offers' side (os_two_pass) a pass simply resumes from the rows' queue positions, and a pass that
      // finds every requirement placed costs a few idle launches (the device decides: RowBits::skip_ptr) — so there are up to
      // three: an EARLY one behind the first chunks that hold four offers per requirement (usually everything is placed there, long
      // before the last chunk lands), one behind all chunks but the last, one over the last.  No host round trip in between.
      // Measured (C5 shapes, decisions cycle, pass after chunk 1 of 4 instead of chunk 2 only): c5 1.54 -> 1.35 ms, MaxPerHostname
      // 4.88 -> 4.54, one region 2.03 -> 1.95; with the matrix 3.02 -> 2.82.  OE_PREFIX_CHUNK = first pass after that chunk.
      // The union commit (more than 32 bit rows cannot happen here, but OE_COMMIT=union) keeps its two passes.
      int early_chunk = h->n_chunks - 2;
      if (os_two_pass) for (int c = 0; c < h->n_chunks - 2; ++c) if (h->chunks[c].r1 >= 4 * h->M) { early_chunk = c; break; }
      static const char* prefix_env = getenv("OE_PREFIX_CHUNK");
      if (prefix_env) early_chunk = std::max(0, std::min(h->n_chunks - 2, atoi(prefix_env)));
      const int64_t wA = h->chunks[h->n_chunks - 2].r1 / 32, wT = (h->N + 31) / 32;
      CU(h->d_dec.ensure(std::max<size_t>((size_t)h->M, 1) * sizeof(oe_decision)));
      CU(h->d_dec_sub.ensure(std::max<size_t>((size_t)h->M, 1) * sizeof(oe_decision)));
      CU(h->d_sub_idx.ensure((size_t)h->M * 4)); CU(h->d_sub_row.ensure((size_t)h->M * 4));
      CU(h->d_tile_open.ensure(((size_t)h->M + 1023) / 1024 * 4));
      if (want_matrix) CU(h->d_matrix.ensure(std::max<size_t>((size_t)h->M * stride, 1) * 4));
      int* h_open = h->h_offer_err + oe_engine::kMaxChunks;
      int64_t w_done = 0;   // words of the class rows the passes enqueued so far have covered
      const auto offers_pass = [&](int64_t w1) -> int {
        RowBits rb{};
        rb.F = h->d_F.as<uint32_t>() + w_done; rb.stride = stride; rb.rows = rows; rb.shard_words = stride; rb.total_words = w1 - w_done;
        if (w_done == 0) {
          k_clear_decisions<<<(unsigned)((h->M + 255) / 256), 256, 0, h->hi_stream>>>(h->d_dec.as<oe_decision>(), h->M);
          CU(cudaMemsetAsync(h->d_nplaced.p, 0, sizeof(int), h->hi_stream));
          if (int rc = queues_ready(h, &launches, rows, h->hi_stream)) return rc;
        } else {
          rb.skip_ptr = h->d_nplaced.as<int>(); rb.skip_at = (int)h->M;
        }
        if (int rc = commit_offers_pass(h, &launches, rb, (int32_t)(w_done * 32), h->M, h->d_dec.as<oe_decision>(), h->hi_stream, false, false, h->d_req_row.as<int32_t>()))
          return rc < 0 ? rc : h->fail(OE_ERR_STATE, "commit: unexpected fallback");
        w_done = w1;
        return 0;
      };
      const std::function<int(int)> prefix_commit = [&](int c) -> int {
        if (c != h->n_chunks - 2 && !(os_two_pass && c == early_chunk)) return 0;
        CU(cudaEventRecord(h->ev_prefix, h->stream));
        CU(cudaStreamWaitEvent(h->hi_stream, h->ev_prefix, 0));
        if (os_two_pass) {
          if (int rc = offers_pass(h->chunks[c].r1 / 32)) return rc;
          tr.mark("prefix_commit_enqueued");
          return 0;
        }
        RowBits rbA{};
        rbA.F = h->d_F.as<uint32_t>(); rbA.stride = stride; rbA.rows = rows; rbA.shard_words = stride; rbA.total_words = wA;
        CU(cudaMemsetAsync(h->d_dec.p, 0, std::max<size_t>((size_t)h->M, 1) * sizeof(oe_decision), h->hi_stream));
        if (int rc = commit_union_pass(h, &launches, rbA, h->bit_row_of_req(), h->M, h->d_dec.as<oe_decision>(), h->hi_stream, true, false, false)) return rc;
        const unsigned n_tiles = (unsigned)((h->M + 1023) / 1024);
        k_count_unplaced<<<n_tiles, 1024, 0, h->hi_stream>>>(h->d_dec.as<oe_decision>(), h->M, h->d_tile_open.as<int>());
        k_collect_unplaced<<<n_tiles, 1024, 0, h->hi_stream>>>(h->d_dec.as<oe_decision>(), h->bit_row_of_req(), h->M, h->d_tile_open.as<int>(),
                                                               h->d_sub_idx.as<int32_t>(), h->d_sub_row.as<int32_t>(), h->d_sub_count.as<int>());
        launches += 2;
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(h_open, h->d_sub_count.p, sizeof(int), cudaMemcpyDeviceToHost, h->hi_stream));
        CU(cudaEventRecord(h->ev_phase_a, h->hi_stream));
        tr.mark("prefix_commit_enqueued");
        return 0;
      };
      if (int rc = run_class_feasibility(h, &launches, want_matrix, true, &prefix_commit)) return rc;
      CU(cudaEventRecord(h->ev[2], h->stream));
      CU(cudaEventRecord(h->ev[3], h->stream));
      CU(cudaEventRecord(h->ev_k1_last, h->stream));
      if (want_matrix) { expand_overlapped = true; CU(cudaEventRecord(h->ev_x1, h->aux_stream)); }
      tr.mark("last_chunk_enqueued");
      if (os_two_pass) {
        CU(cudaStreamWaitEvent(h->hi_stream, h->ev_k1_last, 0));
        if (int rc = offers_pass(wT)) return rc;
      } else {
        CU(cudaEventSynchronize(h->ev_phase_a));
        const int n_open = *h_open;
        tr.mark("prefix_commit_done");
        CU(cudaStreamWaitEvent(h->hi_stream, h->ev_k1_last, 0));
        if (n_open > 0) {
          RowBits rbB{};
          rbB.F = h->d_F.as<uint32_t>() + wA; rbB.stride = stride; rbB.rows = rows; rbB.shard_words = stride; rbB.total_words = wT - wA;
          if (int rc = commit_union_pass(h, &launches, rbB, h->d_sub_row.as<int32_t>(), n_open, h->d_dec_sub.as<oe_decision>(), h->hi_stream, false, false, false)) return rc;
          k_merge_unplaced<<<(unsigned)((n_open + 255) / 256), 256, 0, h->hi_stream>>>(h->d_dec_sub.as<oe_decision>(), h->d_sub_idx.as<int32_t>(), n_open, (int32_t)(wA * 32),
                                                                                    h->d_dec.as<oe_decision>());
          ++launches;
        }
      }
      k_materialize<<<(unsigned)((h->M + 127) / 128), 128, 0, h->hi_stream>>>(h->ctx(), h->d_req_row.as<int32_t>(), h->M, h->d_dec.as<oe_decision>(), h->d_nplaced.as<int>() + 1, decision_push(h));
      ++launches;
      CU(cudaGetLastError());
      CU(cudaEventRecord(h->ev_commit, h->hi_stream));
      CU(cudaStreamWaitEvent(h->stream, h->ev_commit, 0));
    } else {
      if (int rc = run_class_feasibility(h, &launches, want_matrix)) return rc;
      CU(cudaEventRecord(h->ev[2], h->stream));
      CU(cudaEventRecord(h->ev[3], h->stream));
      if (want_matrix) {
        expand_overlapped = true;
        CU(h->d_matrix.ensure(std::max<size_t>((size_t)h->M * h->stride(), 1) * 4));  // before the fork: an allocation would serialise the device
        h->fork_expand = true;
      }
    }
  }
  tr.mark("pass_enqueued");
  if (!two_pass) if (int rc = run_commit(h, &launches)) { h->fork_expand = false; return rc; }
  tr.mark("commit_enqueued");
  CU(cudaEventRecord(h->ev[4], h->stream));
  if (expand_overlapped) {
    if (h->fork_expand) {  // the commit path had no fork point (no requirements, or more bit rows than the union path takes)
      h->fork_expand = false;
      CU(cudaEventRecord(h->ev_fork, h->stream));
      CU(cudaStreamWaitEvent(h->aux_stream, h->ev_fork, 0));
      if (!h->expand_started) CU(cudaEventRecord(h->ev_x0, h->aux_stream));
      if (int rc = run_expand(h, &launches, h->aux_stream, h->expand_from_word)) return rc;
      CU(cudaEventRecord(h->ev_x1, h->aux_stream));
    }
    CU(cudaStreamWaitEvent(h->stream, h->ev_x1, 0));
  }
  int placed_and_mismatch[2] = {0, 0};
  int& n_placed = placed_and_mismatch[0];
  h->decisions_gathered = false;
  if (decision_push(h).n > 0) {
    // every shard's k_materialize has stored its winners' records into all regions: raise this shard's decision flag behind them
    PeerFlags f{};
    f.n = h->world; f.rank = h->rank; f.epoch = h->peer_epoch;
    for (int p = 0; p < h->world; ++p) f.flags[p] = h->peer_base[p] + 16;
    k_peer_signal<<<1, 32, 0, h->stream>>>(f);
    ++launches;
    CU(cudaGetLastError());
  }
  if (h->M > 0 && out) CU(cudaMemcpyAsync(out, h->d_dec.p, (size_t)h->M * sizeof(oe_decision), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(placed_and_mismatch, h->d_nplaced.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  int peer_err = 0;
  if (h->peers_attached) CU(cudaMemcpyAsync(&peer_err, h->d_peer_err.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaEventRecord(h->ev[5], h->stream));
  CU(cudaStreamSynchronize(h->stream));
  tr.mark("synced");
  if (peer_err) {
    CU(cudaMemsetAsync(h->d_peer_err.p, 0, sizeof(int), h->stream));
    h->class_bits_fresh = false;
    return h->fail(OE_ERR_STATE, "oe_evaluate_cycle: shard %d did not publish its class rows of pass %u within %.1f s (every rank must issue the same calls)",
                   peer_err - 1, h->peer_epoch, h->peer_timeout_ns * 1e-9);
  }
  if (placed_and_mismatch[1]) {
    CU(cudaMemsetAsync(h->d_nplaced.as<int>() + 1, 0, sizeof(int), h->stream));
    h->class_bits_fresh = false;
    return h->fail(OE_ERR_STATE, "oe_evaluate_cycle: requirement %d was given an offer whose stage pipeline fails when it is re-run to record the "
                   "ports and volumes (feasibility pass and pipeline disagree): the decisions of this cycle are not usable", placed_and_mismatch[1] - 1);
  }
  h->cycle_done = true;
  if (stats) {
    std::memset(stats, 0, sizeof *stats);
    stats->ms_counters = ms_between(h->ev[0], h->ev[1]);
    stats->ms_feasibility = ms_between(h->ev[1], h->ev[2]);
    stats->ms_expand = expand_overlapped ? ms_between(h->ev_x0, h->ev_x1) : ms_between(h->ev[2], h->ev[3]);  // overlaps ms_commit
    stats->ms_commit = ms_between(h->ev[3], h->ev[4]);
    stats->ms_d2h = ms_between(h->ev[4], h->ev[5]);
    stats->ms_total = ms_between(h->ev[0], h->ev[5]);
    stats->n_placed = n_placed;
    stats->n_pairs = h->N_global * h->M;
    stats->kernel_launches = launches;
  }
  return OE_OK;
}

int32_t oe_stage_failures(oe_engine* h, int64_t req_begin, int64_t req_end, uint64_t* failed) {
  CallerTablesExpire expire_guard{h};
  if (int rc = ready(h, "oe_stage_failures")) return rc;
  if (req_begin < 0 || req_end < req_begin || req_end > h->M) return h->fail(OE_ERR_INVALID, "oe_stage_failures: requirement range");
  if (req_end - req_begin > 65535) return h->fail(OE_ERR_LIMIT, "oe_stage_failures: at most 65535 requirements per call");
  const int64_t n = (req_end - req_begin) * h->N;
  if (n == 0) return OE_OK;
  if (!failed) return h->fail(OE_ERR_INVALID, "oe_stage_failures: null output");
  int64_t launches = 0;
  if (!h->counters_fresh) if (int rc = run_counters(h, &launches)) return rc;
  CU(h->d_dbg.ensure((size_t)n * 8));
  dim3 grid((unsigned)((h->N + 127) / 128), (unsigned)(req_end - req_begin));
  k_stage_failures<<<grid, 128, 0, h->stream>>>(h->ctx(), h->d_req_row.as<int32_t>(), req_begin, h->d_dbg.as<unsigned long long>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(failed, h->d_dbg.p, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return OE_OK;
}

int32_t oe_unexpected_resources(oe_engine* h, const uint32_t* keep_bits, int32_t only_unused_offers, uint8_t* unexpected, int64_t* n_unexpected) {
  CallerTablesExpire expire_guard{h};
  if (int rc = ready(h, "oe_unexpected_resources")) return rc;
  if (n_unexpected) *n_unexpected = 0;
  const int64_t n_res = h->n_reserved;
  if (n_res == 0) return OE_OK;
  if (!keep_bits || !unexpected) return h->fail(OE_ERR_INVALID, "oe_unexpected_resources: null argument");
  const size_t keep_words = (size_t)(h->dict.n_res_ids + 31) / 32;
  CU(h->d_keep.ensure(std::max<size_t>(keep_words, 1) * 4));
  CU(cudaMemcpyAsync(h->d_keep.p, keep_bits, keep_words * 4, cudaMemcpyHostToDevice, h->stream));
  const size_t out_bytes = ((size_t)n_res + 7) / 8 * 8;  // the 8-byte counter sits behind the flags
  CU(h->d_gc_out.ensure(out_bytes + 8));
  CU(cudaMemsetAsync(h->d_gc_out.p, 0, out_bytes + 8, h->stream));
  unsigned long long* d_count = reinterpret_cast<unsigned long long*>(h->d_gc_out.as<uint8_t>() + out_bytes);
  uint32_t* taken = nullptr;
  if (only_unused_offers && h->cycle_done && h->M > 0) {
    const size_t words = (size_t)(h->N + 31) / 32;
    CU(h->d_taken.ensure(std::max<size_t>(words, 1) * 4));
    CU(cudaMemsetAsync(h->d_taken.p, 0, words * 4, h->stream));
    k_mark_taken<<<(unsigned)((h->M + 255) / 256), 256, 0, h->stream>>>(h->d_dec.as<oe_decision>(), h->M, h->row_offset, h->N, h->d_taken.as<uint32_t>());
    taken = h->d_taken.as<uint32_t>();
  }
  k_unexpected<<<(unsigned)((h->N + 127) / 128), 128, 0, h->stream>>>(h->dof, h->d_keep.as<uint32_t>(), taken, h->d_gc_out.as<uint8_t>(), d_count);
  CU(cudaGetLastError());
  unsigned long long count = 0;
  CU(cudaMemcpyAsync(unexpected, h->d_gc_out.p, (size_t)n_res, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(&count, d_count, 8, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (n_unexpected) *n_unexpected = (int64_t)count;
  return OE_OK;
}

namespace {
// Sharded engines: waits (on the device) until every shard's decision records of this pass have landed in this shard's region
// and copies the decisions of winners that live on other shards into the decision table.
int gather_decisions(oe_engine* h, oe_decision* out = nullptr) {
  if (!h->cycle_done) return h->fail(OE_ERR_STATE, "no cycle evaluated since the last load");
  if (!(h->world > 1 && h->peers_attached && h->exch_max_reqs >= h->M && h->exch_rec_words > 0))
    return h->fail(OE_ERR_STATE, "the merged decision table needs attached peers and a decision area sized for the requirement table: load the requirements, then "
                   "oe_peer_export / oe_peer_attach (area for %lld requirements, table has %lld)", (long long)h->exch_max_reqs, (long long)h->M);
  if (h->decisions_gathered || h->M == 0) return 0;
  k_peer_wait<<<1, 32, 0, h->stream>>>(h->peer_base[h->rank] + 16, h->world, h->peer_epoch, h->peer_timeout_ns, h->d_peer_err.as<int>());
  k_gather_decisions<<<(unsigned)((h->M + 255) / 256), 256, 0, h->stream>>>(h->d_dec.as<oe_decision>(), h->M, h->exch_records(h->rank, h->peer_epoch), h->exch_rec_words,
                                                                          h->row_offset, h->N);
  CU(cudaGetLastError());
  int peer_err = 0;
  CU(cudaMemcpyAsync(&peer_err, h->d_peer_err.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  if (out) CU(cudaMemcpyAsync(out, h->d_dec.p, (size_t)h->M * sizeof(oe_decision), cudaMemcpyDeviceToHost, h->stream));   // same round trip
  CU(cudaStreamSynchronize(h->stream));
  if (peer_err) {
    CU(cudaMemsetAsync(h->d_peer_err.p, 0, sizeof(int), h->stream));
    return h->fail(OE_ERR_STATE, "shard %d did not publish its decisions of pass %u within %.1f s (every rank must call oe_evaluate_cycle)", peer_err - 1, h->peer_epoch,
                   h->peer_timeout_ns * 1e-9);
  }
  h->decisions_gathered = true;
  return 0;
}
}  // namespace

int32_t oe_gather_decisions(oe_engine* h, oe_decision* out) {
  if (!h) return OE_ERR_INVALID;
  CU(cudaSetDevice(h->device));
  const bool fresh = !(h->decisions_gathered || h->M == 0);
  if (int rc = gather_decisions(h, out)) return rc;
  if (!fresh && out && h->M > 0) {   // gathered before (oe_evaluate_cycle with a decision buffer, or an earlier call): just the copy
    CU(cudaMemcpyAsync(out, h->d_dec.p, (size_t)h->M * sizeof(oe_decision), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
  }
  return OE_OK;
}

int32_t oe_apply_launches(oe_engine* h) {
  if (int rc = ready(h, "oe_apply_launches")) return rc;
  if (int rc = absorb_host_mirrors(h)) return rc;
  if (!h->cycle_done) return h->fail(OE_ERR_STATE, "oe_apply_launches: no cycle evaluated since the last load");
  const bool from_records = h->world > 1;
  if (from_records) {
    if (!(h->peers_attached && h->exch_max_reqs >= h->M && h->exch_rec_words > 0))
      return h->fail(OE_ERR_STATE, "oe_apply_launches on a sharded engine needs attached peers and a decision area (oe_peer_export after the requirements were loaded): "
                     "the winning offers' placement columns live on other shards");
    if (int rc = gather_decisions(h)) return rc;
  }
  if (h->h_launch_begin.empty()) return h->fail(OE_ERR_STATE, "oe_apply_launches: the requirement table carries no launch rows");
  if (h->dt.TA < h->dof.A) return h->fail(OE_ERR_INVALID, "oe_apply_launches: task table has fewer attribute slots than the offer table");
  std::vector<int32_t> dst, req, name;
  std::vector<uint32_t> filt;
  // the decisions of the last cycle still sit in device memory (no per-cycle host copy is kept for this rarely used entry)
  h->h_last_dec.resize((size_t)h->M);
  if (h->M > 0) CU(cudaMemcpyAsync(h->h_last_dec.data(), h->d_dec.p, (size_t)h->M * sizeof(oe_decision), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (!h->name_map_built) {
    h->name_to_row.reserve(h->h_tname.size() * 2 + 16);
    for (size_t r = 0; r < h->h_tname.size(); ++r) h->name_to_row[h->h_tname[r]] = (int32_t)r;
    h->name_map_built = true;
  }
  int64_t T = h->dt.T;
  const int FW = h->dt.FW;
  for (int64_t j = 0; j < h->M; ++j) {
    if (h->h_last_dec[j].offer_row < 0) continue;
    for (int32_t l = h->h_launch_begin[j]; l < h->h_launch_begin[j + 1]; ++l) {
      const int32_t nm = h->h_launch_name[l];
      auto it = h->name_to_row.find(nm);
      int32_t row;
      if (it == h->name_to_row.end()) {
        row = (int32_t)T++;
        h->name_to_row[nm] = row;
        h->h_tname.push_back(nm); h->h_ttype.push_back(0); h->h_tindex.push_back(0);
      } else {
        row = it->second;
      }
      h->h_ttype[row] = h->h_rtype[j];
      h->task_spans_valid = false;
      h->h_tindex[row] = h->h_rindex[j];
      dst.push_back(row); req.push_back((int32_t)j); name.push_back(nm);
      for (int w = 0; w < FW; ++w) filt.push_back(h->h_launch_filter[(size_t)l * FW + w]);
    }
  }
  if (int rc = reserve_tasks(h, T, h->dt.T)) return rc;
  if (!dst.empty()) {
    if (int rc = upload(h, h->d_l_dst, dst.data(), dst.size())) return rc;
    if (int rc = upload(h, h->d_l_req, req.data(), req.size())) return rc;
    if (int rc = upload(h, h->d_l_name, name.data(), name.size())) return rc;
    if (int rc = upload(h, h->d_l_filter, filt.data(), filt.size())) return rc;
    LaunchArgs a{(int64_t)dst.size(), h->d_l_dst.as<int32_t>(), h->d_l_req.as<int32_t>(), h->d_l_name.as<int32_t>(), h->d_l_filter.as<uint32_t>(),
                 h->d_rtype.as<int32_t>(), h->d_rindex.as<int32_t>()};
    // a pod stores each task name once, and names are unique rows, so no two threads write one row
    if (from_records)
      k_apply_launches_records<<<(unsigned)((a.n + 127) / 128), 128, 0, h->stream>>>(a, h->dt, h->exch_records(h->rank, h->peer_epoch), h->exch_rec_words, h->dof.A, h->dof.Q);
    else
      k_apply_launches<<<(unsigned)((a.n + 127) / 128), 128, 0, h->stream>>>(a, h->dof, h->dt, h->d_dec.as<oe_decision>());
    CU(cudaGetLastError());
  }
  h->dt.T = T;
  CU(cudaStreamSynchronize(h->stream));
  h->counters_fresh = h->class_bits_fresh = h->cycle_done = false;
  if (h->self_provided) {
    h->hasR = false;  // the caller's equivalent-task lists name rows of the old table: requirements must be re-loaded
  } else if (int rc = rebuild_variants(h)) {
    return rc;  // the stored tasks are now "equivalent" to the requirements that placed them
  }
  return OE_OK;
}

int32_t oe_export_tasks(oe_engine* h, int64_t* n_tasks, int32_t* name_id, int32_t* pod_type_id, int32_t* pod_index, int32_t* agent_id,
                        int32_t* hostname_id, int32_t* zone_id, int32_t* region_id, int32_t* attr_id, int32_t* rr_key, uint32_t* filter_bits) {
  if (!h || !h->hasT) return OE_ERR_STATE;
  CU(cudaSetDevice(h->device));
  const DevTasks& t = h->dt;
  if (n_tasks) *n_tasks = t.T;
  if (!name_id) return OE_OK;
  const size_t n = (size_t)t.T;
  int32_t* dst[7] = {name_id, pod_type_id, pod_index, agent_id, hostname_id, zone_id, region_id};
  const int32_t* src[7] = {t.name, t.type, t.index, t.agent, t.host, t.zone, t.region};
  for (int c = 0; c < 7; ++c) CU(cudaMemcpyAsync(dst[c], src[c], n * 4, cudaMemcpyDeviceToHost, h->stream));
  for (int a = 0; a < t.TA; ++a) CU(cudaMemcpyAsync(attr_id + (size_t)a * n, t.attr + (size_t)a * t.cap, n * 4, cudaMemcpyDeviceToHost, h->stream));
  for (int q = 0; q < t.Q; ++q) CU(cudaMemcpyAsync(rr_key + (size_t)q * n, t.rr + (size_t)q * t.cap, n * 4, cudaMemcpyDeviceToHost, h->stream));
  for (int w = 0; w < t.FW; ++w) CU(cudaMemcpyAsync(filter_bits + (size_t)w * n, t.filter + (size_t)w * t.cap, n * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return OE_OK;
}

int32_t oe_peer_export(oe_engine* h, int32_t max_rows, void* handle, void** dev_ptr) {
  if (!h) return OE_ERR_INVALID;
  if (h->world < 2) return h->fail(OE_ERR_STATE, "oe_peer_export: the engine is not sharded (world = 1)");
  if (h->world > OE_PEER_MAX_WORLD) return h->fail(OE_ERR_LIMIT, "oe_peer_export: at most %d shards", OE_PEER_MAX_WORLD);
  if (!h->hasO) return h->fail(OE_ERR_STATE, "oe_peer_export: load the offers first (the row stride comes from shard_rows)");
  if (max_rows < 1) return h->fail(OE_ERR_INVALID, "oe_peer_export: max_rows");
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize(h->stream));
  h->detach_peers();
  h->exch_max_rows = max_rows;
  h->exch_stride = h->stride();
  // decision records (merged decision table, launches on a sharded engine): sized for the requirement table loaded now, with head room
  h->exch_rec_words = ((8 + 4 + h->dof.A + h->dof.Q) + 3) / 4 * 4;
  h->exch_max_reqs = h->hasR ? std::max<int64_t>(1024, h->M + h->M / 4) : 0;
  const size_t bytes = (OE_PEER_HEADER_WORDS + 2 * h->exch_parity_words() + 2 * (size_t)h->exch_max_reqs * h->exch_rec_words) * 4;
  if (h->d_exch.cap < bytes) {   // a region of its own, exactly one cudaMalloc: that is what an IPC handle names
    h->d_exch.release();
    CU(cudaMalloc(&h->d_exch.p, bytes));
    h->d_exch.cap = bytes;
  }
  CU(h->d_peer_err.ensure(sizeof(int)));
  CU(cudaMemsetAsync(h->d_exch.p, 0, OE_PEER_HEADER_WORDS * 4, h->stream));
  CU(cudaMemsetAsync(h->d_peer_err.p, 0, sizeof(int), h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->peer_epoch = 0;
  if (const char* ms = getenv("OE_PEER_TIMEOUT_MS")) {
    const long long v = atoll(ms);
    if (v > 0) h->peer_timeout_ns = (unsigned long long)v * 1000000ull;
  }
  if (handle) {
    static_assert(sizeof(cudaIpcMemHandle_t) == OE_PEER_HANDLE_BYTES, "OE_PEER_HANDLE_BYTES");
    cudaIpcMemHandle_t ih;
    CU(cudaIpcGetMemHandle(&ih, h->d_exch.p));
    std::memcpy(handle, &ih, sizeof ih);
  }
  if (dev_ptr) *dev_ptr = h->d_exch.p;
  return OE_OK;
}

int32_t oe_peer_attach(oe_engine* h, const void* handles, void* const* same_process_ptrs) {
  if (!h) return OE_ERR_INVALID;
  if (!h->d_exch.p || h->exch_max_rows < 1) return h->fail(OE_ERR_STATE, "oe_peer_attach: call oe_peer_export first");
  if (!handles && !same_process_ptrs) return h->fail(OE_ERR_INVALID, "oe_peer_attach: neither handles nor pointers");
  CU(cudaSetDevice(h->device));
  h->detach_peers();
  for (int p = 0; p < h->world; ++p) {
    if (p == h->rank) { h->peer_base[p] = h->d_exch.as<uint32_t>(); continue; }
    if (same_process_ptrs && same_process_ptrs[p]) {
      cudaPointerAttributes at{};
      CU(cudaPointerGetAttributes(&at, same_process_ptrs[p]));
      if (at.type != cudaMemoryTypeDevice) { h->detach_peers(); return h->fail(OE_ERR_INVALID, "oe_peer_attach: pointer of shard %d is not device memory", p); }
      if (at.device != h->device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, h->device, at.device));
        if (!can) { h->detach_peers(); return h->fail(OE_ERR_CUDA, "oe_peer_attach: device %d cannot access device %d", h->device, at.device); }
        cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { h->detach_peers(); return h->cuda_fail(e, "cudaDeviceEnablePeerAccess"); }
        cudaGetLastError();
      }
      h->peer_base[p] = static_cast<uint32_t*>(same_process_ptrs[p]);
      continue;
    }
    if (!handles) { h->detach_peers(); return h->fail(OE_ERR_INVALID, "oe_peer_attach: no handle and no pointer for shard %d", p); }
    cudaIpcMemHandle_t ih;
    std::memcpy(&ih, static_cast<const unsigned char*>(handles) + (size_t)p * OE_PEER_HANDLE_BYTES, sizeof ih);
    void* q = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&q, ih, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); h->detach_peers(); return h->fail(OE_ERR_CUDA, "oe_peer_attach: cudaIpcOpenMemHandle for shard %d: %s", p, cudaGetErrorString(e)); }
    h->peer_base[p] = static_cast<uint32_t*>(q);
    h->peer_opened[p] = true;
  }
  h->peers_attached = true;
  h->class_bits_fresh = false;
  return OE_OK;
}

int32_t oe_peer_detach(oe_engine* h) {
  if (!h) return OE_ERR_INVALID;
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize(h->stream));
  h->detach_peers();
  h->class_bits_fresh = false;
  return OE_OK;
}

int32_t oe_device_buffer(oe_engine* h, int32_t which, void** dev_ptr, int64_t* n_bytes) {
  if (!h || !dev_ptr || !n_bytes) return OE_ERR_INVALID;
  CU(cudaSetDevice(h->device));
  const int64_t stride = h->stride();
  switch (which) {
    case OE_BUF_CLASS_BITS_LOCAL:
      *n_bytes = (int64_t)h->rows_factored() * stride * 4;
      CU(h->d_F.ensure(std::max<int64_t>(*n_bytes, 4)));
      *dev_ptr = h->d_F.p;
      break;
    case OE_BUF_CLASS_BITS_GLOBAL:
      *n_bytes = (int64_t)h->world * h->rows_factored() * stride * 4;
      CU(h->d_Fglobal.ensure(std::max<int64_t>(*n_bytes, 4)));
      *dev_ptr = h->d_Fglobal.p;
      break;
    case OE_BUF_MATRIX:
      *n_bytes = h->M * stride * 4;
      CU(h->d_matrix.ensure(std::max<int64_t>(*n_bytes, 4)));
      *dev_ptr = h->d_matrix.p;
      break;
    case OE_BUF_DECISIONS:
      *n_bytes = h->M * (int64_t)sizeof(oe_decision);
      CU(h->d_dec.ensure(std::max<int64_t>(*n_bytes, 4)));
      *dev_ptr = h->d_dec.p;
      break;
    default: return h->fail(OE_ERR_INVALID, "oe_device_buffer: unknown buffer %d", which);
  }
  return OE_OK;
}

}  // extern "C"
