// oe_kernels_load.cuh — part of the one translation unit oe_engine.cu (included there, in this order; not compiled on its own).
// Kernels of the loads: id / range / offset checks, the packed offer table widened on the device, the range-list dictionary,
// the search for re-deployed pods, small tables fetched from pinned host memory, rows replaced in place.
#pragma once

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
