// oe_kernels_commit_offers.cuh — part of the one translation unit oe_engine.cu (included there, in this order; not compiled on its own).
// Kernels of the offer-side commit: row grouping, per-offer masks, the stable counting sorts, the replay (k_commit_offers), and the
// helpers of the union path's two passes.
#pragma once

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
