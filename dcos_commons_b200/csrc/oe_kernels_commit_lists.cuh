// oe_kernels_commit_lists.cuh — part of the one translation unit oe_engine.cu (included there, in this order; not compiled on its own).
// Kernels of the requirement-side commit (round 1: candidate lists, union lists; kept as the fallback for components of more
// than 32 rows and for OE_COMMIT=union), the bit-row view they share with the offer-side commit, and the row-overlap kernel.
#pragma once

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
