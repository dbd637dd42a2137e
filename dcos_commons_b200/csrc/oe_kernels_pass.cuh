// oe_kernels_pass.cuh — part of the one translation unit oe_engine.cu (included there, in this order; not compiled on its own).
// Kernels of the feasibility pass: K0 (tallies), K0c / K0p / K0t (verdict tables), K1 (general and lean), the peer exchange's
// flags, K1x (the un-factored matrix).
#pragma once

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
