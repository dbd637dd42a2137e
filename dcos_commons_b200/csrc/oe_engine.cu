// oe_engine.cu — B200 (sm_100a) offer-evaluation engine: kernels + the C ABI of include/oe_abi.h, ONE translation unit.
//
// The kernels (all integer / fp64 compare-and-mask work, HBM-, issue- or latency-bound; no tensor cores by design) live in
// headers included below, the two large halves of the host side in fragments included further down:
//   oe_device.cuh                  device-side data model: DevOffers / DevTasks / DevProg, the stage interpreter eval_pair, the
//                                  rule evaluator, FastRow (a folded new-pod pipeline) and eval_fast
//   oe_kernels_load.cuh            loads: k_validate_*, k_unpack_offer_rows (packed wire form), k_tuple_hash_insert /
//                                  k_reintern_tuples (range-list dictionary), k_pod_hash_insert / k_match_tasks (re-deployed
//                                  pods), k_fetch_host, k_scatter_rows
//   oe_kernels_pass.cuh            K0 k_count_tasks / k_rr_summary, K0c k_rule_tables, K0p k_port_tables, K0t k_transpose_tables,
//                                  K1 k_class_feasibility (general) and k_class_feasibility_lean (the PEERS variants also store
//                                  their words into every shard's exchange region; k_peer_signal / k_peer_wait order the readers),
//                                  K1x k_expand_matrix (the HBM-bound pass)
//   oe_kernels_commit_lists.cuh    the bit-row view (RowBits), K3 candidate lists, k_row_overlap, round 1's requirement-side
//                                  commits k_commit / k_commit_union (fallbacks)
//   oe_kernels_commit_offers.cuh   the offer-side commit: k_row_link / k_group_rows, k_offer_masks, k_key_* (stable counting
//                                  sorts), k_commit_offers (the replay), the union path's two-pass helpers
//   oe_kernels_decisions.cuh       K2b k_materialize (+ decision records to the other shards), k_gather_decisions,
//                                  k_stage_failures, k_unexpected, K4 k_apply_launches
//   this file                      oe_engine (state, buffers, streams), create / destroy, and the entry points behind the passes:
//                                  failed-stage masks, the unexpected-resource scan, decisions on sharded engines, launches,
//                                  the peer exchange's set-up, device buffers
//   oe_host_loads.inl              oe_load_* / oe_update_offers
//   oe_host_evaluate.inl           K0 / K1 / commit orchestration, oe_class_feasibility / oe_feasibility / oe_evaluate_cycle
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

#include "oe_kernels_load.cuh"
#include "oe_kernels_pass.cuh"
#include "oe_kernels_commit_lists.cuh"
#include "oe_kernels_commit_offers.cuh"
#include "oe_kernels_decisions.cuh"

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



#include "oe_host_loads.inl"

#include "oe_host_evaluate.inl"

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
