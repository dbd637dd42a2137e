// oe_host_evaluate.inl — part of the one translation unit oe_engine.cu (included there; not compiled on its own).
// The evaluating half of the C ABI: K0 and the folded row descriptors, the feasibility pass (whole table or chunk by chunk while it
// arrives), the commit passes, oe_class_feasibility / oe_requirement_rows / oe_feasibility / oe_evaluate_cycle.

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
