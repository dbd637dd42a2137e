// oe_host_loads.inl — part of the one translation unit oe_engine.cu (included there; not compiled on its own).
// The loading half of the C ABI: oe_load_programs / _tasks / _offers / _offers_packed / oe_update_offers / oe_load_requirements,
// with what they keep on the host (variants of re-deployed pods, launch rows) and what they leave to the device.

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
