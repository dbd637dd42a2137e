// oe_kernels_decisions.cuh — part of the one translation unit oe_engine.cu (included there, in this order; not compiled on its own).
// Kernels behind the replay: K2b (winners re-run with recording, decision records pushed to the other shards), the merged decision
// table, failed-stage masks, the unexpected-resource scan, K4 (launches recorded into the task table).
#pragma once

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
