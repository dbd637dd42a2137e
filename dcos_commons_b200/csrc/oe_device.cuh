// oe_device.cuh — device-side views of the encoded tables and the per-(offer, class) stage/rule evaluation.
//
// The evaluation below is the engine's own formulation of the reference pipeline, NOT a transliteration: placement
// tallies come from per-key counter arrays reduced once per cycle (K0) instead of re-walking every task for every
// offer, the pool is a register/local copy of one SoA offer row, and the rule tree is a postfix program over a
// bit-stack.  Semantics followed (S/ = sdk/scheduler/src/main/java/com/mesosphere/sdk/, P/ = S/offer/evaluate/placement/):
//   S/offer/MesosResourcePool.java:61-70,282-372   sufficientValue / consumeAtomic / consumeReservableMerged
//   S/offer/ValueUtils.java:103-106,144-155        fp64 subtract + compare (no FMA, no reassociation)
//   S/offer/evaluate/PortEvaluationStage.java:72-155,242-298
//   P/MaxPerRule.java:52-87, P/AbstractRoundRobinRule.java:53-151, P/TaskTypeRule.java:98-217
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/oe_abi.h"

namespace oe {

struct DevOffers {
  int64_t N;
  int64_t row_offset;
  int S, K, R, D, A, Q, E;
  const double* scalar;
  const uint32_t* present;
  const uint32_t* port;
  const double* msize;
  const int32_t* mprof;
  const int32_t* alloc;
  const int32_t* agent;
  const int32_t* host;
  const int32_t* zone;
  const int32_t* region;
  const int32_t* attr;
  const int32_t* rr;
  const int32_t* exec;
  const int32_t* ptuple;  // [S][N] id of the (slot, offer)'s `ports` range list among the table's distinct lists, or null
  int32_t n_ptuples;
  const int64_t* rbegin;  // may be null
  const int32_t* rid;
  const int32_t* rkind;
  const double* rval;
};

struct DevTasks {
  int64_t T;
  int64_t cap;  // column stride (rows allocated)
  int TA, Q, FW;
  int32_t* name;
  int32_t* type;
  int32_t* index;
  int32_t* agent;
  int32_t* host;
  int32_t* zone;
  int32_t* region;
  int32_t* attr;    // [TA][cap]
  int32_t* rr;      // [Q][cap]
  uint32_t* filter; // [FW][cap]
};

struct DevProg {
  int C;
  int n_csets, n_tsets;
  int n_agents;
  const int32_t* op_begin;
  const oe_stage_op* ops;
  const oe_rule_node* nodes;
  const int32_t* pod_role;
  const int32_t* excl_begin;
  const uint16_t* excl;
  const int32_t* rs_begin;
  const uint32_t* rs;
  const oe_counter_set* csets;
  const int64_t* cs_off;  // [n_csets+1] offsets into cnt
  const int32_t* ts_type;
  const uint32_t* mbits;
};

// K0 outputs
struct DevCounters {
  const int32_t* cnt;        // concatenated per-counter-set tallies
  const int4* rrsum;         // [n_csets] {min over keys with cnt>0 (0 if none), max, #keys with cnt>0, 0}
  const int32_t* type_total; // [n_tsets]
  const int32_t* type_agent; // [n_tsets][n_agents]
};

// Requirements whose deployed "self" tasks (S/offer/TaskUtils.java:136-145) change the tallies they see.
struct DevVariants {
  int V;
  const int32_t* cls;        // [V]
  const int32_t* self_begin; // [V+1]
  const int32_t* self_rows;  // task rows
  const int4* rrsum;         // [V][n_csets] round-robin summary with the self rows taken out
};

struct DevCtx {
  DevOffers o;
  DevTasks t;
  DevProg p;
  DevCounters c;
  DevVariants v;
};

__device__ __forceinline__ bool mbit(const uint32_t* __restrict__ bits, int32_t off, int32_t v) {
  return v >= 0 && ((__ldg(bits + off + (v >> 5)) >> (v & 31)) & 1u);
}

__device__ __forceinline__ bool task_filter(const DevTasks& t, int32_t row, int f) {
  return (t.filter[(int64_t)(f >> 5) * t.cap + row] >> (f & 31)) & 1u;
}

// How many times key `k` occurs among the keys of the variant's self rows that pass the counter set's filter.
__device__ inline int self_key_count(const DevCtx& cx, int var, const oe_counter_set& cs, int32_t k) {
  int n = 0;
  for (int s = cx.v.self_begin[var]; s < cx.v.self_begin[var + 1]; ++s) {
    int32_t row = cx.v.self_rows[s];
    if (!task_filter(cx.t, row, cs.filter_id)) continue;
    switch (cs.keyspace) {
      case OE_KS_HOSTNAME: n += cx.t.host[row] == k; break;
      case OE_KS_ZONE: n += cx.t.zone[row] == k; break;
      case OE_KS_REGION: n += cx.t.region[row] == k; break;
      case OE_KS_ATTR:
        for (int a = 0; a < cx.t.TA; ++a) n += cx.t.attr[(int64_t)a * cx.t.cap + row] == k;
        break;
      case OE_KS_RR_ATTR: n += cx.t.rr[(int64_t)cs.rr_col * cx.t.cap + row] == k; break;
    }
  }
  return n;
}

__device__ inline int self_type_on_agent(const DevCtx& cx, int var, int32_t pod_type, int32_t agent) {
  int n = 0;
  for (int s = cx.v.self_begin[var]; s < cx.v.self_begin[var + 1]; ++s) {
    int32_t row = cx.v.self_rows[s];
    n += (cx.t.type[row] == pod_type) && (cx.t.agent[row] == agent);
  }
  return n;
}

// Placement columns of one offer row, loaded once per thread.
struct OfferKeys {
  int32_t agent, host, zone, region;
};

// Postfix rule program over a bit-stack (bit 0 = top).
// `nodes` is the node array the program indexes: the global one, or a CTA's shared-memory copy of it.
__device__ inline bool eval_rule(const DevCtx& cx, const oe_rule_node* nodes, int64_t i, const OfferKeys& ok, int var, int32_t begin,
                                 int32_t count) {
  uint64_t st = 0;
  const DevOffers& o = cx.o;
  for (int32_t n = begin; n < begin + count; ++n) {
    const oe_rule_node nd = nodes[n];
    bool v = false;
    switch (nd.op) {
      case OE_R_TRUE: v = true; break;
      case OE_R_FALSE: v = false; break;
      case OE_R_AND: {
        uint64_t m = nd.a >= 64 ? ~0ull : ((1ull << nd.a) - 1);
        v = nd.a > 0 && (st & m) == m;
        st = nd.a >= 64 ? 0 : st >> nd.a;
      } break;
      case OE_R_OR: {
        uint64_t m = nd.a >= 64 ? ~0ull : ((1ull << nd.a) - 1);
        v = (st & m) != 0;
        st = nd.a >= 64 ? 0 : st >> nd.a;
      } break;
      case OE_R_NOT: v = !(st & 1); st >>= 1; break;
      case OE_R_AGENT_EQ: v = ok.agent == nd.a; break;
      case OE_R_ATTR_ANY:
        for (int a = 0; a < o.A; ++a) v = v || mbit(cx.p.mbits, nd.a, __ldg(o.attr + (int64_t)a * o.N + i));
        break;
      case OE_R_HOST_IN: v = mbit(cx.p.mbits, nd.a, ok.host); break;
      case OE_R_ZONE_IN: v = mbit(cx.p.mbits, nd.a, ok.zone); break;
      case OE_R_REGION_IN: v = mbit(cx.p.mbits, nd.a, ok.region); break;
      case OE_R_LOCAL_REGION: v = ok.region < 0 || nd.b == 0 || ok.region == nd.a; break;
      case OE_R_MAXPER_HOST:
      case OE_R_MAXPER_ZONE:
      case OE_R_MAXPER_REGION: {
        const oe_counter_set cs = cx.p.csets[nd.a];
        int32_t k = nd.op == OE_R_MAXPER_HOST ? ok.host : (nd.op == OE_R_MAXPER_ZONE ? ok.zone : ok.region);
        if (k < 0) {
          v = nd.op != OE_R_MAXPER_ZONE;  // no key: zone rule fails outright, the others have nothing to exceed
        } else {
          int c = 1 + __ldg(cx.c.cnt + cx.p.cs_off[nd.a] + k);
          if (var >= 0) c -= self_key_count(cx, var, cs, k);
          v = c <= nd.b;
        }
      } break;
      case OE_R_MAXPER_ATTR: {
        const oe_counter_set cs = cx.p.csets[nd.a];
        v = true;
        for (int a = 0; a < o.A; ++a) {
          int32_t k = __ldg(o.attr + (int64_t)a * o.N + i);
          if (!mbit(cx.p.mbits, nd.c, k)) continue;
          int mult = 0;  // duplicate attribute strings in one offer start the tally above 1 (P/MaxPerRule.java:81-86)
          for (int b = 0; b < o.A; ++b) mult += __ldg(o.attr + (int64_t)b * o.N + i) == k;
          int c = mult + __ldg(cx.c.cnt + cx.p.cs_off[nd.a] + k);
          if (var >= 0) c -= self_key_count(cx, var, cs, k);
          v = v && c <= nd.b;
        }
      } break;
      case OE_R_RR: {
        const oe_counter_set cs = cx.p.csets[nd.a];
        int32_t k;
        switch (cs.keyspace) {
          case OE_KS_HOSTNAME: k = ok.host; break;
          case OE_KS_ZONE: k = ok.zone; break;
          case OE_KS_REGION: k = ok.region; break;
          case OE_KS_RR_ATTR: k = __ldg(o.rr + (int64_t)cs.rr_col * o.N + i); break;
          default: k = -1;
        }
        if (k < 0) { v = false; break; }
        int oc = __ldg(cx.c.cnt + cx.p.cs_off[nd.a] + k);
        int4 sum;
        if (var >= 0) {
          oc -= self_key_count(cx, var, cs, k);
          sum = cx.v.rrsum[(int64_t)var * cx.p.n_csets + nd.a];
        } else {
          sum = cx.c.rrsum[nd.a];
        }
        if (sum.x == sum.y || oc <= sum.x) {
          v = nd.b < 0 || sum.z >= nd.b || oc == 0;
        } else {
          v = false;
        }
      } break;
      case OE_R_TYPE_AVOID:
      case OE_R_TYPE_COLOCATE: {
        if (cx.c.type_total[nd.a] == 0) { v = true; break; }
        int on = __ldg(cx.c.type_agent + (int64_t)nd.a * cx.p.n_agents + ok.agent);
        if (var >= 0) on -= self_type_on_agent(cx, var, cx.p.ts_type[nd.a], ok.agent);
        v = nd.op == OE_R_TYPE_AVOID ? on == 0 : on > 0;
      } break;
      default: v = false;
    }
    st = (st << 1) | (v ? 1ull : 0ull);
  }
  return (st & 1) != 0;
}

#define OE_DEV_MAX_TAKEN 24  // host ports a single pipeline can take (static + dynamic); checked at load
#define OE_DEV_MAX_FREED 4   // ports / MOUNT disks one pipeline returns to the pool (OE_OP_FREE); checked at load

// Mutable view of one offer's pool while one class's pipeline runs over it.
struct PairState {
  double val[OE_MAX_ROLE_SLOTS * OE_MAX_SCALAR_KINDS];
  uint32_t present;
  uint32_t mounts_used;
  // host ports already placed in the pod by earlier stages.  Entries without OE_KEPT were also removed from the
  // pool; entries with it are re-used reservations that never were in the pool.
  uint32_t taken[OE_DEV_MAX_TAKEN];
  int n_taken;
  // orphaned reservations handed back by OE_OP_FREE (MesosResourcePool.free, :374-430): ports as slot << 16 | port, MOUNT
  // disks appended behind the offer's own (reported as slot OE_MAX_MOUNT_SLOTS + k)
  uint32_t freed_port[OE_DEV_MAX_FREED];
  double freed_mount_size[OE_DEV_MAX_FREED];
  int32_t freed_mount_prof[OE_DEV_MAX_FREED];
  int n_freed_ports, n_freed_mounts;
  uint32_t freed_mounts_used;
};
#define OE_KEPT 0x10000u

__device__ __forceinline__ bool port_in_pool(const DevOffers& o, int64_t i, int slot, int nr, int p) {
  for (int r = 0; r < nr; ++r) {
    uint32_t pr = __ldg(o.port + ((int64_t)slot * o.R + r) * o.N + i);
    int b = pr & 0xFFFF, e = pr >> 16;
    if (b <= p && p <= e) return true;
  }
  return false;
}

__device__ __forceinline__ bool port_freed(const PairState& ps, int slot, int p) {
  bool f = false;
  for (int q = 0; q < ps.n_freed_ports; ++q) f = f || ps.freed_port[q] == (((uint32_t)slot << 16) | (uint32_t)p);
  return f;
}

// removed from the pool by an earlier stage
__device__ __forceinline__ bool port_taken(const PairState& ps, int p) {
  bool t = false;
  for (int q = 0; q < ps.n_taken; ++q) t = t || ps.taken[q] == (uint32_t)p;
  return t;
}
// present in the pod's builders (PortEvaluationStage.java:264-269)
__device__ __forceinline__ bool port_placed(const PairState& ps, int p) {
  bool t = false;
  for (int q = 0; q < ps.n_taken; ++q) t = t || (ps.taken[q] & 0xFFFFu) == (uint32_t)p;
  return t;
}

// Lowest port of the slot's ranges that is neither taken by an earlier stage of this pipeline, nor declared
// statically anywhere in the pod, and (when constrained) lies inside the range set.  -1 = none.
__device__ inline bool dynamic_port_ok(const DevCtx& cx, int cls, int32_t rset, const PairState& ps, int p) {
  if (port_placed(ps, p)) return false;
  for (int q = cx.p.excl_begin[cls]; q < cx.p.excl_begin[cls + 1]; ++q) if (cx.p.excl[q] == p) return false;
  if (rset < 0) return true;
  for (int q = cx.p.rs_begin[rset]; q < cx.p.rs_begin[rset + 1]; ++q) {
    const uint32_t rr = cx.p.rs[q];
    if ((int)(rr & 0xFFFF) <= p && p <= (int)(rr >> 16)) return true;
  }
  return false;
}

__device__ inline int pick_dynamic_port_in_ranges(const DevCtx& cx, int64_t i, int cls, int slot, int nr, int32_t rset, const PairState& ps) {
  const DevOffers& o = cx.o;
  const int eb = cx.p.excl_begin[cls], ee = cx.p.excl_begin[cls + 1];
  for (int r = 0; r < nr; ++r) {
    uint32_t pr = __ldg(o.port + ((int64_t)slot * o.R + r) * o.N + i);
    int b = pr & 0xFFFF, e = pr >> 16;
    int p = b;
    while (p <= e) {
      bool blocked = port_placed(ps, p);
      for (int q = eb; q < ee && !blocked; ++q) blocked = cx.p.excl[q] == p;
      if (blocked) { ++p; continue; }
      if (rset >= 0) {
        bool in = false;
        int next = 0x7fffffff;
        for (int q = cx.p.rs_begin[rset]; q < cx.p.rs_begin[rset + 1]; ++q) {
          uint32_t rr = cx.p.rs[q];
          int rb = rr & 0xFFFF, re = rr >> 16;
          if (rb <= p && p <= re) in = true;
          else if (rb > p && rb < next) next = rb;
        }
        if (!in) { p = next; continue; }  // jump to the next constrained range (or past the end)
      }
      return p;
    }
  }
  return -1;
}

// ... over the offered ranges and the ports handed back by earlier OE_OP_FREE stages of this pipeline
__device__ inline int pick_dynamic_port(const DevCtx& cx, int64_t i, int cls, int slot, int nr, int32_t rset, const PairState& ps) {
  int best = pick_dynamic_port_in_ranges(cx, i, cls, slot, nr, rset, ps);
  for (int q = 0; q < ps.n_freed_ports; ++q) {
    if ((int)(ps.freed_port[q] >> 16) != slot) continue;
    const int p = (int)(ps.freed_port[q] & 0xFFFFu);
    if ((best < 0 || p < best) && dynamic_port_ok(cx, cls, rset, ps, p)) best = p;
  }
  return best;
}

__device__ __forceinline__ bool offer_has_exec(const DevOffers& o, int64_t i, int32_t id) {
  bool h = false;
  for (int e = 0; e < o.E; ++e) h = h || __ldg(o.exec + (int64_t)e * o.N + i) == id;
  return h;
}

// Runs the stage program of class `cls` over offer row i.  Returns true iff every stage passes.  With RECORD the
// choices the host needs to materialise the winner (dynamic ports, MOUNT slots, new reservation count) are written.
// With `failed` non-null every stage runs, as in the reference (OfferEvaluator.java:165-171: no short-circuit, a failed
// stage leaves the pool untouched and the later stages still see it), and bit min(stage, 63) is set per failed stage.
template <bool RECORD>
__device__ inline bool eval_pair(const DevCtx& cx, int64_t i, const OfferKeys& ok, int cls, int var, oe_decision* dec,
                                 uint64_t* failed = nullptr) {
  const DevOffers& o = cx.o;
  PairState ps;
  const int SK = o.S * o.K;
  // consumableResource (MesosResourcePool.java:87-97): an offer allocated to another role contributes nothing
  const int32_t pod_role = cx.p.pod_role[cls];
  const int32_t alloc = __ldg(o.alloc + i);
  const bool consumable = !(pod_role >= 0 && alloc >= 0 && alloc != pod_role);
  ps.present = consumable ? __ldg(o.present + i) : 0u;
  for (int q = 0; q < SK; ++q) ps.val[q] = __ldg(o.scalar + (int64_t)q * o.N + i);
  ps.mounts_used = 0;
  ps.n_taken = 0;
  ps.n_freed_ports = 0; ps.n_freed_mounts = 0; ps.freed_mounts_used = 0;
  const int nr = consumable ? o.R : 0;
  const int nd = consumable ? o.D : 0;
  int n_dyn = 0, n_mount = 0, n_res = 0;
  bool all = true;

  const int ob = cx.p.op_begin[cls], oe_ = cx.p.op_begin[cls + 1];
  for (int x = ob; x < oe_; ++x) {
    const oe_stage_op op = cx.p.ops[x];
    bool pass = true;
    if ((op.flags & OE_OPF_SKIP_IF_RUNNING_EXECUTOR) && offer_has_exec(o, i, (int32_t)op.aux)) continue;
    switch (op.opcode) {
      case OE_OP_NOP: break;
      case OE_OP_EXEC_ID: pass = op.arg < 0 || offer_has_exec(o, i, op.arg); break;
      case OE_OP_RULE: pass = eval_rule(cx, cx.p.nodes, i, ok, var, op.arg, (int32_t)op.aux); break;
      case OE_OP_SCALAR_NEW: {
        const int q = op.slot * o.K + op.kind;
        const double avail = ps.val[q];
        const double diff = __dsub_rn(op.amount, avail);
        pass = ((ps.present >> q) & 1u) && !(diff > 0.0);
        if (pass) { ps.val[q] = __dsub_rn(avail, op.amount); ++n_res; }
      } break;
      case OE_OP_PORT_STATIC: {
        pass = (port_in_pool(o, i, op.slot, nr, op.arg) || port_freed(ps, op.slot, op.arg)) && !port_taken(ps, op.arg);
        if (pass) { ps.taken[ps.n_taken++] = (uint32_t)op.arg; ++n_res; }
      } break;
      case OE_OP_PORT_DYNAMIC: {
        int p = pick_dynamic_port(cx, i, cls, op.slot, nr, op.arg, ps);
        pass = p >= 0;
        if (pass) {
          ps.taken[ps.n_taken++] = (uint32_t)p;
          ++n_res;
          if (RECORD && n_dyn < OE_MAX_DYN_PORTS) dec->dyn_port[n_dyn] = (uint16_t)p;
        }
        ++n_dyn;
      } break;
      case OE_OP_VOL_MOUNT: {
        int chosen = -1;
        for (int d = 0; d < nd; ++d) {
          if ((ps.mounts_used >> d) & 1u) continue;
          const double size = __ldg(o.msize + (int64_t)d * o.N + i);
          const int32_t prof = __ldg(o.mprof + (int64_t)d * o.N + i);
          const bool prof_ok = prof >= 0 ? ((op.mask >> prof) & 1ull) != 0 : op.mask == 0;
          if (!(__dsub_rn(op.amount, size) > 0.0) && prof_ok) { chosen = d; break; }
        }
        for (int d = 0; d < ps.n_freed_mounts && chosen < 0; ++d) {
          if ((ps.freed_mounts_used >> d) & 1u) continue;
          const int32_t prof = ps.freed_mount_prof[d];
          const bool prof_ok = prof >= 0 ? ((op.mask >> prof) & 1ull) != 0 : op.mask == 0;
          if (!(__dsub_rn(op.amount, ps.freed_mount_size[d]) > 0.0) && prof_ok) chosen = OE_MAX_MOUNT_SLOTS + d;
        }
        pass = chosen >= 0;
        if (pass) {
          if (chosen >= OE_MAX_MOUNT_SLOTS) ps.freed_mounts_used |= 1u << (chosen - OE_MAX_MOUNT_SLOTS);
          else ps.mounts_used |= 1u << chosen;
          ++n_res;
          if (RECORD && n_mount < OE_MAX_MOUNT_OPS) dec->mount_slot[n_mount] = (int8_t)chosen;
        }
        ++n_mount;
      } break;
      case OE_OP_SCALAR_RESERVED:
      case OE_OP_PORT_RESERVED:
      case OE_OP_MOUNT_RESERVED: {
        // precondition (checked by the host encoder): one resource id is named by at most one op of a class
        double reserved = 0.0;
        bool found = false;
        if (consumable && o.rbegin) {
          for (int64_t q = o.rbegin[i]; q < o.rbegin[i + 1]; ++q)
            if (o.rid[q] == op.arg) { found = true; reserved = o.rval[q]; }
        }
        if (!found) { pass = false; break; }
        if (op.opcode == OE_OP_MOUNT_RESERVED) break;
        if (op.opcode == OE_OP_PORT_RESERVED) {
          const int want = (int)op.aux;
          if ((int)reserved != want) {
            pass = (port_in_pool(o, i, op.slot, nr, want) || port_freed(ps, op.slot, want)) && !port_taken(ps, want);
            if (pass) ps.taken[ps.n_taken++] = (uint32_t)want;
          } else {
            ps.taken[ps.n_taken++] = (uint32_t)want | OE_KEPT;
          }
          break;
        }
        if (reserved == op.amount) break;
        const double diff = __dsub_rn(op.amount, reserved);
        if (diff > 0.0) {
          const int q = op.slot * o.K + op.kind;
          const double avail = ps.val[q];
          pass = ((ps.present >> q) & 1u) && !(__dsub_rn(diff, avail) > 0.0);
          if (pass) ps.val[q] = __dsub_rn(avail, diff);
        }
      } break;
      case OE_OP_FREE: {
        if (op.kind == OE_KIND_PORTS) {  // aux = the port
          if (ps.n_freed_ports < OE_DEV_MAX_FREED) ps.freed_port[ps.n_freed_ports++] = ((uint32_t)op.slot << 16) | (uint32_t)op.aux;
          break;
        }
        if (op.kind == OE_KIND_MOUNT) {  // amount = size, aux = profile id + 1
          if (ps.n_freed_mounts < OE_DEV_MAX_FREED) {
            ps.freed_mount_size[ps.n_freed_mounts] = op.amount;
            ps.freed_mount_prof[ps.n_freed_mounts++] = (int32_t)op.aux - 1;
          }
          break;
        }
        const int q = op.slot * o.K + op.kind;
        const double cur = ((ps.present >> q) & 1u) ? ps.val[q] : 0.0;
        ps.val[q] = __dadd_rn(cur, op.amount);
        ps.present |= 1u << q;
      } break;
      default: pass = false;
    }
    if (!pass) {  // failed stages never deduct and the verdict is the AND of all stages
      all = false;
      if (!failed) break;
      *failed |= 1ull << min(x - ob, 63);
    }
  }
  if (RECORD) dec->n_new_reservations = n_res;
  return all;
}

// ---------------------------------------------------------------------------------------------------------------------
// Folded form of a new-pod pipeline (OfferEvaluator.java:411-522): what oe_load_programs derives from a class whose ops
// are only NOP / EXEC_ID(new) / one RULE / SCALAR_NEW / PORT_STATIC / unconstrained PORT_DYNAMIC on role slot 0.
//   * Scalars.  The stages of one (slot, kind) pool form a chain  d = a_k - v; fail if d > 0; v = v - a_k  (ValueUtils
//     subtract + compare, MesosResourcePool.java:330-372).  Every step is monotone in the starting value, so the set of
//     starting values that pass the whole chain is {v >= thr}; the host finds thr by bisection over the doubles while
//     running the literal chain, hence `!(v < thr)` is exact, rounding included (NaN passes, as it does in the chain).
//   * Ports.  `port` lists the distinct ports that are requested statically or excluded from dynamic selection
//     (PortEvaluationStage.java:252-261).  All static ones must be offered; the dynamic stages succeed iff the offered
//     ports that are not listed number at least n_dyn (each takes the lowest one left, :276-294).
// Anything else (existing-pod ops, MOUNT volumes, constrained dynamic ports, other role slots) runs through eval_pair.
struct alignas(16) FastRow {
  double thr[8];
  uint16_t port[16];
  uint8_t thr_q[8];       // pool index slot * K + kind of each threshold
  uint32_t flags;         // OE_FR_*
  int32_t rule_begin;
  uint16_t rule_count, cls;   // cls: the class itself (a tile's descriptors are copies, one per factored row)
  uint32_t static_mask;   // bit u: port[u] is requested statically
  uint8_t n_thr, n_ports, n_dyn, has_mount;
  int32_t pod_role;
  double mount_amount;    // has_mount: the one MOUNT volume of the pod (consumeAtomic, MesosResourcePool.java:282-328) ...
  uint64_t mount_mask;    // ... and its allowed-profile mask
};
#define OE_FR_U4 9        // a FastRow is staged as nine uint4
static_assert(sizeof(FastRow) == 16 * OE_FR_U4, "FastRow staging");
#define OE_FR_FAST 1u        // the folded form is valid for this class
#define OE_FR_NEVER 2u       // a static port is requested twice: the second stage can never pass
#define OE_FR_NEEDS_POOL 4u  // has scalar or port stages: an offer allocated to another role fails
#define OE_FR_DENSE 16u      // S*K <= 8: thr[] is indexed by pool entry (-inf where the class asks nothing) and the entries
                             // that must be present are the mask in thr_q[0..3]; K1 then tests register copies
#define OE_FR_PORTS_BY_TABLE 32u   // the port stages are only known as a verdict per distinct range list (k_port_tables): a
                                  // range-constrained dynamic port; such a class is folded only while the table has its lists
#define OE_FR_RULE_TAB 8u    // every leaf of the rule reads only the offer's zone / region: the rule is tabulated per
                             // (factored row, zone, region) by k_rule_tables and K1 looks the verdict up

#define OE_FAST_RANGES 8

// Offer row state a thread keeps across all the classes it evaluates.
struct OfferRegs {
  OfferKeys ok;
  int32_t alloc;
  uint32_t present;
  uint32_t rb[OE_FAST_RANGES], rlen[OE_FAST_RANGES];  // slot-0 port ranges: begin and (end - begin); empty = never matches
  int32_t pool_ports;                                 // offered ports of slot 0
  double sv[8];                                       // the pool entries, when S*K <= 8
};

__device__ __forceinline__ void load_offer_regs(const DevOffers& o, int64_t i, bool reg_ranges, OfferRegs& r) {
  r.ok.agent = __ldg(o.agent + i);
  r.ok.host = __ldg(o.host + i);
  r.ok.zone = __ldg(o.zone + i);
  r.ok.region = __ldg(o.region + i);
  r.alloc = __ldg(o.alloc + i);
  r.present = __ldg(o.present + i);
  r.pool_ports = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) r.sv[q] = (o.S * o.K <= 8 && q < o.S * o.K) ? __ldg(o.scalar + (int64_t)q * o.N + i) : 0.0;
  if (reg_ranges) {
#pragma unroll
    for (int q = 0; q < OE_FAST_RANGES; ++q) {
      uint32_t pr = q < o.R ? __ldg(o.port + (int64_t)q * o.N + i) : OE_PORT_EMPTY;
      const uint32_t b = pr & 0xFFFFu, e = pr >> 16;
      const bool some = b <= e;
      r.rb[q] = some ? b : 0x100000u;
      r.rlen[q] = some ? e - b : 0u;
      r.pool_ports += some ? (int32_t)(e - b + 1) : 0;
    }
  }
}

__device__ __forceinline__ bool port_member(const DevOffers& o, int64_t i, bool reg_ranges, const OfferRegs& r, uint32_t p) {
  bool m = false;
  if (reg_ranges) {
#pragma unroll
    for (int q = 0; q < OE_FAST_RANGES; ++q) m = m || (p - r.rb[q]) <= r.rlen[q];
  } else {
    m = port_in_pool(o, i, 0, o.R, (int)p);
  }
  return m;
}

// `bit` if port p lies in one of the eight register-held ranges, else 0.  Written as PTX so that the eight tests stay one
// subtract and one predicate-accumulating compare each (the compiler's own choice for the C++ form was a chain of selects).
__device__ __forceinline__ uint32_t port_member_bit(const OfferRegs& r, uint32_t p, uint32_t bit) {
  static_assert(OE_FAST_RANGES == 8, "port_member_bit tests eight ranges");
  uint32_t out;
  asm("{\n\t.reg .pred q;\n\t.reg .u32 t;\n\t"
      "sub.u32 t, %1, %2;\n\tsetp.le.u32 q, t, %3;\n\t"
      "sub.u32 t, %1, %4;\n\tsetp.le.or.u32 q, t, %5, q;\n\t"
      "sub.u32 t, %1, %6;\n\tsetp.le.or.u32 q, t, %7, q;\n\t"
      "sub.u32 t, %1, %8;\n\tsetp.le.or.u32 q, t, %9, q;\n\t"
      "sub.u32 t, %1, %10;\n\tsetp.le.or.u32 q, t, %11, q;\n\t"
      "sub.u32 t, %1, %12;\n\tsetp.le.or.u32 q, t, %13, q;\n\t"
      "sub.u32 t, %1, %14;\n\tsetp.le.or.u32 q, t, %15, q;\n\t"
      "sub.u32 t, %1, %16;\n\tsetp.le.or.u32 q, t, %17, q;\n\t"
      "selp.u32 %0, %18, 0, q;\n\t}"
      : "=r"(out)
      : "r"(p), "r"(r.rb[0]), "r"(r.rlen[0]), "r"(r.rb[1]), "r"(r.rlen[1]), "r"(r.rb[2]), "r"(r.rlen[2]), "r"(r.rb[3]), "r"(r.rlen[3]),
        "r"(r.rb[4]), "r"(r.rlen[4]), "r"(r.rb[5]), "r"(r.rlen[5]), "r"(r.rb[6]), "r"(r.rlen[6]), "r"(r.rb[7]), "r"(r.rlen[7]), "r"(bit));
  return out;
}

// `sc` = this CTA's shared-memory copy of the scalar columns, [S*K][blockDim.x], or nullptr (read through L1 instead).
// Rule verdicts tabulated over (zone, region): bit (region + 1) * zdim + (zone + 1) of row `row` (absent = -1 -> 0).
struct RuleTab {
  const uint32_t* bits;  // [rows][words]
  int words, zdim;
};

// Port verdicts tabulated over the table's distinct `ports` range lists (oe_load_offers_packed): bit t of row `cls` = the
// class's port stages pass on list t.  Built by k_port_tables, read by K1 instead of testing ranges per offer.
struct PortTab {
  const uint32_t* bits;  // [C][words], or null
  int words;
};

// The port stages of a folded class over one range list given as eight (begin, length) pairs: every statically requested port
// offered, and enough other ports left for the dynamic ones.
__device__ __forceinline__ bool fast_ports_pass(const FastRow& fr, const uint32_t* rb, const uint32_t* rlen, int pool_ports) {
  uint32_t member = 0u;
  for (int u = 0; u < fr.n_ports; ++u) {
    bool m = false;
#pragma unroll
    for (int q = 0; q < OE_FAST_RANGES; ++q) m = m || ((uint32_t)fr.port[u] - rb[q]) <= rlen[q];
    member |= m ? 1u << u : 0u;
  }
  return (member & fr.static_mask) == fr.static_mask && pool_ports - __popc(member) >= (int)fr.n_dyn;
}

// NQ = pool entries kept in registers and tested by the dense form (4 when S*K <= 4, else 8).  tab_word / tab_bit: where the
// tabulated rule verdict of this (row, zone, region) sits — the caller advances the pointer from row to row.
template <int PORTS, int NQ>
__device__ __forceinline__ bool eval_fast(const DevCtx& cx, const oe_rule_node* nodes, const FastRow& fr, const double* sc, int64_t i,
                                          const OfferRegs& r, int var, const uint32_t* tab_word, int tab_bit, const PortTab& ptab, int32_t ptuple) {
  constexpr bool REG_RANGES = PORTS == 1;
  const DevOffers& o = cx.o;
  // every test is evaluated and AND-ed (no short circuit): a branch per test costs more than the test, and a compare that feeds
  // a branch cannot overlap with the next one
  bool pass = !(fr.flags & OE_FR_NEVER);
  uint32_t present = r.present;
  if (fr.flags & OE_FR_NEEDS_POOL) {
    // consumableResource (MesosResourcePool.java:87-97)
    pass &= !((fr.pod_role >= 0) & (r.alloc >= 0) & (r.alloc != fr.pod_role));
  }
  if (fr.flags & OE_FR_DENSE) {
    const uint32_t need = *reinterpret_cast<const uint32_t*>(fr.thr_q);
    pass &= (present & need) == need;
#pragma unroll
    for (int q = 0; q < NQ; ++q) pass &= !(r.sv[q] < fr.thr[q]);
  } else {
    for (int t = 0; t < fr.n_thr; ++t) {
      const int q = fr.thr_q[t];
      const double v = sc ? sc[q * blockDim.x + threadIdx.x] : __ldg(o.scalar + (int64_t)q * o.N + i);
      pass = pass && ((present >> q) & 1u) && !(v < fr.thr[t]);
    }
  }
  if (PORTS == 2) {
    // one lookup: the class's port stages were evaluated once per distinct range list (k_port_tables)
    if (fr.n_ports | fr.n_dyn) pass &= (bool)((__ldg(ptab.bits + (int64_t)fr.cls * ptab.words + (ptuple >> 5)) >> (ptuple & 31)) & 1u);
  } else if (fr.n_ports | fr.n_dyn) {
    int listed = 0;
    bool statics = true;
    int pool_ports = r.pool_ports;
    if (!REG_RANGES) {
      pool_ports = 0;
      for (int q = 0; q < o.R; ++q) {
        const uint32_t pr = __ldg(o.port + (int64_t)q * o.N + i);
        const int b = pr & 0xFFFF, e = pr >> 16;
        pool_ports += b <= e ? e - b + 1 : 0;
      }
    }
    uint32_t member = 0u, bit = 1u;   // bit u: listed port u is offered
    for (int u = 0; u < fr.n_ports; ++u, bit <<= 1) {
      if (REG_RANGES) member |= port_member_bit(r, fr.port[u], bit);
      else if (port_member(o, i, false, r, fr.port[u])) member |= bit;
    }
    listed = __popc(member);
    statics = (member & fr.static_mask) == fr.static_mask;
    pass &= statics & (pool_ports - listed >= (int)fr.n_dyn);
  }
  if (fr.has_mount) {
    // one MOUNT volume: any unreserved MOUNT disk of the offer that is large enough and whose profile the volume allows
    // (MesosResourcePool.java:282-328; with a single such stage "the first one that fits" exists iff any fits)
    bool any = false;
    for (int d = 0; d < o.D; ++d) {
      const double size = __ldg(o.msize + (int64_t)d * o.N + i);
      const int32_t prof = __ldg(o.mprof + (int64_t)d * o.N + i);
      const bool prof_ok = prof >= 0 ? ((fr.mount_mask >> prof) & 1ull) != 0 : fr.mount_mask == 0;
      any |= !(__dsub_rn(fr.mount_amount, size) > 0.0) & prof_ok;
    }
    pass &= any;
  }
  if (fr.flags & OE_FR_RULE_TAB) {
    pass &= (bool)((__ldg(tab_word) >> tab_bit) & 1u);
  } else if (fr.rule_count > 0) {
    pass = pass && eval_rule(cx, nodes, i, r.ok, var, fr.rule_begin, fr.rule_count);
  }
  return pass;
}

}  // namespace oe
