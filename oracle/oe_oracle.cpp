// oe_oracle.cpp — CPU restatement of the reference offer-evaluation path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the parity oracle and the timed CPU baseline.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load it; nothing under dcos_commons_b200/ does.
//
// It consumes the same encoded tables as the engine (include/oe_abi.h) through an identically shaped C ABI (prefix
// oeo_), but evaluates them the way the reference does: one requirement at a time, offers in list order, a fresh
// resource pool per (requirement, offer), every stage run without short-circuit, and every placement rule re-walking
// ALL deployed tasks for every offer.  Nothing is factored by class and no tallies are cached.
//
// Second mode (oeo_set_mode(h, 1, threads)), for inputs the literal walk cannot finish (C5: 10^11 pairs x 10^5 tasks per
// rule): the SAME pair-at-a-time pipeline, but the three rule families that re-walk the task table read tallies built once
// per cycle instead — per-key counts per (key space, task filter), tasks per (pod type, agent) — with the requirement's own
// "equivalent" tasks (S/offer/TaskUtils.java:136-145) subtracted, which is the arithmetic of P/MaxPerRule.java:58-87,
// P/AbstractRoundRobinRule.java:70-112 and P/TaskTypeRule.java:98-217 with the loop over tasks hoisted out of the loop over
// offers.  The cycle keeps one cursor per (class, equivalent-task set): within a cycle the verdict of a pair is a pure
// function of those, and a taken offer never returns (PlanScheduler.java:61-66), so every offer before the cursor stays
// unusable for the requirements that share it.  tests/test_oracle_modes.py pins this mode to the literal one (matrix, masks
// and decisions, byte for byte) on every fuzz table and on config C4; it is NOT how the reference evaluates and is never
// the timed "reference" baseline.
//
// Reference files followed (S/ = sdk/scheduler/src/main/java/com/mesosphere/sdk/, P/ = S/offer/evaluate/placement/):
//   S/scheduler/plan/PlanScheduler.java:50-70            cycle(): steps in order, used offers removed whole
//   S/offer/evaluate/OfferEvaluator.java:113-248         first_fit(): offers in order, all stages, first all-pass wins
//   S/offer/MesosResourcePool.java:53-147,245-455        Pool
//   S/offer/ValueUtils.java:64-67,103-106,144-155        fp64 add / subtract / compare, same operation order
//   S/offer/RangeUtils.java:23-47 (+ antlr4 IntervalSet 4.5.1-1, not vendored: sorted, disjoint, coalescing set)
//   S/offer/evaluate/OfferEvaluationUtils.java:60-227    evaluate_simple_scalar()
//   S/offer/evaluate/PortEvaluationStage.java:72-155,242-298
//   S/offer/evaluate/VolumeEvaluationStage.java:119-263
//   S/offer/evaluate/ExecutorEvaluationStage.java:32-81
//   S/offer/evaluate/UnreserveEvaluationStage.java:24-36
//   P/*.java                                             rule_node()
//   S/offer/TaskUtils.java:136-145                       equivalent()
//   S/offer/evaluate/LaunchEvaluationStage.java:58-68, S/state/PersistentLaunchRecorder.java:74-137   apply_launches()
//
// Parity pin: tests/kat/*.cpp (run by tests/test_kat.py) restate the reference's JUnit vectors against this oracle.  The JVM
// reference itself cannot be built or run in this environment (no JDK, no Maven cache).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <string>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../include/oe_abi.h"

namespace {

template <class T>
std::vector<T> copy_col(const T* p, size_t n) {
  return p ? std::vector<T>(p, p + n) : std::vector<T>();
}

struct Programs {
  int C = 0;
  std::vector<int32_t> op_begin;
  std::vector<oe_stage_op> ops;
  std::vector<oe_rule_node> nodes;
  std::vector<int32_t> pod_role;
  std::vector<int32_t> port_excl_begin;
  std::vector<uint16_t> port_excl;
  std::vector<int32_t> range_set_begin;
  std::vector<uint32_t> range_set;
  std::vector<oe_counter_set> csets;
  std::vector<int32_t> type_set_pod_type;
  std::vector<uint32_t> matcher_bits;
  oe_dict_sizes dict{};
};

struct Tasks {
  int64_t T = 0;
  int TA = 0, Q = 0, FW = 0;
  std::vector<int32_t> name_id, pod_type_id, pod_index, agent_id, hostname_id, zone_id, region_id;
  std::vector<int32_t> attr_id;      // [TA][T]  → stored row-major per task here: [T][TA]
  std::vector<int32_t> rr_key;       // [T][Q]
  std::vector<uint32_t> filter_bits; // [T][FW]
};

struct Offers {
  int64_t N = 0, row_offset = 0, N_global = 0;
  int S = 0, K = 0, R = 0, D = 0, A = 0, Q = 0, E = 0;
  std::vector<double> scalar;          // [S][K][N]
  std::vector<uint32_t> scalar_present;
  std::vector<uint32_t> port_range;    // [S][R][N]
  std::vector<double> mount_size;      // [D][N]
  std::vector<int32_t> mount_profile;
  std::vector<int32_t> alloc_role, agent_id, hostname_id, zone_id, region_id;
  std::vector<int32_t> attr_id, rr_key, exec_id;
  std::vector<int64_t> reserved_begin;
  std::vector<int32_t> reserved_id, reserved_kind;
  std::vector<double> reserved_value;
};

struct Reqs {
  int64_t M = 0;
  std::vector<int32_t> class_id, pod_type_id, pod_index;
  std::vector<int32_t> launch_begin, launch_name_id;
  std::vector<uint32_t> launch_filter_bits;
};

// ---- S/offer/RangeUtils.java + antlr IntervalSet: a sorted, disjoint, adjacency-coalescing set of closed intervals.
struct IntervalSet {
  std::vector<std::pair<int, int>> iv;
  void add(int a, int b) {  // IntervalSet.add(a, b)
    if (b < a) return;
    std::vector<std::pair<int, int>> out;
    bool placed = false;
    for (auto& r : iv) {
      if (r.second < a - 1) {
        out.push_back(r);
      } else if (r.first > b + 1) {
        if (!placed) { out.push_back({a, b}); placed = true; }
        out.push_back(r);
      } else {  // overlapping or adjacent: union
        a = std::min(a, r.first);
        b = std::max(b, r.second);
      }
    }
    if (!placed) out.push_back({a, b});
    iv.swap(out);
  }
  bool contains(int p) const {
    for (auto& r : iv) if (r.first <= p && p <= r.second) return true;
    return false;
  }
  // IntervalSet.subtract(this, other)
  IntervalSet minus(const IntervalSet& o) const {
    IntervalSet res;
    for (auto r : iv) {
      int a = r.first;
      bool alive = true;
      for (auto& s : o.iv) {
        if (s.second < a) continue;
        if (s.first > r.second) break;
        if (s.first > a) res.iv.push_back({a, s.first - 1});
        a = s.second + 1;
        if (a > r.second) { alive = false; break; }
      }
      if (alive && a <= r.second) res.iv.push_back({a, r.second});
    }
    return res;
  }
  bool empty() const { return iv.empty(); }
};

// ---- S/offer/ValueUtils.java:144-155
inline int compare_scalar(double v1, double v2) { return v1 < v2 ? -1 : (v1 > v2 ? 1 : 0); }

struct Mount { double size; int32_t profile; int slot; };
struct Reserved { int32_t kind; double value; };

// ---- S/offer/MesosResourcePool.java
struct Pool {
  int S, K;
  bool role_present[OE_MAX_ROLE_SLOTS];                       // reservableMergedPoolByRole.get(role) != null
  bool present[OE_MAX_ROLE_SLOTS][OE_MAX_SCALAR_KINDS];        // pool.get(name) != null
  double val[OE_MAX_ROLE_SLOTS][OE_MAX_SCALAR_KINDS];
  bool ports_present[OE_MAX_ROLE_SLOTS];
  IntervalSet ports[OE_MAX_ROLE_SLOTS];
  std::vector<Mount> mounts;                                    // unreservedAtomicPool["disk"]
  int n_freed_mounts = 0;                                       // disks handed back by OE_OP_FREE get slots OE_MAX_MOUNT_SLOTS + k
  std::map<int32_t, Reserved> reserved;                         // dynamicallyReservedPoolByResourceId

  // MesosResourcePool.java:61-70 on scalars
  static bool sufficient_scalar(double desired, bool have, double available) {
    if (!have) return false;
    double difference = desired - available;
    return compare_scalar(difference, 0.0) <= 0;
  }
  // :330-372
  bool consume_reservable_merged_scalar(int slot, int kind, double desired) {
    if (!role_present[slot]) return false;
    if (!sufficient_scalar(desired, present[slot][kind], val[slot][kind])) return false;
    val[slot][kind] = val[slot][kind] - desired;
    return true;
  }
  bool consume_reservable_merged_port(int slot, int port) {
    if (!role_present[slot] || !ports_present[slot]) return false;
    IntervalSet desired;
    desired.add(port, port);
    // sufficientValue: subtract(desired, available) must compare <= 0 against the empty range list, i.e. be empty
    if (!desired.minus(ports[slot]).empty()) return false;
    ports[slot] = ports[slot].minus(desired);
    return true;
  }
  // :282-328
  int consume_atomic(double desired, uint64_t profile_mask) {
    std::vector<Mount> filtered;
    int chosen = -1;
    for (auto& m : mounts) {
      bool profile_ok = m.profile >= 0 ? ((profile_mask >> m.profile) & 1) != 0 : profile_mask == 0;
      if (chosen < 0 && sufficient_scalar(desired, true, m.size) && profile_ok) {
        chosen = m.slot;  // do NOT break: the filtered list keeps every other disk
      } else {
        filtered.push_back(m);
      }
    }
    mounts.swap(filtered);
    return chosen;
  }
};

// Tallies of the cached mode (see the file header): what the literal rule functions recompute per pair.
struct TallyCache {
  bool valid = false;
  std::vector<std::unordered_map<int32_t, int>> cnt;   // [counter set] key -> tasks passing the set's filter that carry the key
  std::vector<std::map<int, int>> hist;                // [counter set] tally value -> number of keys with that tally (> 0)
  std::vector<int64_t> type_total;                     // [type set] tasks of that pod type
  std::vector<std::unordered_map<int32_t, int>> type_agent;  // [type set] agent -> tasks of that pod type on it
  std::vector<std::vector<int64_t>> self;              // [requirement] task rows equivalent to it, ascending
};

struct Oracle {
  Programs P;
  Tasks T;
  Offers O;
  Reqs R;
  bool hasP = false, hasT = false, hasO = false, hasR = false;
  int mode = 0, threads = 1;   // mode 1 = tallies cached per cycle
  TallyCache cache;
  std::vector<oe_decision> last_decisions;
  std::string err;
  int64_t pairs_evaluated = 0;

  int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err = buf;
    return code;
  }

  // ---- new MesosResourcePool(offer, role)  (MesosResourcePool.java:53-59, 76-147, 432-455)
  Pool make_pool(int64_t i, int cls) const {
    Pool p;
    p.S = O.S; p.K = O.K;
    // consumableResource (:87-97): resources allocated to a different role than the pod's are dropped
    bool consumable = true;
    int32_t pod_role = P.pod_role[cls];
    if (pod_role >= 0 && O.alloc_role[i] >= 0 && O.alloc_role[i] != pod_role) consumable = false;
    for (int s = 0; s < O.S; ++s) {
      p.role_present[s] = false;
      p.ports_present[s] = false;
      for (int k = 0; k < O.K; ++k) {
        bool have = consumable && ((O.scalar_present[i] >> (s * O.K + k)) & 1);
        p.present[s][k] = have;
        p.val[s][k] = have ? O.scalar[((size_t)s * O.K + k) * O.N + i] : 0.0;
        if (have) p.role_present[s] = true;
      }
      if (consumable) {
        for (int r = 0; r < O.R; ++r) {
          uint32_t pr = O.port_range[((size_t)s * O.R + r) * O.N + i];
          int b = pr & 0xFFFF, e = pr >> 16;
          if (b <= e) { p.ports[s].add(b, e); p.ports_present[s] = true; p.role_present[s] = true; }
        }
      }
    }
    if (consumable) {
      for (int d = 0; d < O.D; ++d) {
        double sz = O.mount_size[(size_t)d * O.N + i];
        if (sz == -std::numeric_limits<double>::infinity()) continue;
        p.mounts.push_back({sz, O.mount_profile[(size_t)d * O.N + i], d});
      }
      if (!O.reserved_begin.empty()) {
        for (int64_t q = O.reserved_begin[i]; q < O.reserved_begin[i + 1]; ++q)
          p.reserved[O.reserved_id[q]] = {O.reserved_kind[q], O.reserved_value[q]};
      }
    }
    return p;
  }

  bool offer_has_executor(int64_t i, int32_t id) const {
    for (int e = 0; e < O.E; ++e) if (O.exec_id[(size_t)e * O.N + i] == id) return true;
    return false;
  }

  bool matcher_bit(int32_t word_off, int32_t v) const {
    if (v < 0) return false;
    return (P.matcher_bits[(size_t)word_off + (v >> 5)] >> (v & 31)) & 1;
  }
  bool task_passes_filter(int64_t t, int f) const { return (T.filter_bits[(size_t)t * T.FW + (f >> 5)] >> (f & 31)) & 1; }
  // S/offer/TaskUtils.java:136-145
  bool equivalent(int64_t t, int64_t j) const {
    return T.pod_type_id[t] >= 0 && T.pod_index[t] != OE_NO_INDEX && T.pod_index[t] == R.pod_index[j] &&
           T.pod_type_id[t] == R.pod_type_id[j];
  }

  // keys of an offer / a task in a counter set's key space
  void offer_keys(const oe_counter_set& cs, int64_t i, int32_t matcher_off, std::vector<int32_t>& keys) const {
    keys.clear();
    switch (cs.keyspace) {
      case OE_KS_HOSTNAME: keys.push_back(O.hostname_id[i]); break;
      case OE_KS_ZONE: if (O.zone_id[i] >= 0) keys.push_back(O.zone_id[i]); break;
      case OE_KS_REGION: if (O.region_id[i] >= 0) keys.push_back(O.region_id[i]); break;
      case OE_KS_ATTR:
        for (int a = 0; a < O.A; ++a) {
          int32_t v = O.attr_id[(size_t)a * O.N + i];
          if (v >= 0 && matcher_bit(matcher_off, v)) keys.push_back(v);
        }
        break;
      case OE_KS_RR_ATTR: { int32_t v = O.rr_key[(size_t)cs.rr_col * O.N + i]; if (v >= 0) keys.push_back(v); } break;
    }
  }
  void task_keys(const oe_counter_set& cs, int64_t t, int32_t matcher_off, std::vector<int32_t>& keys) const {
    keys.clear();
    switch (cs.keyspace) {
      case OE_KS_HOSTNAME: if (T.hostname_id[t] >= 0) keys.push_back(T.hostname_id[t]); break;
      case OE_KS_ZONE: if (T.zone_id[t] >= 0) keys.push_back(T.zone_id[t]); break;
      case OE_KS_REGION: if (T.region_id[t] >= 0) keys.push_back(T.region_id[t]); break;
      case OE_KS_ATTR:
        for (int a = 0; a < T.TA; ++a) {
          int32_t v = T.attr_id[(size_t)t * T.TA + a];
          if (v >= 0 && matcher_bit(matcher_off, v)) keys.push_back(v);
        }
        break;
      case OE_KS_RR_ATTR: { int32_t v = T.rr_key[(size_t)t * T.Q + cs.rr_col]; if (v >= 0) keys.push_back(v); } break;
    }
  }

  // P/MaxPerRule.java:52-87
  bool max_per(const oe_counter_set& cs, int32_t max, int32_t matcher_off, int64_t i, int64_t j) const {
    std::vector<int32_t> okeys, tkeys;
    offer_keys(cs, i, matcher_off, okeys);
    std::map<int32_t, int> counts;
    for (int32_t k : okeys) counts[k] += 1;
    for (int64_t t = 0; t < T.T; ++t) {
      if (!task_passes_filter(t, cs.filter_id)) continue;
      if (equivalent(t, j)) continue;
      task_keys(cs, t, matcher_off, tkeys);
      for (int32_t k : tkeys)
        if (std::find(okeys.begin(), okeys.end(), k) != okeys.end()) counts[k] += 1;
    }
    for (auto& kv : counts) if (kv.second > max) return false;
    return true;
  }

  // P/AbstractRoundRobinRule.java:53-151
  bool round_robin(const oe_counter_set& cs, int32_t distinct, int64_t i, int64_t j) const {
    std::vector<int32_t> okeys, tkeys;
    offer_keys(cs, i, 0, okeys);
    if (okeys.empty()) return false;  // offerKey == null
    int32_t offer_key = okeys[0];
    std::map<int32_t, int> counts;
    for (int64_t t = 0; t < T.T; ++t) {
      if (!task_passes_filter(t, cs.filter_id)) continue;
      if (equivalent(t, j)) continue;
      task_keys(cs, t, 0, tkeys);
      if (tkeys.empty()) continue;
      counts[tkeys[0]] += 1;
    }
    int max_known = 0, min_known = std::numeric_limits<int>::max();
    for (auto& kv : counts) {
      if (kv.second > max_known) max_known = kv.second;
      if (kv.second < min_known) min_known = kv.second;
    }
    if (min_known == std::numeric_limits<int>::max()) min_known = 0;
    auto it = counts.find(offer_key);
    int offer_count = it == counts.end() ? 0 : it->second;
    if (min_known == max_known || offer_count <= min_known) {
      if (distinct < 0) return true;
      if ((int)counts.size() >= distinct) return true;
      return offer_count == 0;
    }
    return false;
  }

  // P/TaskTypeRule.java:98-217
  bool task_type(int32_t pod_type, bool avoid, int64_t i, int64_t j) const {
    std::vector<int64_t> matching;
    for (int64_t t = 0; t < T.T; ++t) if (T.pod_type_id[t] == pod_type) matching.push_back(t);
    if (matching.empty()) return true;
    for (int64_t t : matching) {
      if (equivalent(t, j)) continue;
      if (T.agent_id[t] == O.agent_id[i]) return avoid ? false : true;
    }
    return avoid ? true : false;
  }


  // ---- cached mode ---------------------------------------------------------------------------------------------------
  void build_cache() {
    TallyCache& c = cache;
    c = TallyCache();
    c.cnt.resize(P.csets.size());
    c.hist.resize(P.csets.size());
    std::vector<int32_t> keys;
    for (size_t s = 0; s < P.csets.size(); ++s) {
      const oe_counter_set& cs = P.csets[s];
      for (int64_t t = 0; t < T.T; ++t) {
        if (!task_passes_filter(t, cs.filter_id)) continue;
        if (cs.keyspace == OE_KS_ATTR) {  // every attribute string of the task; the matcher is applied on the offer side
          for (int a = 0; a < T.TA; ++a) { int32_t v = T.attr_id[(size_t)t * T.TA + a]; if (v >= 0) c.cnt[s][v] += 1; }
        } else {
          task_keys(cs, t, 0, keys);
          if (!keys.empty()) c.cnt[s][keys[0]] += 1;
        }
      }
      for (auto& kv : c.cnt[s]) c.hist[s][kv.second] += 1;
    }
    c.type_total.assign(P.type_set_pod_type.size(), 0);
    c.type_agent.resize(P.type_set_pod_type.size());
    for (size_t s = 0; s < P.type_set_pod_type.size(); ++s)
      for (int64_t t = 0; t < T.T; ++t)
        if (T.pod_type_id[t] == P.type_set_pod_type[s]) { c.type_total[s] += 1; c.type_agent[s][T.agent_id[t]] += 1; }
    // equivalent tasks of every requirement (TaskUtils.areEquivalent: same pod type and index)
    std::map<std::pair<int32_t, int32_t>, std::vector<int64_t>> by_pod;
    for (int64_t t = 0; t < T.T; ++t)
      if (T.pod_type_id[t] >= 0 && T.pod_index[t] != OE_NO_INDEX) by_pod[{T.pod_type_id[t], T.pod_index[t]}].push_back(t);
    c.self.assign((size_t)R.M, {});
    for (int64_t j = 0; j < R.M; ++j) {
      auto it = by_pod.find({R.pod_type_id[j], R.pod_index[j]});
      if (it != by_pod.end()) c.self[(size_t)j] = it->second;
    }
    c.valid = true;
  }
  // how often key k occurs among the keys of requirement j's equivalent tasks that pass the set's filter
  int self_key_count(const oe_counter_set& cs, int32_t matcher_off, int64_t j, int32_t k) const {
    int n = 0;
    std::vector<int32_t> tkeys;
    for (int64_t t : cache.self[(size_t)j]) {
      if (!task_passes_filter(t, cs.filter_id)) continue;
      task_keys(cs, t, matcher_off, tkeys);
      if (cs.keyspace == OE_KS_ATTR) { for (int32_t q : tkeys) n += q == k; }
      else if (!tkeys.empty()) n += tkeys[0] == k;
    }
    return n;
  }
  // P/MaxPerRule.java:52-87 with the tallies of :58-76 read from the cache
  bool max_per_cached(int set, int32_t max, int32_t matcher_off, int64_t i, int64_t j) const {
    const oe_counter_set& cs = P.csets[set];
    std::vector<int32_t> okeys;
    offer_keys(cs, i, matcher_off, okeys);
    std::map<int32_t, int> counts;
    for (int32_t k : okeys) counts[k] += 1;
    for (auto& kv : counts) {
      auto it = cache.cnt[set].find(kv.first);
      int tally = it == cache.cnt[set].end() ? 0 : it->second;
      if (tally > 0 && !cache.self[(size_t)j].empty()) tally -= self_key_count(cs, matcher_off, j, kv.first);
      if (kv.second + tally > max) return false;
    }
    return true;
  }
  // P/AbstractRoundRobinRule.java:53-151 with `counts` of :70-89 read from the cache
  bool round_robin_cached(int set, int32_t distinct, int64_t i, int64_t j) const {
    const oe_counter_set& cs = P.csets[set];
    if (cs.keyspace == OE_KS_ATTR) return round_robin(cs, distinct, i, j);  // not a single-key space: no tally kept
    std::vector<int32_t> okeys;
    offer_keys(cs, i, 0, okeys);
    if (okeys.empty()) return false;
    const int32_t offer_key = okeys[0];
    const std::unordered_map<int32_t, int>& cnt = cache.cnt[set];
    auto oc = cnt.find(offer_key);
    int offer_count = oc == cnt.end() ? 0 : oc->second;
    int max_known = 0, min_known = 0, n_keys = 0;
    if (cache.self[(size_t)j].empty()) {
      const std::map<int, int>& h = cache.hist[set];
      if (!h.empty()) { min_known = h.begin()->first; max_known = h.rbegin()->first; }
      n_keys = (int)cnt.size();
    } else {
      // the requirement's own tasks leave the tallies: adjust the histogram of tally values for the keys they touch
      std::map<int32_t, int> minus;
      std::vector<int32_t> tkeys;
      for (int64_t t : cache.self[(size_t)j]) {
        if (!task_passes_filter(t, cs.filter_id)) continue;
        task_keys(cs, t, 0, tkeys);
        if (!tkeys.empty()) minus[tkeys[0]] += 1;
      }
      std::map<int, int> h = cache.hist[set];
      n_keys = (int)cnt.size();
      for (auto& kv : minus) {
        const int before = cnt.at(kv.first), after = before - kv.second;
        if (--h[before] == 0) h.erase(before);
        if (after > 0) h[after] += 1; else --n_keys;
        if (kv.first == offer_key) offer_count = after;
      }
      if (!h.empty()) { min_known = h.begin()->first; max_known = h.rbegin()->first; }
    }
    if (min_known == max_known || offer_count <= min_known) {
      if (distinct < 0) return true;
      if (n_keys >= distinct) return true;
      return offer_count == 0;
    }
    return false;
  }
  // P/TaskTypeRule.java:98-217 with the matching tasks counted per agent once
  bool task_type_cached(int set, bool avoid, int64_t i, int64_t j) const {
    if (cache.type_total[set] == 0) return true;
    auto it = cache.type_agent[set].find(O.agent_id[i]);
    int on = it == cache.type_agent[set].end() ? 0 : it->second;
    for (int64_t t : cache.self[(size_t)j])
      on -= T.pod_type_id[t] == P.type_set_pod_type[set] && T.agent_id[t] == O.agent_id[i];
    return avoid ? on == 0 : on > 0;
  }

  // PlacementRuleEvaluationStage.java:30-42 → PlacementRule.filter; postfix form of the rule tree
  bool rule_program(int32_t begin, int32_t count, int64_t i, int64_t j) const {
    std::vector<char> st;
    for (int32_t n = begin; n < begin + count; ++n) {
      const oe_rule_node& nd = P.nodes[n];
      bool v = false;
      switch (nd.op) {
        case OE_R_TRUE: v = true; break;
        case OE_R_FALSE: v = false; break;
        case OE_R_AND: {  // P/AndRule.java:35-69: every child evaluated; zero children ⇒ fail
          int pass = 0;
          for (int c = 0; c < nd.a; ++c) { pass += st.back() ? 1 : 0; st.pop_back(); }
          v = nd.a > 0 && pass == nd.a;
        } break;
        case OE_R_OR: {  // P/OrRule.java:36-63
          int pass = 0;
          for (int c = 0; c < nd.a; ++c) { pass += st.back() ? 1 : 0; st.pop_back(); }
          v = pass != 0;
        } break;
        case OE_R_NOT: v = !st.back(); st.pop_back(); break;
        case OE_R_AGENT_EQ: v = O.agent_id[i] == nd.a; break;
        case OE_R_ATTR_ANY:  // P/StringMatcherRule.java:48-60 over AttributeRule.getKeys
          for (int a = 0; a < O.A && !v; ++a) v = matcher_bit(nd.a, O.attr_id[(size_t)a * O.N + i]);
          break;
        case OE_R_HOST_IN: v = matcher_bit(nd.a, O.hostname_id[i]); break;
        case OE_R_ZONE_IN: v = O.zone_id[i] >= 0 && matcher_bit(nd.a, O.zone_id[i]); break;
        case OE_R_REGION_IN: v = O.region_id[i] >= 0 && matcher_bit(nd.a, O.region_id[i]); break;
        case OE_R_LOCAL_REGION:  // P/IsLocalRegionRule.java:24-67
          v = O.region_id[i] < 0 || nd.b == 0 || O.region_id[i] == nd.a;
          break;
        case OE_R_MAXPER_HOST: v = mode ? max_per_cached(nd.a, nd.b, 0, i, j) : max_per(P.csets[nd.a], nd.b, 0, i, j); break;
        case OE_R_MAXPER_ZONE: v = O.zone_id[i] >= 0 && (mode ? max_per_cached(nd.a, nd.b, 0, i, j) : max_per(P.csets[nd.a], nd.b, 0, i, j)); break;
        case OE_R_MAXPER_REGION: v = mode ? max_per_cached(nd.a, nd.b, 0, i, j) : max_per(P.csets[nd.a], nd.b, 0, i, j); break;
        case OE_R_MAXPER_ATTR: v = mode ? max_per_cached(nd.a, nd.b, nd.c, i, j) : max_per(P.csets[nd.a], nd.b, nd.c, i, j); break;
        case OE_R_RR: v = mode ? round_robin_cached(nd.a, nd.b, i, j) : round_robin(P.csets[nd.a], nd.b, i, j); break;
        case OE_R_TYPE_AVOID: v = mode ? task_type_cached(nd.a, true, i, j) : task_type(P.type_set_pod_type[nd.a], true, i, j); break;
        case OE_R_TYPE_COLOCATE: v = mode ? task_type_cached(nd.a, false, i, j) : task_type(P.type_set_pod_type[nd.a], false, i, j); break;
        default: v = false;
      }
      st.push_back(v);
    }
    return !st.empty() && st.back();
  }

  // PortEvaluationStage.selectDynamicPort (:242-298)
  int select_dynamic_port(const Pool& pool, int slot, int cls, int32_t range_set, const std::vector<int>& placed) const {
    if (!pool.role_present[slot] || !pool.ports_present[slot]) return -1;
    for (auto& r : pool.ports[slot].iv) {
      for (int p = r.first; p <= r.second; ++p) {
        bool consumed = false;
        for (int32_t q = P.port_excl_begin[cls]; q < P.port_excl_begin[cls + 1] && !consumed; ++q)
          consumed = P.port_excl[q] == p;
        for (int q : placed) consumed = consumed || q == p;
        if (consumed) continue;
        if (range_set >= 0) {
          bool in = false;
          for (int32_t q = P.range_set_begin[range_set]; q < P.range_set_begin[range_set + 1] && !in; ++q) {
            int b = P.range_set[q] & 0xFFFF, e = P.range_set[q] >> 16;
            in = b <= p && p <= e;
          }
          if (!in) continue;
        }
        return p;
      }
    }
    return -1;
  }

  // One (requirement, offer) pass of OfferEvaluator.java:140-184: every stage runs; returns the failed-stage mask.
  // `quick` (cached mode's searches only): stop at the first failed stage — failed stages never deduct, so the verdict
  // "every stage passes" does not depend on the stages behind a failure — and build the pool only when a stage reads it.
  uint64_t evaluate_pair(int64_t i, int64_t j, oe_decision* dec, bool quick = false) const {
    int cls = R.class_id[j];
    if (quick) {   // stages that read nothing but the offer's placement columns come first in every pipeline
      for (int32_t o = P.op_begin[cls]; o < P.op_begin[cls + 1]; ++o) {
        const oe_stage_op& op = P.ops[o];
        if (op.flags) break;
        if (op.opcode == OE_OP_NOP) continue;
        if (op.opcode == OE_OP_EXEC_ID) { if (!(op.arg < 0 || offer_has_executor(i, op.arg))) return 1ull << std::min(o - P.op_begin[cls], 63); continue; }
        if (op.opcode == OE_OP_RULE) { if (!rule_program(op.arg, (int32_t)op.aux, i, j)) return 1ull << std::min(o - P.op_begin[cls], 63); continue; }
        break;
      }
    }
    Pool pool = make_pool(i, cls);
    std::vector<int> placed_ports;  // host ports already put into the pod's builders by earlier stages (:264-269)
    uint64_t failed = 0;
    int n_dyn = 0, n_mount = 0, n_reserve = 0;
    if (dec) {
      for (auto& d : dec->dyn_port) d = 0;
      for (auto& m : dec->mount_slot) m = -1;
    }
    for (int32_t o = P.op_begin[cls]; o < P.op_begin[cls + 1]; ++o) {
      const oe_stage_op& op = P.ops[o];
      bool pass = true;
      if ((op.flags & OE_OPF_SKIP_IF_RUNNING_EXECUTOR) && offer_has_executor(i, (int32_t)op.aux)) {
        continue;  // ResourceEvaluationStage.java:65-84 / VolumeEvaluationStage.java:126-154
      }
      switch (op.opcode) {
        case OE_OP_NOP: break;
        case OE_OP_EXEC_ID: pass = op.arg < 0 || offer_has_executor(i, op.arg); break;
        case OE_OP_RULE: pass = rule_program(op.arg, (int32_t)op.aux, i, j); break;
        case OE_OP_SCALAR_NEW:
          pass = pool.consume_reservable_merged_scalar(op.slot, op.kind, op.amount);
          if (pass) ++n_reserve;
          break;
        case OE_OP_PORT_STATIC:
          pass = pool.consume_reservable_merged_port(op.slot, op.arg);
          if (pass) { placed_ports.push_back(op.arg); ++n_reserve; }
          break;
        case OE_OP_PORT_DYNAMIC: {
          int p = select_dynamic_port(pool, op.slot, cls, op.arg, placed_ports);
          pass = p >= 0 && pool.consume_reservable_merged_port(op.slot, p);
          if (pass) {
            placed_ports.push_back(p);
            ++n_reserve;
            if (dec && n_dyn < OE_MAX_DYN_PORTS) dec->dyn_port[n_dyn] = (uint16_t)p;
          }
          ++n_dyn;
        } break;
        case OE_OP_VOL_MOUNT: {
          int slot = pool.consume_atomic(op.amount, op.mask);
          pass = slot >= 0;
          if (pass) {
            ++n_reserve;
            if (dec && n_mount < OE_MAX_MOUNT_OPS) dec->mount_slot[n_mount] = (int8_t)slot;
          }
          ++n_mount;
        } break;
        case OE_OP_SCALAR_RESERVED: {  // OfferEvaluationUtils.java:60-227 with resourceId present
          auto it = pool.reserved.find(op.arg);
          if (it == pool.reserved.end()) { pass = false; break; }
          double available = it->second.value;
          // consumeReserved (MesosResourcePool.java:259-271)
          if (compare_scalar(available, op.amount) > 0) it->second.value = available - op.amount;
          else pool.reserved.erase(it);
          if (compare_scalar(available, op.amount) == 0) break;  // previously reserved, unchanged
          double difference = op.amount - available;
          if (compare_scalar(difference, 0.0) > 0) {
            pass = pool.consume_reservable_merged_scalar(op.slot, op.kind, difference);  // RESERVE under the SAME id
          }  // else: reservation shrinks, surplus is unreserved by the host: pass
        } break;
        case OE_OP_PORT_RESERVED: {
          auto it = pool.reserved.find(op.arg);
          if (it == pool.reserved.end()) { pass = false; break; }
          int reserved_port = (int)it->second.value, want = (int)op.aux;
          pool.reserved.erase(it);
          if (reserved_port == want) { placed_ports.push_back(want); break; }
          pass = pool.consume_reservable_merged_port(op.slot, want);
          if (pass) placed_ports.push_back(want);
        } break;
        case OE_OP_MOUNT_RESERVED: pass = pool.reserved.count(op.arg) != 0; break;
        case OE_OP_FREE: {  // MesosResourcePool.free (:374-430)
          pool.reserved.erase(op.arg);
          if (op.kind == OE_KIND_PORTS) {
            pool.role_present[op.slot] = true;
            pool.ports_present[op.slot] = true;
            pool.ports[op.slot].add((int)op.aux, (int)op.aux);
          } else if (op.kind == OE_KIND_MOUNT) {
            pool.mounts.push_back({op.amount, (int32_t)op.aux - 1, OE_MAX_MOUNT_SLOTS + pool.n_freed_mounts++});
          } else {
            pool.role_present[op.slot] = true;
            double cur = pool.present[op.slot][op.kind] ? pool.val[op.slot][op.kind] : 0.0;
            pool.present[op.slot][op.kind] = true;
            pool.val[op.slot][op.kind] = cur + op.amount;
          }
        } break;
        default: pass = false;
      }
      if (!pass) failed |= 1ull << std::min(o - P.op_begin[cls], 63);
      if (!pass && quick) return failed;
    }
    if (dec) dec->n_new_reservations = n_reserve;
    return failed;
  }

  int check_loaded() {
    if (!(hasP && hasT && hasO && hasR)) return fail(OE_ERR_STATE, "tables not loaded (programs, tasks, offers, requirements)");
    if (mode && !cache.valid) build_cache();
    return 0;
  }
};

int64_t stride_words(int64_t n) { return std::max<int64_t>(128, ((n + 31) / 32 + 127) / 128 * 128); }

}  // namespace

extern "C" {

int32_t oeo_abi_version(void) { return OE_ABI_VERSION; }

int32_t oeo_create(const oe_config*, Oracle** out) {
  *out = new Oracle();
  return 0;
}
void oeo_destroy(Oracle* h) { delete h; }
const char* oeo_last_error(const Oracle* h) { return h ? h->err.c_str() : ""; }

int32_t oeo_load_programs(Oracle* h, const oe_program_table* t) {
  h->last_decisions.clear();
  h->cache.valid = false;
  Programs& P = h->P;
  P = Programs();
  P.C = t->n_classes;
  P.op_begin = copy_col(t->op_begin, (size_t)t->n_classes + 1);
  P.ops = copy_col(t->ops, (size_t)t->n_ops);
  P.nodes = copy_col(t->rule_nodes, (size_t)t->n_rule_nodes);
  P.pod_role = copy_col(t->pod_role, (size_t)t->n_classes);
  P.port_excl_begin = copy_col(t->port_excl_begin, (size_t)t->n_classes + 1);
  P.port_excl = copy_col(t->port_excl, P.port_excl_begin.empty() ? 0 : (size_t)P.port_excl_begin.back());
  P.range_set_begin = copy_col(t->range_set_begin, (size_t)t->n_range_sets + 1);
  P.range_set = copy_col(t->range_set, P.range_set_begin.empty() ? 0 : (size_t)P.range_set_begin.back());
  P.csets = copy_col(t->counter_sets, (size_t)t->n_counter_sets);
  P.type_set_pod_type = copy_col(t->type_set_pod_type, (size_t)t->n_type_sets);
  P.matcher_bits = copy_col(t->matcher_bits, (size_t)t->n_matcher_words);
  P.dict = t->dict;
  h->hasP = true;
  h->hasT = h->hasR = false;   // a resident offer table survives (include/oe_abi.h, "Loading"): its ids stay valid as dictionaries grow
  return 0;
}

int32_t oeo_load_tasks(Oracle* h, const oe_task_table* t) {
  h->last_decisions.clear();
  h->cache.valid = false;
  Tasks& T = h->T;
  T = Tasks();
  size_t n = (size_t)t->n_tasks;
  T.T = t->n_tasks; T.TA = t->n_attr_slots; T.Q = t->n_rr_cols; T.FW = t->n_filter_words;
  T.name_id = copy_col(t->name_id, n);
  T.pod_type_id = copy_col(t->pod_type_id, n);
  T.pod_index = copy_col(t->pod_index, n);
  T.agent_id = copy_col(t->agent_id, n);
  T.hostname_id = copy_col(t->hostname_id, n);
  T.zone_id = copy_col(t->zone_id, n);
  T.region_id = copy_col(t->region_id, n);
  // transpose the slot-major columns into per-task rows (the oracle appends rows in apply_launches)
  T.attr_id.assign(n * T.TA, -1);
  for (int a = 0; a < T.TA; ++a) for (size_t r = 0; r < n; ++r) T.attr_id[r * T.TA + a] = t->attr_id[(size_t)a * n + r];
  T.rr_key.assign(n * T.Q, -1);
  for (int q = 0; q < T.Q; ++q) for (size_t r = 0; r < n; ++r) T.rr_key[r * T.Q + q] = t->rr_key[(size_t)q * n + r];
  T.filter_bits.assign(n * T.FW, 0);
  for (int w = 0; w < T.FW; ++w) for (size_t r = 0; r < n; ++r) T.filter_bits[r * T.FW + w] = t->filter_bits[(size_t)w * n + r];
  h->hasT = true;
  return 0;
}

int32_t oeo_load_offers(Oracle* h, const oe_offer_table* t) {
  h->last_decisions.clear();
  h->cache.valid = false;
  Offers& O = h->O;
  O = Offers();
  size_t n = (size_t)t->n_offers;
  O.N = t->n_offers; O.row_offset = t->row_offset; O.N_global = t->n_offers_global;
  O.S = t->n_role_slots; O.K = t->n_scalar_kinds; O.R = t->n_port_ranges; O.D = t->n_mount_slots;
  O.A = t->n_attr_slots; O.Q = t->n_rr_cols; O.E = t->n_exec_slots;
  if (O.S > OE_MAX_ROLE_SLOTS || O.K > OE_MAX_SCALAR_KINDS) return h->fail(OE_ERR_LIMIT, "too many role slots / scalar kinds");
  O.scalar = copy_col(t->scalar, n * O.S * O.K);
  O.scalar_present = copy_col(t->scalar_present, n);
  O.port_range = copy_col(t->port_range, n * O.S * O.R);
  O.mount_size = copy_col(t->mount_size, n * O.D);
  O.mount_profile = copy_col(t->mount_profile, n * O.D);
  O.alloc_role = copy_col(t->alloc_role, n);
  O.agent_id = copy_col(t->agent_id, n);
  O.hostname_id = copy_col(t->hostname_id, n);
  O.zone_id = copy_col(t->zone_id, n);
  O.region_id = copy_col(t->region_id, n);
  O.attr_id = copy_col(t->attr_id, n * O.A);
  O.rr_key = copy_col(t->rr_key, n * O.Q);
  O.exec_id = copy_col(t->exec_id, n * O.E);
  if (t->n_reserved > 0) {
    O.reserved_begin = copy_col(t->reserved_begin, n + 1);
    O.reserved_id = copy_col(t->reserved_id, (size_t)t->n_reserved);
    O.reserved_kind = copy_col(t->reserved_kind, (size_t)t->n_reserved);
    O.reserved_value = copy_col(t->reserved_value, (size_t)t->n_reserved);
  }
  h->hasO = true;
  return 0;
}

// oe_update_offers: rows of the resident table replaced in place (no reserved side table on either side, as in the engine).
int32_t oeo_update_offers(Oracle* h, const int64_t* rows, int64_t n_rows, const oe_offer_table* t) {
  h->last_decisions.clear();
  h->cache.valid = false;
  Offers& O = h->O;
  if (!h->hasO) return h->fail(OE_ERR_STATE, "no resident offer table");
  if (!O.reserved_begin.empty() || t->n_reserved > 0) return h->fail(OE_ERR_STATE, "a table with a reserved side table cannot be patched in place");
  if (t->n_offers != n_rows || t->n_role_slots != O.S || t->n_scalar_kinds != O.K || t->n_port_ranges != O.R || t->n_mount_slots != O.D ||
      t->n_attr_slots != O.A || t->n_rr_cols != O.Q || t->n_exec_slots != O.E)
    return h->fail(OE_ERR_INVALID, "slot counts differ from the resident table");
  const size_t N = (size_t)O.N, n = (size_t)n_rows;
  for (size_t k = 0; k < n; ++k) {
    if (rows[k] < 0 || rows[k] >= O.N) return h->fail(OE_ERR_INVALID, "row outside the resident table");
    const size_t i = (size_t)rows[k];
    for (int q = 0; q < O.S * O.K; ++q) O.scalar[(size_t)q * N + i] = t->scalar[(size_t)q * n + k];
    O.scalar_present[i] = t->scalar_present[k];
    for (int q = 0; q < O.S * O.R; ++q) O.port_range[(size_t)q * N + i] = t->port_range[(size_t)q * n + k];
    for (int q = 0; q < O.D; ++q) { O.mount_size[(size_t)q * N + i] = t->mount_size[(size_t)q * n + k]; O.mount_profile[(size_t)q * N + i] = t->mount_profile[(size_t)q * n + k]; }
    O.alloc_role[i] = t->alloc_role[k]; O.agent_id[i] = t->agent_id[k]; O.hostname_id[i] = t->hostname_id[k];
    O.zone_id[i] = t->zone_id[k]; O.region_id[i] = t->region_id[k];
    for (int q = 0; q < O.A; ++q) O.attr_id[(size_t)q * N + i] = t->attr_id[(size_t)q * n + k];
    for (int q = 0; q < O.Q; ++q) O.rr_key[(size_t)q * N + i] = t->rr_key[(size_t)q * n + k];
    for (int q = 0; q < O.E; ++q) O.exec_id[(size_t)q * N + i] = t->exec_id[(size_t)q * n + k];
  }
  return 0;
}

int32_t oeo_load_requirements(Oracle* h, const oe_req_table* t) {
  h->last_decisions.clear();
  h->cache.valid = false;
  Reqs& R = h->R;
  R = Reqs();
  size_t m = (size_t)t->n_reqs;
  R.M = t->n_reqs;
  R.class_id = copy_col(t->class_id, m);
  R.pod_type_id = copy_col(t->pod_type_id, m);
  R.pod_index = copy_col(t->pod_index, m);
  // self_begin / self_rows are deliberately ignored: the oracle finds equivalent tasks by walking the task table
  if (t->launch_begin) {
    R.launch_begin = copy_col(t->launch_begin, m + 1);
    size_t nl = (size_t)R.launch_begin.back();
    R.launch_name_id = copy_col(t->launch_name_id, nl);
    R.launch_filter_bits = copy_col(t->launch_filter_bits, nl * h->T.FW);
  }
  h->hasR = true;
  return 0;
}

int64_t oeo_matrix_stride_words(const Oracle* h) { return stride_words(h->O.N); }

// Full matrix, literally: for every requirement, for every offer, run the whole pipeline.
// `stage_fail` (optional, [M][N] u64) receives the failed-stage mask of every pair.  With threads > 1 (oeo_set_mode) the
// offers of a row are split into word-aligned blocks, one per thread: pairs are independent of each other.
int32_t oeo_feasibility_ex(Oracle* h, uint32_t* bits, uint64_t* stage_fail, int64_t j_begin, int64_t j_end) {
  if (int rc = h->check_loaded()) return rc;
  int64_t stride = stride_words(h->O.N);
  const int64_t N = h->O.N;
  for (int64_t j = j_begin; j < j_end; ++j) {
    uint32_t* row = bits + (size_t)j * stride;
    std::memset(row, 0, (size_t)stride * 4);
    auto span = [&](int64_t i0, int64_t i1) {
      for (int64_t i = i0; i < i1; ++i) {
        uint64_t f = h->evaluate_pair(i, j, nullptr, h->mode && !stage_fail);
        if (stage_fail) stage_fail[(size_t)j * N + i] = f;
        if (f == 0) row[i >> 5] |= 1u << (i & 31);
      }
    };
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(h->threads, N / 4096));
    if (nt <= 1) { span(0, N); continue; }
    const int64_t per = ((N + nt - 1) / nt + 31) / 32 * 32;
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t) {
      const int64_t i0 = std::min<int64_t>(N, t * per), i1 = std::min<int64_t>(N, i0 + per);
      if (i1 > i0) pool.emplace_back(span, i0, i1);
    }
    for (auto& th : pool) th.join();
  }
  h->pairs_evaluated = (j_end - j_begin) * h->O.N;
  return 0;
}

// One matrix row into a buffer of `stride` words (large tables: the caller need not hold M rows).
int32_t oeo_feasibility_row(Oracle* h, int64_t j, uint32_t* row_bits) {
  if (int rc = h->check_loaded()) return rc;
  if (j < 0 || j >= h->R.M) return h->fail(OE_ERR_INVALID, "requirement out of range");
  return oeo_feasibility_ex(h, row_bits - (size_t)j * stride_words(h->O.N), nullptr, j, j + 1);
}

// mode 0 = literal (the reference's own loops), 1 = tallies cached per cycle (see the file header); threads = host threads
// a feasibility row may use.
int32_t oeo_set_mode(Oracle* h, int32_t mode, int32_t threads) {
  h->mode = mode ? 1 : 0;
  h->threads = std::max(1, threads);
  h->cache.valid = false;
  return 0;
}

int32_t oeo_feasibility(Oracle* h, uint32_t* bits, oe_stats* stats) {
  int rc = oeo_feasibility_ex(h, bits, nullptr, 0, h->R.M);
  if (stats) { std::memset(stats, 0, sizeof *stats); stats->n_pairs = h->pairs_evaluated; }
  return rc;
}

// PlanScheduler.resourceOffers (:50-70) over OfferEvaluator.evaluate (:113-248).
int32_t oeo_evaluate_cycle(Oracle* h, oe_decision* out, oe_stats* stats) {
  if (int rc = h->check_loaded()) return rc;
  if (h->mode) {
    // One cursor per (class, equivalent-task set): the pairs' verdicts depend on nothing else within a cycle and taken offers
    // never return, so the offers before a cursor stay unusable for every requirement that shares it.
    const int64_t N = h->O.N;
    std::vector<char> taken((size_t)N, 0);
    std::map<std::pair<int32_t, std::vector<int64_t>>, int64_t> cursor;
    int64_t pairs = 0;
    int placed = 0;
    for (int64_t j = 0; j < h->R.M; ++j) {
      oe_decision dec;
      std::memset(&dec, 0, sizeof dec);
      dec.offer_row = -1;
      for (auto& m : dec.mount_slot) m = -1;
      int64_t& cur = cursor[{h->R.class_id[j], h->cache.self[(size_t)j]}];
      // first offer at or after the cursor that is free and passes: the first 512 one by one, then in blocks that the host
      // threads search side by side (the lowest hit wins, so the answer is that of the plain scan)
      auto first_hit = [&](int64_t a, int64_t b) -> int64_t {
        for (int64_t i = a; i < b; ++i)
          if (!taken[(size_t)i] && h->evaluate_pair(i, j, nullptr, true) == 0) return i;
        return b;
      };
      int64_t i = cur;
      {
        const int64_t head_end = std::min<int64_t>(N, i + 512);
        const int64_t hit = first_hit(i, head_end);
        pairs += std::min(hit + 1, head_end) - i;
        i = hit;
        if (hit == head_end) {
          int64_t block = 4096;
          while (i < N) {
            const int nt = h->threads;
            const int64_t span_end = std::min<int64_t>(N, i + block * nt);
            std::vector<int64_t> hits((size_t)nt, span_end);
            std::vector<std::thread> pool;
            for (int t = 0; t < nt; ++t) {
              const int64_t a = std::min(span_end, i + t * block), b = std::min(span_end, a + block);
              if (b > a) pool.emplace_back([&, t, a, b] { const int64_t x = first_hit(a, b); hits[(size_t)t] = x < b ? x : span_end; });
            }
            for (auto& th : pool) th.join();
            const int64_t best = *std::min_element(hits.begin(), hits.end());
            pairs += std::min(best + 1, span_end) - i;
            i = best;
            if (best < span_end) break;
            block = std::min<int64_t>(block * 2, 65536);
          }
        }
      }
      if (i < N) {
        oe_decision trial;
        std::memset(&trial, 0, sizeof trial);
        h->evaluate_pair(i, j, &trial);
        dec = trial;
        dec.offer_row = (int32_t)(i + h->O.row_offset);
        dec.reserved0 = 0;
        taken[(size_t)i] = 1;
        ++placed;
      }
      cur = i;
      if (dec.offer_row < 0) dec.n_new_reservations = 0;
      out[j] = dec;
    }
    h->last_decisions.assign(out, out + h->R.M);
    h->pairs_evaluated = pairs;
    if (stats) { std::memset(stats, 0, sizeof *stats); stats->n_pairs = pairs; stats->n_placed = placed; }
    return 0;
  }
  std::vector<int64_t> available((size_t)h->O.N);  // availableOffers: list of rows, shrinking
  for (int64_t i = 0; i < h->O.N; ++i) available[i] = i;
  int64_t pairs = 0;
  int placed = 0;
  for (int64_t j = 0; j < h->R.M; ++j) {
    oe_decision dec;
    std::memset(&dec, 0, sizeof dec);
    dec.offer_row = -1;
    for (auto& m : dec.mount_slot) m = -1;
    size_t hit = available.size();
    for (size_t a = 0; a < available.size(); ++a) {
      oe_decision trial;
      std::memset(&trial, 0, sizeof trial);
      ++pairs;
      if (h->evaluate_pair(available[a], j, &trial) == 0) {
        dec = trial;
        dec.offer_row = (int32_t)(available[a] + h->O.row_offset);
        dec.reserved0 = 0;
        hit = a;
        break;
      }
    }
    if (hit < available.size()) { available.erase(available.begin() + hit); ++placed; }
    else { dec.n_new_reservations = 0; }
    out[j] = dec;
  }
  h->last_decisions.assign(out, out + h->R.M);
  h->pairs_evaluated = pairs;
  if (stats) { std::memset(stats, 0, sizeof *stats); stats->n_pairs = pairs; stats->n_placed = placed; }
  return 0;
}

// LaunchEvaluationStage.java:58-68 + PersistentLaunchRecorder.record: every task of a placed pod is stored under its
// name with the winning offer's hostname / attributes / zone / region / agent.
int32_t oeo_apply_launches(Oracle* h) {
  if (int rc = h->check_loaded()) return rc;
  if (h->R.launch_begin.empty()) return h->fail(OE_ERR_STATE, "requirement table carries no launch rows");
  if ((int64_t)h->last_decisions.size() != h->R.M) return h->fail(OE_ERR_STATE, "no cycle evaluated");
  Tasks& T = h->T;
  const Offers& O = h->O;
  if (T.TA < O.A) return h->fail(OE_ERR_INVALID, "task table has fewer attribute slots than the offer table");
  h->cache.valid = false;
  for (int64_t j = 0; j < h->R.M; ++j) {
    int64_t row = h->last_decisions[j].offer_row;
    if (row < 0) continue;
    int64_t i = row - O.row_offset;
    for (int32_t l = h->R.launch_begin[j]; l < h->R.launch_begin[j + 1]; ++l) {
      int32_t name = h->R.launch_name_id[l];
      int64_t t = -1;
      for (int64_t x = 0; x < T.T; ++x) if (T.name_id[x] == name) { t = x; break; }
      if (t < 0) {
        t = T.T++;
        T.name_id.push_back(name);
        T.pod_type_id.push_back(0); T.pod_index.push_back(0); T.agent_id.push_back(0);
        T.hostname_id.push_back(0); T.zone_id.push_back(0); T.region_id.push_back(0);
        T.attr_id.resize((size_t)T.T * T.TA, -1);
        T.rr_key.resize((size_t)T.T * T.Q, -1);
        T.filter_bits.resize((size_t)T.T * T.FW, 0);
      }
      T.pod_type_id[t] = h->R.pod_type_id[j];
      T.pod_index[t] = h->R.pod_index[j];
      T.agent_id[t] = O.agent_id[i];
      T.hostname_id[t] = O.hostname_id[i];
      T.zone_id[t] = O.zone_id[i];
      T.region_id[t] = O.region_id[i];
      for (int a = 0; a < T.TA; ++a) T.attr_id[(size_t)t * T.TA + a] = a < O.A ? O.attr_id[(size_t)a * O.N + i] : -1;
      for (int q = 0; q < T.Q; ++q) T.rr_key[(size_t)t * T.Q + q] = O.rr_key[(size_t)q * O.N + i];
      for (int w = 0; w < T.FW; ++w) T.filter_bits[(size_t)t * T.FW + w] = h->R.launch_filter_bits[(size_t)l * T.FW + w];
    }
  }
  return 0;
}

// DefaultScheduler.getUnexpectedResources (S/scheduler/DefaultScheduler.java:483-538): for every unused offer, every resource
// with a resource id that is not in the keep set.
int32_t oeo_unexpected_resources(Oracle* h, const uint32_t* keep_bits, int32_t only_unused_offers, uint8_t* unexpected, int64_t* n_unexpected) {
  if (int rc = h->check_loaded()) return rc;
  int64_t count = 0;
  std::vector<char> taken((size_t)h->O.N, 0);
  if (only_unused_offers)
    for (auto& d : h->last_decisions) if (d.offer_row >= 0 && d.offer_row < h->O.N) taken[(size_t)d.offer_row] = 1;
  for (int64_t i = 0; i < h->O.N; ++i) {
    for (int64_t q = h->O.reserved_begin.empty() ? 0 : h->O.reserved_begin[i]; !h->O.reserved_begin.empty() && q < h->O.reserved_begin[i + 1]; ++q) {
      unexpected[q] = 0;
      if (taken[(size_t)i]) continue;
      const int32_t id = h->O.reserved_id[q];
      const bool keep = (keep_bits[id >> 5] >> (id & 31)) & 1u;
      unexpected[q] = !keep;
      count += !keep;
    }
  }
  if (n_unexpected) *n_unexpected = count;
  return OE_OK;
}

int64_t oeo_task_count(const Oracle* h) { return h->T.T; }
int64_t oeo_pairs_evaluated(const Oracle* h) { return h->pairs_evaluated; }

// Copies the oracle's current task table back out (slot-major columns), so tests can compare it with the engine's.
int32_t oeo_export_tasks(const Oracle* h, int32_t* name_id, int32_t* pod_type_id, int32_t* pod_index, int32_t* agent_id,
                         int32_t* hostname_id, int32_t* zone_id, int32_t* region_id, int32_t* attr_id, int32_t* rr_key,
                         uint32_t* filter_bits) {
  const Tasks& T = h->T;
  size_t n = (size_t)T.T;
  for (size_t t = 0; t < n; ++t) {
    name_id[t] = T.name_id[t]; pod_type_id[t] = T.pod_type_id[t]; pod_index[t] = T.pod_index[t];
    agent_id[t] = T.agent_id[t]; hostname_id[t] = T.hostname_id[t]; zone_id[t] = T.zone_id[t]; region_id[t] = T.region_id[t];
    for (int a = 0; a < T.TA; ++a) attr_id[(size_t)a * n + t] = T.attr_id[t * T.TA + a];
    for (int q = 0; q < T.Q; ++q) rr_key[(size_t)q * n + t] = T.rr_key[t * T.Q + q];
    for (int w = 0; w < T.FW; ++w) filter_bits[(size_t)w * n + t] = T.filter_bits[t * T.FW + w];
  }
  return 0;
}

}  // extern "C"
