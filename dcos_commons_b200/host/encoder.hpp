// encoder.hpp — turns one cycle's Offers, TaskInfos and PodInstanceRequirements into the columnar, dictionary-encoded
// tables of include/oe_abi.h.  This is the host half of the drop-in: everything string-shaped is resolved here
// (dictionaries, StringMatchers, attribute formatting), and the stage pipeline the reference would build per
// requirement (S/offer/evaluate/OfferEvaluator.java:411-522) is flattened into a class program.
//
// Reference semantics restated (S/ = sdk/scheduler/src/main/java/com/mesosphere/sdk/):
//   S/offer/MesosResourcePool.java:53-147,432-455   pool partition: merged by role, reserved by id, atomic MOUNT disks
//   S/offer/MesosResource.java:26-62                isAtomic, getRole
//   S/offer/ResourceUtils.java:82-117               getReservation / getResourceId
//   S/offer/RangeUtils.java:23-47 (antlr IntervalSet) merged port ranges: sorted, disjoint, adjacency-coalesced
//   S/offer/taskdata/TaskLabelReader.java:47-92     what a deployed task contributes to the task table
//   S/offer/evaluate/OfferEvaluator.java:380-403,411-536   stage order of a new pod
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <set>
#include <stdexcept>
#include <unordered_map>

#include "specification.hpp"

namespace sdk {

// S/offer/InvalidRequirementException.java
struct InvalidRequirementException : std::runtime_error { using std::runtime_error::runtime_error; };

namespace labels {  // S/offer/taskdata/LabelConstants.java
constexpr const char* OFFER_ATTRIBUTES = "offer_attributes";
constexpr const char* OFFER_HOSTNAME = "offer_hostname";
constexpr const char* OFFER_ZONE = "offer_zone";
constexpr const char* OFFER_REGION = "offer_region";
constexpr const char* TASK_TYPE = "task_type";
constexpr const char* TASK_INDEX = "index";
constexpr const char* RESOURCE_ID = "resource_id";
}  // namespace labels

// S/offer/taskdata/TaskLabelWriter.java (the setters LaunchEvaluationStage.java:58-68 uses)
struct TaskLabelWriter {
  mesos::TaskInfo& task;
  explicit TaskLabelWriter(mesos::TaskInfo& t) : task(t) {}
  TaskLabelWriter& setType(const std::string& t) { task.labels[labels::TASK_TYPE] = t; return *this; }
  TaskLabelWriter& setIndex(int i) { task.labels[labels::TASK_INDEX] = std::to_string(i); return *this; }
  TaskLabelWriter& setHostname(const mesos::Offer& o) { task.labels[labels::OFFER_HOSTNAME] = o.hostname; return *this; }
  TaskLabelWriter& setOfferAttributes(const mesos::Offer& o) { task.labels[labels::OFFER_ATTRIBUTES] = AttributeStringUtils::toString(o.attributes); return *this; }
  TaskLabelWriter& setZone(const std::string& z) { task.labels[labels::OFFER_ZONE] = z; return *this; }
  TaskLabelWriter& setRegion(const std::string& r) { task.labels[labels::OFFER_REGION] = r; return *this; }
};

struct ResourceUtils {
  // S/offer/ResourceUtils.java:82-117: the effective reservation is the last entry of the refinement stack
  static std::optional<std::string> getResourceId(const mesos::Resource& r) {
    if (r.reservations.empty()) return std::nullopt;
    auto it = r.reservations.back().labels.find(labels::RESOURCE_ID);
    if (it == r.reservations.back().labels.end()) return std::nullopt;
    return it->second;
  }
  static std::string getRole(const mesos::Resource& r) { return r.reservations.empty() ? ANY_ROLE : r.reservations.back().role; }  // MesosResource.java:52-62
  static bool isAtomic(const mesos::Resource& r) { return r.disk && r.disk->source == mesos::DiskInfo::MOUNT; }                    // MesosResource.java:26-30
};

// antlr4 IntervalSet as S/offer/RangeUtils.java uses it: closed intervals kept sorted, disjoint and adjacency-coalesced.
struct IntervalSet {
  std::vector<std::pair<int64_t, int64_t>> iv;
  void add(int64_t a, int64_t b) {
    if (b < a) return;
    std::vector<std::pair<int64_t, int64_t>> out;
    bool placed = false;
    for (auto& x : iv) {
      if (x.second + 1 < a) out.push_back(x);
      else if (b + 1 < x.first) { if (!placed) { out.push_back({a, b}); placed = true; } out.push_back(x); }
      else { a = std::min(a, x.first); b = std::max(b, x.second); }
    }
    if (!placed) out.push_back({a, b});
    std::sort(out.begin(), out.end());
    iv.swap(out);
  }
};

class Dictionary {
 public:
  int32_t intern(const std::string& s) {
    auto it = ids_.find(s);
    if (it != ids_.end()) return it->second;
    const int32_t id = (int32_t)values_.size();
    ids_.emplace(s, id);
    values_.push_back(s);
    return id;
  }
  int32_t find(const std::string& s) const { auto it = ids_.find(s); return it == ids_.end() ? -1 : it->second; }
  int32_t size() const { return (int32_t)values_.size(); }
  const std::vector<std::string>& values() const { return values_; }
 private:
  std::unordered_map<std::string, int32_t> ids_;
  std::vector<std::string> values_;
};

// The dictionaries an encoded offer row refers to.  One CycleEncoder owns a fresh set; an adapter that keeps the offer table on
// the device from cycle to cycle (ResidentOffers below) owns one set for as long as the table lives: ids handed out once stay
// valid, dictionaries only grow, and the engine keeps the resident table across program re-loads on exactly that condition.
struct OfferDictionaries {
  Dictionary agents, hosts, zones, regions, attrs, rrValues, rrAttrs, execs, roles, slots, kinds, resids, profiles;
};

// Host-side shadow of one OfferEvaluationStage of a class, index-aligned with the class's oe_stage_op list.  It is what
// the adapter needs to turn the engine's decision for the WINNING offer back into recommendations.
struct StageInfo {
  enum Kind { EXECUTOR, PLACEMENT, RESOURCE, PORT, VOLUME, TLS, LAUNCH, UNRESERVE, DESTROY };
  Kind kind = EXECUTOR;
  std::string resourceName;             // RESOURCE / PORT / VOLUME
  double value = 0.0;
  int port = 0;                         // PORT: requested port (0 = dynamic)
  bool hostPort = true;                 // PORT: false = overlay network, no reservation
  VolumeSpec::Type volumeType = VolumeSpec::ROOT;
  std::string containerPath;
  std::vector<std::string> taskNames;   // task spec names that receive the resource; empty = executor level
  std::string taskSpecName;             // LAUNCH
  bool shouldLaunch = false;            // LAUNCH
  std::string role, preReservedRole, principal;
  std::string portName;                 // PORT: PortSpec.getPortName()
  // existing-pod path: the reservation this stage expects to find in the offer (ResourceEvaluationStage.requiredResourceId)
  std::optional<std::string> resourceId;
  std::optional<std::string> persistenceId;
  bool skipIfRunningExecutor = false;   // executor-level resource of a pod whose executor still runs
  std::string executorId;
  mesos::Resource orphan;               // UNRESERVE / DESTROY: the resource no spec claims any more
};

// What OfferEvaluator.getEvaluationPipeline (OfferEvaluator.java:250-310) derives from the state store for a pod that has
// launched before: its stored tasks and the executor to re-use, if one is still alive.
struct ExistingPod {
  std::map<std::string, mesos::TaskInfo> podTasks;  // by task name
  std::optional<std::string> executorId;            // a task of the pod that is not being relaunched still runs in it
  mesos::ExecutorInfo executorInfo;
};

struct EncodedCycle {
  // programs
  std::vector<int32_t> op_begin{0};
  std::vector<oe_stage_op> ops;
  std::vector<oe_rule_node> nodes;
  std::vector<int32_t> pod_role;
  std::vector<int32_t> excl_begin{0};
  std::vector<uint16_t> excl;
  std::vector<int32_t> rs_begin{0};
  std::vector<uint32_t> rs;
  std::vector<oe_counter_set> csets;
  std::vector<int32_t> tsets;
  std::vector<uint32_t> mbits;
  // tasks
  int TA = 1, Q = 0, FW = 1;
  std::vector<int32_t> t_name, t_type, t_index, t_agent, t_host, t_zone, t_region, t_attr, t_rr;
  std::vector<uint32_t> t_filter;
  // offers
  int S = 1, K = 4, R = 1, D = 0, A = 1, E = 0;
  std::vector<double> o_scalar, o_msize, o_rval;
  std::vector<uint32_t> o_present, o_port;
  std::vector<int32_t> o_mprof, o_alloc, o_agent, o_host, o_zone, o_region, o_attr, o_rr, o_exec, o_rid, o_rkind;
  std::vector<int64_t> o_rbegin;
  // Resident offer table (ResidentOffers): when `resident` is set the o_* columns above hold only the rows in `delta_rows` (row k
  // of the columns = resident row delta_rows[k]) unless `full_reload` says the whole table is in them; N_resident = rows of the
  // table on the device.
  bool resident = false, full_reload = true;
  std::vector<int64_t> delta_rows;
  int64_t N_resident = 0;
  // requirements
  std::vector<int32_t> r_class, r_type, r_index, l_begin{0}, l_name;
  std::vector<uint32_t> l_filter;
  // host-side shadows
  std::vector<std::vector<StageInfo>> classStages;
  int64_t N = 0, T = 0, M = 0;
  oe_dict_sizes dict{};

  oe_program_table programTable() const {
    oe_program_table t{};
    t.n_classes = (int32_t)pod_role.size(); t.n_ops = (int32_t)ops.size(); t.n_rule_nodes = (int32_t)nodes.size();
    t.n_counter_sets = (int32_t)csets.size(); t.n_type_sets = (int32_t)tsets.size(); t.n_range_sets = (int32_t)rs_begin.size() - 1;
    t.n_matcher_words = (int64_t)mbits.size();
    t.op_begin = op_begin.data(); t.ops = ops.data(); t.rule_nodes = nodes.data(); t.pod_role = pod_role.data();
    t.port_excl_begin = excl_begin.data(); t.port_excl = excl.data(); t.range_set_begin = rs_begin.data(); t.range_set = rs.data();
    t.counter_sets = csets.data(); t.type_set_pod_type = tsets.data(); t.matcher_bits = mbits.data(); t.dict = dict;
    return t;
  }
  oe_task_table taskTable() const {
    oe_task_table t{};
    t.n_tasks = T; t.n_attr_slots = TA; t.n_rr_cols = Q; t.n_filter_words = FW;
    t.name_id = t_name.data(); t.pod_type_id = t_type.data(); t.pod_index = t_index.data(); t.agent_id = t_agent.data();
    t.hostname_id = t_host.data(); t.zone_id = t_zone.data(); t.region_id = t_region.data(); t.attr_id = t_attr.data();
    t.rr_key = t_rr.data(); t.filter_bits = t_filter.data();
    return t;
  }
  oe_offer_table offerTable() const {
    oe_offer_table t{};
    t.n_offers = N; t.row_offset = 0; t.n_offers_global = N; t.shard_rows = 0;
    t.n_role_slots = S; t.n_scalar_kinds = K; t.n_port_ranges = R; t.n_mount_slots = D; t.n_attr_slots = A; t.n_rr_cols = Q; t.n_exec_slots = E;
    t.scalar = o_scalar.data(); t.scalar_present = o_present.data(); t.port_range = o_port.data(); t.mount_size = o_msize.data();
    t.mount_profile = o_mprof.data(); t.alloc_role = o_alloc.data(); t.agent_id = o_agent.data(); t.hostname_id = o_host.data();
    t.zone_id = o_zone.data(); t.region_id = o_region.data(); t.attr_id = o_attr.data(); t.rr_key = o_rr.data(); t.exec_id = o_exec.data();
    t.n_reserved = (int64_t)o_rid.size();
    if (t.n_reserved) { t.reserved_begin = o_rbegin.data(); t.reserved_id = o_rid.data(); t.reserved_kind = o_rkind.data(); t.reserved_value = o_rval.data(); }
    return t;
  }
  oe_req_table reqTable() const {
    oe_req_table t{};
    t.n_reqs = M; t.class_id = r_class.data(); t.pod_type_id = r_type.data(); t.pod_index = r_index.data();
    t.launch_begin = l_begin.data(); t.launch_name_id = l_name.data(); t.launch_filter_bits = l_filter.data();
    return t;
  }
};

// The compact wire form of an encoded offer table (oe_offer_packing, include/oe_abi.h): what goes over the bus instead of
// EncodedCycle's full-width offer columns.  Decided per column from the values themselves — narrow unsigned ids when the
// dictionary fits (all-ones = -1), fp32 scalars only when every amount survives the round trip, no hostname column when every
// row's hostname id is its agent id, no allocation-role column when all rows agree, and the `ports` ranges as a dictionary of
// the table's distinct range lists plus one list id per (slot, offer).  The engine widens all of it on the device, exactly.
class PackedOffers {
 public:
  explicit PackedOffers(const EncodedCycle& c) : c_(c) {
    const size_t N = (size_t)c.N;
    pk_ = oe_offer_packing{};
    // scalars
    bool f32 = true;
    for (double v : c.o_scalar) if (!((double)(float)v == v || v != v)) { f32 = false; break; }
    pk_.scalar_is_f32 = f32 ? 1 : 0;
    if (f32) { scalarF_.resize(c.o_scalar.size()); for (size_t k = 0; k < c.o_scalar.size(); ++k) scalarF_[k] = (float)c.o_scalar[k]; }
    const int bits = c.S * c.K;
    pk_.present_bytes = bits <= 8 ? 1 : bits <= 16 ? 2 : 4;
    narrow(c.o_present, pk_.present_bytes, present_, false);
    pk_.mount_profile_bytes = idBytes(64); narrow(c.o_mprof, pk_.mount_profile_bytes, mprof_, true);
    pk_.zone_bytes = idBytes(c.dict.n_zones); narrow(c.o_zone, pk_.zone_bytes, zone_, true);
    pk_.region_bytes = idBytes(c.dict.n_regions); narrow(c.o_region, pk_.region_bytes, region_, true);
    pk_.attr_bytes = idBytes(c.dict.n_attrs); narrow(c.o_attr, pk_.attr_bytes, attr_, true);
    pk_.rr_bytes = idBytes(c.dict.n_rr_values); narrow(c.o_rr, pk_.rr_bytes, rr_, true);
    pk_.exec_bytes = idBytes(c.dict.n_exec_ids); narrow(c.o_exec, pk_.exec_bytes, exec_, true);
    pk_.hostname_is_agent = c.o_host == c.o_agent ? 1 : 0;
    bool constRole = true;
    for (size_t i = 1; i < N && constRole; ++i) constRole = c.o_alloc[i] == c.o_alloc[0];
    pk_.alloc_role_is_const = constRole ? 1 : 0;
    pk_.alloc_role_const = constRole && N ? c.o_alloc[0] : -1;
    // the distinct range lists, numbered in the order the rows bring them
    pk_.port_tuple_bytes = 0;
    const int R = c.R, S = c.S;
    if (R > 0 && N > 0) {
      std::map<std::vector<uint32_t>, uint32_t> seen;
      std::vector<uint32_t> ids((size_t)S * N), list((size_t)R), last;
      uint32_t lastId = 0;
      bool have = false;
      const size_t limit = std::max<size_t>(N, 4096) / 4;   // beyond this the dictionary is no shorter than the column
      bool fits = true;
      for (int s = 0; s < S && fits; ++s)
        for (size_t i = 0; i < N; ++i) {
          for (int r = 0; r < R; ++r) list[(size_t)r] = c.o_port[((size_t)s * R + r) * N + i];
          if (have && list == last) { ids[(size_t)s * N + i] = lastId; continue; }   // most agents offer the list their neighbour offers
          auto it = seen.find(list);
          if (it == seen.end()) {
            if (seen.size() >= limit) { fits = false; break; }
            it = seen.emplace(list, (uint32_t)seen.size()).first;
            tuples_.insert(tuples_.end(), list.begin(), list.end());
          }
          ids[(size_t)s * N + i] = it->second;
          last = list; lastId = it->second; have = true;
        }
      if (fits) {
        pk_.port_tuple_bytes = seen.size() <= 65536 ? 2 : 4;
        pk_.n_port_tuples = (int32_t)seen.size();
        narrow(ids, pk_.port_tuple_bytes, tupleId_, false);
      } else {
        tuples_.clear();
      }
    }
  }
  // the same oe_offer_table as EncodedCycle::offerTable(), its column pointers re-aimed at the narrow columns
  oe_offer_table table() const {
    oe_offer_table t = c_.offerTable();
    if (pk_.scalar_is_f32) t.scalar = reinterpret_cast<const double*>(scalarF_.data());
    t.scalar_present = reinterpret_cast<const uint32_t*>(present_.data());
    t.mount_profile = reinterpret_cast<const int32_t*>(mprof_.data());
    t.zone_id = reinterpret_cast<const int32_t*>(zone_.data());
    t.region_id = reinterpret_cast<const int32_t*>(region_.data());
    t.attr_id = reinterpret_cast<const int32_t*>(attr_.data());
    t.rr_key = reinterpret_cast<const int32_t*>(rr_.data());
    t.exec_id = reinterpret_cast<const int32_t*>(exec_.data());
    if (pk_.hostname_is_agent) t.hostname_id = nullptr;
    if (pk_.alloc_role_is_const) t.alloc_role = nullptr;
    if (pk_.port_tuple_bytes) t.port_range = nullptr;
    return t;
  }
  oe_offer_packing packing() const {
    oe_offer_packing p = pk_;
    p.port_tuples = tuples_.data();
    p.port_tuple_id = tupleId_.data();
    return p;
  }
  size_t bytes() const {   // what crosses the bus for the offer columns
    const size_t N = (size_t)c_.N;
    return (pk_.scalar_is_f32 ? scalarF_.size() * 4 : c_.o_scalar.size() * 8) + present_.size() + c_.o_msize.size() * 8 + mprof_.size() + zone_.size() +
           region_.size() + attr_.size() + rr_.size() + exec_.size() + N * 4 + (pk_.hostname_is_agent ? 0 : N * 4) + (pk_.alloc_role_is_const ? 0 : N * 4) +
           (pk_.port_tuple_bytes ? tuples_.size() * 4 + tupleId_.size() : c_.o_port.size() * 4);
  }

 private:
  static uint8_t idBytes(int64_t dictSize) { return dictSize < 0xff ? 1 : dictSize < 0xffff ? 2 : 4; }
  template <class T>
  static void narrow(const std::vector<T>& src, int bytes, std::vector<uint8_t>& out, bool absentIsAllOnes) {
    out.resize(src.size() * (size_t)bytes + 16);   // (+16: an empty column still has an address)
    for (size_t k = 0; k < src.size(); ++k) {
      const int64_t v = (int64_t)src[k];
      if (bytes == 1) out[k] = (uint8_t)((absentIsAllOnes && v < 0) ? 0xff : v);
      else if (bytes == 2) { const uint16_t w = (uint16_t)((absentIsAllOnes && v < 0) ? 0xffff : v); std::memcpy(&out[k * 2], &w, 2); }
      else { const uint32_t w = (uint32_t)v; std::memcpy(&out[k * 4], &w, 4); }
    }
    out.resize(src.size() * (size_t)bytes);
    if (out.empty()) out.reserve(16);
  }
  const EncodedCycle& c_;
  oe_offer_packing pk_{};
  std::vector<float> scalarF_;
  std::vector<uint8_t> present_, mprof_, zone_, region_, attr_, rr_, exec_, tupleId_;
  std::vector<uint32_t> tuples_;
};

// One requirement as the encoder sees it: either a full pod pipeline or (for rule-level tests, the counterpart of
// calling PlacementRule.filter directly) just a rule.
struct EncoderRequirement {
  const PodInstanceRequirement* requirement = nullptr;
  std::shared_ptr<const ExistingPod> existing;  // set: getExistingEvaluationPipeline; null: getNewEvaluationPipeline
  // PodInfoBuilder.portsByTask (PodInfoBuilder.java:130-140): task spec name -> port name -> the port the stored task of that name
  // (unless permanently failed) advertises; a dynamic port that had a value keeps it (PortEvaluationStage.java:79-87), also when
  // the pod goes through the new-pod pipeline (recovery type PERMANENT without the failed label)
  std::map<std::string, std::map<std::string, int64_t>> priorPorts;
  // Multi-service cycles (S/scheduler/multi/MultiServiceEventClient.java:312-372): the service the requirement belongs to.  Its
  // placement rules count, and its pod is compared with, the tasks of the SAME service only (every service has its own state
  // store), so pod types, task names and task filters are namespaced by it.  "" = the single-service scheduler.
  std::string service;
  const SchedulerConfig* schedulerConfig = nullptr;  // the service's own executor overhead; null = the encoder's
  RulePtr ruleOnly;          // when set: the class is [PlacementRuleEvaluationStage(ruleOnly)]
  PodInstance podInstance;   // used with ruleOnly
};

class CycleEncoder : private RuleDictionary, private RuleEmitter {
 public:
  // `persistent`: dictionaries that outlive this cycle (ResidentOffers); null = the encoder's own, fresh ones
  explicit CycleEncoder(SchedulerConfig cfg = SchedulerConfig(), OfferDictionaries* persistent = nullptr)
      : cfg_(cfg), od_(persistent ? *persistent : ownDicts_), agents_(od_.agents), hosts_(od_.hosts), zones_(od_.zones), regions_(od_.regions), attrs_(od_.attrs),
        rrValues_(od_.rrValues), rrAttrs_(od_.rrAttrs), execs_(od_.execs), roles_(od_.roles), slots_(od_.slots), kinds_(od_.kinds), resids_(od_.resids),
        profiles_(od_.profiles) {}

  // `taskServices`: empty, or one service name per task (the state store the task came from)
  // The rows of a resident offer table as ResidentOffers hands them over: `slots` in row order (null = withdrawn offer), `dirty` =
  // the rows whose offer changed since the table was last sent, `shape` = {S, K, R, D, A, E, Q, has side table} of the table on the
  // device (all -1: nothing loaded yet).
  struct ResidentRows {
    const std::vector<const mesos::Offer*>* slots = nullptr;
    const std::vector<int64_t>* dirty = nullptr;
    int shape[8] = {-1, -1, -1, -1, -1, -1, -1, 0};
  };

  EncodedCycle encode(const std::vector<mesos::Offer>& offers, const std::vector<mesos::TaskInfo>& tasks,
                      const std::vector<EncoderRequirement>& reqs, const std::vector<std::string>& taskServices = {}) {
    std::vector<const mesos::Offer*> rows;
    rows.reserve(offers.size());
    for (auto& o : offers) rows.push_back(&o);
    return encodeRows(rows, nullptr, tasks, reqs, taskServices);
  }
  // The same cycle over a resident offer table: only the dirty rows are encoded (and later sent) when the table on the device
  // still has room for them; otherwise every row is, and EncodedCycle::full_reload says so.
  EncodedCycle encodeResident(const ResidentRows& resident, const std::vector<mesos::TaskInfo>& tasks, const std::vector<EncoderRequirement>& reqs,
                              const std::vector<std::string>& taskServices = {}) {
    return encodeRows(*resident.slots, &resident, tasks, reqs, taskServices);
  }

 private:
  EncodedCycle encodeRows(const std::vector<const mesos::Offer*>& allRows, const ResidentRows* resident, const std::vector<mesos::TaskInfo>& tasks,
                          const std::vector<EncoderRequirement>& reqs, const std::vector<std::string>& taskServices) {
    if (used_) throw std::logic_error("CycleEncoder: one encoder per cycle");
    used_ = true;
    // the offers whose strings this cycle has to look at: all of them, or — resident table — the ones that changed
    const bool tableLoaded = resident && resident->shape[0] >= 0;
    std::vector<const mesos::Offer*> offers;
    if (tableLoaded) { for (int64_t r : *resident->dirty) if (allRows[(size_t)r]) offers.push_back(allRows[(size_t)r]); }
    else for (auto* o : allRows) if (o) offers.push_back(o);
    if (!taskServices.empty() && taskServices.size() != tasks.size()) throw std::logic_error("CycleEncoder: one service name per task");
    auto serviceOf = [&](size_t t) -> const std::string& { static const std::string none; return taskServices.empty() ? none : taskServices[t]; };
    roles_.intern(ANY_ROLE);  // role slot 0
    slots_.intern(ANY_ROLE);
    for (const char* k : {"cpus", "mem", "disk", "gpus"}) kinds_.intern(k);

    // ---- pass 1: dictionaries
    for (auto* op : offers) {
      const mesos::Offer& o = *op;
      agents_.intern(o.slave_id);
      hosts_.intern(o.hostname);
      if (o.domain) { zones_.intern(o.domain->zone); regions_.intern(o.domain->region); }
      for (auto& a : o.attributes) attrs_.intern(AttributeStringUtils::toString(a));
      for (auto& e : o.executor_ids) execs_.intern(e);
      for (auto& r : o.resources) {
        if (r.allocation_role) roles_.intern(*r.allocation_role);
        if (r.value.type == mesos::Value::SCALAR && !ResourceUtils::isAtomic(r)) kinds_.intern(r.name);
        if (auto id = ResourceUtils::getResourceId(r)) resids_.intern(*id);
        else if (!ResourceUtils::isAtomic(r)) slots_.intern(ResourceUtils::getRole(r));
        if (r.disk && r.disk->profile) profiles_.intern(*r.disk->profile);
      }
    }
    for (size_t ti = 0; ti < tasks.size(); ++ti) {
      const mesos::TaskInfo& t = tasks[ti];
      names_.intern(ns(serviceOf(ti), t.name));
      agents_.intern(t.slave_id);
      if (auto v = label(t, labels::OFFER_HOSTNAME)) hosts_.intern(*v);
      if (auto v = label(t, labels::OFFER_ZONE)) zones_.intern(*v);
      if (auto v = label(t, labels::OFFER_REGION)) regions_.intern(*v);
      if (auto v = label(t, labels::TASK_TYPE)) types_.intern(ns(serviceOf(ti), *v));
      if (auto v = label(t, labels::OFFER_ATTRIBUTES)) for (auto& a : AttributeStringUtils::toStringList(*v)) attrs_.intern(a);
    }
    for (auto& r : reqs) {
      const PodInstance& pi = r.ruleOnly ? r.podInstance : r.requirement->podInstance;
      curService_ = r.service;
      types_.intern(ns(r.service, pi.pod->type));
      if (r.existing) {
        if (r.existing->executorId) execs_.intern(*r.existing->executorId);
        for (auto& res : r.existing->executorInfo.resources) if (auto id = ResourceUtils::getResourceId(res)) resids_.intern(*id);
        for (auto& [n, t] : r.existing->podTasks) for (auto& res : t.resources) if (auto id = ResourceUtils::getResourceId(res)) resids_.intern(*id);
      }
      if (r.ruleOnly) r.ruleOnly->declare(*this);
      else {
        if (pi.pod->placementRule) (*pi.pod->placementRule)->declare(*this);
        for (auto& ts : pi.pod->tasks) {
          names_.intern(ns(r.service, taskInstanceName(pi, ts.name)));
          for (auto& rs : ts.resourceSet.resources) {
            roles_.intern(rs.role); slots_.intern(rs.preReservedRole);
            if (!rs.isPort) kinds_.intern(rs.name);
          }
          for (auto& v : ts.resourceSet.volumes) { slots_.intern(v.preReservedRole); for (auto& p : v.profiles) profiles_.intern(p); }
        }
        for (auto& v : pi.pod->volumes) { slots_.intern(v.preReservedRole); for (auto& p : v.profiles) profiles_.intern(p); }
      }
    }
    if (slots_.size() > OE_MAX_ROLE_SLOTS) throw InvalidRequirementException("more reservable roles in one cycle than OE_MAX_ROLE_SLOTS");
    if (kinds_.size() > OE_MAX_SCALAR_KINDS) throw InvalidRequirementException("more scalar resource names than OE_MAX_SCALAR_KINDS");
    if (profiles_.size() > OE_MAX_PROFILES) throw InvalidRequirementException("more volume profiles than OE_MAX_PROFILES");
    if (rrAttrs_.size() > OE_MAX_RR_COLS) throw InvalidRequirementException("more round-robin attribute names than OE_MAX_RR_COLS");
    // round-robin value dictionary: the values of the named attributes, on offers and in task labels
    // (a round-robin attribute name that is new this cycle adds a column to the table: every row is looked at again, and re-sent)
    if (tableLoaded && rrAttrs_.size() != resident->shape[6]) { offers.clear(); for (auto* x : allRows) if (x) offers.push_back(x); }
    for (auto* o : offers) for (auto& a : o->attributes) if (rrColumn(a.name) >= 0) rrValues_.intern(AttributeStringUtils::toString(a.value));
    for (auto& t : tasks)
      if (auto v = label(t, labels::OFFER_ATTRIBUTES))
        for (auto& a : AttributeStringUtils::toStringList(*v)) {
          const size_t k = a.find(':');
          if (k != std::string::npos && rrColumn(a.substr(0, k)) >= 0) rrValues_.intern(a.substr(k + 1));
        }

    encodeOffers(allRows, resident);
    encodeTasks(tasks, taskServices);

    // ---- programs and requirements
    std::map<std::string, int> classOf;
    for (auto& r : reqs) {
      const PodInstance& pi = r.ruleOnly ? r.podInstance : r.requirement->podInstance;
      curService_ = r.service;
      curCfg_ = r.schedulerConfig ? r.schedulerConfig : &cfg_;
      std::string key;
      if (r.ruleOnly) key = "rule@" + std::to_string((uintptr_t)r.ruleOnly.get()) + "|" + r.service;
      else if (r.existing) key = "existing@" + std::to_string(out_.r_class.size());   // reservations are per pod instance
      else if (!r.priorPorts.empty()) key = "sticky@" + std::to_string(out_.r_class.size());  // remembered ports are per pod instance
      else {
        key = "pod@" + std::to_string((uintptr_t)pi.pod.get()) + "|" + r.service + "|" + std::to_string((int)r.requirement->recoveryType);
        for (auto& t : r.requirement->tasksToLaunch) key += "|" + t;
      }
      auto it = classOf.find(key);
      int cls;
      if (it != classOf.end()) cls = it->second;
      else {
        cls = r.ruleOnly ? addRuleOnlyClass(*r.ruleOnly) : r.existing ? addExistingPodClass(*r.requirement, *r.existing) : addNewPodClass(*r.requirement, r.priorPorts);
        classOf[key] = cls;
      }
      out_.r_class.push_back(cls);
      out_.r_type.push_back(types_.find(ns(r.service, pi.pod->type)));
      out_.r_index.push_back(pi.index);
      if (!r.ruleOnly)
        for (auto& ts : pi.pod->tasks) pendingLaunchNames_.push_back({r.service, taskInstanceName(pi, ts.name)});
      pendingLaunchEnd_.push_back((int32_t)pendingLaunchNames_.size());
    }
    out_.M = (int64_t)reqs.size();
    // launch rows (LaunchEvaluationStage stores every task of the pod, launched or not: :49-54)
    const int FW = std::max(1, (filters_.size() + 31) / 32);
    out_.FW = FW;
    for (size_t j = 0, l = 0; j < reqs.size(); ++j) {
      for (; l < (size_t)pendingLaunchEnd_[j]; ++l) {
        out_.l_name.push_back(names_.find(ns(pendingLaunchNames_[l].first, pendingLaunchNames_[l].second)));
        for (int w = 0; w < FW; ++w) out_.l_filter.push_back(filterWord(pendingLaunchNames_[l].first, pendingLaunchNames_[l].second, w));
      }
      out_.l_begin.push_back((int32_t)out_.l_name.size());
    }
    // task filter bits (filters are only known after the rules were emitted)
    out_.t_filter.assign((size_t)FW * std::max<int64_t>(out_.T, 1), 0u);
    for (int64_t t = 0; t < out_.T; ++t)
      for (int w = 0; w < FW; ++w) out_.t_filter[(size_t)w * out_.T + t] = filterWord(serviceOf((size_t)t), tasks[(size_t)t].name, w);

    oe_dict_sizes& d = out_.dict;
    d.n_agents = std::max(agents_.size(), 1); d.n_hostnames = std::max(hosts_.size(), 1); d.n_zones = zones_.size(); d.n_regions = regions_.size();
    d.n_attrs = attrs_.size(); d.n_rr_values = rrValues_.size(); d.n_pod_types = std::max(types_.size(), 1); d.n_task_names = names_.size();
    d.n_filters = std::max(filters_.size(), 1); d.n_exec_ids = execs_.size(); d.n_roles = roles_.size(); d.n_res_ids = resids_.size();
    return std::move(out_);
  }

 public:
  const Dictionary& hostnames() const { return hosts_; }
  const Dictionary& resourceIds() const { return resids_; }

 private:
  static std::optional<std::string> label(const mesos::TaskInfo& t, const char* key) {
    auto it = t.labels.find(key);
    if (it == t.labels.end()) return std::nullopt;
    return it->second;
  }
  static std::string lower(std::string s) { for (auto& c : s) c = (char)std::tolower((unsigned char)c); return s; }
  int rrColumn(const std::string& attrName) const { return rrAttrs_.find(lower(attrName)); }  // equalsIgnoreCase

  // a name inside a service's own namespace (the unit separator cannot occur in a Mesos id)
  static std::string ns(const std::string& service, const std::string& x) { return service.empty() ? x : service + '\x1f' + x; }
  uint32_t filterWord(const std::string& service, const std::string& taskName, int w) const {
    uint32_t bits = 0;
    for (int f = w * 32; f < std::min(filters_.size(), (w + 1) * 32); ++f)
      if (filterService_[(size_t)f] == service && filterMatchers_[(size_t)f]->matches(taskName)) bits |= 1u << (f & 31);
    return bits;
  }

  // ---- offers -------------------------------------------------------------------------------------------------------
  // One offer, reduced to what its table row holds — `new MesosResourcePool(offer, role)` (MesosResourcePool.java:53-147) plus the
  // placement getters — before the table's slot counts are known.
  struct OfferRow {
    std::vector<double> scalar;                       // [S * K]
    uint32_t present = 0;
    std::vector<IntervalSet> ports;                   // [S]
    std::vector<std::pair<double, int32_t>> mounts;   // (size, profile id)
    int32_t alloc = -1, agent = 0, host = 0, zone = -1, region = -1;
    std::vector<int32_t> attrs, rr, execs;
    std::vector<int32_t> rid, rkind;                  // id-carrying resources (dynamicallyReservedPoolByResourceId)
    std::vector<double> rval;
  };
  OfferRow rowOf(const mesos::Offer* ofp, int S, int K, int Q) const {
    OfferRow row;
    row.scalar.assign((size_t)S * K, 0.0);
    row.ports.resize((size_t)S);
    row.rr.assign((size_t)Q, -1);
    if (!ofp) return row;   // a withdrawn offer: no pool entry, no port, nothing fits it
    const mesos::Offer& of = *ofp;
    row.agent = agents_.find(of.slave_id);
    row.host = hosts_.find(of.hostname);
    if (of.domain) { row.zone = zones_.find(of.domain->zone); row.region = regions_.find(of.domain->region); }
    std::optional<std::string> alloc;
    for (auto& r : of.resources) {
      if (r.allocation_role) {
        if (alloc && *alloc != *r.allocation_role)
          throw InvalidRequirementException("offer " + of.id + " mixes allocation roles; the table carries one per offer");
        alloc = r.allocation_role;
      }
      if (auto id = ResourceUtils::getResourceId(r)) {  // dynamicallyReservedPoolByResourceId (MesosResourcePool.java:119-131)
        row.rid.push_back(resids_.find(*id));
        if (ResourceUtils::isAtomic(r)) { row.rkind.push_back(OE_KIND_MOUNT); row.rval.push_back(r.value.scalar); }
        else if (r.value.type == mesos::Value::RANGES) { row.rkind.push_back(OE_KIND_PORTS); row.rval.push_back(r.value.ranges.empty() ? 0.0 : (double)r.value.ranges[0].first); }
        else { row.rkind.push_back(kinds_.find(r.name)); row.rval.push_back(r.value.scalar); }
        continue;
      }
      if (ResourceUtils::isAtomic(r)) {  // unreservedAtomicPool (MesosResourcePool.java:99-117)
        if (r.name == "disk") row.mounts.push_back({r.value.scalar, r.disk->profile ? profiles_.find(*r.disk->profile) : -1});
        continue;
      }
      const int slot = slots_.find(ResourceUtils::getRole(r));  // reservableMergedPoolByRole (MesosResourcePool.java:432-455)
      if (r.value.type == mesos::Value::SCALAR) {
        const int q = slot * K + kinds_.find(r.name);
        const bool had = (row.present >> q) & 1u;
        // ValueUtils.add(zero, v) then + v2 ...: the sum runs in offer-resource order (MesosResourcePool.java:133-147)
        row.scalar[(size_t)q] = (had ? row.scalar[(size_t)q] : 0.0) + r.value.scalar;
        row.present |= 1u << q;
      } else if (r.value.type == mesos::Value::RANGES && r.name == "ports") {
        for (auto& iv : r.value.ranges) row.ports[(size_t)slot].add((int)iv.first, (int)iv.second);  // RangeUtils.java:65-67 narrows with (int)
      }
    }
    if (alloc) row.alloc = roles_.find(*alloc);
    for (auto& a : of.attributes) row.attrs.push_back(attrs_.find(AttributeStringUtils::toString(a)));
    for (int q = 0; q < Q; ++q)
      for (auto& a : of.attributes)
        if (rrColumn(a.name) == q) { row.rr[(size_t)q] = rrValues_.find(AttributeStringUtils::toString(a.value)); break; }  // first match wins
    for (auto& e : of.executor_ids) row.execs.push_back(execs_.find(e));
    return row;
  }

  // `resident` null: the plain table of all rows.  Otherwise: the dirty rows only, in the shape of the table on the device, when
  // they fit it; else all rows in a shape that holds everything (full_reload).
  void encodeOffers(const std::vector<const mesos::Offer*>& allRows, const ResidentRows* resident) {
    EncodedCycle& o = out_;
    o.S = std::max(1, slots_.size());
    o.K = std::max(4, kinds_.size());
    o.Q = rrAttrs_.size();
    o.resident = resident != nullptr;
    o.N_resident = (int64_t)allRows.size();
    const bool tableLoaded = resident && resident->shape[0] >= 0;
    std::vector<OfferRow> rows;
    std::vector<int64_t> which;   // table row of rows[k]
    int R = 1, D = 0, A = 1, E = 0;
    bool side = false;
    auto take = [&](int64_t r) {
      rows.push_back(rowOf(allRows[(size_t)r], o.S, o.K, o.Q));
      which.push_back(r);
      const OfferRow& x = rows.back();
      for (auto& p : x.ports) R = std::max<int>(R, (int)p.iv.size());
      D = std::max<int>(D, (int)x.mounts.size());
      A = std::max<int>(A, (int)x.attrs.size());
      E = std::max<int>(E, (int)x.execs.size());
      side = side || !x.rid.empty();
    };
    bool delta = tableLoaded && resident->shape[0] == o.S && resident->shape[1] == o.K && resident->shape[6] == o.Q && !resident->shape[7];
    if (delta) {
      for (int64_t r : *resident->dirty) take(r);
      // a changed offer with more port ranges / MOUNT disks / attributes / executor ids than the table has slots for, or with
      // reserved resources (oe_update_offers cannot patch the reserved side table): everything is sent again
      delta = R <= resident->shape[2] && D <= resident->shape[3] && A <= resident->shape[4] && E <= resident->shape[5] && !side;
      if (delta) { R = resident->shape[2]; D = resident->shape[3]; A = resident->shape[4]; E = resident->shape[5]; }
    }
    if (!delta) {
      rows.clear(); which.clear();
      R = tableLoaded ? std::max(1, resident->shape[2]) : 1; D = tableLoaded ? resident->shape[3] : 0;
      A = tableLoaded ? std::max(1, resident->shape[4]) : 1; E = tableLoaded ? resident->shape[5] : 0;
      for (int64_t r = 0; r < (int64_t)allRows.size(); ++r) take(r);
    }
    if (A > OE_MAX_ATTR_SLOTS || E > OE_MAX_EXEC_SLOTS) throw InvalidRequirementException("offer exceeds the attribute / executor-id slot limits");
    if (R > OE_MAX_PORT_RANGES || D > OE_MAX_MOUNT_SLOTS) throw InvalidRequirementException("offer exceeds the port-range / MOUNT-disk slot limits");
    o.full_reload = !delta;
    o.delta_rows = delta ? which : std::vector<int64_t>();
    const int64_t N = (int64_t)rows.size();
    o.N = N;
    o.R = R; o.D = D; o.A = A; o.E = E;
    const size_t n = (size_t)std::max<int64_t>(N, 1);
    o.o_scalar.assign((size_t)o.S * o.K * n, 0.0);
    o.o_present.assign(n, 0u);
    o.o_alloc.assign(n, -1); o.o_agent.assign(n, 0); o.o_host.assign(n, 0); o.o_zone.assign(n, -1); o.o_region.assign(n, -1);
    o.o_rbegin.assign((size_t)N + 1, 0);
    o.o_port.assign((size_t)o.S * R * n, OE_PORT_EMPTY);
    o.o_msize.assign((size_t)std::max(D, 1) * n, -std::numeric_limits<double>::infinity());
    o.o_mprof.assign((size_t)std::max(D, 1) * n, -1);
    o.o_attr.assign((size_t)A * n, -1);
    o.o_rr.assign((size_t)std::max(o.Q, 1) * n, -1);
    o.o_exec.assign((size_t)std::max(E, 1) * n, -1);
    for (int64_t i = 0; i < N; ++i) {
      const OfferRow& x = rows[(size_t)i];
      for (int q = 0; q < o.S * o.K; ++q) o.o_scalar[(size_t)q * n + (size_t)i] = x.scalar[(size_t)q];
      o.o_present[(size_t)i] = x.present;
      o.o_alloc[(size_t)i] = x.alloc; o.o_agent[(size_t)i] = std::max(x.agent, 0); o.o_host[(size_t)i] = std::max(x.host, 0);
      o.o_zone[(size_t)i] = x.zone; o.o_region[(size_t)i] = x.region;
      for (int s = 0; s < o.S; ++s)
        for (size_t r = 0; r < x.ports[(size_t)s].iv.size(); ++r) {
          auto iv = x.ports[(size_t)s].iv[r];
          o.o_port[((size_t)s * R + r) * n + (size_t)i] = OE_PORT_PACK(iv.first, iv.second);
        }
      for (size_t d = 0; d < x.mounts.size(); ++d) { o.o_msize[d * n + (size_t)i] = x.mounts[d].first; o.o_mprof[d * n + (size_t)i] = x.mounts[d].second; }
      for (size_t a = 0; a < x.attrs.size(); ++a) o.o_attr[a * n + (size_t)i] = x.attrs[a];
      for (int q = 0; q < o.Q; ++q) o.o_rr[(size_t)q * n + (size_t)i] = x.rr[(size_t)q];
      for (size_t e = 0; e < x.execs.size(); ++e) o.o_exec[e * n + (size_t)i] = x.execs[e];
      o.o_rid.insert(o.o_rid.end(), x.rid.begin(), x.rid.end());
      o.o_rkind.insert(o.o_rkind.end(), x.rkind.begin(), x.rkind.end());
      o.o_rval.insert(o.o_rval.end(), x.rval.begin(), x.rval.end());
      o.o_rbegin[(size_t)i + 1] = (int64_t)o.o_rid.size();
    }
  }

  // ---- tasks ---------------------------------------------------------------------------------------------------------
  void encodeTasks(const std::vector<mesos::TaskInfo>& tasks, const std::vector<std::string>& taskServices) {
    EncodedCycle& o = out_;
    auto serviceOf = [&](size_t t) -> const std::string& { static const std::string none; return taskServices.empty() ? none : taskServices[t]; };
    const int64_t T = (int64_t)tasks.size();
    o.T = T;
    int TA = std::max(1, o.A);
    std::vector<std::vector<std::string>> attrStrings((size_t)T);
    for (int64_t t = 0; t < T; ++t) {
      if (auto v = label(tasks[(size_t)t], labels::OFFER_ATTRIBUTES)) attrStrings[(size_t)t] = AttributeStringUtils::toStringList(*v);
      TA = std::max<int>(TA, (int)attrStrings[(size_t)t].size());
    }
    if (TA > OE_MAX_ATTR_SLOTS) throw InvalidRequirementException("task exceeds the attribute slot limit");
    o.TA = TA;
    const size_t n = (size_t)std::max<int64_t>(T, 1);
    o.t_name.assign(n, 0); o.t_type.assign(n, -1); o.t_index.assign(n, OE_NO_INDEX); o.t_agent.assign(n, 0);
    o.t_host.assign(n, -1); o.t_zone.assign(n, -1); o.t_region.assign(n, -1);
    o.t_attr.assign((size_t)TA * n, -1);
    o.t_rr.assign((size_t)std::max(o.Q, 1) * n, -1);
    for (int64_t t = 0; t < T; ++t) {
      const mesos::TaskInfo& ti = tasks[(size_t)t];
      o.t_name[(size_t)t] = names_.find(ns(serviceOf((size_t)t), ti.name));
      o.t_agent[(size_t)t] = agents_.find(ti.slave_id);
      if (auto v = label(ti, labels::TASK_TYPE)) o.t_type[(size_t)t] = types_.find(ns(serviceOf((size_t)t), *v));
      if (auto v = label(ti, labels::TASK_INDEX)) {  // Integer.parseInt (TaskLabelReader.java:57-59); unparsable = not equivalent to anything
        char* end = nullptr;
        const long x = std::strtol(v->c_str(), &end, 10);
        if (!v->empty() && end && *end == 0 && x >= INT32_MIN + 1 && x <= INT32_MAX) o.t_index[(size_t)t] = (int32_t)x;
      }
      if (auto v = label(ti, labels::OFFER_HOSTNAME)) o.t_host[(size_t)t] = hosts_.find(*v);
      if (auto v = label(ti, labels::OFFER_ZONE)) o.t_zone[(size_t)t] = zones_.find(*v);
      if (auto v = label(ti, labels::OFFER_REGION)) o.t_region[(size_t)t] = regions_.find(*v);
      for (size_t a = 0; a < attrStrings[(size_t)t].size(); ++a) o.t_attr[a * n + (size_t)t] = attrs_.find(attrStrings[(size_t)t][a]);
      for (int q = 0; q < o.Q; ++q)
        for (auto& a : attrStrings[(size_t)t]) {
          const size_t k = a.find(':');
          if (k == std::string::npos) throw std::invalid_argument("Unable to split attribute into name:value elements: " + a);  // AttributeStringUtils.split
          if (rrColumn(a.substr(0, k)) == q) { o.t_rr[(size_t)q * n + (size_t)t] = rrValues_.find(a.substr(k + 1)); break; }
        }
    }
  }

  // ---- classes -------------------------------------------------------------------------------------------------------
  oe_stage_op op(int opcode, int kind = 0, int slot = 0, int arg = 0, double amount = 0.0, uint64_t mask = 0, uint64_t aux = 0) {
    oe_stage_op x{};
    x.opcode = (uint8_t)opcode; x.kind = (uint8_t)kind; x.slot = (uint8_t)slot; x.arg = arg; x.amount = amount; x.mask = mask; x.aux = aux;
    return x;
  }
  void pushRule(const PlacementRule& rule, std::vector<oe_stage_op>& ops) {
    const int begin = (int)out_.nodes.size();
    rule.emit(*this);
    const int count = (int)out_.nodes.size() - begin;
    if (count > OE_MAX_RULE_NODES) throw InvalidRequirementException("placement rule larger than OE_MAX_RULE_NODES");
    ops.push_back(op(OE_OP_RULE, 0, 0, begin, 0.0, 0, (uint64_t)count));
  }
  int finishClass(std::vector<oe_stage_op>& ops, std::vector<StageInfo>& stages, int podRole, const std::vector<int>& staticPorts) {
    if (ops.size() > OE_MAX_STAGE_OPS) throw InvalidRequirementException("pipeline longer than OE_MAX_STAGE_OPS");
    out_.ops.insert(out_.ops.end(), ops.begin(), ops.end());
    out_.op_begin.push_back((int32_t)out_.ops.size());
    out_.pod_role.push_back(podRole);
    std::set<int> excl(staticPorts.begin(), staticPorts.end());
    for (int p : excl) out_.excl.push_back((uint16_t)p);
    out_.excl_begin.push_back((int32_t)out_.excl.size());
    out_.classStages.push_back(stages);
    return (int)out_.pod_role.size() - 1;
  }
  int addRuleOnlyClass(const PlacementRule& rule) {
    std::vector<oe_stage_op> ops;
    std::vector<StageInfo> stages(1);
    stages[0].kind = StageInfo::PLACEMENT;
    pushRule(rule, ops);
    return finishClass(ops, stages, -1, {});
  }

  // PortEvaluationStage.requireHostPorts (PortEvaluationStage.java:325-334): no network = host network; otherwise only
  // networks that map ports (S/dcos/DcosConstants.java: the "mesos-bridge" network) need host ports
  static bool requireHostPorts(const std::vector<std::string>& networkNames) {
    if (networkNames.empty()) return true;
    for (auto& n : networkNames) if (n == "mesos-bridge") return true;
    return false;
  }

  // PodInfoBuilder's assignedOverlayPorts (PodInfoBuilder.java:118-122): every explicitly requested port of the pod is taken
  void resetOverlayPorts(const PodSpec& pod) {
    overlayAssigned_.clear();
    for (auto& ts : pod.tasks) for (auto& rs : ts.resourceSet.resources) if (rs.isPort && rs.port > 0) overlayAssigned_.insert(rs.port);
  }
  // PortEvaluationStage.selectOverlayPort (:300-313): the next free port of 1025..2025 (S/dcos/DcosConstants.java), offer-independent
  int selectOverlayPort() {
    for (int p = 1025; p <= 2025; ++p) if (overlayAssigned_.insert(p).second) return p;
    throw InvalidRequirementException("no overlay port left in 1025..2025");  // the stage would fail on every offer
  }
  using PriorPorts = std::map<std::string, std::map<std::string, int64_t>>;
  static int priorPortFor(const PriorPorts& prior, const std::vector<std::string>& taskNames, const std::string& portName) {
    for (auto& n : taskNames) {  // getTaskNames().stream().map(getPriorPortForTask)...findAny() (PortEvaluationStage.java:80-84)
      auto pt = prior.find(n);
      if (pt == prior.end()) continue;
      auto pp = pt->second.find(portName);
      if (pp != pt->second.end()) return (int)pp->second;
    }
    return 0;
  }

  // getNewEvaluationPipeline (OfferEvaluator.java:411-522)
  int addNewPodClass(const PodInstanceRequirement& req, const PriorPorts& priorPorts = {}) {
    const PodSpec& pod = *req.podInstance.pod;
    std::vector<oe_stage_op> ops;
    std::vector<StageInfo> stages;
    resetOverlayPorts(pod);
    auto stage = [&](StageInfo s, oe_stage_op o) { stages.push_back(std::move(s)); ops.push_back(o); };
    { StageInfo s; s.kind = StageInfo::EXECUTOR; stage(s, op(OE_OP_EXEC_ID, 0, 0, -1)); }                                  // :417
    if (pod.placementRule) { StageInfo s; s.kind = StageInfo::PLACEMENT; stages.push_back(s); pushRule(**pod.placementRule, ops); }   // :419-422
    for (auto& v : pod.volumes) volumeStage(v, {}, stage);                                                                    // :424-427
    for (auto& ts : pod.tasks) if (ts.usesTLS) { StageInfo s; s.kind = StageInfo::TLS; s.taskSpecName = ts.name; stage(s, op(OE_OP_NOP)); }  // :431-437
    // static ports declared anywhere in the pod are never picked dynamically (PortEvaluationStage.java:252-261)
    std::vector<int> staticPorts;
    for (auto& ts : pod.tasks) for (auto& rs : ts.resourceSet.resources) if (rs.isPort && rs.port != 0) staticPorts.push_back(rs.port);

    std::map<std::string, const TaskSpec*> byName;  // TreeMap: task spec name order (:439-449)
    for (auto& ts : pod.tasks) {
      if (!byName.emplace(ts.name, &ts).second) throw std::logic_error("Duplicate key " + ts.name);
    }
    bool addedExecutorResources = false;
    std::set<std::string> addedResourceSets;
    for (auto& [taskSpecName, ts] : byName) {
      const ResourceSet& rset = ts->resourceSet;
      // getOrderedResourceSpecs (:380-403): static ports, dynamic ports, everything else
      std::vector<const ResourceSpec*> ordered;
      for (auto& r : rset.resources) if (r.isPort && r.port != 0) ordered.push_back(&r);
      for (auto& r : rset.resources) if (r.isPort && r.port == 0) ordered.push_back(&r);
      for (auto& r : rset.resources) if (!r.isPort) ordered.push_back(&r);
      if (!addedExecutorResources) {  // :460-478
        addedExecutorResources = true;
        if (ordered.empty()) throw InvalidRequirementException("resource set " + rset.id + " has no resources");  // resourceSpecs.get(0)
        for (auto& [name, value] : curCfg_->getExecutorResources()) {
          StageInfo s;
          s.kind = StageInfo::RESOURCE; s.resourceName = name; s.value = value;
          s.role = ordered[0]->role; s.preReservedRole = ordered[0]->preReservedRole; s.principal = ordered[0]->principal;
          stage(s, op(OE_OP_SCALAR_NEW, kinds_.find(name), slots_.find(ordered[0]->preReservedRole), 0, value));
        }
      }
      if (!addedResourceSets.count(rset.id)) {  // :480-513
        addedResourceSets.insert(rset.id);
        std::vector<std::string> taskNames;
        for (auto& [n2, t2] : byName) if (t2->resourceSet.id == rset.id) taskNames.push_back(n2);
        for (const ResourceSpec* r : ordered) {
          StageInfo s;
          s.taskNames = taskNames; s.role = r->role; s.preReservedRole = r->preReservedRole; s.principal = r->principal;
          const int slot = slots_.find(r->preReservedRole);
          if (r->isPort) {
            s.kind = StageInfo::PORT; s.resourceName = "ports"; s.hostPort = requireHostPorts(r->networkNames);
            s.portName = r->portName;
            int port = r->port;
            if (port == 0) port = priorPortFor(priorPorts, taskNames, r->portName);           // PortEvaluationStage.java:79-87
            if (port == 0 && !s.hostPort) port = selectOverlayPort();
            s.port = port;
            if (!s.hostPort) stage(s, op(OE_OP_NOP));                                           // PortEvaluationStage.java:145-154
            else if (port != 0) {
              if (port < 0 || port > 65535) throw InvalidRequirementException("port out of range");
              stage(s, op(OE_OP_PORT_STATIC, 0, slot, port));
            } else {
              int rangeSet = -1;
              if (!r->ranges.empty()) {
                for (auto& rg : r->ranges) out_.rs.push_back(OE_PORT_PACK(std::clamp(rg.begin, 0, 65535), std::clamp(rg.end, 0, 65535)));
                out_.rs_begin.push_back((int32_t)out_.rs.size());
                rangeSet = (int)out_.rs_begin.size() - 2;
              }
              stage(s, op(OE_OP_PORT_DYNAMIC, 0, slot, rangeSet));
            }
          } else {
            s.kind = StageInfo::RESOURCE; s.resourceName = r->name; s.value = r->value;
            stage(s, op(OE_OP_SCALAR_NEW, kinds_.find(r->name), slot, 0, r->value));
          }
        }
        for (auto& v : rset.volumes) volumeStage(v, taskNames, stage);  // :509-512
      }
      StageInfo s;  // :515-518
      s.kind = StageInfo::LAUNCH; s.taskSpecName = taskSpecName;
      s.shouldLaunch = std::find(req.tasksToLaunch.begin(), req.tasksToLaunch.end(), taskSpecName) != req.tasksToLaunch.end();
      stage(s, op(OE_OP_NOP));
    }
    // OfferEvaluationUtils.getRole(podSpec) (OfferEvaluationUtils.java:229-235): role of the first resource spec
    int podRole = -1;
    for (auto& ts : pod.tasks) { if (!ts.resourceSet.resources.empty()) { podRole = roles_.find(ts.resourceSet.resources[0].role); break; } }
    return finishClass(ops, stages, podRole, staticPorts);
  }

  // ---- getExistingEvaluationPipeline (OfferEvaluator.java:538-728) with ExecutorResourceMapper.java:66-129,
  // TaskResourceMapper.java:80-194 and ResourceMapperUtils.java:26-72
  struct SpecRef {  // a ResourceSpec or a VolumeSpec ("disk"), as the mappers' `remainingResourceSpecs` hold them
    const ResourceSpec* r = nullptr;
    const VolumeSpec* v = nullptr;
    const std::string name() const { return v ? std::string("disk") : r->name; }
  };
  int slotOfPreviousRole(const mesos::Resource& r) {  // MesosResource.getPreviousRole (MesosResource.java:64-83)
    const std::string role = r.reservations.size() >= 2 ? r.reservations[r.reservations.size() - 2].role : std::string(ANY_ROLE);
    const int s = slots_.find(role);
    if (s < 0) throw InvalidRequirementException("orphaned reservation returns to role '" + role + "', which no table of this cycle carries");
    return s;
  }
  template <class StageFn>
  void unreserveStage(const mesos::Resource& r, StageFn& stage) {  // UnreserveEvaluationStage.java:24-36 -> MesosResourcePool.free
    StageInfo s;
    s.kind = StageInfo::UNRESERVE; s.orphan = r; s.resourceName = r.name;
    const int resid = resids_.find(*ResourceUtils::getResourceId(r));
    if (ResourceUtils::isAtomic(r)) {  // freeAtomicResource (:408-430): the bare disk rejoins the MOUNT list
      stage(s, op(OE_OP_FREE, OE_KIND_MOUNT, 0, resid, r.value.scalar, 0, (uint64_t)((r.disk->profile ? profiles_.intern(*r.disk->profile) : -1) + 1)));
    } else if (r.value.type == mesos::Value::RANGES) {
      if (r.name != "ports" || r.value.ranges.size() != 1 || r.value.ranges[0].first != r.value.ranges[0].second)
        throw InvalidRequirementException("returning an orphaned multi-port reservation to the pool is not supported");
      stage(s, op(OE_OP_FREE, OE_KIND_PORTS, slotOfPreviousRole(r), resid, 0.0, 0, (uint64_t)r.value.ranges[0].first));
    } else if (r.value.type == mesos::Value::SCALAR) {
      stage(s, op(OE_OP_FREE, kinds_.find(r.name), slotOfPreviousRole(r), resid, r.value.scalar));
    } else {
      throw InvalidRequirementException("orphaned " + r.name + " reservation of an unsupported value type");
    }
  }
  template <class StageFn>
  void updateStage(const SpecRef& spec, const mesos::Resource& prior, const std::vector<std::string>& taskNames, const ExistingPod& ex,
                   const std::map<std::string, int64_t>& priorPorts, StageFn& stage) {
    StageInfo s;
    s.taskNames = taskNames;
    s.resourceId = ResourceUtils::getResourceId(prior);
    const int resid = resids_.find(*s.resourceId);
    const bool executorLevel = taskNames.empty();
    uint8_t flags = 0;
    uint64_t aux = 0;
    if (executorLevel && ex.executorId) { s.skipIfRunningExecutor = true; s.executorId = *ex.executorId; flags = OE_OPF_SKIP_IF_RUNNING_EXECUTOR; aux = (uint64_t)execs_.find(*ex.executorId); }
    if (spec.v) {  // VolumeEvaluationStage.getExisting: the size is the one already reserved (ResourceMapperUtils.java:40-48)
      s.kind = StageInfo::VOLUME; s.resourceName = "disk"; s.volumeType = spec.v->type; s.containerPath = spec.v->containerPath;
      s.value = prior.disk ? prior.value.scalar : spec.v->size;
      s.role = spec.v->role; s.preReservedRole = spec.v->preReservedRole; s.principal = spec.v->principal;
      if (prior.disk) s.persistenceId = prior.disk->persistence_id;
      // the running-executor shortcut of volumes additionally needs a persistence id (VolumeEvaluationStage.java:126-154)
      if (!(s.persistenceId && executorLevel && ex.executorId)) { flags = 0; aux = 0; s.skipIfRunningExecutor = false; }
      oe_stage_op o = spec.v->type == VolumeSpec::ROOT ? op(OE_OP_SCALAR_RESERVED, kinds_.find("disk"), slots_.find(spec.v->preReservedRole), resid, s.value)
                                                       : op(OE_OP_MOUNT_RESERVED, 0, 0, resid);
      o.flags = flags; o.aux = aux;
      stage(s, o);
    } else if (spec.r->isPort) {  // PortEvaluationStage with a resource id (PortEvaluationStage.java:75-127)
      s.kind = StageInfo::PORT; s.resourceName = "ports"; s.hostPort = requireHostPorts(spec.r->networkNames); s.portName = spec.r->portName;
      s.role = spec.r->role; s.preReservedRole = spec.r->preReservedRole; s.principal = spec.r->principal;
      int port = spec.r->port;
      if (port == 0) { auto it = priorPorts.find(spec.r->portName); if (it != priorPorts.end()) port = (int)it->second; }
      if (port == 0 && !s.hostPort) port = selectOverlayPort();
      s.port = port;
      if (!s.hostPort) { stage(s, op(OE_OP_NOP)); return; }
      stage(s, op(OE_OP_PORT_RESERVED, 0, slots_.find(spec.r->preReservedRole), resid, 0.0, 0, (uint64_t)port));
    } else {  // ResourceEvaluationStage with a resource id (ResourceEvaluationStage.java:62-122)
      s.kind = StageInfo::RESOURCE; s.resourceName = spec.r->name; s.value = spec.r->value;
      s.role = spec.r->role; s.preReservedRole = spec.r->preReservedRole; s.principal = spec.r->principal;
      oe_stage_op o = op(OE_OP_SCALAR_RESERVED, kinds_.find(spec.r->name), slots_.find(spec.r->preReservedRole), resid, spec.r->value);
      o.flags = flags; o.aux = aux;
      stage(s, o);
    }
  }
  template <class StageFn>
  void createStage(const SpecRef& spec, const std::vector<std::string>& taskNames, const std::map<std::string, int64_t>& priorPorts, StageFn& stage) {
    if (spec.v) { volumeStage(*spec.v, taskNames, stage); return; }
    StageInfo s;
    s.taskNames = taskNames; s.role = spec.r->role; s.preReservedRole = spec.r->preReservedRole; s.principal = spec.r->principal;
    const int slot = slots_.find(spec.r->preReservedRole);
    if (spec.r->isPort) {
      s.kind = StageInfo::PORT; s.resourceName = "ports"; s.port = spec.r->port; s.hostPort = requireHostPorts(spec.r->networkNames); s.portName = spec.r->portName;
      int port = spec.r->port;
      if (port == 0) { auto it = priorPorts.find(spec.r->portName); if (it != priorPorts.end()) port = (int)it->second; }  // PortEvaluationStage.java:79-87
      if (port == 0 && !s.hostPort) port = selectOverlayPort();
      if (!s.hostPort) { s.port = port; stage(s, op(OE_OP_NOP)); }
      else if (port != 0) { s.port = port; stage(s, op(OE_OP_PORT_STATIC, 0, slot, port)); }
      else {
        int rangeSet = -1;
        if (!spec.r->ranges.empty()) {
          for (auto& rg : spec.r->ranges) out_.rs.push_back(OE_PORT_PACK(std::clamp(rg.begin, 0, 65535), std::clamp(rg.end, 0, 65535)));
          out_.rs_begin.push_back((int32_t)out_.rs.size());
          rangeSet = (int)out_.rs_begin.size() - 2;
        }
        stage(s, op(OE_OP_PORT_DYNAMIC, 0, slot, rangeSet));
      }
    } else {
      s.kind = StageInfo::RESOURCE; s.resourceName = spec.r->name; s.value = spec.r->value;
      stage(s, op(OE_OP_SCALAR_NEW, kinds_.find(spec.r->name), slot, 0, spec.r->value));
    }
  }

  int addExistingPodClass(const PodInstanceRequirement& req, const ExistingPod& ex) {
    const PodSpec& pod = *req.podInstance.pod;
    std::vector<oe_stage_op> ops;
    std::vector<StageInfo> stages;
    auto stage = [&](StageInfo s, oe_stage_op o) { stages.push_back(std::move(s)); ops.push_back(o); };
    resetOverlayPorts(pod);
    { StageInfo s; s.kind = StageInfo::EXECUTOR; if (ex.executorId) s.executorId = *ex.executorId;
      stage(s, op(OE_OP_EXEC_ID, 0, 0, ex.executorId ? execs_.find(*ex.executorId) : -1)); }                     // OfferEvaluator.java:295-304
    for (auto& ts : pod.tasks) if (ts.usesTLS) { StageInfo s; s.kind = StageInfo::TLS; s.taskSpecName = ts.name; stage(s, op(OE_OP_NOP)); }  // :548-554
    // (a PERMANENT recovery never comes here: it takes the new-pod pipeline, :268-271)
    std::vector<int> staticPorts;
    for (auto& ts : pod.tasks) for (auto& rs : ts.resourceSet.resources) if (rs.isPort && rs.port != 0) staticPorts.push_back(rs.port);

    // ---- executor resources (ExecutorResourceMapper.java:66-129): volumes first, then the overhead specs
    const ResourceSpec* any = nullptr;
    for (auto& ts : pod.tasks) if (!ts.resourceSet.resources.empty()) { any = &ts.resourceSet.resources[0]; break; }
    if (!any) throw InvalidRequirementException("pod " + pod.type + " has no resource spec to take role and principal from");
    std::vector<ResourceSpec> executorSpecs;
    for (auto& [name, value] : curCfg_->getExecutorResources()) executorSpecs.push_back(ResourceSpec::scalar(name, value, any->role, any->preReservedRole, any->principal));
    {
      std::vector<SpecRef> remaining;
      for (auto& v : pod.volumes) { SpecRef s; s.v = &v; remaining.push_back(s); }
      for (auto& r : executorSpecs) { SpecRef s; s.r = &r; remaining.push_back(s); }
      std::vector<std::pair<SpecRef, const mesos::Resource*>> matched;
      std::vector<const mesos::Resource*> orphaned;
      for (auto& res : ex.executorInfo.resources) {
        int hit = -1;
        if (ResourceUtils::getResourceId(res)) {
          for (size_t k = 0; k < remaining.size() && hit < 0; ++k) {
            if (res.name == "disk" && res.disk) { if (remaining[k].v && remaining[k].v->containerPath == res.disk->container_path) hit = (int)k; }
            else if (remaining[k].name() == res.name) hit = (int)k;
          }
        }
        if (hit >= 0) { matched.push_back({remaining[(size_t)hit], &res}); remaining.erase(remaining.begin() + hit); }
        else if (res.disk) orphaned.push_back(&res);
      }
      for (auto* o : orphaned) { StageInfo s; s.kind = StageInfo::DESTROY; s.orphan = *o; stage(s, op(OE_OP_NOP)); }   // DestroyEvaluationStage
      for (auto* o : orphaned) unreserveStage(*o, stage);
      for (auto& [spec, res] : matched) updateStage(spec, *res, {}, ex, {}, stage);
      for (auto& spec : remaining) createStage(spec, {}, {}, stage);
    }

    // ---- task resources (getExistingTaskEvaluationPipeline, OfferEvaluator.java:598-668)
    std::map<std::string, const TaskSpec*> byName;
    for (auto& ts : pod.tasks) if (!byName.emplace(ts.name, &ts).second) throw std::logic_error("Duplicate key " + ts.name);
    std::set<std::string> updatedResourceSets;
    for (auto& launchName : req.tasksToLaunch) {
      auto it = byName.find(launchName);
      if (it == byName.end()) throw std::logic_error("Unable to find task to launch " + launchName + " among the defined tasks of pod " + pod.type);
      const ResourceSet& rset = it->second->resourceSet;
      if (!updatedResourceSets.insert(rset.id).second) continue;
      std::vector<std::string> namesInSet;  // TreeSet order
      for (auto& [n2, t2] : byName) if (t2->resourceSet.id == rset.id) namesInSet.push_back(n2);
      const mesos::TaskInfo* prior = nullptr;
      for (auto& n2 : namesInSet) {
        auto pt = ex.podTasks.find(taskInstanceName(req.podInstance, n2));
        if (pt != ex.podTasks.end()) { prior = &pt->second; break; }
      }
      if (prior) {  // TaskResourceMapper.java:80-139; without a stored TaskInfo the reference adds no resource stages at all (:690-697)
        std::vector<SpecRef> remaining;
        for (auto& r : rset.resources) { SpecRef s; s.r = &r; remaining.push_back(s); }
        for (auto& v : rset.volumes) { SpecRef s; s.v = &v; remaining.push_back(s); }
        std::vector<std::pair<SpecRef, const mesos::Resource*>> matched;
        std::vector<const mesos::Resource*> orphaned;
        for (auto& res : prior->resources) {
          int hit = -1;
          const bool hasId = ResourceUtils::getResourceId(res).has_value();
          for (size_t k = 0; k < remaining.size() && hit < 0 && hasId; ++k) {
            const SpecRef& sp = remaining[k];
            if (res.name == "disk") { if (sp.v && res.disk && sp.v->containerPath == res.disk->container_path) hit = (int)k; }
            else if (res.name == "ports") {  // findMatchingPortSpec (:141-194): single-port resources only
              if (res.value.ranges.size() != 1 || res.value.ranges[0].first != res.value.ranges[0].second || !sp.r || !sp.r->isPort) continue;
              const int64_t p = res.value.ranges[0].first;
              if (sp.r->port == 0) { auto pp = prior->discovery_ports.find(sp.r->portName); if (pp != prior->discovery_ports.end() && pp->second == p) hit = (int)k; }
              else if (sp.r->port == p) hit = (int)k;
            } else if (sp.name() == res.name) hit = (int)k;
          }
          if (hit >= 0) { matched.push_back({remaining[(size_t)hit], &res}); remaining.erase(remaining.begin() + hit); }
          else orphaned.push_back(&res);
        }
        for (auto* o : orphaned) unreserveStage(*o, stage);
        for (auto& [spec, res] : matched) updateStage(spec, *res, namesInSet, ex, prior->discovery_ports, stage);
        for (auto& spec : remaining) createStage(spec, namesInSet, prior->discovery_ports, stage);
      }
      for (auto& n2 : namesInSet) {
        StageInfo s;
        s.kind = StageInfo::LAUNCH; s.taskSpecName = n2;
        s.shouldLaunch = std::find(req.tasksToLaunch.begin(), req.tasksToLaunch.end(), n2) != req.tasksToLaunch.end();
        stage(s, op(OE_OP_NOP));
      }
    }
    int podRole = -1;
    for (auto& ts : pod.tasks) { if (!ts.resourceSet.resources.empty()) { podRole = roles_.find(ts.resourceSet.resources[0].role); break; } }
    return finishClass(ops, stages, podRole, staticPorts);
  }

  template <class StageFn>
  void volumeStage(const VolumeSpec& v, const std::vector<std::string>& taskNames, StageFn& stage) {
    StageInfo s;
    s.kind = StageInfo::VOLUME; s.resourceName = "disk"; s.value = v.size; s.volumeType = v.type; s.containerPath = v.containerPath;
    s.taskNames = taskNames; s.role = v.role; s.preReservedRole = v.preReservedRole; s.principal = v.principal;
    if (v.type == VolumeSpec::ROOT) {  // VolumeEvaluationStage.java:155-181
      stage(s, op(OE_OP_SCALAR_NEW, kinds_.find("disk"), slots_.find(v.preReservedRole), 0, v.size));
    } else {                           // consumeAtomic (:182-186, MesosResourcePool.java:282-328)
      uint64_t mask = 0;
      for (auto& p : v.profiles) mask |= 1ull << profiles_.find(p);
      stage(s, op(OE_OP_VOL_MOUNT, 0, 0, 0, v.size, mask));
    }
  }

  // ---- RuleDictionary ------------------------------------------------------------------------------------------------
  void needAgent(const std::string& a) override { agents_.intern(a); }
  void needRegion(const std::string& r) override { regions_.intern(r); }
  void needPodType(const std::string& t) override { types_.intern(ns(curService_, t)); }
  void needRoundRobinAttribute(const std::string& name) override { rrAttrs_.intern(lower(name)); }
  void needTaskFilter(const StringMatcher&) override {}

  // ---- RuleEmitter ---------------------------------------------------------------------------------------------------
  void node(int opc, int a, int b, int c) override { out_.nodes.push_back(oe_rule_node{opc, a, b, c}); }
  int matcherOver(const Dictionary& dict, const char* tag, const StringMatcher& m) {
    const std::string key = std::string(tag) + m.toString();
    auto it = matcherCache_.find(key);
    if (it != matcherCache_.end()) return it->second;
    const int off = (int)out_.mbits.size();
    const int words = std::max(1, (dict.size() + 31) / 32);
    out_.mbits.resize(out_.mbits.size() + (size_t)words, 0u);
    for (int v = 0; v < dict.size(); ++v)
      if (m.matches(dict.values()[(size_t)v])) out_.mbits[(size_t)off + (v >> 5)] |= 1u << (v & 31);
    matcherCache_[key] = off;
    return off;
  }
  int matcherOverAttributes(const StringMatcher& m) override { return matcherOver(attrs_, "attr:", m); }
  int matcherOverHostnames(const StringMatcher& m) override { return matcherOver(hosts_, "host:", m); }
  int matcherOverZones(const StringMatcher& m) override { return matcherOver(zones_, "zone:", m); }
  int matcherOverRegions(const StringMatcher& m) override { return matcherOver(regions_, "region:", m); }
  int counterSet(int keyspace, const StringMatcher& taskFilter, const std::string& rrAttributeName) override {
    const std::string fkey = ns(curService_, taskFilter.toString());
    int f = filters_.find(fkey);
    if (f < 0) {
      f = filters_.intern(fkey);
      // keep our own copy of the matcher: the rule object may not outlive the encoder call in every caller
      filterMatchers_.push_back(cloneMatcher(taskFilter));
      filterService_.push_back(curService_);
    }
    const int col = keyspace == OE_KS_RR_ATTR ? rrColumn(rrAttributeName) : 0;
    for (size_t c = 0; c < out_.csets.size(); ++c)
      if (out_.csets[c].keyspace == keyspace && out_.csets[c].rr_col == col && out_.csets[c].filter_id == f) return (int)c;
    out_.csets.push_back(oe_counter_set{keyspace, col, f, 0});
    return (int)out_.csets.size() - 1;
  }
  static MatcherPtr cloneMatcher(const StringMatcher& m) {
    if (auto* r = dynamic_cast<const RegexMatcher*>(&m)) return RegexMatcher::create(r->pattern);
    if (auto* e = dynamic_cast<const ExactMatcher*>(&m)) { auto p = std::make_shared<ExactMatcher>(""); p->str = e->str; return p; }
    return AnyMatcher::create();
  }
  int typeSet(const std::string& podType) override {
    const int t = types_.find(ns(curService_, podType));
    for (size_t s = 0; s < out_.tsets.size(); ++s) if (out_.tsets[s] == t) return (int)s;
    out_.tsets.push_back(t);
    return (int)out_.tsets.size() - 1;
  }
  int agentId(const std::string& slaveId) override { return agents_.find(slaveId); }
  int regionId(const std::string& region) override { return regions_.find(region); }

  SchedulerConfig cfg_;
  EncodedCycle out_;
  bool used_ = false;
  OfferDictionaries ownDicts_;
  OfferDictionaries& od_;
  Dictionary &agents_, &hosts_, &zones_, &regions_, &attrs_, &rrValues_, &rrAttrs_, &execs_, &roles_, &slots_, &kinds_, &resids_, &profiles_;
  Dictionary types_, names_, filters_;
  std::vector<MatcherPtr> filterMatchers_;
  std::vector<std::string> filterService_;  // the service whose tasks filter f looks at
  std::string curService_;                  // the service of the requirement being declared / emitted
  const SchedulerConfig* curCfg_ = &cfg_;
  std::unordered_map<std::string, int> matcherCache_;
  std::set<int> overlayAssigned_;
  std::vector<std::pair<std::string, std::string>> pendingLaunchNames_;  // (service, task name)
  std::vector<int32_t> pendingLaunchEnd_;
};

}  // namespace sdk
