// offer_evaluator.hpp — the reference-facing entry points, backed by the C ABI of include/oe_abi.h:
//   OfferEvaluator::evaluate(requirement, offers)        S/offer/evaluate/OfferEvaluator.java:113-248
//   PlanScheduler::resourceOffers(offers, steps)         S/scheduler/plan/PlanScheduler.java:50-70
//   PlacementRuleFilter::filter(rule, offer, pod, tasks)  P/PlacementRule.java:38 (as the unit tests call it)
// Same names, argument meaning and error behaviour as the reference (S/ = sdk/scheduler/src/main/java/com/mesosphere/
// sdk/).  Matching is done by the engine; this layer encodes the inputs (encoder.hpp) and, for the ONE winning offer per
// requirement, turns the engine's decision back into the recommendation list the reference would have produced
// (RESERVE / CREATE / LAUNCH_GROUP / StoreTaskInfo, in stage order).
//
// The library behind the ABI is liboe_b200.so — the CUDA engine; there is no CPU implementation in this layer and no way to
// ask this header for one: EngineLibrary::open() binds the "oe_" entry points only.  (The known-answer tests derive their own
// loader from EngineLibrary in tests/kat/kat.hpp to point the same host code at the oracle and so pin the ORACLE against the
// reference's JUnit vectors; that subclass lives under tests/.)
#pragma once
#include <algorithm>
#include <dlfcn.h>
#include <cstdlib>

#include <functional>
#include <set>
#include <random>

#include "encoder.hpp"

namespace sdk {

struct EngineError : std::runtime_error {  // surfaces as IOException in the Java adapter (PlanScheduler.java:99-104)
  int code;
  EngineError(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

class EngineLibrary {
 public:
  // The CUDA engine (liboe_b200.so, built from dcos_commons_b200/csrc): the entry points of include/oe_abi.h, by their own names.
  static EngineLibrary open(const std::string& path) {
    void* h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) throw std::runtime_error(std::string("cannot load ") + path + ": " + dlerror());
    EngineLibrary lib;
    lib.path = path;
    lib.bind([&](const char* name) {
      void* s = dlsym(h, name);
      if (!s) throw std::runtime_error(path + " does not export " + name);
      return s;
    });
    return lib;
  }

  int32_t (*create)(const oe_config*, oe_engine**) = nullptr;
  void (*destroy)(oe_engine*) = nullptr;
  const char* (*last_error)(const oe_engine*) = nullptr;
  int32_t (*load_programs)(oe_engine*, const oe_program_table*) = nullptr;
  int32_t (*load_tasks)(oe_engine*, const oe_task_table*) = nullptr;
  int32_t (*load_offers)(oe_engine*, const oe_offer_table*) = nullptr;
  int32_t (*update_offers)(oe_engine*, const int64_t*, int64_t, const oe_offer_table*) = nullptr;
  int32_t (*load_offers_packed)(oe_engine*, const oe_offer_table*, const oe_offer_packing*) = nullptr;   // optional: null when the library has only the plain load
  int32_t (*load_requirements)(oe_engine*, const oe_req_table*) = nullptr;
  int64_t (*matrix_stride_words)(const oe_engine*) = nullptr;
  int32_t (*feasibility)(oe_engine*, uint32_t*, oe_stats*) = nullptr;
  int32_t (*evaluate_cycle)(oe_engine*, oe_decision*, oe_stats*) = nullptr;
  int32_t (*apply_launches)(oe_engine*) = nullptr;
  int32_t (*unexpected_resources)(oe_engine*, const uint32_t*, int32_t, uint8_t*, int64_t*) = nullptr;
  std::string path;

 protected:
  EngineLibrary() = default;
  // `symbol` resolves an entry point of include/oe_abi.h by its declared name ("oe_create", ...) or throws.
  void bind(const std::function<void*(const char*)>& symbol) {
    create = (decltype(create))symbol("oe_create"); destroy = (decltype(destroy))symbol("oe_destroy");
    last_error = (decltype(last_error))symbol("oe_last_error");
    load_programs = (decltype(load_programs))symbol("oe_load_programs"); load_tasks = (decltype(load_tasks))symbol("oe_load_tasks");
    load_offers = (decltype(load_offers))symbol("oe_load_offers"); update_offers = (decltype(update_offers))symbol("oe_update_offers");
    load_requirements = (decltype(load_requirements))symbol("oe_load_requirements");
    matrix_stride_words = (decltype(matrix_stride_words))symbol("oe_matrix_stride_words");
    feasibility = (decltype(feasibility))symbol("oe_feasibility"); evaluate_cycle = (decltype(evaluate_cycle))symbol("oe_evaluate_cycle");
    apply_launches = (decltype(apply_launches))symbol("oe_apply_launches");
    unexpected_resources = (decltype(unexpected_resources))symbol("oe_unexpected_resources");
    try { load_offers_packed = (decltype(load_offers_packed))symbol("oe_load_offers_packed"); } catch (const std::runtime_error&) { load_offers_packed = nullptr; }
  }
};

// One engine handle (one per service, like OfferEvaluator: S/scheduler/DefaultScheduler.java:198-210).
class Engine {
 public:
  explicit Engine(const EngineLibrary& lib, int device = 0) : lib_(lib) {
    oe_config cfg{};
    cfg.device = device; cfg.rank = 0; cfg.world = 1;
    const int rc = lib_.create(&cfg, &h_);
    if (rc != 0) throw EngineError(rc, std::string("oe_create: ") + lib_.last_error(nullptr));
  }
  ~Engine() { if (h_) lib_.destroy(h_); }
  Engine(const Engine&) = delete;
  Engine& operator=(const Engine&) = delete;

  // A cycle over a resident offer table (EncodedCycle::resident) sends only the rows that changed (oe_update_offers) unless the
  // encoder had to re-encode the whole table; the programs are re-loaded either way (the engine keeps the table across that).
  void load(const EncodedCycle& c) {
    auto p = c.programTable(); check(lib_.load_programs(h_, &p), "oe_load_programs");
    auto t = c.taskTable(); check(lib_.load_tasks(h_, &t), "oe_load_tasks");
    auto o = c.offerTable();
    if (c.resident && !c.full_reload) {
      if (!c.delta_rows.empty()) check(lib_.update_offers(h_, c.delta_rows.data(), (int64_t)c.delta_rows.size(), &o), "oe_update_offers");
      ++deltaLoads; rowsSent += (int64_t)c.delta_rows.size();
    } else if (lib_.load_offers_packed && packedWire) {
      // a whole table goes over the bus in its compact form (PackedOffers); rows replaced later arrive at full width
      const PackedOffers packed(c);
      const oe_offer_table po = packed.table();
      const oe_offer_packing pk = packed.packing();
      check(lib_.load_offers_packed(h_, &po, &pk), "oe_load_offers_packed");
      ++fullLoads; rowsSent += c.N; offerBytesSent += (int64_t)packed.bytes();
    } else {
      check(lib_.load_offers(h_, &o), "oe_load_offers");
      ++fullLoads; rowsSent += c.N;
    }
    auto r = c.reqTable(); check(lib_.load_requirements(h_, &r), "oe_load_requirements");
    M_ = c.M; N_ = c.resident ? c.N_resident : c.N;
  }
  int64_t fullLoads = 0, deltaLoads = 0, rowsSent = 0;   // what crossed the boundary so far (tests, diagnostics)
  int64_t offerBytesSent = 0;                            // ... bytes of the packed whole-table loads among them
  // OE_HOST_WIRE=plain in the environment: whole tables at full width (oe_load_offers) although the library can take them packed
  bool packedWire = [] { const char* w = std::getenv("OE_HOST_WIRE"); return !(w && std::string(w) == "plain"); }();
  std::vector<oe_decision> evaluateCycle() {
    std::vector<oe_decision> d((size_t)std::max<int64_t>(M_, 1));
    oe_stats st{};
    check(lib_.evaluate_cycle(h_, d.data(), &st), "oe_evaluate_cycle");
    d.resize((size_t)M_);
    return d;
  }
  // flags per reserved side-table entry (offers in order, their id-carrying resources in order)
  std::vector<uint8_t> unexpectedResources(const std::vector<uint32_t>& keepBits, size_t nReserved, bool onlyUnusedOffers) {
    std::vector<uint8_t> out(std::max<size_t>(nReserved, 1));
    int64_t n = 0;
    check(lib_.unexpected_resources(h_, keepBits.data(), onlyUnusedOffers ? 1 : 0, out.data(), &n), "oe_unexpected_resources");
    out.resize(nReserved);
    return out;
  }
  // feasibility bit of (requirement j, offer i)
  std::vector<std::vector<bool>> feasibility() {
    const int64_t stride = lib_.matrix_stride_words(h_);
    std::vector<uint32_t> bits((size_t)std::max<int64_t>(M_ * stride, 1));
    oe_stats st{};
    check(lib_.feasibility(h_, bits.data(), &st), "oe_feasibility");
    std::vector<std::vector<bool>> out((size_t)M_, std::vector<bool>((size_t)N_));
    for (int64_t j = 0; j < M_; ++j)
      for (int64_t i = 0; i < N_; ++i) out[(size_t)j][(size_t)i] = (bits[(size_t)(j * stride + (i >> 5))] >> (i & 31)) & 1u;
    return out;
  }

 private:
  void check(int rc, const char* what) {
    if (rc != 0) throw EngineError(rc, std::string(what) + ": " + lib_.last_error(h_));
  }
  const EngineLibrary& lib_;
  oe_engine* h_ = nullptr;
  int64_t M_ = 0, N_ = 0;
};

// ---- recommendations (S/offer/*OfferRecommendation.java) ---------------------------------------------------------------
struct OfferRecommendation {
  enum Type { RESERVE, UNRESERVE, CREATE, DESTROY, LAUNCH_GROUP, STORE_TASK };
  Type type = RESERVE;
  std::string offerId, agentId;
  mesos::Resource resource;               // RESERVE / CREATE
  std::optional<mesos::TaskInfo> task;    // LAUNCH_GROUP / STORE_TASK
  bool hasOperation() const { return type != STORE_TASK; }   // StoreTaskInfoRecommendation carries no Mesos operation
};

// S/offer/OfferUtils.java:26-38: the offers no recommendation WITH an operation refers to (an offer that only led to internal
// bookkeeping — a StoreTaskInfoRecommendation — is not "used").  PlanScheduler.resourceOffers (:61-66) and the multi-service
// fan-out narrow the offer list with it between steps / services; the engine's ordered commit does the same on the device.
struct OfferUtils {
  static std::vector<mesos::Offer> filterOutAcceptedOffers(const std::vector<mesos::Offer>& offers, const std::vector<OfferRecommendation>& recommendations) {
    std::vector<std::string> used;
    for (auto& r : recommendations) if (r.hasOperation()) used.push_back(r.offerId);
    std::vector<mesos::Offer> out;
    for (auto& o : offers) if (std::find(used.begin(), used.end(), o.id) == used.end()) out.push_back(o);
    return out;
  }
};

// S/state/StateStore.java:340 fetchTasks / storeTasks, reduced to an in-memory map (the MemPersister role in the
// reference's tests, T/offer/evaluate/OfferEvaluatorTestBase.java:32-50).
class StateStore {
 public:
  void storeTasks(const std::vector<mesos::TaskInfo>& tasks) { for (auto& t : tasks) tasks_[t.name] = t; }
  std::vector<mesos::TaskInfo> fetchTasks() const {
    std::vector<mesos::TaskInfo> out;
    for (auto& [n, t] : tasks_) out.push_back(t);
    return out;
  }
  // S/state/StateStore.java storeStatus / fetchStatus: the TaskState name ("TASK_RUNNING", ...) last reported for a task
  void storeStatus(const std::string& taskName, const std::string& state) { statuses_[taskName] = state; }
  std::optional<std::string> fetchStatus(const std::string& taskName) const {
    auto it = statuses_.find(taskName);
    if (it == statuses_.end()) return std::nullopt;
    return it->second;
  }
  // S/state/StateStore.java storeGoalOverrideStatus / fetchGoalOverrideStatus (the decommission plan sets "DECOMMISSIONING")
  void storeGoalOverrideStatus(const std::string& taskName, const std::string& status) { overrides_[taskName] = status; }
  std::string fetchGoalOverrideStatus(const std::string& taskName) const {
    auto it = overrides_.find(taskName);
    return it == overrides_.end() ? std::string("NONE") : it->second;
  }
  void clear() { tasks_.clear(); statuses_.clear(); overrides_.clear(); }
 private:
  std::map<std::string, mesos::TaskInfo> tasks_;
  std::map<std::string, std::string> statuses_, overrides_;
};

// Direct counterpart of calling `rule.filter(offer, podInstance, tasks).isPassing()` in the reference's rule tests.
struct PlacementRuleFilter {
  const EngineLibrary& lib;
  bool filter(const RulePtr& rule, const mesos::Offer& offer, const PodInstance& podInstance, const std::vector<mesos::TaskInfo>& tasks) const {
    CycleEncoder enc;
    EncoderRequirement r;
    r.ruleOnly = rule; r.podInstance = podInstance;
    EncodedCycle c = enc.encode({offer}, tasks, {r});
    Engine e(lib);
    e.load(c);
    return e.feasibility()[0][0];
  }
};

// The offers a scheduler holds between cycles, kept as ONE table on the device.  The reference's queue takes what the master
// sends and drops what it rescinds (S/framework/OfferQueue.java:46-85: offer(), remove(offerId)); every cycle the evaluator sees
// the current list (S/framework/OfferProcessor.java:240-262) and the offers it accepted leave it (PlanScheduler.java:61-66).
// Here an agent keeps its ROW: a new offer of an agent replaces the agent's row in place, an agent seen for the first time
// appends a row, and an offer that was rescinded or accepted leaves an empty row behind (no pool entry, no port: nothing fits
// it).  The list the evaluator is given is the rows in row order — list() — so first fit (OfferEvaluator.java:137) means the same
// thing on both sides; only the rows that changed since the last cycle are encoded and sent (oe_update_offers).
class ResidentOffers {
 public:
  void offer(const mesos::Offer& o) {
    auto it = rowOfAgent_.find(o.slave_id);
    size_t row;
    if (it == rowOfAgent_.end()) { row = slots_.size(); slots_.emplace_back(); rowOfAgent_[o.slave_id] = row; }
    else { row = it->second; if (slots_[row]) rowOfOffer_.erase(slots_[row]->id); }
    slots_[row] = o;
    rowOfOffer_[o.id] = row;
    dirty_.insert((int64_t)row);
  }
  // OfferQueue.remove(offerId) — rescinded by the master — and what PlanScheduler does with an accepted offer
  void remove(const std::string& offerId) {
    auto it = rowOfOffer_.find(offerId);
    if (it == rowOfOffer_.end()) return;
    slots_[it->second].reset();
    dirty_.insert((int64_t)it->second);
    rowOfOffer_.erase(it);
  }
  // the evaluator's `offers` argument: the standing offers in row order
  std::vector<mesos::Offer> list() const {
    std::vector<mesos::Offer> out;
    for (auto& s : slots_) if (s) out.push_back(*s);
    return out;
  }
  size_t rows() const { return slots_.size(); }
  const std::optional<mesos::Offer>& row(size_t r) const { return slots_[r]; }

 private:
  friend class OfferEvaluator;
  std::vector<std::optional<mesos::Offer>> slots_;
  std::map<std::string, size_t> rowOfAgent_, rowOfOffer_;
  std::set<int64_t> dirty_;                 // rows whose offer changed since the table was last sent
  OfferDictionaries dicts_;                 // ids live as long as the table
  int shape_[8] = {-1, -1, -1, -1, -1, -1, -1, 0};
  size_t capacity_ = 0;                     // rows of the table on the device (the offers' rows and spare ones behind them)
  std::unique_ptr<Engine> engine_;          // the handle that holds the table
};

class OfferEvaluator {
 public:
  OfferEvaluator(const EngineLibrary& lib, StateStore& stateStore, std::string serviceName = "test-service", SchedulerConfig cfg = SchedulerConfig())
      : lib_(lib), stateStore_(stateStore), serviceName_(std::move(serviceName)), cfg_(cfg) {}

  // S/offer/evaluate/OfferEvaluator.java:113-248: recommendations of the first offer (list order) all of whose stages
  // pass; empty when none does.
  std::vector<OfferRecommendation> evaluate(const PodInstanceRequirement& requirement, const std::vector<mesos::Offer>& offers) {
    auto all = evaluateCycle({requirement}, offers);
    return all.empty() ? std::vector<OfferRecommendation>() : all[0];
  }

  // The whole-cycle form PlanScheduler uses: one engine call for every requirement; result[j] belongs to requirement j.
  std::vector<std::vector<OfferRecommendation>> evaluateCycle(const std::vector<PodInstanceRequirement>& requirements,
                                                              const std::vector<mesos::Offer>& offers) {
    std::vector<std::vector<OfferRecommendation>> out(requirements.size());
    if (requirements.empty()) return out;
    const std::vector<mesos::TaskInfo> allTasks = stateStore_.fetchTasks();  // OfferEvaluator.java:124 (once per cycle here)
    CycleEncoder enc(cfg_);
    std::vector<EncoderRequirement> ers = encoderRequirements(requirements, allTasks);
    EncodedCycle c = enc.encode(offers, allTasks, ers);
    Engine e(lib_);
    e.packedWire = e.packedWire && packedWire;
    e.load(c);
    offerBytesSent += e.offerBytesSent;
    const std::vector<oe_decision> dec = e.evaluateCycle();
    lastDecisions_ = dec;
    for (size_t j = 0; j < requirements.size(); ++j)
      if (dec[j].offer_row >= 0)
        out[j] = materialise(requirements[j], offers[(size_t)dec[j].offer_row], c.classStages[(size_t)c.r_class[j]], dec[j], ers[j].existing.get());
    return out;
  }

  // The same cycle over the scheduler's standing offers: result[j] names offers of `resident.list()`; accepted offers leave the
  // table (PlanScheduler.java:61-66), as they leave the reference's list.
  std::vector<std::vector<OfferRecommendation>> evaluateCycle(const std::vector<PodInstanceRequirement>& requirements, ResidentOffers& resident) {
    std::vector<std::vector<OfferRecommendation>> out(requirements.size());
    if (requirements.empty()) return out;
    const std::vector<mesos::TaskInfo> allTasks = stateStore_.fetchTasks();
    CycleEncoder enc(cfg_, &resident.dicts_);
    std::vector<EncoderRequirement> ers = encoderRequirements(requirements, allTasks);
    // the table on the device has spare (empty) rows at its end, so that an agent seen for the first time is a changed row like
    // any other; only when they run out is the table re-sent, with an eighth more room
    if (resident.slots_.size() > resident.capacity_) {
      resident.capacity_ = resident.slots_.size() + std::max<size_t>(16, resident.slots_.size() / 8);
      resident.shape_[0] = -1;
    }
    std::vector<const mesos::Offer*> slots;
    for (auto& s : resident.slots_) slots.push_back(s ? &*s : nullptr);
    slots.resize(resident.capacity_, nullptr);
    const std::vector<int64_t> dirty(resident.dirty_.begin(), resident.dirty_.end());
    CycleEncoder::ResidentRows rr;
    rr.slots = &slots; rr.dirty = &dirty;
    std::copy(resident.shape_, resident.shape_ + 8, rr.shape);
    EncodedCycle c = enc.encodeResident(rr, allTasks, ers);
    if (!resident.engine_) resident.engine_ = std::make_unique<Engine>(lib_);
    resident.engine_->load(c);
    if (c.full_reload) {
      const int shape[8] = {c.S, c.K, c.R, c.D, c.A, c.E, c.Q, c.o_rid.empty() ? 0 : 1};
      std::copy(shape, shape + 8, resident.shape_);
    }
    resident.dirty_.clear();
    const std::vector<oe_decision> dec = resident.engine_->evaluateCycle();
    lastDecisions_ = dec;
    for (size_t j = 0; j < requirements.size(); ++j)
      if (dec[j].offer_row >= 0) {
        const mesos::Offer winner = *resident.slots_[(size_t)dec[j].offer_row];
        out[j] = materialise(requirements[j], winner, c.classStages[(size_t)c.r_class[j]], dec[j], ers[j].existing.get());
        resident.remove(winner.id);
      }
    return out;
  }
  const Engine* residentEngine(const ResidentOffers& r) const { return r.engine_.get(); }
  bool packedWire = true;        // false: whole offer tables cross the boundary at full width (oe_load_offers)
  int64_t offerBytesSent = 0;    // offer-column bytes of the packed whole-table loads of evaluateCycle(requirements, offers)

  // The two halves of evaluateCycle, for callers that put several services' requirements into ONE engine cycle
  // (MultiServiceEventClient below).  `requirements` must outlive the returned vector.
  std::vector<EncoderRequirement> encoderRequirements(const std::vector<PodInstanceRequirement>& requirements,
                                                      const std::vector<mesos::TaskInfo>& allTasks, const std::string& service = "") const {
    std::vector<EncoderRequirement> ers(requirements.size());
    for (size_t j = 0; j < requirements.size(); ++j) {
      ers[j].requirement = &requirements[j];
      ers[j].service = service;
      ers[j].schedulerConfig = &cfg_;
      ers[j].existing = existingPod(requirements[j], allTasks);
      for (auto& ts : requirements[j].podInstance.pod->tasks) {  // PodInfoBuilder.portsByTask (PodInfoBuilder.java:130-140)
        const std::string name = taskInstanceName(requirements[j].podInstance, ts.name);
        for (auto& t : allTasks)
          if (t.name == name && !isPermanentlyFailed(t) && !t.discovery_ports.empty()) ers[j].priorPorts[ts.name] = t.discovery_ports;
      }
    }
    return ers;
  }
  std::vector<OfferRecommendation> recommendationsFor(const PodInstanceRequirement& requirement, const mesos::Offer& winner,
                                                      const std::vector<StageInfo>& stages, const oe_decision& decision, const ExistingPod* existing) {
    return materialise(requirement, winner, stages, decision, existing);
  }
  StateStore& stateStore() { return stateStore_; }
  const std::string& serviceName() const { return serviceName_; }
  const std::vector<oe_decision>& lastDecisions() const { return lastDecisions_; }

 private:
  static bool isPermanentlyFailed(const mesos::TaskInfo& t) {  // S/scheduler/recovery/FailureUtils.java: label "permanently-failed"
    auto it = t.labels.find("permanently-failed");
    return it != t.labels.end() && it->second == "true";
  }
  // getEvaluationPipeline's choice (OfferEvaluator.java:250-310): null = new / failed pod, else the existing-pod context.
  std::shared_ptr<const ExistingPod> existingPod(const PodInstanceRequirement& req, const std::vector<mesos::TaskInfo>& allTasks) const {
    auto ex = std::make_shared<ExistingPod>();
    for (auto& ts : req.podInstance.pod->tasks) {  // thisPodTasks (:127-131)
      const std::string name = taskInstanceName(req.podInstance, ts.name);
      for (auto& t : allTasks) if (t.name == name) ex->podTasks[name] = t;
    }
    bool noLaunchedTasksExist = true, allFailed = !ex->podTasks.empty();
    for (auto& [n, t] : ex->podTasks) {
      for (auto& r : t.resources) if (auto id = ResourceUtils::getResourceId(r); id && !id->empty()) noLaunchedTasksExist = false;
      allFailed = allFailed && isPermanentlyFailed(t);
    }
    if (req.recoveryType == RecoveryType::PERMANENT || allFailed || noLaunchedTasksExist) return nullptr;
    // getExecutorInfo (:312-345): a task of the pod that is NOT being relaunched and still runs keeps its executor alive
    std::set<std::string> launching;
    for (auto& n : req.tasksToLaunch) launching.insert(taskInstanceName(req.podInstance, n));
    for (auto& [n, t] : ex->podTasks) {
      if (launching.count(n) || isPermanentlyFailed(t)) continue;
      auto st = stateStore_.fetchStatus(n);
      if (st && (*st == "TASK_STAGING" || *st == "TASK_STARTING" || *st == "TASK_RUNNING")) {
        ex->executorId = t.executor.executor_id;
        ex->executorInfo = t.executor;
        return ex;
      }
    }
    ex->executorInfo = ex->podTasks.begin()->second.executor;
    ex->executorInfo.executor_id.clear();
    return ex;
  }

  std::string newResourceId() {  // ResourceBuilder.java:318,343 mints UUID.randomUUID(): 36 characters
    static const char* hex = "0123456789abcdef";
    std::string s(36, '-');
    for (int k = 0; k < 36; ++k) if (k != 8 && k != 13 && k != 18 && k != 23) s[(size_t)k] = hex[rng_() & 15];
    return s;
  }

  mesos::Resource reservedResource(const StageInfo& s, const mesos::Value& v) {
    mesos::Resource r;
    r.name = s.resourceName; r.value = v;
    if (s.preReservedRole != ANY_ROLE) {  // refined reservation on top of the pre-reserved role (MesosResourcePool.java:349-357)
      mesos::ReservationInfo st;
      st.type = mesos::ReservationInfo::STATIC; st.role = s.preReservedRole;
      r.reservations.push_back(st);
    }
    mesos::ReservationInfo dyn;
    dyn.type = mesos::ReservationInfo::DYNAMIC; dyn.role = s.role; dyn.principal = s.principal;
    dyn.labels[labels::RESOURCE_ID] = newResourceId();
    r.reservations.push_back(dyn);
    return r;
  }

  static const mesos::Resource* reservedInOffer(const mesos::Offer& offer, const std::string& id) {
    for (auto& r : offer.resources) if (auto rid = ResourceUtils::getResourceId(r); rid && *rid == id) return &r;
    return nullptr;
  }
  mesos::Resource withId(const StageInfo& s, const mesos::Value& v, const std::string& id) {
    mesos::Resource r = reservedResource(s, v);
    r.reservations.back().labels[labels::RESOURCE_ID] = id;
    return r;
  }

  std::vector<OfferRecommendation> materialise(const PodInstanceRequirement& req, const mesos::Offer& offer,
                                               const std::vector<StageInfo>& stages, const oe_decision& d, const ExistingPod* existing) {
    std::vector<OfferRecommendation> recs;
    const PodInstance& pi = req.podInstance;
    std::map<std::string, mesos::TaskInfo> taskBuilders;  // PodInfoBuilder: one TaskInfo per task spec
    mesos::ExecutorInfo executor;
    for (auto& ts : pi.pod->tasks) {
      mesos::TaskInfo t;
      t.name = taskInstanceName(pi, ts.name);
      taskBuilders[ts.name] = t;
    }
    int dyn = 0, mount = 0, newIds = 0;
    auto rec = [&](OfferRecommendation::Type type) -> OfferRecommendation& {
      OfferRecommendation r;
      r.type = type; r.offerId = offer.id; r.agentId = offer.slave_id;
      recs.push_back(r);
      return recs.back();
    };
    auto attach = [&](const StageInfo& s, const mesos::Resource& r) {
      if (s.taskNames.empty()) executor.resources.push_back(r);
      else for (auto& n : s.taskNames) taskBuilders[n].resources.push_back(r);
    };
    const bool runningExecutor = existing && existing->executorId &&
                                 std::find(offer.executor_ids.begin(), offer.executor_ids.end(), *existing->executorId) != offer.executor_ids.end();
    if (existing && existing->executorId) executor.executor_id = *existing->executorId;
    for (auto& s : stages) {
      if (s.resourceId && s.kind != StageInfo::UNRESERVE && s.kind != StageInfo::DESTROY) {
        // ---- a stage that expects an earlier reservation (evaluateSimpleResource with a resource id,
        // OfferEvaluationUtils.java:105-225; VolumeEvaluationStage.java:126-226)
        mesos::Value want = s.kind == StageInfo::PORT ? mesos::Value::Ranges({{s.port, s.port}}) : mesos::Value::Scalar(s.value);
        mesos::Resource r = withId(s, want, *s.resourceId);
        if (s.kind == StageInfo::VOLUME) {
          mesos::DiskInfo disk;
          disk.container_path = s.containerPath; disk.persistence_id = s.persistenceId;
          if (s.volumeType != VolumeSpec::ROOT) disk.source = mesos::DiskInfo::MOUNT;
          r.disk = disk;
        }
        if (s.kind == StageInfo::PORT && !s.hostPort) { attach(s, r); continue; }
        if (!(s.skipIfRunningExecutor && runningExecutor)) {
          const mesos::Resource* have = reservedInOffer(offer, *s.resourceId);
          if (!have) throw std::logic_error("engine accepted an offer that lacks reservation " + *s.resourceId);
          if (s.kind == StageInfo::PORT) {
            const int64_t had = have->value.ranges.empty() ? 0 : have->value.ranges[0].first;
            if (had != s.port) rec(OfferRecommendation::RESERVE).resource = r;   // the port moved: reserve the new one under the same id
          } else if (!(s.kind == StageInfo::VOLUME && s.volumeType != VolumeSpec::ROOT)) {
            const double had = have->value.scalar;
            if (s.value > had) rec(OfferRecommendation::RESERVE).resource = withId(s, mesos::Value::Scalar(s.value - had), *s.resourceId);
            else if (s.value < had) rec(OfferRecommendation::UNRESERVE).resource = withId(s, mesos::Value::Scalar(had - s.value), *s.resourceId);
          }
          if (s.kind == StageInfo::VOLUME && !s.persistenceId) { r.disk->persistence_id = newResourceId(); rec(OfferRecommendation::CREATE).resource = r; }
        }
        if (s.kind == StageInfo::PORT) for (auto& n : s.taskNames) taskBuilders[n].discovery_ports[s.portName] = s.port;
        attach(s, r);
        continue;
      }
      switch (s.kind) {
        case StageInfo::EXECUTOR: case StageInfo::PLACEMENT: case StageInfo::TLS: break;
        case StageInfo::DESTROY: rec(OfferRecommendation::DESTROY).resource = s.orphan; break;
        case StageInfo::UNRESERVE: rec(OfferRecommendation::UNRESERVE).resource = s.orphan; break;
        case StageInfo::RESOURCE: {
          mesos::Resource r = reservedResource(s, mesos::Value::Scalar(s.value));
          ++newIds;
          rec(OfferRecommendation::RESERVE).resource = r;
          attach(s, r);
        } break;
        case StageInfo::PORT: {
          int port = s.port;
          if (s.hostPort && port == 0) port = d.dyn_port[dyn < OE_MAX_DYN_PORTS ? dyn : OE_MAX_DYN_PORTS - 1];
          if (s.hostPort && s.port == 0) ++dyn;
          if (!s.hostPort) {  // overlay: no reservation, the port is only advertised (PortEvaluationStage.java:145-154)
            for (auto& n : s.taskNames) taskBuilders[n].discovery_ports[s.portName] = s.port;
            break;
          }
          mesos::Resource r = reservedResource(s, mesos::Value::Ranges({{port, port}}));
          ++newIds;
          rec(OfferRecommendation::RESERVE).resource = r;
          for (auto& n : s.taskNames) taskBuilders[n].discovery_ports[s.portName] = port;
          attach(s, r);
        } break;
        case StageInfo::VOLUME: {
          mesos::Value v = mesos::Value::Scalar(s.value);
          mesos::DiskInfo disk;
          disk.container_path = s.containerPath;
          disk.persistence_id = newResourceId();
          if (s.volumeType != VolumeSpec::ROOT) {
            // the whole MOUNT disk is taken (VolumeEvaluationStage.java:205-215: setValue(mesosResource.getValue()))
            const int slot = d.mount_slot[mount < OE_MAX_MOUNT_OPS ? mount : OE_MAX_MOUNT_OPS - 1];
            ++mount;
            int seen = 0;
            for (auto& or_ : offer.resources)
              if (ResourceUtils::isAtomic(or_) && !ResourceUtils::getResourceId(or_) && or_.name == "disk" && seen++ == slot) {
                v = or_.value; disk.source = mesos::DiskInfo::MOUNT; disk.profile = or_.disk->profile;
              }
          }
          mesos::Resource r = reservedResource(s, v);
          r.disk = disk;
          ++newIds;
          rec(OfferRecommendation::RESERVE).resource = r;
          rec(OfferRecommendation::CREATE).resource = r;   // no persistence id yet: createsVolume() (:222-226)
          attach(s, r);
        } break;
        case StageInfo::LAUNCH: {  // LaunchEvaluationStage.java:42-90
          mesos::TaskInfo& t = taskBuilders[s.taskSpecName];
          t.task_id = s.shouldLaunch ? t.name + "__" + newResourceId() : "";
          t.slave_id = offer.slave_id;
          TaskLabelWriter w(t);
          w.setOfferAttributes(offer).setType(pi.pod->type).setIndex(pi.index).setHostname(offer);
          if (offer.domain) w.setRegion(offer.domain->region).setZone(offer.domain->zone);
          t.executor = executor;
          if (s.shouldLaunch) rec(OfferRecommendation::LAUNCH_GROUP).task = t;
          rec(OfferRecommendation::STORE_TASK).task = t;
        } break;
      }
    }
    if (newIds != d.n_new_reservations)
      throw std::logic_error("engine and adapter disagree on the number of new reservations (" + std::to_string(d.n_new_reservations) +
                             " vs " + std::to_string(newIds) + ")");
    return recs;
  }

  const EngineLibrary& lib_;
  StateStore& stateStore_;
  std::string serviceName_;
  SchedulerConfig cfg_;
  std::mt19937_64 rng_{20240607};
  std::vector<oe_decision> lastDecisions_;
};

// S/scheduler/DefaultScheduler.java:483-538 getUnexpectedResources: reservations found in the unused offers that no stored
// task (that is neither permanently failed nor being decommissioned) still claims.  The keep set is built on the host, the
// scan over every reserved resource of every unused offer runs on the engine.
struct OfferResources {  // S/offer/OfferResources.java
  mesos::Offer offer;
  std::vector<mesos::Resource> resources;
};
class UnexpectedResourceScanner {
 public:
  UnexpectedResourceScanner(const EngineLibrary& lib, StateStore& stateStore) : lib_(lib), stateStore_(stateStore) {}
  std::vector<OfferResources> getUnexpectedResources(const std::vector<mesos::Offer>& unusedOffers) {
    std::vector<OfferResources> out;
    if (unusedOffers.empty()) return out;
    std::set<std::string> resourceIdsToKeep;  // :487-499
    for (auto& t : stateStore_.fetchTasks()) {
      auto failed = t.labels.find("permanently-failed");
      if (failed != t.labels.end() && failed->second == "true") continue;
      if (stateStore_.fetchGoalOverrideStatus(t.name) == "DECOMMISSIONING") continue;
      for (auto& r : t.resources) if (auto id = ResourceUtils::getResourceId(r)) resourceIdsToKeep.insert(*id);
      for (auto& r : t.executor.resources) if (auto id = ResourceUtils::getResourceId(r)) resourceIdsToKeep.insert(*id);  // getAllResources
    }
    CycleEncoder enc;
    EncodedCycle c = enc.encode(unusedOffers, {}, {});
    const Dictionary& ids = enc.resourceIds();
    std::vector<uint32_t> keep((size_t)std::max(1, (ids.size() + 31) / 32), 0u);
    for (int v = 0; v < ids.size(); ++v)
      if (resourceIdsToKeep.count(ids.values()[(size_t)v])) keep[(size_t)v >> 5] |= 1u << (v & 31);
    Engine e(lib_);
    e.load(c);
    const std::vector<uint8_t> flags = e.unexpectedResources(keep, c.o_rid.size(), false);
    size_t q = 0;
    for (auto& offer : unusedOffers) {  // side-table order: offers in order, their id-carrying resources in order
      OfferResources found{offer, {}};
      for (auto& r : offer.resources)
        if (ResourceUtils::getResourceId(r) && flags[q++]) found.resources.push_back(r);
      if (!found.resources.empty()) out.push_back(found);
    }
    return out;
  }
 private:
  const EngineLibrary& lib_;
  StateStore& stateStore_;
};

// S/scheduler/plan/Step.java reduced to what PlanScheduler reads.
struct Step {
  std::string name;
  std::optional<PodInstanceRequirement> requirement;
  enum Status { PENDING, PREPARED, STARTING, COMPLETE } status = PENDING;
  bool isPending() const { return status == PENDING; }
  bool isPrepared() const { return status == PREPARED; }
  bool isPendingOrPrepared() const { return isPending() || isPrepared(); }
  void start() {}  // DeploymentStep.start (S/scheduler/plan/DeploymentStep.java:106-108) does nothing
  // DeploymentStep.updateOfferStatus (:122-136): no operations -> PREPARED, waiting for another offer cycle; operations sent -> STARTING
  void updateOfferStatus(size_t nOperations) { status = nOperations == 0 ? PREPARED : STARTING; }
};

// S/scheduler/plan/PlanScheduler.java:50-124
class PlanScheduler {
 public:
  explicit PlanScheduler(OfferEvaluator& e) : evaluator_(e) {}
  // Steps in order; each takes the first offer that fits among the offers no earlier step of this cycle used (:56-67).
  // Steps that are not pending / prepared, or carry no requirement, are skipped (:73-92).  `steps` get STARTING on success.
  std::vector<OfferRecommendation> resourceOffers(const std::vector<mesos::Offer>& offers, std::vector<Step>& steps) {
    std::vector<OfferRecommendation> all;
    if (steps.empty()) return all;
    std::vector<PodInstanceRequirement> reqs;
    std::vector<size_t> which;
    for (size_t k = 0; k < steps.size(); ++k) {
      if (!steps[k].isPendingOrPrepared()) continue;                                   // :73-78
      steps[k].start();                                                                // :81
      if (!steps[k].requirement) { steps[k].updateOfferStatus(0); continue; }          // :84-88
      reqs.push_back(*steps[k].requirement);
      which.push_back(k);
    }
    // no offers: every evaluate() comes back empty (OfferEvaluator.java:136 loops over nothing); the steps still move to PREPARED
    auto per = offers.empty() ? std::vector<std::vector<OfferRecommendation>>(reqs.size()) : evaluator_.evaluateCycle(reqs, offers);
    for (size_t j = 0; j < per.size(); ++j) {
      size_t nOperations = 0;  // only what is actually sent to Mesos counts (:115-118)
      for (auto& r : per[j]) nOperations += r.hasOperation();
      steps[which[j]].updateOfferStatus(nOperations);
      all.insert(all.end(), per[j].begin(), per[j].end());
    }
    return all;
  }
 private:
  OfferEvaluator& evaluator_;
};

// S/scheduler/multi/MultiServiceEventClient.java:312-372 offers(): the services take their turn in order, each seeing the offers
// the earlier ones left (OfferUtils.filterOutAcceptedOffers).  Every service runs PlanScheduler.resourceOffers over its own
// candidate steps against its own state store, so the whole fan-out is ONE ordered first-fit over the concatenation of the
// services' requirements — one engine cycle instead of one per service, the services' tasks told apart by namespace.
class MultiServiceEventClient {
 public:
  struct Service {
    OfferEvaluator* evaluator;   // carries the service's name, state store and scheduler config
    std::vector<Step>* steps;    // the candidate steps of this cycle (PlanCoordinator.getCandidates), in order
  };
  explicit MultiServiceEventClient(const EngineLibrary& lib) : lib_(lib) {}

  std::vector<OfferRecommendation> offers(const std::vector<mesos::Offer>& offers, const std::vector<Service>& services) {
    std::vector<OfferRecommendation> all;
    struct Pending { size_t service, step; };
    std::vector<Pending> which;
    std::vector<std::vector<PodInstanceRequirement>> reqs(services.size());
    std::vector<mesos::TaskInfo> allTasks;
    std::vector<std::string> taskServices;
    std::vector<EncoderRequirement> ers;
    for (size_t s = 0; s < services.size(); ++s) {
      std::vector<Step>& steps = *services[s].steps;
      for (size_t k = 0; k < steps.size(); ++k) {  // PlanScheduler.java:72-92
        if (!steps[k].isPendingOrPrepared()) continue;
        steps[k].start();
        if (!steps[k].requirement) { steps[k].updateOfferStatus(0); continue; }
        reqs[s].push_back(*steps[k].requirement);
        which.push_back({s, k});
      }
    }
    for (size_t s = 0; s < services.size(); ++s) {
      const std::vector<mesos::TaskInfo> tasks = services[s].evaluator->stateStore().fetchTasks();
      auto e = services[s].evaluator->encoderRequirements(reqs[s], tasks, services[s].evaluator->serviceName());
      ers.insert(ers.end(), e.begin(), e.end());
      allTasks.insert(allTasks.end(), tasks.begin(), tasks.end());
      taskServices.insert(taskServices.end(), tasks.size(), services[s].evaluator->serviceName());
    }
    std::vector<oe_decision> dec(ers.size());
    EncodedCycle c;
    if (!offers.empty() && !ers.empty()) {  // with no offers left the services are still cranked (:327-328): nothing can match
      CycleEncoder enc;
      c = enc.encode(offers, allTasks, ers, taskServices);
      Engine e(lib_);
      e.load(c);
      dec = e.evaluateCycle();
    } else {
      for (auto& d : dec) d.offer_row = -1;
    }
    for (size_t j = 0; j < ers.size(); ++j) {
      Step& step = (*services[which[j].service].steps)[which[j].step];
      if (dec[j].offer_row < 0) { step.updateOfferStatus(0); continue; }
      auto recs = services[which[j].service].evaluator->recommendationsFor(*ers[j].requirement, offers[(size_t)dec[j].offer_row],
                                                                           c.classStages[(size_t)c.r_class[j]], dec[j], ers[j].existing.get());
      size_t nOperations = 0;
      for (auto& r : recs) nOperations += r.hasOperation();
      step.updateOfferStatus(nOperations);
      all.insert(all.end(), recs.begin(), recs.end());
    }
    return all;
  }

 private:
  const EngineLibrary& lib_;
};

}  // namespace sdk
