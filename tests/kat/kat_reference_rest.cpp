// kat_reference_rest.cpp — the remaining reference vectors on the path: pods with several tasks and shared resource sets
// (T/offer/evaluate/OfferEvaluatorTest.java:349-741 over src/test/resources/resource-set-seq.yml), refined reservations
// (:806-915 over resource-refinement.yml), agent placement (OfferEvaluatorPlacementTest.java), the remaining port / volume /
// pool / range vectors, the stage-level tests (PortEvaluationStageTest, VolumeEvaluationStageTest, ExecutorEvaluationStageTest,
// LaunchEvaluationStageTest, OfferEvaluationUtilsTest) restated through the whole evaluator, and PlanSchedulerTest.
//
// The `...Refined` variants of the reference re-run a test with the RESERVATION_REFINEMENT capability on, i.e. with
// reservations carried in Resource.reservations[] instead of the legacy role + reservation pair.  mesos.hpp models only the
// stack form (the legacy pair is the same data flattened), so a Refined variant re-runs the same body and additionally
// looks at the shape of the stack.
//
// Not restated, because they assert on things outside the matching path: OfferEvaluatorTest.testGetTargetConfig* (:922-1117,
// which config UUID a relaunch is labelled with), testLogOutcome* (:1119-1145, log text), PortEvaluationStageTest's
// *Envvar* tests (:296-450, environment variables of health / readiness checks).
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct Fixture {  // T/offer/evaluate/OfferEvaluatorTestBase.java:32-94
  StateStore stateStore;
  OfferEvaluator evaluator;
  explicit Fixture(SchedulerConfig cfg = SchedulerConfig()) : evaluator(*kat::library(), stateStore, TestConstants::SERVICE_NAME, cfg) {}
  void recordOperations(const std::vector<Rec>& recs) {  // :76-94: every StoreTaskInfoRecommendation goes to the state store
    std::vector<mesos::TaskInfo> store;
    for (auto& r : recs) if (r.type == Rec::STORE_TASK) store.push_back(*r.task);
    stateStore.storeTasks(store);
  }
  std::vector<mesos::Resource> recordLaunchWithCompleteOfferedResources(const PodInstanceRequirement& req, std::vector<mesos::Resource> offered) {
    auto recs = evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer(offered)});
    recordOperations(recs);
    std::vector<mesos::Resource> reserved;
    for (auto& r : recs) if (r.type == Rec::RESERVE) reserved.push_back(r.resource);
    return reserved;
  }
};

std::string types(const std::vector<Rec>& recs) {  // the reference's operation-type lists; '-' = StoreTaskInfoRecommendation (null)
  std::string s;
  for (auto& r : recs)
    s += r.type == Rec::RESERVE ? 'R' : r.type == Rec::UNRESERVE ? 'U' : r.type == Rec::CREATE ? 'C' : r.type == Rec::DESTROY ? 'D'
       : r.type == Rec::LAUNCH_GROUP ? 'L' : '-';
  return s;
}
std::string resourceId(const mesos::Resource& r) { return ResourceUtils::getResourceId(r).value_or(""); }

mesos::Resource reserved(mesos::Resource r, const std::string& id) {  // ResourceTestUtils.addReservation (:168-185)
  mesos::ReservationInfo dyn;
  dyn.role = TestConstants::ROLE; dyn.principal = TestConstants::PRINCIPAL; dyn.labels[labels::RESOURCE_ID] = id;
  r.reservations.push_back(dyn);
  return r;
}
mesos::Resource getReservedCpus(double v, const std::string& id) { return reserved(ResourceTestUtils::getUnreservedCpus(v), id); }
mesos::Resource getReservedMem(double v, const std::string& id) { return reserved(ResourceTestUtils::getUnreservedMem(v), id); }
mesos::Resource getReservedDisk(double v, const std::string& id) { return reserved(ResourceTestUtils::getUnreservedDisk(v), id); }
mesos::Resource getReservedRootVolume(double v, const std::string& id, const std::string& persistenceId) {
  mesos::Resource r = getReservedDisk(v, id);
  mesos::DiskInfo d;
  d.persistence_id = persistenceId; d.container_path = TestConstants::CONTAINER_PATH;
  r.disk = d;
  return r;
}
// getExpectedExecutorResources (OfferEvaluatorTest.java:64-92): the executor's three reservations as the agent re-offers them
std::vector<mesos::Resource> getExpectedExecutorResources(const mesos::ExecutorInfo& executor) {
  auto idOf = [&](const char* name) {
    for (auto& r : executor.resources) if (r.name == name) if (auto id = ResourceUtils::getResourceId(r)) return *id;
    KAT_FAIL("executor has no reserved " << name);
  };
  return {getReservedCpus(0.1, idOf("cpus")), getReservedMem(32, idOf("mem")), getReservedDisk(256, idOf("disk"))};
}

// src/test/resources/resource-set-seq.yml: pod "name" (count 2); resource sets name-resources {cpus 1.0, ROOT volume name-path 50}
// and sidecar-resources {cpus 2.0}; tasks node, format, bootstrap on name-resources and backup on sidecar-resources
PodInstance resourceSetSeqPod(int index = 0) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = "name"; pod->count = 2;
  const ResourceSet name = ResourceSetBuilder("test-role", ANY_ROLE, "test-principal").id("name-resources").cpus(1.0).addRootVolume(50, "name-path").build();
  const ResourceSet sidecar = ResourceSetBuilder("test-role", ANY_ROLE, "test-principal").id("sidecar-resources").cpus(2.0).build();
  for (const char* n : {"node", "format", "bootstrap"}) { TaskSpec t; t.name = n; t.resourceSet = name; pod->tasks.push_back(t); }
  TaskSpec backup; backup.name = "backup"; backup.resourceSet = sidecar; pod->tasks.push_back(backup);
  return PodInstance{pod, index};
}
mesos::Offer sufficientSeqOffer() { return OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(3.0), ResourceTestUtils::getUnreservedDisk(500.0)}); }

}  // namespace

// ---- T/offer/evaluate/OfferEvaluatorTest.java
KAT(OfferEvaluatorTest_testReserveLaunchScalarRefined) {  // :94-102 (body :30-62)
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))});
  assertEquals(6u, recs.size());
  const mesos::Resource& reserve = recs[3].resource;
  assertEquals(1.0, reserve.value.scalar);
  assertEquals(1u, reserve.reservations.size());   // validateRole: no pre-reserved role, one DYNAMIC entry
  assertTrue(reserve.reservations[0].type == mesos::ReservationInfo::DYNAMIC);
  assertEquals(TestConstants::ROLE, reserve.reservations[0].role);
  assertEquals(TestConstants::PRINCIPAL, reserve.reservations[0].principal);
  assertEquals(36u, resourceId(reserve).size());
  assertEquals(resourceId(reserve), resourceId(recs[4].task->resources[0]));
}
KAT(OfferEvaluatorTest_testLaunchMultipleTasksPerExecutor) {  // :348-438
  Fixture f;
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  TaskSpec a; a.name = "taskA"; a.resourceSet = ResourceSetBuilder(TestConstants::ROLE, ANY_ROLE, TestConstants::PRINCIPAL).cpus(1.0).id("resourceSetA").build();
  TaskSpec b; b.name = "taskB"; b.resourceSet = ResourceSetBuilder(TestConstants::ROLE, ANY_ROLE, TestConstants::PRINCIPAL).cpus(2.0).id("resourceSetB").build();
  pod->tasks = {a, b};
  auto req = PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"taskA", "taskB"});
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(3.0))});
  assertEquals(std::string("RRRRL-RL-"), types(recs));
  assertEquals(std::string("pod-type-0-taskA"), recs[4].task->name);
  assertEquals(std::string("pod-type-0-taskB"), recs[7].task->name);
  assertEquals(1.0, recs[3].resource.value.scalar);
  assertEquals(2.0, recs[6].resource.value.scalar);
  // one executor for both launches, and the stored TaskInfos carry it too
  auto executorIds = [](const mesos::TaskInfo& t) { std::string s; for (auto& r : t.executor.resources) s += resourceId(r) + ","; return s; };
  assertEquals(3u, recs[4].task->executor.resources.size());
  assertEquals(executorIds(*recs[4].task), executorIds(*recs[7].task));
  assertEquals(executorIds(*recs[4].task), executorIds(*recs[5].task));
  assertEquals(executorIds(*recs[7].task), executorIds(*recs[8].task));
}
KAT(OfferEvaluatorTest_testLaunchSequencedTasksInPod) {  // :464-576
  Fixture f;
  PodInstance podInstance = resourceSetSeqPod();
  auto formatReq = PodInstanceRequirement::newBuilder(podInstance, {"format"});
  auto recs = f.evaluator.evaluate(formatReq, {sufficientSeqOffer()});
  // executor x3 | backup (sidecar-resources): RESERVE + store | bootstrap, format (+launch), node (name-resources): RESERVE RESERVE CREATE
  assertEquals(std::string("RRRR-RRC-L--"), types(recs));
  assertEquals(std::string("name-0-backup"), recs[4].task->name);
  assertEquals(std::string("name-0-bootstrap"), recs[8].task->name);
  assertEquals(std::string("name-0-format"), recs[9].task->name);
  assertEquals(std::string("name-0-format"), recs[10].task->name);
  assertEquals(std::string("name-0-node"), recs[11].task->name);
  const mesos::TaskInfo formatTaskToReoffer = *recs[10].task;
  f.recordOperations(recs);

  // name-0-node later: sufficient but unreserved resources give nothing
  auto nodeReq = PodInstanceRequirement::newBuilder(podInstance, {"node"});
  assertEquals(0u, f.evaluator.evaluate(nodeReq, {sufficientSeqOffer()}).size());

  // the format task shares "name-resources" with node, so its reservations are what node needs
  auto expected = getExpectedExecutorResources(formatTaskToReoffer.executor);
  expected.push_back(getReservedCpus(1.0, resourceId(formatTaskToReoffer.resources[0])));
  mesos::Resource volume = getReservedRootVolume(50.0, resourceId(formatTaskToReoffer.resources[1]), *formatTaskToReoffer.resources[1].disk->persistence_id);
  volume.disk->container_path = "name-path";
  expected.push_back(volume);
  mesos::Offer offer = OfferTestUtils::getOffer(expected);
  offer.executor_ids = {"name__some-executor"};
  recs = f.evaluator.evaluate(nodeReq, {offer});
  // every task of the resource set gets a TaskInfo update, node also a launch; backup (other resource set) is left alone
  assertEquals(std::string("--L-"), types(recs));
  assertEquals(std::string("name-0-bootstrap"), recs[0].task->name);
  assertEquals(std::string("name-0-format"), recs[1].task->name);
  assertEquals(std::string("name-0-node"), recs[2].task->name);
  assertEquals(std::string("name-0-node"), recs[3].task->name);
  assertEquals(resourceId(formatTaskToReoffer.resources[0]), resourceId(recs[2].task->resources[0]));
}
KAT(OfferEvaluatorTest_testRelaunchFailedPod) {  // :578-668
  Fixture f;
  PodInstance podInstance = resourceSetSeqPod();
  auto recs = f.evaluator.evaluate(PodInstanceRequirement::newBuilder(podInstance, {"format"}), {sufficientSeqOffer()});
  assertEquals(std::string("RRRR-RRC-L--"), types(recs));
  f.recordOperations(recs);
  // again as non-failed: the pod is fine according to the state store, it wants its reservations, not fresh ones
  auto nodeReq = PodInstanceRequirement::newBuilder(podInstance, {"node"});
  assertEquals(0u, f.evaluator.evaluate(nodeReq, {sufficientSeqOffer()}).size());
  // flagged as permanently failed: a new deployment replaces the prior one; only node is launched
  nodeReq.recoveryType = RecoveryType::PERMANENT;
  recs = f.evaluator.evaluate(nodeReq, {sufficientSeqOffer()});
  assertEquals(std::string("RRRR-RRC--L-"), types(recs));
  assertEquals(std::string("name-0-backup"), recs[4].task->name);
  assertEquals(std::string("name-0-bootstrap"), recs[8].task->name);
  assertEquals(std::string("name-0-format"), recs[9].task->name);
  assertEquals(std::string("name-0-node"), recs[10].task->name);
  assertEquals(std::string("name-0-node"), recs[11].task->name);
}
KAT(OfferEvaluatorTest_testTransientToPermanentFailure) {  // :670-741
  Fixture f;
  PodInstance podInstance = resourceSetSeqPod();
  auto req = PodInstanceRequirement::newBuilder(podInstance, {"node"});
  auto recs = f.evaluator.evaluate(req, {sufficientSeqOffer()});
  assertEquals(std::string("RRRR-RRC--L-"), types(recs));
  f.recordOperations(recs);
  f.stateStore.storeStatus("name-0-node", "TASK_LOST");   // the agent is gone
  auto tasks = f.stateStore.fetchTasks();                  // FailureUtils.setPermanentlyFailed(stateStore, podInstance)
  for (auto& t : tasks) t.labels["permanently-failed"] = "true";
  f.stateStore.storeTasks(tasks);
  assertEquals(std::string("RRRR-RRC--L-"), types(f.evaluator.evaluate(req, {sufficientSeqOffer()})));
}
// :743-802 (not annotated @Test in the reference, restated all the same): src/test/resources/valid-minimal-volume.yml
KAT(OfferEvaluatorTest_testReplaceDeployStep) {
  Fixture f;
  auto pod = std::make_shared<PodSpec>();
  pod->type = "pod-type";
  TaskSpec t; t.name = "task-name";
  t.resourceSet = ResourceSetBuilder("hello-world-role", ANY_ROLE, "hello-world-principal").id("task-name-resource-set").cpus(1.0).memory(512).addRootVolume(10, "container-path").build();
  pod->tasks = {t};
  auto req = PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"task-name"});
  mesos::Offer sufficient = OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(3.0), ResourceTestUtils::getUnreservedMem(1024), ResourceTestUtils::getUnreservedDisk(500.0)});
  auto recs = f.evaluator.evaluate(req, {sufficient});
  assertEquals(9u, recs.size());
  assertTrue(recs[7].type == Rec::LAUNCH_GROUP);
  f.recordOperations(recs);
  auto tasks = f.stateStore.fetchTasks();   // `pod replace`: everything in the pod is marked permanently failed
  for (auto& x : tasks) x.labels["permanently-failed"] = "true";
  f.stateStore.storeTasks(tasks);
  assertEquals(std::string("RRRRRRCL-"), types(f.evaluator.evaluate(req, {sufficient})));
}

// src/test/resources/resource-refinement.yml: service hello-world, pod-type with pre-reserved-role base-role, test-task cpus 1.0
namespace {
PodInstanceRequirement refinementRequirement() {
  auto pod = std::make_shared<PodSpec>();
  pod->type = "pod-type";
  pod->preReservedRole = TestConstants::PRE_RESERVED_ROLE;
  TaskSpec t; t.name = "test-task";
  t.resourceSet = ResourceSetBuilder(TestConstants::PRE_RESERVED_ROLE + "/hello-world-role", TestConstants::PRE_RESERVED_ROLE, "hello-world-principal").id("test-task-resource-set").cpus(1.0).build();
  pod->tasks = {t};
  return PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"test-task"});
}
mesos::Resource staticallyReservedTo(mesos::Resource r, const std::string& role) {
  mesos::ReservationInfo st;
  st.type = mesos::ReservationInfo::STATIC; st.role = role;
  r.reservations.push_back(st);
  return r;
}
}  // namespace
KAT(OfferEvaluatorTest_testResourceRefinementSucceeds) {  // :804-855
  Fixture f;
  const std::string base = TestConstants::PRE_RESERVED_ROLE;
  mesos::Offer sufficient = OfferTestUtils::getCompleteOffer({staticallyReservedTo(ResourceTestUtils::getUnreservedCpus(0.1), base), staticallyReservedTo(ResourceTestUtils::getUnreservedMem(256), base),
                                                              staticallyReservedTo(ResourceTestUtils::getUnreservedDisk(512), base), staticallyReservedTo(ResourceTestUtils::getUnreservedCpus(3.0), base)});
  auto recs = f.evaluator.evaluate(refinementRequirement(), {sufficient});
  assertEquals(6u, recs.size());
  const mesos::Resource& reserve = recs[0].resource;
  assertEquals(2u, reserve.reservations.size());
  assertTrue(reserve.reservations[0].type == mesos::ReservationInfo::STATIC);
  assertEquals(base, reserve.reservations[0].role);
  assertTrue(reserve.reservations[0].labels.empty());
  assertTrue(reserve.reservations[1].type == mesos::ReservationInfo::DYNAMIC);
  assertEquals(base + "/hello-world-role", reserve.reservations[1].role);
  assertFalse(reserve.reservations[1].labels.empty());
}
KAT(OfferEvaluatorTest_testResourceRefinementFailsForMissingPreReservation) {  // :857-881
  Fixture f;
  assertEquals(0u, f.evaluator.evaluate(refinementRequirement(), {OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(3.0)})}).size());
}
KAT(OfferEvaluatorTest_testResourceRefinementFailsForDifferentPreReservation) {  // :883-915
  Fixture f;
  mesos::Offer bad = OfferTestUtils::getOffer({staticallyReservedTo(ResourceTestUtils::getUnreservedCpus(3.0), "different-role")});
  assertEquals(0u, f.evaluator.evaluate(refinementRequirement(), {bad}).size());
  // (and with everything the pod needs under that other role it still fails)
  mesos::Offer alsoBad = OfferTestUtils::getOffer({staticallyReservedTo(ResourceTestUtils::getUnreservedCpus(3.1), "different-role"), staticallyReservedTo(ResourceTestUtils::getUnreservedMem(256), "different-role"),
                                                   staticallyReservedTo(ResourceTestUtils::getUnreservedDisk(512), "different-role")});
  assertEquals(0u, f.evaluator.evaluate(refinementRequirement(), {alsoBad}).size());
}
KAT(OfferEvaluatorTest_testEvaluationPipelineGeneratesSingleTLSEvaluationPerTask) {  // :1147-1167
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  for (int i = 0; i < 2; ++i) {
    TaskSpec t; t.name = TestConstants::TASK_NAME + std::to_string(i); t.resourceSet = PodInstanceRequirementTestUtils::getCpuResourceSet(1.0); t.usesTLS = true;
    pod->tasks.push_back(t);
  }
  auto req = PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {pod->tasks[0].name, pod->tasks[1].name});
  CycleEncoder enc;
  EncoderRequirement er;
  er.requirement = &req;
  EncodedCycle c = enc.encode({OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))}, {}, {er});
  std::vector<std::string> tlsTasks;
  for (auto& s : c.classStages[(size_t)c.r_class[0]]) if (s.kind == StageInfo::TLS) tlsTasks.push_back(s.taskSpecName);
  std::sort(tlsTasks.begin(), tlsTasks.end());
  assertEquals(2u, tlsTasks.size());
  assertEquals(pod->tasks[0].name, tlsTasks[0]);
  assertEquals(pod->tasks[1].name, tlsTasks[1]);
}

// ---- T/offer/evaluate/OfferEvaluatorPlacementTest.java
namespace {
size_t withAgentRule(std::optional<RulePtr> rule) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getRequirement(PodInstanceRequirementTestUtils::getCpuResourceSet(1.0), TestConstants::POD_TYPE, 0, std::move(rule));
  return f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))}).size();
}
}  // namespace
KAT(OfferEvaluatorPlacementTest_testAvoidAgents) {  // :27-71
  assertEquals(0u, withAgentRule(PlacementUtils::getAgentPlacementRule({TestConstants::AGENT_ID}, {})));
  assertEquals(6u, withAgentRule(PlacementUtils::getAgentPlacementRule({"some-random-agent"}, {})));
}
KAT(OfferEvaluatorPlacementTest_testColocateAgents) {  // :73-116
  assertEquals(0u, withAgentRule(PlacementUtils::getAgentPlacementRule({}, {"some-random-agent"})));
  assertEquals(6u, withAgentRule(PlacementUtils::getAgentPlacementRule({}, {TestConstants::AGENT_ID})));
}

// ---- T/offer/evaluate/OfferEvaluatorPortsTest.java
KAT(OfferEvaluatorPortsTest_testReserveTaskDynamicPort) {  // :129-149
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({0}), {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(10000, 10000))});
  assertEquals(6u, recs.size());
  assertTrue(recs[4].type == Rec::LAUNCH_GROUP);
  assertFalse(recs[5].hasOperation());
  assertFalse(resourceId(recs[4].task->resources[0]).empty());
  assertEquals((int64_t)10000, recs[4].task->resources[0].value.ranges[0].first);   // what PORT_..._0 would be set to
  assertEquals((int64_t)10000, recs[4].task->discovery_ports.at("test-port-0"));
}
// :309-343 and :345-379: a NamedVIPSpec is a PortSpec plus discovery labels (host-side protobuf); the port part is the path
namespace {
void vipPort(int taskPort) {
  Fixture f;
  auto b = PodInstanceRequirementTestUtils::builder();
  b.port("test-vip-port", taskPort);   // getVIPResourceSet (:130-160): the resource set holds nothing but the VIP port
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(b.build(), 0), {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(10000, 10000))});
  assertEquals(6u, recs.size());
  assertTrue(recs[4].type == Rec::LAUNCH_GROUP);
  assertFalse(recs[5].hasOperation());
  const mesos::Resource& fulfilled = recs[4].task->resources[0];
  assertEquals((int64_t)10000, fulfilled.value.ranges[0].first);
  assertFalse(resourceId(fulfilled).empty());
  assertEquals((int64_t)10000, recs[4].task->discovery_ports.at("test-vip-port"));
}
}  // namespace
KAT(OfferEvaluatorPortsTest_testReserveTaskNamedVIPPort) { vipPort(10000); }
KAT(OfferEvaluatorPortsTest_testReserveTaskDynamicVIPPort) { vipPort(0); }

// ---- T/offer/evaluate/OfferEvaluatorVolumesTest.java
KAT(OfferEvaluatorVolumesTest_testCreateMultipleRootVolumes) {  // :320-374
  Fixture f;
  auto rs = PodInstanceRequirementTestUtils::builder().cpus(1.0).addRootVolume(1.0, TestConstants::CONTAINER_PATH + "-a").addRootVolume(2.0, TestConstants::CONTAINER_PATH + "-b").build();
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(rs, 0), {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedDisk(3)})});
  assertEquals(std::string("RRRRRCRCL-"), types(recs));
  assertEquals(TestConstants::CONTAINER_PATH + "-a", recs[5].resource.disk->container_path);
  assertEquals(TestConstants::CONTAINER_PATH + "-b", recs[7].resource.disk->container_path);
  for (auto& r : recs[8].task->resources) assertFalse(resourceId(r).empty());
}
namespace {
PodInstanceRequirement executorVolumeRequirement() {  // PodInstanceRequirementTestUtils.getExecutorRequirement (:191-217)
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  TaskSpec t; t.name = TestConstants::TASK_NAME; t.resourceSet = PodInstanceRequirementTestUtils::getCpuResourceSet(1.0);
  pod->tasks = {t};
  VolumeSpec v;
  v.type = VolumeSpec::MOUNT; v.size = 1000; v.containerPath = TestConstants::CONTAINER_PATH; v.role = TestConstants::ROLE; v.principal = TestConstants::PRINCIPAL;
  pod->volumes = {v};
  return PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {TestConstants::TASK_NAME});
}
}  // namespace
KAT(OfferEvaluatorVolumesTest_testReserveCreateExecutorVolumeRefined) {  // :533-541 (body :486-531)
  Fixture f;
  auto recs = f.evaluator.evaluate(executorVolumeRequirement(), {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedMountVolume(2000), ResourceTestUtils::getUnreservedCpus(1.0)})});
  assertEquals(8u, recs.size());
  const mesos::Resource& reserve = recs[0].resource;
  assertTrue(recs[0].type == Rec::RESERVE);
  assertEquals(2000.0, reserve.value.scalar);     // the whole MOUNT disk is taken
  assertEquals(1u, reserve.reservations.size());
  assertEquals(TestConstants::ROLE, ResourceUtils::getRole(reserve));
  assertTrue(reserve.disk->source == mesos::DiskInfo::MOUNT);
  assertEquals(TestConstants::PRINCIPAL, reserve.reservations[0].principal);
  assertEquals(36u, resourceId(reserve).size());
  assertTrue(recs[1].type == Rec::CREATE);
  assertEquals(resourceId(reserve), resourceId(recs[1].resource));
  assertEquals(36u, recs[1].resource.disk->persistence_id->size());
  assertTrue(recs[1].resource.disk->source == mesos::DiskInfo::MOUNT);
}
KAT(OfferEvaluatorVolumesTest_testRelaunchExecutorVolumeFailure) {  // :543-599
  Fixture f;
  auto req = executorVolumeRequirement();
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedMountVolume(2000), ResourceTestUtils::getUnreservedCpus(1.0)})});
  assertEquals(std::string("RCRRRRL-"), types(recs));
  const std::string id = resourceId(recs[0].resource);
  const std::string persistenceId = *recs[1].resource.disk->persistence_id;
  // nothing was recorded, so this is a NEW pod again: the volume that now carries a reservation is not an unreserved MOUNT
  // disk any more, and the reserved cpus are not in the "*" pool
  mesos::Resource mount = reserved(ResourceTestUtils::getUnreservedMountVolume(2000), id);
  mount.disk->persistence_id = persistenceId; mount.disk->container_path = TestConstants::CONTAINER_PATH;
  assertEquals(0u, f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer({mount, getReservedCpus(1.0, id)})}).size());
}

// ---- T/offer/evaluate/PortEvaluationStageTest.java, stage by stage through a pod that holds nothing but the port
namespace {
struct PortStage {
  Fixture f{[] { SchedulerConfig c; c.executorResources.clear(); return c; }()};   // the stage alone: no executor overhead around it
  std::vector<Rec> evaluate(const ResourceSet& rs, const mesos::Offer& offer, const std::string& podPreReservedRole = ANY_ROLE) {
    auto pod = std::make_shared<PodSpec>();
    pod->type = TestConstants::POD_TYPE; pod->preReservedRole = podPreReservedRole;
    TaskSpec t; t.name = TestConstants::TASK_NAME; t.resourceSet = rs;
    pod->tasks = {t};
    return f.evaluator.evaluate(PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {TestConstants::TASK_NAME}), {offer});
  }
};
ResourceSetBuilder portOnly(const std::string& preReservedRole = ANY_ROLE) {
  ResourceSetBuilder b(TestConstants::ROLE, preReservedRole, TestConstants::PRINCIPAL);
  b.id("resourceSet");
  return b;
}
}  // namespace
KAT(PortEvaluationStageTest_testPortResourceIsIgnoredOnOverlay) {  // :138-179: port 80 is not on offer, and need not be
  PortStage s;
  auto recs = s.evaluate(portOnly().port("overlay-port-name", 80, {}, {"dcos"}).build(), OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10000, 10000)}));
  assertEquals(std::string("L-"), types(recs));     // passes with no recommendation of its own
  assertEquals(0u, recs[0].task->resources.size());
  assertEquals((int64_t)80, recs[0].task->discovery_ports.at("overlay-port-name"));
}
KAT(PortEvaluationStageTest_testDynamicPortResourceOnOverlay) {  // :181-223: DcosConstants.OVERLAY_DYNAMIC_PORT_RANGE_START
  PortStage s;
  auto recs = s.evaluate(portOnly().port("dyn-port-name", 0, {}, {"dcos"}).build(), OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10000, 10000)}));
  assertEquals(std::string("L-"), types(recs));
  assertEquals(0u, recs[0].task->resources.size());
  assertEquals((int64_t)1025, recs[0].task->discovery_ports.at("dyn-port-name"));
}
KAT(PortEvaluationStageTest_testDynamicPortResourceOnOverlayAndRequestedPort) {  // :225-293: 1025 is asked for explicitly, the dynamic one moves on
  PortStage s;
  auto recs = s.evaluate(portOnly().port("explicit-port", 1025, {}, {"dcos"}).port("dynamic-port", 0, {}, {"dcos"}).build(),
                         OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10000, 10000)}));
  assertEquals(std::string("L-"), types(recs));
  assertEquals(0u, recs[0].task->resources.size());
  assertEquals((int64_t)1025, recs[0].task->discovery_ports.at("explicit-port"));
  assertEquals((int64_t)1026, recs[0].task->discovery_ports.at("dynamic-port"));
}
KAT(PortEvaluationStageTest_testDynamicPortNotStickyAfterReplacement) {  // :451-531
  PortStage s;
  const ResourceSet rs = portOnly().port("TEST", 0).build();
  auto recs = s.evaluate(rs, OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10000, 10050)}));
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)10000, recs[1].task->discovery_ports.at("TEST"));   // the lowest port on offer
  // The stage is then re-run with a PodInfoBuilder that knows the old task and WITHOUT a resource id, i.e. as the new-pod
  // pipeline does for a PERMANENT recovery: the port sticks, and an offer without 10000 fails ...
  s.f.recordOperations(recs);
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  TaskSpec t; t.name = TestConstants::TASK_NAME; t.resourceSet = rs;
  pod->tasks = {t};
  auto replace = PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {TestConstants::TASK_NAME});
  replace.recoveryType = RecoveryType::PERMANENT;
  mesos::Offer without10000 = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10001, 10050)});
  assertEquals(0u, s.f.evaluator.evaluate(replace, {without10000}).size());
  mesos::Offer with10000 = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(9000, 10050)});
  recs = s.f.evaluator.evaluate(replace, {without10000, with10000});
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)10000, recs[0].resource.value.ranges[0].first);     // not 9000
  // ... but once the task is marked permanently failed the old port is forgotten and the dynamic port re-rolls
  auto tasks = s.f.stateStore.fetchTasks();
  for (auto& x : tasks) x.labels["permanently-failed"] = "true";
  s.f.stateStore.storeTasks(tasks);
  recs = s.f.evaluator.evaluate(replace, {without10000});
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)10001, recs[0].resource.value.ranges[0].first);
}
KAT(PortEvaluationStageTest_testDynamicPortWithPrereservedRoleFails) {  // :533-572: no slave_public ports on offer
  PortStage s;
  assertEquals(0u, s.evaluate(portOnly("slave_public").port("TEST", 0).build(), OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(10000, 10050)}), "slave_public").size());
}
KAT(PortEvaluationStageTest_testDynamicPortWithPrereservedRole) {  // :574-612
  PortStage s;
  auto recs = s.evaluate(portOnly("slave_public").port("TEST", 0).build(), OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedPorts(23, 5050, "slave_public")}, "slave_public"), "slave_public");
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)23, recs[0].resource.value.ranges[0].first);
  assertEquals(2u, recs[0].resource.reservations.size());   // refined on top of slave_public
}
namespace {
std::vector<Rec> dynamicWithin(std::vector<RangeSpec> ranges, int offeredBegin, int offeredEnd) {
  PortStage s;
  return s.evaluate(portOnly("slave_public").port("TEST", 0, std::move(ranges)).build(),
                    OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedPorts(offeredBegin, offeredEnd, "slave_public")}, "slave_public"), "slave_public");
}
}  // namespace
KAT(PortEvaluationStageTest_testDynamicPortWithMultipleRangesFails) {  // :706-750: 23..5050 touches neither 6000-8000 nor 2-21
  assertEquals(0u, dynamicWithin({RangeSpec(6000, 8000), RangeSpec(2, 21)}, 23, 5050).size());
  auto recs = dynamicWithin({RangeSpec(6000, 8000), RangeSpec(2, 23)}, 23, 5050);   // one port of overlap is enough
  assertEquals((int64_t)23, recs[0].resource.value.ranges[0].first);
}
KAT(PortEvaluationStageTest_testDynamicPortWithUnboundedRanges) {  // :752-795: RangeSpec(1024, null) has no upper bound
  auto recs = dynamicWithin({RangeSpec(1024)}, 3000, 5050);
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)3000, recs[0].resource.value.ranges[0].first);
}
KAT(PortEvaluationStageTest_testDynamicPortWithRangeInclusiveUpperbound) {  // :797-840
  auto recs = dynamicWithin({RangeSpec(1024, 3000)}, 3000, 5050);
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)3000, recs[0].resource.value.ranges[0].first);
}
KAT(PortEvaluationStageTest_testDynamicPortWithRangeInclusiveLowerbound) {  // :842-886
  auto recs = dynamicWithin({RangeSpec(5050, 6000)}, 3000, 5050);
  assertEquals(std::string("RL-"), types(recs));
  assertEquals((int64_t)5050, recs[0].resource.value.ranges[0].first);
}

// ---- T/offer/MesosResourcePoolTest.java: what lands in which pool, seen through what a pod can then take from the offer
namespace {
bool takes(const PodInstanceRequirement& req, const std::vector<mesos::Resource>& offered) {
  PortStage s;
  return !s.f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)}).empty();
}
}  // namespace
KAT(MesosResourcePoolTest_testEmptyUnreservedAtomicPool) {  // :24-29: cpus are not a MOUNT disk
  assertFalse(takes(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1000), {ResourceTestUtils::getUnreservedCpus(1.0)}));
}
KAT(MesosResourcePoolTest_testCreateSingleUnreservedAtomicPool) {  // :31-38
  assertTrue(takes(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1000), {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(1000)}));
}
KAT(MesosResourcePoolTest_testCreateSingleReservedAtomicPool) {  // :40-51: a reserved MOUNT disk is in the by-id pool, not the atomic one
  assertFalse(takes(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1000), {ResourceTestUtils::getUnreservedCpus(1.0), reserved(ResourceTestUtils::getUnreservedMountVolume(1000), "some-resource-id")}));
}
KAT(MesosResourcePoolTest_testMultipleUnreservedAtomicPool) {  // :53-61: two disks, two volumes; a third finds nothing
  auto two = PodInstanceRequirementTestUtils::builder().cpus(1.0).addMountVolume(1000, "a").addMountVolume(1000, "b").build();
  auto three = PodInstanceRequirementTestUtils::builder().cpus(1.0).addMountVolume(1000, "a").addMountVolume(1000, "b").addMountVolume(1000, "c").build();
  std::vector<mesos::Resource> offered = {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(1000), ResourceTestUtils::getUnreservedMountVolume(1000)};
  assertTrue(takes(PodInstanceRequirementTestUtils::getRequirement(two, 0), offered));
  assertFalse(takes(PodInstanceRequirementTestUtils::getRequirement(three, 0), offered));
}
KAT(MesosResourcePoolTest_testConsumeUnreservedMergedResource) {  // :165-179: 1.0 of 1.0 leaves exactly zero
  auto one = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  assertTrue(takes(one, {ResourceTestUtils::getUnreservedCpus(1.0)}));
  auto twoTasks = std::make_shared<PodSpec>();
  twoTasks->type = TestConstants::POD_TYPE;
  for (const char* n : {"a", "b"}) {
    TaskSpec t; t.name = n;
    t.resourceSet = ResourceSetBuilder(TestConstants::ROLE, ANY_ROLE, TestConstants::PRINCIPAL).id(std::string("rs-") + n).cpus(n[0] == 'a' ? 1.0 : 0.0).build();
    twoTasks->tasks.push_back(t);
  }
  // after task a's 1.0 the pool holds the SCALAR zero: a second request of 0.0 still fits, anything above it does not
  assertTrue(takes(PodInstanceRequirement::newBuilder(PodInstance{twoTasks, 0}, {"a", "b"}), {ResourceTestUtils::getUnreservedCpus(1.0)}));
}
KAT(MesosResourcePoolTest_testConsumeReservedMergedResource) {  // :152-163: consumeReserved finds the resource by its id, whole
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  auto res = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(1.0)});
  auto offered = getExpectedExecutorResources(f.stateStore.fetchTasks().front().executor);
  offered.push_back(getReservedCpus(1.0, resourceId(res[3])));
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(resourceId(res[3]), resourceId(recs[0].task->resources[0]));
}

// ---- T/offer/RangeUtilsTest.java:19-96.  mergeRanges is pinned in kat_evaluator.cpp (RangeUtilsTest_merging); subtractRanges and
// isInAny are what the static-port and dynamic-port stages compute on the device, so they are restated through ports.
KAT(RangeUtilsTest_testMergingOverlappingRanges) { IntervalSet a; a.add(1, 3); a.add(2, 4); assertEquals(1u, a.iv.size()); assertEquals((int64_t)1, a.iv[0].first); assertEquals((int64_t)4, a.iv[0].second); }
KAT(RangeUtilsTest_testMergingNonOverlappingRanges) {
  IntervalSet b; b.add(1, 3); b.add(5, 7);
  assertEquals(2u, b.iv.size());
  assertEquals((int64_t)1, b.iv[0].first); assertEquals((int64_t)3, b.iv[0].second); assertEquals((int64_t)5, b.iv[1].first); assertEquals((int64_t)7, b.iv[1].second);
}
KAT(RangeUtilsTest_testMergingInternallyOverlappingRanges) { IntervalSet c; c.add(1, 5); c.add(2, 4); assertEquals(1u, c.iv.size()); assertEquals((int64_t)1, c.iv[0].first); assertEquals((int64_t)5, c.iv[0].second); }
KAT(RangeUtilsTest_testSubtractRanges) {  // {[1,3],[5,7]} - {[1,3]} = {[5,7]}: after static 1,2,3 the dynamic ports are 5,6,7 and then none
  PortStage s;
  mesos::Offer offer = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(1, 3), ResourceTestUtils::getUnreservedPorts(5, 7)});
  auto b = portOnly();
  for (int p : {1, 2, 3}) b.port("s" + std::to_string(p), p);
  for (int k = 0; k < 3; ++k) b.port("d" + std::to_string(k), 0);
  auto recs = s.evaluate(b.build(), offer);
  assertEquals(std::string("RRRRRRL-"), types(recs));
  for (int k = 0; k < 3; ++k) assertEquals((int64_t)(5 + k), recs[(size_t)(3 + k)].resource.value.ranges[0].first);
  b.port("d3", 0);
  assertEquals(0u, s.evaluate(b.build(), offer).size());
}
KAT(RangeUtilsTest_testSubtractRangesForNilResult) {  // {[2,3]} - {[1,5]} is empty: ports 2 and 3 are both on offer
  PortStage s;
  auto recs = s.evaluate(portOnly().port("two", 2).port("three", 3).build(), OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(1, 5)}));
  assertEquals(std::string("RRL-"), types(recs));
}
KAT(RangeUtilsTest_testRangeIsInAny) {
  PortStage s;
  mesos::Offer r1 = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(1, 3), ResourceTestUtils::getUnreservedPorts(5, 7)});
  for (int p : {1, 2, 3, 5, 6, 7}) assertEquals(3u, s.evaluate(portOnly().port("p", p).build(), r1).size());
  for (int p : {4, 8, 65535}) assertEquals(0u, s.evaluate(portOnly().port("p", p).build(), r1).size());
  mesos::Offer r2 = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedPorts(2, 2)});
  assertEquals(0u, s.evaluate(portOnly().port("p", 1).build(), r2).size());
  assertEquals(3u, s.evaluate(portOnly().port("p", 2).build(), r2).size());
  assertEquals(0u, s.evaluate(portOnly().port("p", 3).build(), r2).size());
}

// ---- T/offer/evaluate/OfferEvaluationUtilsTest.java:49-187 (evaluateSimpleResource against a mocked pool): the same five
// situations with a real pool, desired 5 / current 4 as there
namespace {
struct Blocks {  // the reference uses a made-up scalar called "blocks"; any scalar name takes the same code path
  Fixture f{[] { SchedulerConfig c; c.executorResources.clear(); return c; }()};
  PodInstanceRequirement want(double v) { return PodInstanceRequirementTestUtils::getRequirement(PodInstanceRequirementTestUtils::builder().scalar("blocks", v).build(), 0); }
  mesos::Resource blocks(double v) { return ResourceTestUtils::unreservedScalar("blocks", v); }
};
}  // namespace
KAT(OfferEvaluationUtilsTest_testResourceInsufficient) {
  Blocks b;
  assertEquals(0u, b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({b.blocks(4)})}).size());          // new reservation
  auto recs = b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({b.blocks(5)})});
  b.f.recordOperations(recs);
  assertEquals(0u, b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({b.blocks(5)})}).size());          // with a resource id: the id is not in the offer
}
KAT(OfferEvaluationUtilsTest_testResourceSufficient) {
  Blocks b;
  auto recs = b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({b.blocks(5)})});
  assertEquals(std::string("RL-"), types(recs));                                                                 // no id yet: a RESERVE of the desired value
  assertEquals(5.0, recs[0].resource.value.scalar);
  assertFalse(resourceId(recs[0].resource).empty());
  b.f.recordOperations(recs);
  auto again = b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({reserved(b.blocks(5), resourceId(recs[0].resource))})});
  assertEquals(std::string("L-"), types(again));                                                                 // with the id and the same value: nothing to do
  assertEquals(resourceId(recs[0].resource), resourceId(again[0].task->resources[0]));
}
KAT(OfferEvaluationUtilsTest_testResourceIncreaseSufficient) {
  Blocks b;
  auto recs = b.f.evaluator.evaluate(b.want(4), {OfferTestUtils::getOffer({b.blocks(4)})});
  b.f.recordOperations(recs);
  const std::string id = resourceId(recs[0].resource);
  auto grown = b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({reserved(b.blocks(4), id), b.blocks(1)})});
  assertEquals(std::string("RL-"), types(grown));
  assertEquals(1.0, grown[0].resource.value.scalar);       // toAdd = desired - current
  assertEquals(id, resourceId(grown[0].resource));
}
KAT(OfferEvaluationUtilsTest_testResourceIncreaseInsufficient) {
  Blocks b;
  auto recs = b.f.evaluator.evaluate(b.want(4), {OfferTestUtils::getOffer({b.blocks(4)})});
  b.f.recordOperations(recs);
  const std::string id = resourceId(recs[0].resource);
  assertEquals(0u, b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({reserved(b.blocks(4), id), b.blocks(0.5)})}).size());
  assertEquals(0u, b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({reserved(b.blocks(4), id)})}).size());
}
KAT(OfferEvaluationUtilsTest_testResourceDecrease) {
  Blocks b;
  auto recs = b.f.evaluator.evaluate(b.want(5), {OfferTestUtils::getOffer({b.blocks(5)})});
  b.f.recordOperations(recs);
  const std::string id = resourceId(recs[0].resource);
  auto shrunk = b.f.evaluator.evaluate(b.want(4), {OfferTestUtils::getOffer({reserved(b.blocks(5), id)})});
  assertEquals(std::string("UL-"), types(shrunk));
  assertEquals(1.0, shrunk[0].resource.value.scalar);      // the surplus goes back
  assertEquals(id, resourceId(shrunk[0].resource));
  assertEquals(4.0, shrunk[1].task->resources[0].value.scalar);
}

// ---- T/offer/evaluate/VolumeEvaluationStageTest.java
KAT(VolumeEvaluationStageTest_testCreateSucceeds) {  // :20-70
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1000), {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(2000)})});
  assertEquals(std::string("RRRRRCL-"), types(recs));
  assertEquals(std::string("disk"), recs[4].resource.name);
  assertEquals(2000.0, recs[4].resource.value.scalar);     // the offered disk's size, not the requested 1000
  assertFalse(resourceId(recs[4].resource).empty());
  assertEquals(std::string("disk"), recs[5].resource.name);
  assertEquals(2000.0, recs[5].resource.value.scalar);
  assertEquals(resourceId(recs[4].resource), resourceId(recs[5].resource));
  assertFalse(recs[5].resource.disk->persistence_id->empty());
}
KAT(VolumeEvaluationStageTest_testCreateFails) {  // :72-100
  Fixture f;
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 2000), {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(1000)})}).size());
}

// ---- T/offer/evaluate/ExecutorEvaluationStageTest.java:25-108.  The reference builds the stage by hand around the executor id
// of a RUNNING task; through the evaluator that is a pod with a second task that keeps running while the first is relaunched.
namespace {
struct RunningExecutor : Fixture {
  std::shared_ptr<PodSpec> pod = std::make_shared<PodSpec>();
  std::vector<mesos::Resource> taskAResources;
  RunningExecutor() {
    pod->type = TestConstants::POD_TYPE;
    for (const char* n : {"a", "b"}) {
      TaskSpec t; t.name = n;
      t.resourceSet = ResourceSetBuilder(TestConstants::ROLE, ANY_ROLE, TestConstants::PRINCIPAL).id(std::string("rs-") + n).cpus(1.0).build();
      pod->tasks.push_back(t);
    }
    auto recs = evaluator.evaluate(PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"a", "b"}), {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.1))});
    recordOperations(recs);
    auto tasks = stateStore.fetchTasks();
    for (auto& t : tasks) { t.executor.executor_id = "pod-type__the-executor"; if (t.name == "pod-type-0-a") taskAResources = t.resources; }
    stateStore.storeTasks(tasks);
    stateStore.storeStatus("pod-type-0-b", "TASK_RUNNING");
  }
  std::vector<Rec> relaunchA(const mesos::Offer& offer) { return evaluator.evaluate(PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"a"}), {offer}); }
};
}  // namespace
KAT(ExecutorEvaluationStageTest_testRejectOfferWithoutExpectedExecutorId) {
  RunningExecutor f;
  assertEquals(0u, f.relaunchA(OfferTestUtils::getCompleteOffer(f.taskAResources)).size());
}
KAT(ExecutorEvaluationStageTest_testAcceptOfferWithExpectedExecutorId) {
  RunningExecutor f;
  mesos::Offer offer = OfferTestUtils::getCompleteOffer(f.taskAResources);
  offer.executor_ids = {"pod-type__the-executor"};
  auto recs = f.relaunchA(offer);
  assertEquals(std::string("L-"), types(recs));
  assertEquals(std::string("pod-type__the-executor"), recs[0].task->executor.executor_id);
}

// ---- T/offer/evaluate/LaunchEvaluationStageTest.java: the stage always passes and writes the labels the NEXT cycle's placement
// rules read back (S/offer/evaluate/LaunchEvaluationStage.java:58-68).  label1 / target_configuration come from the task spec and
// the config store, outside the path.
namespace {
mesos::TaskInfo launchedOn(const mesos::Offer& offer) {
  PortStage s;
  auto recs = s.f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {offer});
  assertEquals(std::string("RL-"), types(recs));   // isPassing
  return *recs[1].task;
}
}  // namespace
KAT(LaunchEvaluationStageTest_isPassing) { launchedOn(OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(2.0)})); }
KAT(LaunchEvaluationStageTest_labelsAreCorrect) {
  mesos::TaskInfo t = launchedOn(OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(2.0)}));
  assertEquals(4u, t.labels.size());
  assertEquals(std::string("0"), t.labels.at("index"));
  assertEquals(std::string(""), t.labels.at("offer_attributes"));
  assertEquals(TestConstants::HOSTNAME, t.labels.at("offer_hostname"));
  assertEquals(TestConstants::POD_TYPE, t.labels.at("task_type"));
}
KAT(LaunchEvaluationStageTest_regionAndZoneInjected) {
  mesos::Offer offer = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(2.0)});
  offer.domain = TestConstants::LOCAL_DOMAIN_INFO();
  mesos::TaskInfo t = launchedOn(offer);
  assertEquals(TestConstants::LOCAL_REGION, t.labels.at("offer_region"));
  assertEquals(TestConstants::ZONE, t.labels.at("offer_zone"));
}
KAT(LaunchEvaluationStageTest_regionAndZoneNotInjected) {
  mesos::TaskInfo t = launchedOn(OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(2.0)}));
  assertTrue(t.labels.find("offer_region") == t.labels.end());
  assertTrue(t.labels.find("offer_zone") == t.labels.end());
}

// ---- T/scheduler/plan/PlanSchedulerTest.java:69-96
namespace {
mesos::Offer planSchedulerOffer() {
  mesos::Offer o;
  o.id = "offerid"; o.framework_id = "frameworkid"; o.slave_id = "slaveid"; o.hostname = "hello";
  return o;
}
}  // namespace
KAT(PlanSchedulerTest_testNonPendingStep) {
  Fixture f;
  PlanScheduler scheduler(f.evaluator);
  std::vector<Step> steps(1);
  steps[0].status = Step::PREPARED;
  assertTrue(scheduler.resourceOffers({planSchedulerOffer()}, steps).empty());
  assertTrue(steps[0].isPrepared());
  steps[0].status = Step::COMPLETE;   // "Ignoring resource offers for step" (:73-78)
  steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  assertTrue(scheduler.resourceOffers({OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))}, steps).empty());
  assertTrue(steps[0].status == Step::COMPLETE);
}
KAT(PlanSchedulerTest_testStartNoRequirement) {
  Fixture f;
  PlanScheduler scheduler(f.evaluator);
  std::vector<Step> steps(1);
  steps[0].status = Step::PENDING;
  assertTrue(scheduler.resourceOffers({planSchedulerOffer()}, steps).empty());
  assertTrue(steps[0].isPrepared());
}
KAT(PlanSchedulerTest_testEvaluateNoRecommendations) {  // the reference mocks the evaluator to return nothing; this offer really has nothing
  Fixture f;
  PlanScheduler scheduler(f.evaluator);
  std::vector<Step> steps(1);
  steps[0].status = Step::PENDING;
  steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  assertTrue(scheduler.resourceOffers({planSchedulerOffer()}, steps).empty());
  assertTrue(steps[0].isPrepared());
}

// ---- T/offer/evaluate/PlacementRuleEvaluationStageTest.java:25-104: AgentRule.require(agent) as the pod's placement rule; the
// offer from that agent passes (and the stage leaves the pool alone: the launch that follows still finds its 1.0 cpus + 0.1 of the
// executor in the 1.1 offered), the offer from another agent fails.
namespace {
size_t requireAgent(const std::string& offerAgent) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getRequirement(PodInstanceRequirementTestUtils::getCpuResourceSet(1.0), TestConstants::POD_TYPE, 0,
                                                             std::optional<RulePtr>(AgentRule::require(std::vector<std::string>{"test-agent"})));
  mesos::Offer offer = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(1.1));
  offer.slave_id = offerAgent;
  return f.evaluator.evaluate(req, {offer}).size();
}
}  // namespace
KAT(PlacementRuleEvaluationStageTest_testOfferPassesPlacementRule) { assertEquals(6u, requireAgent("test-agent")); }
KAT(PlacementRuleEvaluationStageTest_testOfferFailsPlacementRule) { assertEquals(0u, requireAgent("other-agent")); }

// ---- T/offer/OfferUtilsTest.java:24-113: filterOutAcceptedOffers keeps the offers that no recommendation with an operation names
namespace {
struct OfferUtilsFixture {
  std::vector<mesos::Offer> OFFERS;
  std::vector<Rec> RECS;
  static Rec rec(const std::string& offerId, bool withOperation) {
    Rec r;
    r.type = withOperation ? Rec::RESERVE : Rec::STORE_TASK;   // toOfferRecommendation(offer, withOperation), :114-135
    r.offerId = offerId;
    return r;
  }
  OfferUtilsFixture() {
    for (const char* id : {"no-operation", "with-without-operation", "with-operation", "no-recommendation"}) {
      mesos::Offer o = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(2), ResourceTestUtils::getUnreservedMem(1000), ResourceTestUtils::getUnreservedDisk(10000)});
      o.id = id;
      OFFERS.push_back(o);
    }
    RECS = {rec("no-operation", false), rec("with-without-operation", false), rec("with-without-operation", true), rec("with-operation", true)};
  }
  std::vector<std::string> left(const std::vector<Rec>& recs) const {
    std::vector<std::string> ids;
    for (auto& o : OfferUtils::filterOutAcceptedOffers(OFFERS, recs)) ids.push_back(o.id);
    return ids;
  }
};
using Ids = std::vector<std::string>;
}  // namespace
KAT(OfferUtilsTest_testFilterAcceptedOffers) {
  OfferUtilsFixture f;
  assertTrue(f.left({f.RECS[2]}) == (Ids{"no-operation", "with-operation", "no-recommendation"}));
}
KAT(OfferUtilsTest_testFilterAcceptedOffersNoAccepted) {
  OfferUtilsFixture f;
  assertTrue(f.left({}) == (Ids{"no-operation", "with-without-operation", "with-operation", "no-recommendation"}));
}
KAT(OfferUtilsTest_testFilterAcceptedOffersAllAccepted) {
  OfferUtilsFixture f;
  assertTrue(f.left(f.RECS) == (Ids{"no-operation", "no-recommendation"}));
}
KAT(OfferUtilsTest_testFilterAcceptedOffersAcceptedInvalidId) {
  OfferUtilsFixture f;
  assertTrue(f.left({OfferUtilsFixture::rec("abc", false), OfferUtilsFixture::rec("def", true)}) ==
             (Ids{"no-operation", "with-without-operation", "with-operation", "no-recommendation"}));
}
