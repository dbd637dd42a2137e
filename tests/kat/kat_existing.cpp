// kat_existing.cpp — relaunching a pod that already holds reservations (S/offer/evaluate/OfferEvaluator.java:538-728,
// TaskResourceMapper, ExecutorResourceMapper, evaluateSimpleResource with a resource id), restated from
// T/offer/evaluate/OfferEvaluatorTest.java:105-305, OfferEvaluatorPortsTest.java:72-259 and OfferEvaluatorVolumesTest.java.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct Fixture {
  StateStore stateStore;
  OfferEvaluator evaluator;
  Fixture() : evaluator(*kat::library(), stateStore, TestConstants::SERVICE_NAME) {}
  // T/offer/evaluate/OfferEvaluatorTestBase.java:52-94: evaluate, store the TaskInfos, hand back the reserved resources
  std::vector<mesos::Resource> recordLaunch(const mesos::Offer& offer, const PodInstanceRequirement& req) {
    auto recs = evaluator.evaluate(req, {offer});
    std::vector<mesos::Resource> reserved;
    std::vector<mesos::TaskInfo> store;
    for (auto& r : recs) {
      if (r.type == Rec::RESERVE) reserved.push_back(r.resource);
      if (r.type == Rec::STORE_TASK) store.push_back(*r.task);
    }
    stateStore.storeTasks(store);
    return reserved;
  }
  std::vector<mesos::Resource> recordLaunchWithCompleteOfferedResources(const PodInstanceRequirement& req, std::vector<mesos::Resource> offered,
                                                                        const std::string& preReservedRole = ANY_ROLE) {
    return recordLaunch(OfferTestUtils::getCompleteOffer(offered, preReservedRole), req);
  }
  // getExpectedExecutorResources (OfferEvaluatorTest.java:64-99): the executor's reservations as the agent would re-offer them
  std::vector<mesos::Resource> expectedExecutorResources() { return stateStore.fetchTasks().front().executor.resources; }
};

std::string types(const std::vector<Rec>& recs) {
  std::string s;
  for (auto& r : recs)
    s += r.type == Rec::RESERVE ? 'R' : r.type == Rec::UNRESERVE ? 'U' : r.type == Rec::CREATE ? 'C' : r.type == Rec::DESTROY ? 'D'
       : r.type == Rec::LAUNCH_GROUP ? 'L' : '-';
  return s;
}
std::string resourceId(const mesos::Resource& r) { return ResourceUtils::getResourceId(r).value_or(""); }
mesos::Resource withValue(mesos::Resource r, double v) { r.value = mesos::Value::Scalar(v); return r; }

}  // namespace

KAT(OfferEvaluatorTest_testRelaunchExpectedScalar) {  // :105-129
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(2.0)});
  const std::string id = resourceId(reserved[3]);
  auto expected = f.expectedExecutorResources();
  expected.push_back(reserved[3]);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(expected)});
  assertEquals(2u, recs.size());
  assertEquals(std::string("L-"), types(recs));
  assertEquals(id, resourceId(recs[0].task->resources[0]));
}
KAT(OfferEvaluatorTest_testIncreaseReservationScalar) {  // :131-177
  Fixture f;
  auto reserved = f.recordLaunchWithCompleteOfferedResources(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {ResourceTestUtils::getUnreservedCpus(2.0)});
  const std::string id = resourceId(reserved[3]);
  auto expected = f.expectedExecutorResources();
  expected.push_back(reserved[3]);                            // 1.0 cpus reserved
  expected.push_back(ResourceTestUtils::getUnreservedCpus(1.0));
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(2.0), {OfferTestUtils::getOffer(expected)});
  assertEquals(std::string("RL-"), types(recs));
  assertEquals(1.0, recs[0].resource.value.scalar);           // only the difference is reserved, under the same id
  assertEquals(id, resourceId(recs[0].resource));
  assertEquals(TestConstants::ROLE, ResourceUtils::getRole(recs[0].resource));
  assertEquals(id, resourceId(recs[1].task->resources[0]));
  assertEquals(2.0, recs[1].task->resources[0].value.scalar);
}
// :180-188, :282-290: the same two with the refinement capability on (see the note at the top of kat_reference_rest.cpp)
KAT(OfferEvaluatorTest_testIncreaseReservationScalarRefined) { kat_OfferEvaluatorTest_testIncreaseReservationScalar(); }
KAT(OfferEvaluatorTest_testIncreasePreReservedReservationScalar) {  // :190-236
  Fixture f;
  const std::string role = "slave_public";
  auto reserved = f.recordLaunchWithCompleteOfferedResources(PodInstanceRequirementTestUtils::getCpuRequirement(1.0, role),
                                                             {ResourceTestUtils::getUnreservedCpus(2.0, role)}, role);
  const std::string id = resourceId(reserved[3]);
  auto expected = f.expectedExecutorResources();
  expected.push_back(reserved[3]);
  expected.push_back(ResourceTestUtils::getUnreservedCpus(1.0, role));
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(2.0, role), {OfferTestUtils::getOffer(expected)});
  assertEquals(std::string("RL-"), types(recs));
  assertEquals(1.0, recs[0].resource.value.scalar);
  assertEquals(id, resourceId(recs[0].resource));
  // ... and the extra cpu must come from the pre-reserved role's pool, not from "*"
  expected.back() = ResourceTestUtils::getUnreservedCpus(1.0);
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(2.0, role), {OfferTestUtils::getOffer(expected)}).size());
}
KAT(OfferEvaluatorTest_testDecreaseReservationScalar) {  // :238-280
  Fixture f;
  auto reserved = f.recordLaunchWithCompleteOfferedResources(PodInstanceRequirementTestUtils::getCpuRequirement(2.0), {ResourceTestUtils::getUnreservedCpus(2.0)});
  const std::string id = resourceId(reserved[3]);
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);                             // 2.0 cpus reserved
  offered.push_back(ResourceTestUtils::getUnreservedCpus(1.0));
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("UL-"), types(recs));
  assertEquals(1.0, recs[0].resource.value.scalar);
  assertEquals(id, resourceId(recs[0].resource));
  assertEquals(id, resourceId(recs[1].task->resources[0]));
  assertEquals(1.0, recs[1].task->resources[0].value.scalar);
}
KAT(OfferEvaluatorTest_testDecreaseReservationScalarRefined) { kat_OfferEvaluatorTest_testDecreaseReservationScalar(); }
KAT(OfferEvaluatorTest_testFailIncreaseReservationScalar) {  // :292-305
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(2.0);
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(2.0)});
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer(withValue(reserved[3], 1.0))});
  assertEquals(0u, recs.size());
  // the same with the executor's reservations offered back and nothing unreserved to grow into
  auto offered = f.expectedExecutorResources();
  offered.push_back(withValue(reserved[3], 1.0));
  assertEquals(0u, f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)}).size());
  offered.push_back(ResourceTestUtils::getUnreservedCpus(0.5));
  assertEquals(0u, f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)}).size());
  offered.back() = ResourceTestUtils::getUnreservedCpus(1.0);
  assertEquals(std::string("RL-"), types(f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)})));
}
// an offer that lacks one of the pod's reservations is refused ("Offer lacks previously reserved ...", OfferEvaluationUtils.java:90-99),
// and the first offer that carries them all wins
KAT(OfferEvaluator_relaunchNeedsEveryReservation) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(2.0)});
  auto all = f.expectedExecutorResources();
  all.push_back(reserved[3]);
  auto missingExecutorMem = all;
  missingExecutorMem.erase(missingExecutorMem.begin() + 2);
  mesos::Offer a = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(8.0));  // plenty of room, but not OUR reservations
  a.id = "fresh";
  mesos::Offer b = OfferTestUtils::getOffer(missingExecutorMem);
  b.id = "partial";
  mesos::Offer c = OfferTestUtils::getOffer(all);
  c.id = "home";
  auto recs = f.evaluator.evaluate(req, {a, b, c});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(std::string("home"), recs[0].offerId);
}
// OfferEvaluatorPortsTest.java:72-118 (testLaunchExpectedStaticPort) and :120-165 (testLaunchExpectedDynamicPort)
KAT(OfferEvaluatorPortsTest_testLaunchExpectedStaticPort) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getPortRequirement({555});
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedPorts(555, 555)});
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(resourceId(reserved[3]), resourceId(recs[0].task->resources[0]));
  assertEquals((int64_t)555, recs[0].task->resources[0].value.ranges[0].first);
}
KAT(OfferEvaluatorPortsTest_testLaunchExpectedDynamicPort) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getPortRequirement({0});
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedPorts(10000, 10000)});
  assertEquals((int64_t)10000, reserved[3].value.ranges[0].first);
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});     // the port chosen last time is kept
  assertEquals(std::string("L-"), types(recs));
  assertEquals((int64_t)10000, recs[0].task->resources[0].value.ranges[0].first);
  assertEquals(resourceId(reserved[3]), resourceId(recs[0].task->resources[0]));
}
// OfferEvaluatorPortsTest.java:150-186 (testUpdateStaticToStaticPort) and :190-222 (testUpdateDynamicToStaticPort): the old
// port's reservation matches no spec any more -> UNRESERVE (and it is back in the pool), the new port is a fresh RESERVE
namespace {
void movesToPort666(int firstPort) {
  Fixture f;
  auto reserved = f.recordLaunchWithCompleteOfferedResources(PodInstanceRequirementTestUtils::getPortRequirement({firstPort}), {ResourceTestUtils::getUnreservedPorts(555, 555)});
  assertEquals((int64_t)555, reserved[3].value.ranges[0].first);
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);
  offered.push_back(ResourceTestUtils::getUnreservedPorts(666, 666));
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({666}), {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("URL-"), types(recs));
  assertEquals((int64_t)555, recs[0].resource.value.ranges[0].first);
  assertEquals(resourceId(reserved[3]), resourceId(recs[0].resource));
  assertEquals((int64_t)666, recs[1].resource.value.ranges[0].first);
  assertEquals((int64_t)666, recs[2].task->resources[0].value.ranges[0].first);
  // without 666 on offer the move is impossible ...
  offered.pop_back();
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({666}), {OfferTestUtils::getOffer(offered)}).size());
  // ... but the port that was just handed back is usable by a new spec of the same pod (MesosResourcePool.free)
  auto b = PodInstanceRequirementTestUtils::builder();
  b.port("renamed-port", 555);
  auto renamed = PodInstanceRequirementTestUtils::getRequirement(b.build(), 0);
  if (firstPort != 0) assertEquals(std::string("L-"), types(f.evaluator.evaluate(renamed, {OfferTestUtils::getOffer(offered)})));  // same number: matched, kept
}
}  // namespace
KAT(OfferEvaluatorPortsTest_testUpdateStaticToStaticPort) { movesToPort666(555); }
KAT(OfferEvaluatorPortsTest_testUpdateDynamicToStaticPort) { movesToPort666(0); }
KAT(OfferEvaluatorPortsTest_testLaunchExpectedMultiplePorts) {  // :224-258
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getPortRequirement({10000, 10001});
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedPorts(10000, 10001)});
  assertEquals(5u, reserved.size());
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);
  offered.push_back(reserved[4]);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(2u, recs[0].task->resources.size());
  assertEquals(resourceId(reserved[3]), resourceId(recs[0].task->resources[0]));
  assertEquals(resourceId(reserved[4]), resourceId(recs[0].task->resources[1]));
}
// a volume whose container path no spec names any more: DESTROY is the host's business, the disk space returns to the pool
KAT(OfferEvaluator_orphanedRootVolumeIsUnreserved) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500);
  f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedDisk(2000)});
  auto stored = f.stateStore.fetchTasks().front();
  auto offered = f.expectedExecutorResources();
  for (auto& r : stored.resources) offered.push_back(r);
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {OfferTestUtils::getOffer(offered)});  // the volume is gone from the spec
  assertEquals(std::string("UL-"), types(recs));
  assertEquals(std::string("disk"), recs[0].resource.name);
  assertEquals(1u, recs[1].task->resources.size());
}
KAT(OfferEvaluator_addResourceToExistingPod) {
  Fixture f;
  auto req1 = PodInstanceRequirementTestUtils::getPortRequirement({555});
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req1, {ResourceTestUtils::getUnreservedPorts(555, 555)});
  auto b = PodInstanceRequirementTestUtils::builder();
  b.port("test-port-555", 555).cpus(1.0);
  auto req2 = PodInstanceRequirementTestUtils::getRequirement(b.build(), 0);
  auto offered = f.expectedExecutorResources();
  offered.push_back(reserved[3]);
  assertEquals(0u, f.evaluator.evaluate(req2, {OfferTestUtils::getOffer(offered)}).size());  // no cpus to be had
  offered.push_back(ResourceTestUtils::getUnreservedCpus(1.5));
  auto recs = f.evaluator.evaluate(req2, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("RL-"), types(recs));    // matched port: nothing to do; missing cpus spec: new reservation
  assertEquals(std::string("cpus"), recs[0].resource.name);
  assertFalse(resourceId(recs[0].resource) == resourceId(reserved[3]));
}
// OfferEvaluatorVolumesTest.java:95-182 (testExpectedRootVolume) and :244-300 (testExpectedMountVolume)
KAT(OfferEvaluatorVolumesTest_testExpectedRootVolume) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500);
  auto reserved = f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedDisk(2000)});
  assertEquals(5u, reserved.size());
  auto stored = f.stateStore.fetchTasks().front();
  auto offered = f.expectedExecutorResources();
  for (auto& r : stored.resources) offered.push_back(r);   // cpus + the created volume, as the agent re-offers them
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(*stored.resources[1].disk->persistence_id, *recs[0].task->resources[1].disk->persistence_id);
  assertEquals(resourceId(stored.resources[1]), resourceId(recs[0].task->resources[1]));
}
KAT(OfferEvaluatorVolumesTest_testExpectedMountVolume) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1500);
  f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(2000)});
  auto stored = f.stateStore.fetchTasks().front();
  auto offered = f.expectedExecutorResources();
  for (auto& r : stored.resources) offered.push_back(r);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(2000.0, recs[0].task->resources[1].value.scalar);
  // without the volume's reservation in the offer the pod cannot come back there
  offered.pop_back();
  offered.push_back(ResourceTestUtils::getUnreservedMountVolume(2000));
  assertEquals(0u, f.evaluator.evaluate(req, {OfferTestUtils::getOffer(offered)}).size());
}
// OfferEvaluatorVolumesTest.java:173-181: despite its name, testReserveLaunchScalarRefined re-runs testExpectedRootVolume
KAT(OfferEvaluatorVolumesTest_testReserveLaunchScalarRefined) { kat_OfferEvaluatorVolumesTest_testExpectedRootVolume(); }
KAT(OfferEvaluatorVolumesTest_testExpectedMountVolumeRefined) { kat_OfferEvaluatorVolumesTest_testExpectedMountVolume(); }  // :310-318
// OfferEvaluator.java:268-293: a permanently failed pod (or `pod replace`) starts over with the new-pod pipeline
KAT(OfferEvaluator_permanentRecoveryUsesTheNewPodPipeline) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  f.recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(2.0)});
  auto fresh = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0));
  assertEquals(0u, f.evaluator.evaluate(req, {fresh}).size());     // existing pod: wants its reservations back
  auto replace = req;
  replace.recoveryType = RecoveryType::PERMANENT;
  assertEquals(std::string("RRRRL-"), types(f.evaluator.evaluate(replace, {fresh})));
  auto failed = f.stateStore.fetchTasks();
  for (auto& t : failed) t.labels["permanently-failed"] = "true";   // FailureUtils.setPermanentlyFailed
  f.stateStore.storeTasks(failed);
  assertEquals(std::string("RRRRL-"), types(f.evaluator.evaluate(req, {fresh})));
}
// ExecutorEvaluationStage.java:32-81 + ResourceEvaluationStage.java:65-84: a second task of a pod whose executor still runs must
// land on the agent that lists that executor id, and the executor's own reservations are not looked for in the offer
KAT(ExecutorEvaluationStage_runningExecutorIsReused) {
  Fixture f;
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  for (const char* name : {"a", "b"}) {
    TaskSpec t;
    t.name = name;
    t.resourceSet = ResourceSetBuilder(TestConstants::ROLE, ANY_ROLE, TestConstants::PRINCIPAL).id(std::string("rs-") + name).cpus(1.0).build();
    pod->tasks.push_back(t);
  }
  PodInstance pi{pod, 0};
  auto both = PodInstanceRequirement::newBuilder(pi, {"a", "b"});
  f.recordLaunch(OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(4.0)}), both);
  auto tasks = f.stateStore.fetchTasks();
  assertEquals(2u, tasks.size());
  for (auto& t : tasks) { t.executor.executor_id = "executor-1"; }
  f.stateStore.storeTasks(tasks);
  f.stateStore.storeStatus(tasks[0].name, "TASK_RUNNING");   // task a keeps running, task b is relaunched
  auto relaunchB = PodInstanceRequirement::newBuilder(pi, {"b"});
  mesos::Offer wrongAgent = OfferTestUtils::getOffer(tasks[1].resources);      // b's reservation, but the executor is not there
  wrongAgent.id = "wrong";
  mesos::Offer home = wrongAgent;
  home.id = "home";
  home.executor_ids = {"executor-1"};
  auto recs = f.evaluator.evaluate(relaunchB, {wrongAgent, home});
  assertEquals(std::string("L-"), types(recs));
  assertEquals(std::string("home"), recs[0].offerId);
  assertEquals(std::string("executor-1"), recs[0].task->executor.executor_id);
}

// ---- S/scheduler/DefaultScheduler.java:483-538, T/scheduler/DefaultSchedulerTest.java:697-752
namespace {
struct Installed : Fixture {
  mesos::TaskInfo task;
  Installed() {
    auto req = PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500);
    recordLaunchWithCompleteOfferedResources(req, {ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedDisk(2000)});
    task = stateStore.fetchTasks().front();
  }
  std::vector<OfferResources> scan(const std::vector<mesos::Offer>& offers) { return UnexpectedResourceScanner(*kat::library(), stateStore).getUnexpectedResources(offers); }
};
}  // namespace
KAT(DefaultSchedulerTest_testUnexpectedPermanentlyFailedResources) {
  Installed f;
  assertFalse(f.task.resources.empty());
  std::vector<mesos::Offer> offers = {OfferTestUtils::getOffer(f.task.resources)};
  assertTrue(f.scan(offers).empty());                       // currently expected
  f.task.labels["permanently-failed"] = "true";
  f.stateStore.storeTasks({f.task});
  auto found = f.scan(offers);
  assertEquals(1u, found.size());
  assertEquals(f.task.resources.size(), found[0].resources.size());
  for (size_t k = 0; k < f.task.resources.size(); ++k) assertEquals(resourceId(f.task.resources[k]), resourceId(found[0].resources[k]));
}
KAT(DefaultSchedulerTest_testUnexpectedDecommissioningResources) {
  Installed f;
  std::vector<mesos::Offer> offers = {OfferTestUtils::getOffer(f.task.resources)};
  assertTrue(f.scan(offers).empty());
  f.stateStore.storeGoalOverrideStatus(f.task.name, "DECOMMISSIONING");
  auto found = f.scan(offers);
  assertEquals(1u, found.size());
  assertEquals(f.task.resources.size(), found[0].resources.size());
}
KAT(DefaultScheduler_unexpectedResourcesAcrossManyOffers) {
  Installed f;
  mesos::Resource stray = ResourceTestUtils::getUnreservedCpus(0.5);
  mesos::ReservationInfo dyn;
  dyn.role = TestConstants::ROLE; dyn.labels[labels::RESOURCE_ID] = "left-behind-by-an-old-task";
  stray.reservations.push_back(dyn);
  std::vector<mesos::Offer> offers;
  offers.push_back(OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(4.0)));  // nothing reserved: not listed
  offers.push_back(OfferTestUtils::getOffer({f.task.resources[0], stray, ResourceTestUtils::getUnreservedMem(64)}));
  offers.push_back(OfferTestUtils::getOffer(f.task.executor.resources));                            // executor resources are kept too
  offers.push_back(OfferTestUtils::getOffer({stray}));
  for (size_t k = 0; k < offers.size(); ++k) offers[k].id = "offer-" + std::to_string(k);
  auto found = f.scan(offers);
  assertEquals(2u, found.size());
  assertEquals(std::string("offer-1"), found[0].offer.id);
  assertEquals(1u, found[0].resources.size());
  assertEquals(std::string("left-behind-by-an-old-task"), resourceId(found[0].resources[0]));
  assertEquals(std::string("offer-3"), found[1].offer.id);
}
