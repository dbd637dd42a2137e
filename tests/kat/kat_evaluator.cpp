// kat_evaluator.cpp — the reference's evaluator-level vectors (T/offer/evaluate/OfferEvaluator*Test.java,
// T/offer/MesosResourcePoolTest.java, T/offer/RangeUtilsTest.java, T/offer/evaluate/PortEvaluationStageTest.java),
// PlanScheduler's greedy commit and the helloworld ServiceTest trace, restated through sdk::OfferEvaluator.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct Fixture {  // T/offer/evaluate/OfferEvaluatorTestBase.java:32-50
  StateStore stateStore;
  OfferEvaluator evaluator;
  explicit Fixture(SchedulerConfig cfg = SchedulerConfig()) : evaluator(*kat::library(), stateStore, TestConstants::SERVICE_NAME, cfg) {}
};

std::string types(const std::vector<Rec>& recs) {  // "RRRRL-" like the reference's operation-type lists (null = store)
  std::string s;
  for (auto& r : recs)
    s += r.type == Rec::RESERVE ? 'R' : r.type == Rec::CREATE ? 'C' : r.type == Rec::LAUNCH_GROUP ? 'L' : r.type == Rec::STORE_TASK ? '-' : '?';
  return s;
}
std::string resourceId(const mesos::Resource& r) { return ResourceUtils::getResourceId(r).value_or(""); }

}  // namespace

// ---- T/offer/evaluate/OfferEvaluatorTest.java
KAT(OfferEvaluatorTest_testReserveLaunchScalar) {  // :30-62
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))});
  assertEquals(6u, recs.size());
  assertEquals(std::string("RRRRL-"), types(recs));
  const mesos::Resource& reserve = recs[3].resource;
  assertEquals(1.0, reserve.value.scalar);
  assertEquals(std::string("cpus"), reserve.name);
  assertEquals(TestConstants::ROLE, ResourceUtils::getRole(reserve));
  assertEquals(TestConstants::PRINCIPAL, reserve.reservations.back().principal);
  assertEquals(36u, resourceId(reserve).size());
  assertFalse(reserve.disk.has_value());
  const mesos::TaskInfo& launched = *recs[4].task;
  assertEquals(resourceId(reserve), resourceId(launched.resources[0]));
  // executor overhead: cpus 0.1, disk 256, mem 32 in TreeMap order (:64-99)
  assertEquals(0.1, recs[0].resource.value.scalar);
  assertEquals(std::string("disk"), recs[1].resource.name);
  assertEquals(256.0, recs[1].resource.value.scalar);
  assertEquals(32.0, recs[2].resource.value.scalar);
  assertEquals(3u, launched.executor.resources.size());
}
KAT(OfferEvaluatorTest_testLaunchNotOnFirstOffer) {  // :441-462
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  mesos::Offer insufficient = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedMem(2.0));
  insufficient.id = "first";
  mesos::Offer sufficient = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0));
  sufficient.id = "second";
  auto recs = f.evaluator.evaluate(req, {insufficient, sufficient});
  assertEquals(std::string("RRRRL-"), types(recs));
  for (auto& r : recs) assertEquals(std::string("second"), r.offerId);
}
KAT(OfferEvaluatorTest_insufficientEverywhereGivesNothing) {
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(0.5)), OfferTestUtils::getOffer({})});
  assertEquals(0u, recs.size());
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {}).size());
}
// OfferEvaluatorTest.java:307-340 (testLaunchAttributesEmbedded) — what the next cycle's placement rules will read back
KAT(OfferEvaluatorTest_testLaunchAttributesEmbedded) {
  Fixture f;
  mesos::Offer offer = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0));
  offer.attributes.push_back(mesos::Attribute{"rack", mesos::Value::Text("foo")});
  offer.attributes.push_back(mesos::Attribute{"diskspeed", mesos::Value::Scalar(1234.5678)});
  offer.domain = TestConstants::LOCAL_DOMAIN_INFO();
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {offer});
  assertEquals(6u, recs.size());
  const mesos::TaskInfo& t = *recs[4].task;
  assertEquals(std::string("rack:foo;diskspeed:1234.568"), t.labels.at(labels::OFFER_ATTRIBUTES));
  assertEquals(offer.hostname, t.labels.at(labels::OFFER_HOSTNAME));
  assertEquals(TestConstants::ZONE, t.labels.at(labels::OFFER_ZONE));
  assertEquals(TestConstants::LOCAL_REGION, t.labels.at(labels::OFFER_REGION));
  assertEquals(TestConstants::POD_TYPE, t.labels.at(labels::TASK_TYPE));
  assertEquals(std::string("0"), t.labels.at(labels::TASK_INDEX));
}
// pre-reserved role: resources come out of that role's pool only (OfferEvaluatorTest.java testReserveLaunchScalar w/ role,
// MesosResourcePool.java:330-372)
KAT(OfferEvaluatorTest_preReservedRolePool) {
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0, TestConstants::PRE_RESERVED_ROLE);
  auto unreservedOnly = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0));
  assertEquals(0u, f.evaluator.evaluate(req, {unreservedOnly}).size());
  auto preReserved = OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(2.0, TestConstants::PRE_RESERVED_ROLE)}, TestConstants::PRE_RESERVED_ROLE);
  auto recs = f.evaluator.evaluate(req, {preReserved});
  assertEquals(std::string("RRRRL-"), types(recs));
  assertEquals(2u, recs[3].resource.reservations.size());  // STATIC base-role + DYNAMIC refinement
  assertEquals(TestConstants::PRE_RESERVED_ROLE, recs[3].resource.reservations[0].role);
}

// ---- T/offer/evaluate/OfferEvaluatorPortsTest.java
KAT(OfferEvaluatorPortsTest_testReserveStaticPort) {  // :35-60
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({555}),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(555, 555))});
  assertEquals(std::string("RRRRL-"), types(recs));
  assertFalse(resourceId(recs[4].task->resources[0]).empty());
  assertEquals((int64_t)555, recs[3].resource.value.ranges[0].first);
}
KAT(OfferEvaluatorPortsTest_testReserveStaticPortFailure) {  // :62-70
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({555}),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(666, 666))});
  assertEquals(0u, recs.size());
}
KAT(OfferEvaluatorPortsTest_testReserveTaskMultipleDynamicPorts) {  // :261-306
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({0, 0}),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(10000, 10001))});
  assertEquals(std::string("RRRRRL-"), types(recs));
  assertEquals((int64_t)10000, recs[3].resource.value.ranges[0].first);
  assertEquals((int64_t)10000, recs[3].resource.value.ranges[0].second);
  assertEquals((int64_t)10001, recs[4].resource.value.ranges[0].first);
  const mesos::TaskInfo& t = *recs[5].task;
  assertEquals(resourceId(recs[3].resource), resourceId(t.resources[0]));
  assertEquals(resourceId(recs[4].resource), resourceId(t.resources[1]));
}
KAT(OfferEvaluatorPortsTest_threeDynamicPortsFromTwoFail) {
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({0, 0, 0}),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(10000, 10001))});
  assertEquals(0u, recs.size());
}
// PortEvaluationStage.java:252-261: a dynamic port never takes a port that the pod requests statically elsewhere
KAT(PortEvaluationStage_dynamicSkipsStaticallyClaimedPorts) {
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({10000, 0}),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedPorts(10000, 10002))});
  assertEquals(std::string("RRRRRL-"), types(recs));
  assertEquals((int64_t)10000, recs[3].resource.value.ranges[0].first);  // static ports are evaluated first (:380-403)
  assertEquals((int64_t)10001, recs[4].resource.value.ranges[0].first);
}
// T/offer/evaluate/PortEvaluationStageTest.java:615-700 (dynamic port constrained to ranges, pre-reserved role)
namespace {
std::vector<Rec> dynamicInRange(int begin, int end) {
  Fixture f;
  mesos::Resource offered = ResourceTestUtils::getUnreservedPorts(23, 5050, "slave_public");
  mesos::Offer offer = OfferTestUtils::getCompleteOffer({offered}, "slave_public");
  auto b = PodInstanceRequirementTestUtils::builder("slave_public");
  b.port("TEST", 0, {RangeSpec(begin, end)});
  return f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(b.build(), 0), {offer});
}
}  // namespace
KAT(PortEvaluationStageTest_testDynamicPortWithRangesPasses) {
  auto recs = dynamicInRange(25, 600);
  assertEquals(std::string("RRRRL-"), types(recs));
  assertEquals((int64_t)25, recs[3].resource.value.ranges[0].first);
}
KAT(PortEvaluationStageTest_testDynamicPortWithRangesFails) { assertEquals(0u, dynamicInRange(6000, 8000).size()); }
KAT(PortEvaluationStageTest_testDynamicPortRangeBoundsAreInclusive) {
  assertEquals((int64_t)5050, dynamicInRange(5050, 6000)[3].resource.value.ranges[0].first);
  assertEquals((int64_t)23, dynamicInRange(1, 23)[3].resource.value.ranges[0].first);
  assertEquals(0u, dynamicInRange(5051, 6000).size());
  assertEquals(0u, dynamicInRange(1, 22).size());
}
KAT(PortEvaluationStage_overlayPortsNeedNoHostPort) {  // PortEvaluationStage.java:145-154,325-334
  Fixture f;
  auto b = PodInstanceRequirementTestUtils::builder();
  b.cpus(1.0).port("overlay", 8080, {}, {"dcos"});
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(b.build(), 0),
                                   {OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(1.0))});
  assertEquals(std::string("RRRRL-"), types(recs));  // no ports offered at all, and no RESERVE for the overlay port
}

// ---- T/offer/evaluate/OfferEvaluatorVolumesTest.java
KAT(OfferEvaluatorVolumesTest_testReserveCreateLaunchRootVolume) {  // :35-93
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500),
                                   {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedDisk(2000), ResourceTestUtils::getUnreservedCpus(1.0)})});
  assertEquals(8u, recs.size());
  assertEquals(std::string("RRRRRCL-"), types(recs));
  assertEquals(1.0, recs[3].resource.value.scalar);
  assertEquals(1500.0, recs[4].resource.value.scalar);
  assertEquals(resourceId(recs[4].resource), resourceId(recs[5].resource));
  assertEquals(36u, recs[5].resource.disk->persistence_id->size());
  const mesos::Resource& launched = recs[6].task->resources[1];
  assertEquals(resourceId(recs[4].resource), resourceId(launched));
  assertEquals(*recs[5].resource.disk->persistence_id, *launched.disk->persistence_id);
}
// :94-102 re-runs the test above with the refinement capability on (note in kat_reference_rest.cpp)
KAT(OfferEvaluatorVolumesTest_testReserveCreateLaunchRootVolumeRefined) { kat_OfferEvaluatorVolumesTest_testReserveCreateLaunchRootVolume(); }
KAT(OfferEvaluatorVolumesTest_testReserveCreateLaunchMountVolume) {  // :184-242
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, 1500),
                                   {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(2000)})});
  assertEquals(std::string("RRRRRCL-"), types(recs));
  assertEquals(2000.0, recs[4].resource.value.scalar);  // the whole MOUNT disk is reserved
  assertTrue(recs[4].resource.disk->source == mesos::DiskInfo::MOUNT);
  assertEquals(2000.0, recs[6].task->resources[1].value.scalar);
}
KAT(OfferEvaluatorVolumesTest_testCreateMultipleProfileMountVolumes) {  // :377-445
  Fixture f;
  auto rs = PodInstanceRequirementTestUtils::builder()
                .cpus(1.0)
                .addMountVolume(1.0, TestConstants::CONTAINER_PATH + "-a", {"x"})
                .addMountVolume(1.0, TestConstants::CONTAINER_PATH + "-b", {"x", "y"})
                .addMountVolume(1.0, TestConstants::CONTAINER_PATH + "-c", {})
                .build();
  mesos::Offer offer = OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(1.0),
                                                         ResourceTestUtils::getUnreservedMountVolume(1.0, "x"), ResourceTestUtils::getUnreservedMountVolume(1.0, "y")});
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(rs, 0), {offer});
  assertEquals(std::string("RRRRRCRCRCL-"), types(recs));
  assertEquals(TestConstants::CONTAINER_PATH + "-a", recs[5].resource.disk->container_path);
  assertEquals(std::string("x"), *recs[5].resource.disk->profile);
  assertEquals(TestConstants::CONTAINER_PATH + "-b", recs[7].resource.disk->container_path);
  assertEquals(std::string("y"), *recs[7].resource.disk->profile);
  assertEquals(TestConstants::CONTAINER_PATH + "-c", recs[9].resource.disk->container_path);
  assertFalse(recs[9].resource.disk->profile.has_value());
}
KAT(OfferEvaluatorVolumesTest_testConsumeMultipleMountVolumesFailure) {  // :447-462
  Fixture f;
  auto rs = PodInstanceRequirementTestUtils::builder().cpus(1.0).addMountVolume(1000.0, "A", {}).addMountVolume(1000.0, "B", {}).build();
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRequirement(rs, 0),
                                   {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(2000)})});
  assertEquals(0u, recs.size());
}
KAT(OfferEvaluatorVolumesTest_testFailCreateRootVolume) {  // :464-472
  Fixture f;
  // complete offer: disk 512 + 1000 = 1512; executor takes 256; 1256 < 1500
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500),
                                   {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedDisk(1000)})});
  assertEquals(0u, recs.size());
}
KAT(OfferEvaluatorVolumesTest_testFailToCreateVolumeWithWrongResource) {  // :474-484: a MOUNT disk is not root disk
  Fixture f;
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getRootVolumeRequirement(1.0, 1500),
                                   {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(2000)})});
  assertEquals(0u, recs.size());
}
KAT(OfferEvaluatorVolumesTest_testReserveCreateExecutorVolume) {  // :486-520: a pod-level volume is evaluated before the task stages
  Fixture f;
  auto req = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  auto pod = std::make_shared<PodSpec>(*req.podInstance.pod);
  VolumeSpec v;
  v.type = VolumeSpec::MOUNT; v.size = 1000; v.containerPath = TestConstants::CONTAINER_PATH; v.role = TestConstants::ROLE; v.principal = TestConstants::PRINCIPAL;
  pod->volumes.push_back(v);
  req.podInstance.pod = pod;
  auto recs = f.evaluator.evaluate(req, {OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedMountVolume(2000), ResourceTestUtils::getUnreservedCpus(1.0)})});
  assertEquals(std::string("RCRRRRL-"), types(recs));
  assertEquals(2000.0, recs[0].resource.value.scalar);
  assertEquals(1u + 3u, recs[6].task->executor.resources.size());  // the volume belongs to the executor
}

// ---- T/offer/MesosResourcePoolTest.java:63-195, through a minimal pipeline (cpus + the volume under test)
namespace {
bool mountFits(double wanted, std::vector<std::string> profiles, double offered, std::optional<std::string> profile) {
  SchedulerConfig cfg;
  cfg.executorResources.clear();
  Fixture f(cfg);
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getMountVolumeRequirement(1.0, wanted, std::move(profiles)),
                                   {OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(1.0), ResourceTestUtils::getUnreservedMountVolume(offered, profile)})});
  return !recs.empty();
}
}  // namespace
KAT(MesosResourcePoolTest_testConsumeUnreservedAtomicResource) { assertTrue(mountFits(1000, {}, 1000, std::nullopt)); }
KAT(MesosResourcePoolTest_testConsumeUnreservedAtomicResourceMatchedProfile) { assertTrue(mountFits(1000, {"foo", "bar"}, 1000, "bar")); }
KAT(MesosResourcePoolTest_testConsumeUnreservedAtomicResourceUnmatchedProfile1) { assertFalse(mountFits(1000, {"foo"}, 1000, "bar")); }
KAT(MesosResourcePoolTest_testConsumeUnreservedAtomicResourceUnmatchedProfile2) { assertFalse(mountFits(1000, {"foo", "bar"}, 1000, std::nullopt)); }
KAT(MesosResourcePoolTest_testConsumeUnreservedAtomicResourceUnmatchedProfile3) { assertFalse(mountFits(1000, {}, 1000, "bar")); }
KAT(MesosResourcePoolTest_atomicTooSmall) { assertFalse(mountFits(1000.5, {}, 1000, std::nullopt)); }
KAT(MesosResourcePoolTest_testConsumeInsufficientUnreservedMergedResource) {
  SchedulerConfig cfg;
  cfg.executorResources.clear();
  Fixture f(cfg);
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(2.0), {OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(1.0)})}).size());
  assertEquals(3u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(1.0)})}).size());
}
KAT(MesosResourcePoolTest_testConsumeUnreservedWhenNoUnreservedResources) {  // a reserved resource is not in the merged pool
  SchedulerConfig cfg;
  cfg.executorResources.clear();
  Fixture f(cfg);
  mesos::Resource reserved = ResourceTestUtils::getUnreservedCpus(1.0);
  mesos::ReservationInfo dyn;
  dyn.role = TestConstants::ROLE; dyn.principal = TestConstants::PRINCIPAL; dyn.labels[labels::RESOURCE_ID] = "some-other-pods-reservation";
  reserved.reservations.push_back(dyn);
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {OfferTestUtils::getOffer({reserved})}).size());
}
// MesosResourcePool.java:133-147 + ValueUtils.java:64-67: the pool is the fp64 sum of the offer's entries in order, and the
// stages subtract in order; 0.1 + 0.2 is not 0.3
KAT(MesosResourcePool_sumsAndDeductsInOrder) {
  SchedulerConfig cfg;
  cfg.executorResources = {{"cpus", 0.1}};
  Fixture f(cfg);
  auto offerOf = [](std::vector<double> cpus) {
    std::vector<mesos::Resource> rs;
    for (double c : cpus) rs.push_back(ResourceTestUtils::getUnreservedCpus(c));
    return OfferTestUtils::getOffer(rs);
  };
  // pool 0.2 + 0.1 = 0.30000000000000004; executor 0.1 leaves 0.20000000000000004 >= 0.2
  assertEquals(4u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(0.2), {offerOf({0.2, 0.1})}).size());  // R(executor) R(cpus) L -
  // pool 0.3 exactly; executor 0.1 leaves 0.19999999999999998 < 0.2
  assertEquals(0u, f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(0.2), {offerOf({0.3})}).size());
}
// consumableResource (MesosResourcePool.java:87-97): resources allocated to another role are invisible to the pod
KAT(MesosResourcePool_allocationRoleFilter) {
  Fixture f;
  mesos::Offer mine = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0));
  mesos::Offer theirs = mine;
  for (auto& r : mine.resources) r.allocation_role = TestConstants::ROLE;
  for (auto& r : theirs.resources) r.allocation_role = "some-other-role";
  theirs.id = "theirs"; mine.id = "mine";
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getCpuRequirement(1.0), {theirs, mine});
  assertEquals(6u, recs.size());
  assertEquals(std::string("mine"), recs[0].offerId);
}

// ---- T/offer/RangeUtilsTest.java:19-96 (antlr IntervalSet behaviour the encoder restates)
KAT(RangeUtilsTest_merging) {
  IntervalSet a; a.add(1, 3); a.add(2, 4);
  assertEquals(1u, a.iv.size()); assertEquals((int64_t)1, a.iv[0].first); assertEquals((int64_t)4, a.iv[0].second);
  IntervalSet b; b.add(1, 3); b.add(5, 7);
  assertEquals(2u, b.iv.size()); assertEquals((int64_t)5, b.iv[1].first);
  IntervalSet c; c.add(1, 5); c.add(2, 4);
  assertEquals(1u, c.iv.size()); assertEquals((int64_t)5, c.iv[0].second);
  IntervalSet d; d.add(5, 7); d.add(1, 3); d.add(4, 4);  // adjacency coalesces (IntervalSet.add)
  assertEquals(1u, d.iv.size()); assertEquals((int64_t)1, d.iv[0].first); assertEquals((int64_t)7, d.iv[0].second);
}
KAT(RangeUtilsTest_portsSplitOverSeveralResourcesMerge) {
  Fixture f;
  mesos::Offer offer = OfferTestUtils::getCompleteOffer({ResourceTestUtils::getUnreservedPorts(5, 7), ResourceTestUtils::getUnreservedPorts(1, 3),
                                                         ResourceTestUtils::getUnreservedPorts(4, 4)});
  auto recs = f.evaluator.evaluate(PodInstanceRequirementTestUtils::getPortRequirement({0, 4, 0}), {offer});
  assertEquals(std::string("RRRRRRL-"), types(recs));
  assertEquals((int64_t)4, recs[3].resource.value.ranges[0].first);  // static first
  assertEquals((int64_t)1, recs[4].resource.value.ranges[0].first);
  assertEquals((int64_t)2, recs[5].resource.value.ranges[0].first);
}

// ---- S/scheduler/plan/PlanScheduler.java:50-124 (T/scheduler/plan/PlanSchedulerTest.java covers 0 / 1 step only; the
// multi-step expectations follow the code: steps in order, first fit, used offers removed whole)
KAT(PlanSchedulerTest_emptyInputs) {
  Fixture f;
  PlanScheduler ps(f.evaluator);
  std::vector<Step> steps(1);
  steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  assertEquals(0u, ps.resourceOffers({}, steps).size());
  std::vector<Step> none;
  assertEquals(0u, ps.resourceOffers({OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(2.0))}, none).size());
}
KAT(PlanScheduler_greedyCommitRemovesUsedOffers) {
  Fixture f;
  PlanScheduler ps(f.evaluator);
  auto offer = [](const std::string& id, double cpus) {
    mesos::Offer o = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(cpus));
    o.id = id; o.slave_id = "agent-" + id; o.hostname = "host-" + id;
    return o;
  };
  std::vector<mesos::Offer> offers = {offer("small", 1.0), offer("big", 4.0), offer("medium", 2.0)};
  std::vector<Step> steps(5);
  steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(2.0, 0);  // skips small -> big
  steps[1].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(2.0, 1);  // big is gone although it had room -> medium
  steps[2].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(2.0, 2);  // only small left -> nothing
  steps[3].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(0.5, 3);  // small
  steps[3].status = Step::COMPLETE;                                                   // ... but it is not pending (:73-78)
  steps[4].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(1.0, 4);  // small
  auto recs = ps.resourceOffers(offers, steps);
  assertEquals(18u, recs.size());
  assertEquals(std::string("big"), recs[0].offerId);
  assertEquals(std::string("medium"), recs[6].offerId);
  assertEquals(std::string("small"), recs[12].offerId);
  assertTrue(steps[0].status == Step::STARTING);
  assertTrue(steps[1].status == Step::STARTING);
  assertTrue(steps[2].status == Step::PREPARED);   // evaluated, nothing found: updateOfferStatus(empty) (:106-112)
  assertTrue(steps[3].status == Step::COMPLETE);
  assertTrue(steps[4].status == Step::STARTING);
}

// ---- frameworks/helloworld/src/test/java/com/mesosphere/sdk/helloworld/scheduler/ServiceTest.java:898-950 — the default
// deployment trace.  hello: 0.1 cpu / 252 mem / 25 disk ROOT; world: 0.2 cpu / 512 mem / 2 x 25 disk
// ROOT; both [["hostname","UNIQUE"]]; executor overhead = cpus 0.1 only (sdk/testing ServiceTestRunner.java:43-48,307-308);
// offers are SendOffer's exact-fit rows, resources NOT merged (sdk/testing SendOffer.java:127-171).
namespace {
struct HelloWorld {
  std::shared_ptr<PodSpec> hello = std::make_shared<PodSpec>(), world = std::make_shared<PodSpec>();
  SchedulerConfig cfg;
  HelloWorld() {
    cfg.executorResources = {{"cpus", 0.1}};
    hello->type = "hello";
    hello->placementRule = MarathonConstraintParser::parse("hello", "[[\"hostname\", \"UNIQUE\"]]");
    TaskSpec hs;
    hs.name = "server";
    hs.resourceSet = ResourceSetBuilder("hello-world-role", ANY_ROLE, "hello-world-principal").id("server-resource-set").cpus(0.1).memory(252).addRootVolume(25, "hello-container-path").build();
    hello->tasks = {hs};
    world->type = "world";
    world->placementRule = MarathonConstraintParser::parse("world", "[[\"hostname\", \"UNIQUE\"]]");
    TaskSpec ws;
    ws.name = "server";
    ws.resourceSet = ResourceSetBuilder("hello-world-role", ANY_ROLE, "hello-world-principal").id("server-resource-set").cpus(0.2).memory(512)
                         .addRootVolume(25, "world-container-path1").addRootVolume(25, "world-container-path2").build();
    world->tasks = {ws};
  }
  static int n() { static int k = 0; return ++k; }
  mesos::Offer sendOffer(const std::shared_ptr<PodSpec>& pod, const std::string& hostname = "test-hostname") {
    mesos::Offer o;
    o.id = "offer-" + std::to_string(n()); o.framework_id = "test-framework-id"; o.slave_id = "test-agent-" + std::to_string(n()); o.hostname = hostname;
    for (auto& t : pod->tasks) {
      for (auto& r : t.resourceSet.resources) o.resources.push_back(ResourceTestUtils::unreservedScalar(r.name, r.value));
      for (auto& v : t.resourceSet.volumes) o.resources.push_back(ResourceTestUtils::unreservedScalar("disk", v.size));
    }
    o.resources.push_back(ResourceTestUtils::getUnreservedCpus(0.1));
    return o;
  }
};
}  // namespace
KAT(HelloWorldServiceTest_defaultDeploymentTrace) {
  HelloWorld hw;
  Fixture f(hw.cfg);
  PlanScheduler ps(f.evaluator);
  auto cycle = [&](const mesos::Offer& offer, const std::shared_ptr<PodSpec>& pod, int index) {  // one candidate step per cycle (serial plan)
    std::vector<Step> steps(1);
    steps[0].requirement = PodInstanceRequirement::newBuilder(PodInstance{pod, index}, {"server"});
    auto recs = ps.resourceOffers({offer}, steps);
    std::vector<mesos::TaskInfo> store;
    std::vector<std::string> launched;
    for (auto& r : recs) {
      if (r.type == Rec::STORE_TASK) store.push_back(*r.task);   // PersistentLaunchRecorder.record (after the cycle)
      if (r.type == Rec::LAUNCH_GROUP) launched.push_back(r.task->name);
    }
    f.stateStore.storeTasks(store);
    return launched;
  };
  using V = std::vector<std::string>;
  auto eq = [](const V& a, const V& b) { return a == b; };
  assertTrue(eq(V{"hello-0-server"}, cycle(hw.sendOffer(hw.hello), hw.hello, 0)));              // cycle 1
  assertTrue(eq(V{"world-0-server"}, cycle(hw.sendOffer(hw.world), hw.world, 0)));              // cycle 3: hello-0 on the host is not counted ("world-.*")
  assertTrue(eq(V{}, cycle(hw.sendOffer(hw.world), hw.world, 1)));                              // cycle 5: 1 + 1 > max 1 on test-hostname
  assertTrue(eq(V{"world-1-server"}, cycle(hw.sendOffer(hw.world, "host-foo"), hw.world, 1)));  // cycle 6
  assertEquals(3u, f.stateStore.fetchTasks().size());
  // BASELINE.json's variant (hello x 2): hello-1 is refused on test-hostname and accepted elsewhere
  assertTrue(eq(V{}, cycle(hw.sendOffer(hw.hello), hw.hello, 1)));
  assertTrue(eq(V{"hello-1-server"}, cycle(hw.sendOffer(hw.hello, "host-bar"), hw.hello, 1)));
  // a world offer does not fit a hello pod the other way round: cpus 0.1 + 0.1 < 0.2 + 0.1
  assertTrue(eq(V{}, cycle(hw.sendOffer(hw.hello, "host-baz"), hw.world, 2)));
}
// the same service with every step offered in ONE cycle: tallies are the cycle-start snapshot (DefaultScheduler.java:440,455),
// so both world pods may land on the same hostname — the engine must reproduce that, not "fix" it
KAT(HelloWorld_oneCycleUsesTheCycleStartSnapshot) {
  HelloWorld hw;
  Fixture f(hw.cfg);
  PlanScheduler ps(f.evaluator);
  std::vector<Step> steps(2);
  steps[0].requirement = PodInstanceRequirement::newBuilder(PodInstance{hw.world, 0}, {"server"});
  steps[1].requirement = PodInstanceRequirement::newBuilder(PodInstance{hw.world, 1}, {"server"});
  mesos::Offer a = hw.sendOffer(hw.world), b = hw.sendOffer(hw.world);  // two agents that report the same hostname
  auto recs = ps.resourceOffers({a, b}, steps);
  assertEquals(2, countType(recs, Rec::LAUNCH_GROUP));
}
