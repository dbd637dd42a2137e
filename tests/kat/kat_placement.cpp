// kat_placement.cpp — the reference's placement-rule unit vectors, restated.  Each KAT names the JUnit method it follows
// (T/ = sdk/scheduler/src/test/java/com/mesosphere/sdk/, PT/ = T/offer/evaluate/placement/).  Where the reference builds
// tasks with a random pod index (T/testutils/TaskTestUtils.java:80-84) the restatement uses fixed distinct indexes.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;

namespace {

bool filter(const RulePtr& rule, const mesos::Offer& offer, const PodInstance& pod, const std::vector<mesos::TaskInfo>& tasks) {
  return PlacementRuleFilter{*kat::library()}.filter(rule, offer, pod, tasks);
}

// PodInstance of type "type", index 0, as PT/AttributeRuleTest.java:26-28 and the round-robin tests build it
PodInstance typePod(const std::string& type = "type", int index = 0) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = type;
  TaskSpec t;
  t.name = "test-task-name";
  t.resourceSet = PodInstanceRequirementTestUtils::getCpuResourceSet(1.0);
  pod->tasks.push_back(t);
  return PodInstance{pod, index};
}

mesos::Attribute textAttr(const std::string& name, const std::string& value) { return mesos::Attribute{name, mesos::Value::Text(value)}; }

// ---- PT/MaxPerTest.java:52-236.  TestMaxPerRule hands arbitrary key lists to MaxPerRule; the same keys are expressed
// here as attribute strings "key0:x" under a MaxPerAttributeRule whose matcher accepts everything.
int nextIndex = 5000;
mesos::TaskInfo keyedTask(const std::vector<std::string>& keys, const std::string& type = "different-type", int index = 100) {
  mesos::TaskInfo t = getTaskInfo();
  TaskLabelWriter w(t);
  w.setType(type).setIndex(index);
  std::string joined;
  for (size_t k = 0; k < keys.size(); ++k) joined += (k ? ";" : "") + keys[k] + ":x";
  if (!keys.empty()) t.labels[labels::OFFER_ATTRIBUTES] = joined;
  return t;
}
mesos::Offer keyedOffer(const std::vector<std::string>& keys) {
  mesos::Offer o = OfferTestUtils::getEmptyOfferBuilder();
  for (auto& k : keys) o.attributes.push_back(textAttr(k, "x"));
  return o;
}
bool maxPer(int max, const std::vector<std::string>& taskKeys, const std::vector<std::string>& offerKeys, bool withTask = true, bool self = false) {
  RulePtr rule = std::make_shared<MaxPerAttributeRule>(max, RegexMatcher::create(".*"), AnyMatcher::create());
  PodInstance pod = PodTestUtils::getPodInstance(0);
  std::vector<mesos::TaskInfo> tasks;
  if (withTask) tasks.push_back(self ? keyedTask(taskKeys, pod.pod->type, pod.index) : keyedTask(taskKeys));
  return filter(rule, keyedOffer(offerKeys), pod, tasks);
}

}  // namespace

KAT(MaxPerTest_limitZero) { assertThrows(MaxPerHostnameRule(0), std::invalid_argument); }
KAT(MaxPerTest_acceptEmptyAll) { assertTrue(maxPer(1, {}, {}, false)); }
KAT(MaxPerTest_acceptEmptyKeysNonEmptyTasks) { assertTrue(maxPer(1, {}, {})); }
KAT(MaxPerTest_acceptOfferWithoutKeys) { assertTrue(maxPer(1, {"key0"}, {})); }
KAT(MaxPerTest_acceptOfferWithDifferentKey) { assertTrue(maxPer(1, {"key0"}, {"key1"})); }
KAT(MaxPerTest_rejectSingleTask) { assertFalse(maxPer(1, {"key0"}, {"key0"})); }
KAT(MaxPerTest_ignoreSelf) { assertTrue(maxPer(1, {"key0"}, {"key0"}, true, true)); }
KAT(MaxPerTest_ignoreSelfLargerLimit) { assertTrue(maxPer(2, {"key0", "key0"}, {"key0"}, true, true)); }
KAT(MaxPerTest_rejectOneOfMultipleOverLimit) { assertFalse(maxPer(1, {"key0", "key1"}, {"key1"})); }
KAT(MaxPerTest_rejectOneOfMultipleOverLargerLimit) { assertFalse(maxPer(2, {"key0", "key1", "key1"}, {"key1"})); }
KAT(MaxPerTest_rejectMultipleOverLimit) { assertFalse(maxPer(1, {"key0", "key1"}, {"key1", "key0"})); }
KAT(MaxPerTest_rejectMultipleOverLargerLimit) { assertFalse(maxPer(2, {"key0", "key0", "key1", "key1"}, {"key1", "key0"})); }
KAT(MaxPerTest_acceptOfferWithDisjointKeys) { assertTrue(maxPer(1, {"key0", "key1"}, {"key2", "key3"})); }
// MaxPerRule.java:66-67,81-86: the offer's own keys are tallied first, so a key listed twice in the offer starts at 2
KAT(MaxPerRule_duplicateOfferKeysStartAtTwo) {
  RulePtr one = std::make_shared<MaxPerAttributeRule>(1, RegexMatcher::create(".*"), AnyMatcher::create());
  RulePtr two = std::make_shared<MaxPerAttributeRule>(2, RegexMatcher::create(".*"), AnyMatcher::create());
  assertFalse(filter(one, keyedOffer({"key0", "key0"}), PodTestUtils::getPodInstance(0), {}));
  assertTrue(filter(two, keyedOffer({"key0", "key0"}), PodTestUtils::getPodInstance(0), {}));
}

// ---- PT/MaxPerHostnameRuleTest.java / MaxPerZoneRuleTest.java / MaxPerRegionRuleTest.java (key providers)
namespace {
mesos::TaskInfo placedTask(const std::string& name, const std::string& type, int index, const std::string& host, const std::string& agent,
                           std::optional<mesos::DomainInfo> domain = std::nullopt, const std::string& attrs = "") {
  mesos::TaskInfo t = getTaskInfo(name);
  t.slave_id = agent;
  TaskLabelWriter w(t);
  w.setType(type).setIndex(index);
  t.labels[labels::OFFER_HOSTNAME] = host;
  if (domain) w.setZone(domain->zone).setRegion(domain->region);
  if (!attrs.empty()) t.labels[labels::OFFER_ATTRIBUTES] = attrs;
  return t;
}
mesos::Offer hostOffer(const std::string& host, std::optional<mesos::DomainInfo> domain = std::nullopt) {
  mesos::Offer o = OfferTestUtils::getEmptyOfferBuilder();
  o.hostname = host; o.slave_id = "agent-" + host; o.domain = domain;
  return o;
}
}  // namespace

KAT(MaxPerHostnameRule_countsOnlyFilteredTasksOnThatHost) {
  MatcherPtr filterA = RegexMatcher::create("a-.*");
  RulePtr rule = std::make_shared<MaxPerHostnameRule>(2, filterA);
  std::vector<mesos::TaskInfo> tasks = {placedTask("a-0-t", "a", 0, "host1", "agent-host1"), placedTask("a-1-t", "a", 1, "host1", "agent-host1"),
                                        placedTask("b-0-t", "b", 0, "host2", "agent-host2"), placedTask("b-1-t", "b", 1, "host2", "agent-host2"),
                                        placedTask("a-2-t", "a", 2, "host2", "agent-host2")};
  PodInstance fresh = typePod("a", 7);
  assertFalse(filter(rule, hostOffer("host1"), fresh, tasks));  // 1 + 2 > 2
  assertTrue(filter(rule, hostOffer("host2"), fresh, tasks));   // 1 + 1 (b-* not counted)
  assertTrue(filter(rule, hostOffer("host3"), fresh, tasks));
  assertTrue(filter(rule, hostOffer("host1"), typePod("a", 1), tasks));  // a-1 redeploying: its own task is not counted
}
KAT(MaxPerZoneRule_failsWithoutZoneAndCountsPerZone) {
  MatcherPtr any = AnyMatcher::create();
  RulePtr rule = std::make_shared<MaxPerZoneRule>(1, any);
  mesos::DomainInfo z1{"r", "z1"}, z2{"r", "z2"};
  std::vector<mesos::TaskInfo> tasks = {placedTask("x-0-t", "x", 0, "h1", "a1", z1)};
  assertFalse(filter(rule, hostOffer("h9"), typePod("x", 5), tasks));  // MaxPerZoneRule.java:58-59: no zone in the offer
  assertFalse(filter(rule, hostOffer("h2", z1), typePod("x", 5), tasks));
  assertTrue(filter(rule, hostOffer("h2", z2), typePod("x", 5), tasks));
}
KAT(MaxPerRegionRule_offerWithoutDomainHasNoKeys) {
  RulePtr rule = std::make_shared<MaxPerRegionRule>(1, AnyMatcher::create());
  mesos::DomainInfo r1{"r1", "z"}, r2{"r2", "z"};
  std::vector<mesos::TaskInfo> tasks = {placedTask("x-0-t", "x", 0, "h1", "a1", r1)};
  assertTrue(filter(rule, hostOffer("h9"), typePod("x", 5), tasks));   // MaxPerRegionRule.java:33-45: no keys, nothing to exceed
  assertFalse(filter(rule, hostOffer("h2", r1), typePod("x", 5), tasks));
  assertTrue(filter(rule, hostOffer("h2", r2), typePod("x", 5), tasks));
}

// ---- PT/RoundRobinByAttributeRuleTest.java:94-216
namespace {
const char* ATTRIB_NAME = "rack_id";
mesos::TaskInfo rrTask(const std::string& taskName, const std::string& attrName, const std::string& attrVal) {
  mesos::TaskInfo t = getTaskInfo(taskName);
  TaskLabelWriter(t).setType(TestConstants::TASK_TYPE).setIndex(nextIndex++);
  t.labels[labels::OFFER_ATTRIBUTES] = attrName + ":" + attrVal;
  return t;
}
mesos::TaskInfo rrTask(const std::string& taskName, const std::string& attrVal) { return rrTask(taskName, ATTRIB_NAME, attrVal); }
mesos::Offer offerWithAttribute(const std::string& value, const std::string& name = ATTRIB_NAME) {
  mesos::Offer o = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(1.0));
  o.attributes.push_back(textAttr(name, value));
  return o;
}
PodInstance podOf(const mesos::TaskInfo& t) { return typePod(t.labels.at(labels::TASK_TYPE), std::stoi(t.labels.at(labels::TASK_INDEX))); }
std::vector<mesos::TaskInfo> ignoredAttrTasks() {
  return {rrTask("ignored1", "value1"), rrTask("ignored2", "value2"), rrTask("ignored3", "value3"), rrTask("ignored4", "value4"),
          rrTask("ignored5", "value5"), rrTask("ignored6", "foo", "value6"), rrTask("ignored7", "bar", "value7"), rrTask("ignored8", "baz", "value8")};
}
}  // namespace

KAT(RoundRobinByAttributeRuleTest_testRolloutWithAgentCount) {
  RulePtr rule = std::make_shared<RoundRobinByAttributeRule>(ATTRIB_NAME, 3, RegexMatcher::create("[0-9]"));
  const PodInstance POD_INSTANCE = typePod();
  auto tasks = ignoredAttrTasks();
  auto pass = [&](const char* v, const PodInstance& p) { return filter(rule, offerWithAttribute(v), p, tasks); };
  assertTrue(pass("value1", POD_INSTANCE));
  mesos::TaskInfo taskInfo1 = rrTask("1", "value1");
  PodInstance req1 = podOf(taskInfo1);
  tasks.push_back(taskInfo1);  // value1:1
  assertFalse(pass("value1", POD_INSTANCE));
  assertTrue(pass("value2", POD_INSTANCE));
  assertTrue(pass("value3", POD_INSTANCE));
  mesos::TaskInfo taskInfo2 = rrTask("2", "value3");
  PodInstance req2 = podOf(taskInfo2);
  tasks.push_back(taskInfo2);  // value1:1, value3:1
  assertTrue(pass("value1", req1));
  assertTrue(pass("value3", req2));
  assertFalse(pass("value1", POD_INSTANCE));
  assertTrue(pass("value2", POD_INSTANCE));
  assertFalse(pass("value3", POD_INSTANCE));
  tasks.push_back(rrTask("3", "value2"));  // 1,1,1
  assertTrue(pass("value1", POD_INSTANCE));
  assertTrue(pass("value2", POD_INSTANCE));
  assertTrue(pass("value3", POD_INSTANCE));
  tasks.push_back(rrTask("4", "value2"));  // value1:1, value2:2, value3:1
  assertTrue(pass("value1", POD_INSTANCE));
  assertFalse(pass("value2", POD_INSTANCE));
  assertTrue(pass("value3", POD_INSTANCE));
  tasks.push_back(rrTask("5", "value3"));  // 1,2,2
  assertTrue(pass("value4", POD_INSTANCE));
  tasks.push_back(rrTask("6", "value4"));  // 1,2,2,1
  assertTrue(pass("value4", POD_INSTANCE));
  tasks.push_back(rrTask("7", "value4"));  // 1,2,2,2
  assertFalse(pass("value2", POD_INSTANCE));
  assertFalse(pass("value3", POD_INSTANCE));
  assertFalse(pass("value4", POD_INSTANCE));
  assertTrue(pass("value1", POD_INSTANCE));
  tasks.push_back(rrTask("8", "value1"));  // 2,2,2,2
  for (const char* v : {"value1", "value2", "value3", "value4"}) assertTrue(pass(v, POD_INSTANCE));
}

KAT(RoundRobinByAttributeRuleTest_testRolloutWithoutAgentCount) {
  RulePtr rule = std::make_shared<RoundRobinByAttributeRule>(ATTRIB_NAME, std::nullopt, RegexMatcher::create("[0-9]"));
  const PodInstance POD_INSTANCE = typePod();
  auto tasks = ignoredAttrTasks();
  auto pass = [&](const char* v, const PodInstance& p) { return filter(rule, offerWithAttribute(v), p, tasks); };
  assertTrue(pass("value1", POD_INSTANCE));
  mesos::TaskInfo taskInfo1 = rrTask("1", "value1");
  PodInstance req1 = podOf(taskInfo1);
  tasks.push_back(taskInfo1);
  for (const char* v : {"value1", "value2", "value3"}) assertTrue(pass(v, POD_INSTANCE));
  mesos::TaskInfo taskInfo2 = rrTask("2", "value3");
  PodInstance req2 = podOf(taskInfo2);
  tasks.push_back(taskInfo2);
  assertTrue(pass("value1", req1));
  assertTrue(pass("value3", req2));
  for (const char* v : {"value1", "value2", "value3"}) assertTrue(pass(v, POD_INSTANCE));
  tasks.push_back(rrTask("3", "value2"));
  for (const char* v : {"value1", "value2", "value3"}) assertTrue(pass(v, POD_INSTANCE));
  tasks.push_back(rrTask("4", "value2"));  // 1,2,1
  assertTrue(pass("value1", POD_INSTANCE));
  assertFalse(pass("value2", POD_INSTANCE));
  assertTrue(pass("value3", POD_INSTANCE));
  tasks.push_back(rrTask("5", "value3"));  // 1,2,2
  assertTrue(pass("value4", POD_INSTANCE));
  tasks.push_back(rrTask("6", "value4"));
  assertTrue(pass("value4", POD_INSTANCE));
  tasks.push_back(rrTask("7", "value4"));  // 1,2,2,2
  assertFalse(pass("value2", POD_INSTANCE));
  assertFalse(pass("value3", POD_INSTANCE));
  assertFalse(pass("value4", POD_INSTANCE));
  assertTrue(pass("value1", POD_INSTANCE));
  tasks.push_back(rrTask("8", "value1"));
  for (const char* v : {"value1", "value2", "value3", "value4"}) assertTrue(pass(v, POD_INSTANCE));
}

// RoundRobinByAttributeRule.java:62-81: the attribute NAME is compared case-insensitively; an offer without it fails
KAT(RoundRobinByAttributeRule_nameIgnoresCaseAndMissingAttributeFails) {
  RulePtr rule = std::make_shared<RoundRobinByAttributeRule>("Rack_ID", std::nullopt, AnyMatcher::create());
  std::vector<mesos::TaskInfo> tasks = {rrTask("1", "value1"), rrTask("2", "value1"), rrTask("3", "value2")};
  assertFalse(filter(rule, offerWithAttribute("value1"), typePod(), tasks));   // 2 > min 1
  assertTrue(filter(rule, offerWithAttribute("value2"), typePod(), tasks));
  assertFalse(filter(rule, offerWithAttribute("value2", "other"), typePod(), tasks));
}

// ---- PT/RoundRobinByHostnameRuleTest.java (same rollout table keyed by hostname)
namespace {
mesos::TaskInfo hostTask(const std::string& taskName, const std::string& host) {
  mesos::TaskInfo t = getTaskInfo(taskName);
  TaskLabelWriter(t).setType(TestConstants::TASK_TYPE).setIndex(nextIndex++);
  t.labels[labels::OFFER_HOSTNAME] = host;
  return t;
}
}  // namespace
KAT(RoundRobinByHostnameRuleTest_testRolloutWithAgentCount) {
  RulePtr rule = std::make_shared<RoundRobinByHostnameRule>(3, RegexMatcher::create("[0-9]"));
  const PodInstance POD_INSTANCE = typePod();
  std::vector<mesos::TaskInfo> tasks = {hostTask("ignored1", "host1"), hostTask("ignored2", "host2"), hostTask("ignored3", "host3"),
                                        hostTask("ignored4", "host4"), hostTask("ignored5", "host5")};
  auto pass = [&](const char* h, const PodInstance& p) { return filter(rule, hostOffer(h), p, tasks); };
  assertTrue(pass("host1", POD_INSTANCE));
  mesos::TaskInfo t1 = hostTask("1", "host1");
  PodInstance req1 = podOf(t1);
  tasks.push_back(t1);
  assertFalse(pass("host1", POD_INSTANCE));
  assertTrue(pass("host2", POD_INSTANCE));
  assertTrue(pass("host3", POD_INSTANCE));
  mesos::TaskInfo t2 = hostTask("2", "host3");
  PodInstance req2 = podOf(t2);
  tasks.push_back(t2);
  assertTrue(pass("host1", req1));
  assertTrue(pass("host3", req2));
  assertFalse(pass("host1", POD_INSTANCE));
  assertTrue(pass("host2", POD_INSTANCE));
  assertFalse(pass("host3", POD_INSTANCE));
  tasks.push_back(hostTask("3", "host2"));
  for (const char* h : {"host1", "host2", "host3"}) assertTrue(pass(h, POD_INSTANCE));
  tasks.push_back(hostTask("4", "host2"));
  assertTrue(pass("host1", POD_INSTANCE));
  assertFalse(pass("host2", POD_INSTANCE));
  assertTrue(pass("host3", POD_INSTANCE));
  tasks.push_back(hostTask("5", "host3"));
  assertTrue(pass("host4", POD_INSTANCE));
  tasks.push_back(hostTask("6", "host4"));
  assertTrue(pass("host4", POD_INSTANCE));
  tasks.push_back(hostTask("7", "host4"));
  assertFalse(pass("host2", POD_INSTANCE));
  assertFalse(pass("host3", POD_INSTANCE));
  assertFalse(pass("host4", POD_INSTANCE));
  assertTrue(pass("host1", POD_INSTANCE));
}
KAT(RoundRobinByZoneRule_offerWithoutZoneFails) {
  RulePtr rule = std::make_shared<RoundRobinByZoneRule>(std::nullopt, AnyMatcher::create());
  assertFalse(filter(rule, hostOffer("h"), typePod(), {}));
  assertTrue(filter(rule, hostOffer("h", mesos::DomainInfo{"r", "z"}), typePod(), {}));
}

// ---- PT/TaskTypeRuleTest.java:112-196
namespace {
mesos::TaskInfo typedTask(const std::string& type, const std::string& name, const std::string& agent, int index) {
  mesos::TaskInfo t = getTaskInfo(name);
  t.slave_id = agent;
  TaskLabelWriter(t).setType(type).setIndex(index);
  return t;
}
mesos::Offer agentOffer(const std::string& agent) {
  mesos::Offer o = OfferTestUtils::getEmptyOfferBuilder();
  o.slave_id = agent;
  return o;
}
struct TaskTypeFixture {
  mesos::TaskInfo MATCH_1 = typedTask("match", "matchtask-1", "agent1", 11), MATCH_3 = typedTask("match", "matchtask-3", "agent3", 13),
                  MISMATCH_1 = typedTask("mismatch", "othertask-1", "agent1", 21), MISMATCH_2 = typedTask("mismatch", "othertask-2", "agent2", 22),
                  MISMATCH_3 = typedTask("mismatch", "othertask-3", "agent3", 23);
  std::vector<mesos::TaskInfo> TASKS{MISMATCH_1, MATCH_1, MISMATCH_2, MISMATCH_3, MATCH_3}, MISMATCH_TASKS{MISMATCH_1, MISMATCH_2, MISMATCH_3};
  mesos::Offer OFFER_1 = agentOffer("agent1"), OFFER_2 = agentOffer("agent2"), OFFER_3 = agentOffer("agent3");
  PodInstance POD = PodInstanceRequirementTestUtils::getCpuRequirement(1.0).podInstance;
  std::array<bool, 3> run(const RulePtr& rule, const PodInstance& pod, const std::vector<mesos::TaskInfo>& tasks) {
    return {filter(rule, OFFER_1, pod, tasks), filter(rule, OFFER_2, pod, tasks), filter(rule, OFFER_3, pod, tasks)};
  }
};
std::ostream& operator<<(std::ostream& o, const std::array<bool, 3>& a) { return o << a[0] << a[1] << a[2]; }
using B3 = std::array<bool, 3>;
}  // namespace
KAT(TaskTypeRuleTest_testColocate) { TaskTypeFixture f; assertEquals((B3{true, false, true}), f.run(TaskTypeRule::colocateWith("match"), f.POD, f.TASKS)); }
KAT(TaskTypeRuleTest_testColocateWithSamePresent) {
  TaskTypeFixture f;
  RulePtr rule = TaskTypeRule::colocateWith("match");
  assertEquals((B3{false, false, true}), f.run(rule, podOf(f.MATCH_1), f.TASKS));
  assertEquals((B3{true, false, false}), f.run(rule, podOf(f.MATCH_3), f.TASKS));
  assertEquals((B3{true, false, true}), f.run(rule, podOf(f.MISMATCH_1), f.TASKS));
  assertEquals((B3{true, false, true}), f.run(rule, podOf(f.MISMATCH_2), f.TASKS));
  assertEquals((B3{true, false, true}), f.run(rule, podOf(f.MISMATCH_3), f.TASKS));
}
KAT(TaskTypeRuleTest_testAvoid) { TaskTypeFixture f; assertEquals((B3{false, true, false}), f.run(TaskTypeRule::avoid("match"), f.POD, f.TASKS)); }
KAT(TaskTypeRuleTest_testAvoidWithSamePresent) {
  TaskTypeFixture f;
  RulePtr rule = TaskTypeRule::avoid("match");
  assertEquals((B3{true, true, false}), f.run(rule, podOf(f.MATCH_1), f.TASKS));
  assertEquals((B3{false, true, true}), f.run(rule, podOf(f.MATCH_3), f.TASKS));
  assertEquals((B3{false, true, false}), f.run(rule, podOf(f.MISMATCH_1), f.TASKS));
  assertEquals((B3{false, true, false}), f.run(rule, podOf(f.MISMATCH_2), f.TASKS));
  assertEquals((B3{false, true, false}), f.run(rule, podOf(f.MISMATCH_3), f.TASKS));
}
KAT(TaskTypeRuleTest_testColocateNotFound) { TaskTypeFixture f; assertEquals((B3{true, true, true}), f.run(TaskTypeRule::colocateWith("match"), f.POD, f.MISMATCH_TASKS)); }
KAT(TaskTypeRuleTest_testAvoidNotFound) { TaskTypeFixture f; assertEquals((B3{true, true, true}), f.run(TaskTypeRule::avoid("match"), f.POD, f.MISMATCH_TASKS)); }

// ---- PT/AndRuleTest.java, OrRuleTest.java, NotRuleTest.java
namespace {
RulePtr PASS() { return std::make_shared<PassthroughRule>(); }
RulePtr FAIL() { return std::make_shared<InvalidPlacementRule>("", ""); }
bool tree(const RulePtr& r) { return filter(r, OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}); }
RulePtr And(std::vector<RulePtr> r) { return std::make_shared<AndRule>(std::move(r)); }
RulePtr Or(std::vector<RulePtr> r) { return std::make_shared<OrRule>(std::move(r)); }
RulePtr Not(RulePtr r) { return std::make_shared<NotRule>(std::move(r)); }
}  // namespace
KAT(AndRuleTest_truthTable) {
  assertFalse(tree(And({})));  // AndRule.java:40-44
  assertTrue(tree(And({PASS()})));
  assertFalse(tree(And({FAIL()})));
  assertTrue(tree(And({PASS(), PASS()})));
  assertFalse(tree(And({PASS(), FAIL()})));
  assertFalse(tree(And({FAIL(), PASS()})));
  assertFalse(tree(And({FAIL(), FAIL()})));
  assertFalse(tree(And({PASS(), PASS(), FAIL()})));
  assertTrue(tree(And({PASS(), PASS(), PASS()})));
}
KAT(OrRuleTest_truthTable) {
  assertFalse(tree(Or({})));  // OrRule.java:36-58
  assertTrue(tree(Or({PASS()})));
  assertFalse(tree(Or({FAIL()})));
  assertTrue(tree(Or({PASS(), FAIL()})));
  assertTrue(tree(Or({FAIL(), PASS()})));
  assertFalse(tree(Or({FAIL(), FAIL()})));
  assertTrue(tree(Or({FAIL(), FAIL(), PASS()})));
}
KAT(NotRuleTest_invertsAndNests) {
  assertFalse(tree(Not(PASS())));
  assertTrue(tree(Not(FAIL())));
  assertTrue(tree(Not(Not(PASS()))));
  assertTrue(tree(And({Not(FAIL()), Or({FAIL(), Not(And({PASS(), FAIL()}))})})));
}

// ---- PT/AttributeRuleTest.java:54-190
namespace {
mesos::Attribute ATTR_TEXT() { return textAttr("footext", "bar"); }
mesos::Attribute ATTR_SCALAR() { return mesos::Attribute{"fooscalar", mesos::Value::Scalar(123.456)}; }
mesos::Attribute ATTR_RANGES() { return mesos::Attribute{"fooranges", mesos::Value::Ranges({{234, 345}, {456, 567}})}; }
mesos::Attribute ATTR_SET() { return mesos::Attribute{"fooset", mesos::Value::Set({"bar", "baz"})}; }
mesos::Offer attrOffer(std::vector<mesos::Attribute> attrs) {
  mesos::Offer o = OfferTestUtils::getEmptyOfferBuilder();
  o.attributes = std::move(attrs);
  return o;
}
bool exactAttr(const mesos::Offer& o, const mesos::Attribute& a) {
  return filter(AttributeRuleFactory::require(ExactMatcher::create(AttributeStringUtils::toString(a))), o, typePod(), {});
}
bool regexAttr(const mesos::Offer& o, const char* re) { return filter(AttributeRuleFactory::require(RegexMatcher::create(re)), o, typePod(), {}); }
}  // namespace
KAT(AttributeStringUtils_formats) {
  assertEquals(std::string("footext:bar"), AttributeStringUtils::toString(ATTR_TEXT()));
  assertEquals(std::string("fooscalar:123.456"), AttributeStringUtils::toString(ATTR_SCALAR()));
  assertEquals(std::string("fooranges:[234-345,456-567]"), AttributeStringUtils::toString(ATTR_RANGES()));
  assertEquals(std::string("fooset:{bar,baz}"), AttributeStringUtils::toString(ATTR_SET()));
  assertEquals(std::string("a:1.000"), AttributeStringUtils::toString(mesos::Attribute{"a", mesos::Value::Scalar(1)}));
}
KAT(AttributeRuleTest_testExactMatchesString) {
  for (auto a : {ATTR_TEXT(), ATTR_SCALAR(), ATTR_RANGES(), ATTR_SET()}) assertTrue(exactAttr(attrOffer({a}), a));
}
KAT(AttributeRuleTest_testAnyMatchesString) {
  mesos::Offer o = attrOffer({ATTR_TEXT(), ATTR_SCALAR(), ATTR_RANGES(), ATTR_SET()});
  for (auto a : {ATTR_TEXT(), ATTR_SCALAR(), ATTR_RANGES(), ATTR_SET()}) assertTrue(exactAttr(o, a));
}
KAT(AttributeRuleTest_testMismatchesString) {
  mesos::Offer o = attrOffer({ATTR_SCALAR(), ATTR_SET()});
  assertFalse(exactAttr(o, ATTR_TEXT()));
  assertTrue(exactAttr(o, ATTR_SCALAR()));
  assertFalse(exactAttr(o, ATTR_RANGES()));
  assertTrue(exactAttr(o, ATTR_SET()));
  o = attrOffer({ATTR_RANGES(), ATTR_TEXT()});
  assertTrue(exactAttr(o, ATTR_TEXT()));
  assertFalse(exactAttr(o, ATTR_SCALAR()));
  assertTrue(exactAttr(o, ATTR_RANGES()));
  assertFalse(exactAttr(o, ATTR_SET()));
}
KAT(AttributeRuleTest_testRegexes) {
  const char *T = "footext:...", *S = ".*:[0-9.]+", *R = ".*ranges:.+", *E = ".*:\\{bar,baz\\}";
  assertTrue(regexAttr(attrOffer({ATTR_TEXT()}), T));
  assertTrue(regexAttr(attrOffer({ATTR_SCALAR()}), S));
  assertTrue(regexAttr(attrOffer({ATTR_RANGES()}), R));
  assertTrue(regexAttr(attrOffer({ATTR_SET()}), E));
  mesos::Offer o = attrOffer({ATTR_SCALAR(), ATTR_SET()});
  assertFalse(regexAttr(o, T));
  assertTrue(regexAttr(o, S));
  assertFalse(regexAttr(o, R));
  assertTrue(regexAttr(o, E));
  assertFalse(regexAttr(attrOffer({}), ".*"));  // no attributes, nothing to match
}

// ---- PT/HostnameRuleTest.java, ZoneRuleTest.java, RegionRuleTest.java, IsLocalRegionRuleTest.java, AgentRuleTest.java
KAT(HostnameRuleTest_requireAvoidExactAndRegex) {
  mesos::Offer o = hostOffer("host1");
  assertTrue(filter(HostnameRuleFactory::require(ExactMatcher::create("host1")), o, typePod(), {}));
  assertFalse(filter(HostnameRuleFactory::require(ExactMatcher::create("host2")), o, typePod(), {}));
  assertFalse(filter(HostnameRuleFactory::avoid(ExactMatcher::create("host1")), o, typePod(), {}));
  assertTrue(filter(HostnameRuleFactory::avoid(ExactMatcher::create("host2")), o, typePod(), {}));
  assertTrue(filter(HostnameRuleFactory::require(RegexMatcher::create("host[0-9]")), o, typePod(), {}));
  assertFalse(filter(HostnameRuleFactory::require(RegexMatcher::create("host")), o, typePod(), {}));  // full match, not find
  assertTrue(filter(HostnameRuleFactory::require(AnyMatcher::create()), o, typePod(), {}));
}
KAT(ZoneRuleTest_offersWithAndWithoutZone) {
  RulePtr rule = ZoneRuleFactory::require(ExactMatcher::create(TestConstants::ZONE));
  assertFalse(filter(rule, OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}));  // ZoneRule.java:39-40
  assertTrue(filter(rule, hostOffer("h", TestConstants::LOCAL_DOMAIN_INFO()), typePod(), {}));
  assertFalse(filter(ZoneRuleFactory::require(ExactMatcher::create("elsewhere")), hostOffer("h", TestConstants::LOCAL_DOMAIN_INFO()), typePod(), {}));
}
KAT(RegionRuleTest_offersWithAndWithoutRegion) {
  RulePtr rule = RegionRuleFactory::require(ExactMatcher::create(TestConstants::LOCAL_REGION));
  assertFalse(filter(rule, OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}));
  assertTrue(filter(rule, hostOffer("h", TestConstants::LOCAL_DOMAIN_INFO()), typePod(), {}));
  assertFalse(filter(rule, hostOffer("h", TestConstants::REMOTE_DOMAIN_INFO()), typePod(), {}));
}
KAT(IsLocalRegionRuleTest_all) {
  RulePtr rule = std::make_shared<IsLocalRegionRule>();
  IsLocalRegionRule::clearLocalDomain();
  assertTrue(filter(rule, hostOffer("h", TestConstants::REMOTE_DOMAIN_INFO()), typePod(), {}));  // local domain unknown
  IsLocalRegionRule::setLocalDomain(TestConstants::LOCAL_DOMAIN_INFO());
  assertTrue(filter(rule, OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}));              // offer without a domain
  assertTrue(filter(rule, hostOffer("h", TestConstants::LOCAL_DOMAIN_INFO()), typePod(), {}));
  assertFalse(filter(rule, hostOffer("h", TestConstants::REMOTE_DOMAIN_INFO()), typePod(), {}));
  IsLocalRegionRule::clearLocalDomain();
}
KAT(AgentRuleTest_requireAvoidAndPlacementUtils) {
  assertTrue(filter(AgentRule::require({"agent1"}), agentOffer("agent1"), typePod(), {}));
  assertFalse(filter(AgentRule::require({"agent1"}), agentOffer("agent2"), typePod(), {}));
  assertTrue(filter(AgentRule::require({"agent1", "agent2"}), agentOffer("agent2"), typePod(), {}));
  assertFalse(filter(AgentRule::avoid({"agent1", "agent2"}), agentOffer("agent2"), typePod(), {}));
  assertTrue(filter(AgentRule::avoid({"agent1", "agent2"}), agentOffer("agent3"), typePod(), {}));
  auto both = PlacementUtils::getAgentPlacementRule({"agent1"}, {"agent2", "agent3"});
  assertTrue(both.has_value());
  assertFalse(filter(*both, agentOffer("agent1"), typePod(), {}));
  assertTrue(filter(*both, agentOffer("agent3"), typePod(), {}));
  assertFalse(filter(*both, agentOffer("agent4"), typePod(), {}));
  assertFalse(PlacementUtils::getAgentPlacementRule({}, {}).has_value());
}

// ---- PT/ExactMatcherTest.java (number normalisation through DecimalFormat("0.###"))
KAT(ExactMatcherTest_numbersAndStrings) {
  assertTrue(ExactMatcher::create("1")->matches("1.0"));
  assertTrue(ExactMatcher::create("1.000")->matches("1"));
  assertTrue(ExactMatcher::create("0.5")->matches(".5"));
  assertTrue(ExactMatcher::create("1.2345")->matches("1.234"));  // both round to 1.234 (HALF_EVEN on the binary value)
  assertFalse(ExactMatcher::create("1.2")->matches("1.3"));
  assertTrue(ExactMatcher::create("foo")->matches("foo"));
  assertFalse(ExactMatcher::create("foo")->matches("Foo"));
  assertFalse(ExactMatcher::create("foo")->matches("foo "));
  assertEquals(std::string("ExactMatcher{str='rack:1'}"), ExactMatcher::createAttribute("rack", "1")->toString());  // "rack:1" is not a number
  assertEquals(std::string("ExactMatcher{str='1000'}"), ExactMatcher::create("1e3")->toString());
}

// ---- PT/AbstractRoundRobinRuleTest.java:15-157.  The reference's TestRoundRobinRule hands arbitrary offer / task keys to
// AbstractRoundRobinRule.filter; here the keys are the values of attribute "rack_id" under a RoundRobinByAttributeRule (a null
// key = the attribute is missing).  TASK_INFO carries no type / index labels there, so no task is "the same pod" (:60-68 of
// P/AbstractRoundRobinRule.java); the restated tasks have distinct indexes of another type.
namespace {
std::vector<mesos::TaskInfo> keyTasks(const std::vector<std::string>& keys) {
  std::vector<mesos::TaskInfo> tasks;
  for (size_t k = 0; k < keys.size(); ++k) tasks.push_back(rrTask("dummy" + std::to_string(k), keys[k]));
  return tasks;
}
bool roundRobin(MatcherPtr taskFilter, std::optional<int> distinct, const char* offerKey, const std::vector<mesos::TaskInfo>& tasks) {
  RulePtr rule = std::make_shared<RoundRobinByAttributeRule>(ATTRIB_NAME, distinct, std::move(taskFilter));
  mesos::Offer offer = offerKey ? offerWithAttribute(offerKey) : OfferTestUtils::getEmptyOfferBuilder();
  return filter(rule, offer, PodTestUtils::getPodInstance(0), tasks);
}
mesos::TaskInfo taskWithoutKey(const std::string& name) {   // getKey(task) == null: the attribute label is absent
  mesos::TaskInfo t = getTaskInfo(name);
  TaskLabelWriter(t).setType(TestConstants::TASK_TYPE).setIndex(nextIndex++);
  return t;
}
}  // namespace
KAT(AbstractRoundRobinRuleTest_offerKeyNull) { assertFalse(roundRobin(AnyMatcher::create(), 1, nullptr, {})); }                                     // :15-31
KAT(AbstractRoundRobinRuleTest_noTasksMatch) { assertTrue(roundRobin(ExactMatcher::create("banana"), 1, "key0", {taskWithoutKey("t")})); }          // :33-47
KAT(AbstractRoundRobinRuleTest_allTasksMatch) { assertTrue(roundRobin(AnyMatcher::create(), 1, "key0", {taskWithoutKey("t")})); }                   // :49-63
KAT(AbstractRoundRobinRuleTest_rejectDueToInsufficientSpread) { assertFalse(roundRobin(AnyMatcher::create(), 2, "key0", keyTasks({"key0"}))); }     // :69-83
KAT(AbstractRoundRobinRuleTest_acceptSufficientSpread) { assertTrue(roundRobin(AnyMatcher::create(), 2, "key1", keyTasks({"key0"}))); }             // :89-103
KAT(AbstractRoundRobinRuleTest_acceptSecondLayerOfRoundRobining) { assertTrue(roundRobin(AnyMatcher::create(), 2, "key0", keyTasks({"key0", "key1"}))); }   // :109-123
KAT(AbstractRoundRobinRuleTest_rejectFullKey) { assertFalse(roundRobin(AnyMatcher::create(), 2, "key0", keyTasks({"key0", "key1", "key0"}))); }     // :129-143

// ---- PT/RoundRobinByZoneRuleTest.java:24-62 and RoundRobinByRegionRuleTest.java:24-62 assert on getKey(offer) / getKey(task); the
// same four facts seen through filter(): an offer without a fault domain has no key (FAIL, AbstractRoundRobinRule.java:56-59), an
// offer's key is its zone (region) name, a task without the label has no key (it is not tallied), a task's key is its zone
// (region) label.  TestConstants.LOCAL_DOMAIN_INFO = region "test-region", zone "test-zone".
namespace {
const mesos::DomainInfo LOCAL{"test-region", "test-zone"}, REMOTE{"other-region", "other-zone"};
mesos::TaskInfo domainTask(const std::string& name, std::optional<mesos::DomainInfo> d) {
  return placedTask(name, TestConstants::TASK_TYPE, nextIndex++, "h-" + name, "a-" + name, d);
}
}  // namespace
KAT(RoundRobinByZoneRuleTest_getNullOfferKey) {
  assertFalse(filter(std::make_shared<RoundRobinByZoneRule>(2), OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}));
}
KAT(RoundRobinByZoneRuleTest_getZoneOfferKey_getZoneTaskKey) {
  RulePtr rule = std::make_shared<RoundRobinByZoneRule>(2);
  // one task in the local zone: the offer's zone key equals the task's, 1 of 2 zones used -> this zone is full for now
  std::vector<mesos::TaskInfo> tasks = {domainTask("t0", LOCAL)};
  assertFalse(filter(rule, hostOffer("h", LOCAL), typePod(), tasks));
  assertTrue(filter(rule, hostOffer("h", REMOTE), typePod(), tasks));
  // the key is the ZONE name: same region, other zone is a different key
  assertTrue(filter(rule, hostOffer("h", mesos::DomainInfo{"test-region", "zone-b"}), typePod(), tasks));
}
KAT(RoundRobinByZoneRuleTest_getNullTaskKey) {
  RulePtr rule = std::make_shared<RoundRobinByZoneRule>(2);
  assertTrue(filter(rule, hostOffer("h", LOCAL), typePod(), {domainTask("t0", std::nullopt)}));   // an unlabelled task is not tallied
}
KAT(RoundRobinByRegionRuleTest_getNullOfferKey) {
  assertFalse(filter(std::make_shared<RoundRobinByRegionRule>(2), OfferTestUtils::getEmptyOfferBuilder(), typePod(), {}));
}
KAT(RoundRobinByRegionRuleTest_getRegionOfferKey_getRegionTaskKey) {
  RulePtr rule = std::make_shared<RoundRobinByRegionRule>(2);
  std::vector<mesos::TaskInfo> tasks = {domainTask("t0", LOCAL)};
  assertFalse(filter(rule, hostOffer("h", LOCAL), typePod(), tasks));
  assertFalse(filter(rule, hostOffer("h", mesos::DomainInfo{"test-region", "zone-b"}), typePod(), tasks));   // the key is the REGION name
  assertTrue(filter(rule, hostOffer("h", REMOTE), typePod(), tasks));
}
KAT(RoundRobinByRegionRuleTest_getNullTaskKey) {
  RulePtr rule = std::make_shared<RoundRobinByRegionRule>(2);
  assertTrue(filter(rule, hostOffer("h", LOCAL), typePod(), {domainTask("t0", std::nullopt)}));
}

// ---- PT/MaxPerAttributeRuleTest.java:57-97 assert on getKeys(task) / getKeys(offer) of MaxPerAttributeRule(2, matcher[, taskFilter]);
// through filter(): keys that exist are counted (offer 1 + tasks), keys the matcher rejects or tasks the filter rejects are not.
namespace {
mesos::Offer fooOffer() {
  mesos::Offer o = OfferTestUtils::getEmptyOfferBuilder();
  o.attributes.push_back(textAttr("footext", "123"));
  return o;
}
mesos::TaskInfo fooTask(const std::string& name) {   // getTask("match-1__abc", OFFER_ATTR_MATCH_1): the offer's attributes as the task's label
  mesos::TaskInfo t = getTaskInfo(name);
  TaskLabelWriter(t).setType("different-type").setIndex(nextIndex++);
  t.labels[labels::OFFER_ATTRIBUTES] = "footext:123";
  return t;
}
}  // namespace
KAT(MaxPerAttributeRuleTest_getAllTaskKeys_getAllOfferKeys) {
  RulePtr rule = std::make_shared<MaxPerAttributeRule>(2, AnyMatcher::create());
  assertTrue(filter(rule, fooOffer(), typePod(), {fooTask("match-1")}));                       // "footext:123": offer 1 + task 1 <= 2
  assertFalse(filter(rule, fooOffer(), typePod(), {fooTask("match-1"), fooTask("match-2")}));  // 3 > 2
}
KAT(MaxPerAttributeRuleTest_getTaskKeysMatchNoAttributes_getOfferKeysMatchNoAttributes) {
  RulePtr rule = std::make_shared<MaxPerAttributeRule>(2, RegexMatcher::create("banana"));
  assertTrue(filter(rule, fooOffer(), typePod(), {fooTask("match-1"), fooTask("match-2"), fooTask("match-3")}));   // no key matches: nothing to exceed
}
KAT(MaxPerAttributeRuleTest_getTaskKeysMatchNoTasks) {
  RulePtr rule = std::make_shared<MaxPerAttributeRule>(2, AnyMatcher::create(), RegexMatcher::create("banana"));
  assertTrue(filter(rule, fooOffer(), typePod(), {fooTask("match-1"), fooTask("match-2"), fooTask("match-3")}));   // no task passes the filter
}

// ---- PT/PlacementUtilsTest.java:40-112: which rule trees read the offer's region (PlacementUtils.java:126-141 over every rule's
// getPlacementFields()).  getPodSpec(rule) there only wraps the rule into a PodSpec; the restatement hands PodSpec::placementRule over.
namespace {
RulePtr regionRule() { return std::make_shared<RegionRule>(ExactMatcher::create("region")); }
RulePtr zoneRule() { return std::make_shared<ZoneRule>(ExactMatcher::create("zone")); }
RulePtr attributeRule() { return std::make_shared<AttributeRule>(ExactMatcher::create("attribute")); }
bool refsRegion(std::optional<RulePtr> rule) {
  PodSpec pod;
  pod.placementRule = std::move(rule);
  return PlacementUtils::placementRuleReferencesRegion(pod.placementRule);
}
}  // namespace
KAT(PlacementUtilsTest_emptyRuleNotRegion) { assertFalse(refsRegion(std::nullopt)); }
KAT(PlacementUtilsTest_passThroughRuleNotRegion) { assertFalse(refsRegion(RulePtr(std::make_shared<PassthroughRule>()))); }
KAT(PlacementUtilsTest_simpleRegionRuleSucceeds) { assertTrue(refsRegion(regionRule())); }
KAT(PlacementUtilsTest_simpleNotRegionRuleFails) { assertFalse(refsRegion(zoneRule())); }
KAT(PlacementUtilsTest_orRegionRuleSucceeds) { assertTrue(refsRegion(RulePtr(std::make_shared<OrRule>(std::vector<RulePtr>{regionRule(), zoneRule()})))); }
KAT(PlacementUtilsTest_orNotRegionRuleFails) { assertFalse(refsRegion(RulePtr(std::make_shared<OrRule>(std::vector<RulePtr>{attributeRule(), zoneRule()})))); }
KAT(PlacementUtilsTest_nestedOrRegionRuleSucceeds) {
  assertTrue(refsRegion(RulePtr(std::make_shared<AndRule>(attributeRule(), std::make_shared<OrRule>(std::vector<RulePtr>{regionRule(), zoneRule()})))));
}
KAT(PlacementUtilsTest_nestedOrNotRegionRuleFails) {
  assertFalse(refsRegion(RulePtr(std::make_shared<AndRule>(attributeRule(), std::make_shared<OrRule>(std::vector<RulePtr>{attributeRule(), zoneRule()})))));
  assertTrue(PlacementUtils::placementRuleReferencesZone(std::optional<RulePtr>(zoneRule())));
}
