// kat_services.cpp — BASELINE.json configs 2 and 3, built from the shapes of the reference's own service
// definitions and driven through sdk::PlanScheduler.  The reference holds NO offer-simulation trace for these two services
// (frameworks/cassandra/src/test/.../ServiceTest.java:20-82 and frameworks/hdfs/src/test/.../ServiceTest.java:27-123 only
// render specs), so the expected placements below are derived by hand from the reference's rules; they are run against the
// oracle on CPU and against the engine on GPU, which must therefore agree with each other.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct Cluster {
  StateStore store;
  OfferEvaluator evaluator;
  PlanScheduler scheduler;
  Cluster() : evaluator(*kat::library(), store, "svc"), scheduler(evaluator) {}
  // one cycle: offers from all agents, the given steps; launches are recorded afterwards (PersistentLaunchRecorder)
  std::vector<std::pair<std::string, std::string>> cycle(const std::vector<mesos::Offer>& offers, std::vector<Step>& steps) {
    auto recs = scheduler.resourceOffers(offers, steps);
    std::vector<mesos::TaskInfo> toStore;
    std::vector<std::pair<std::string, std::string>> placed;  // (task, hostname)
    for (auto& r : recs) {
      if (r.type == Rec::STORE_TASK) toStore.push_back(*r.task);
      if (r.type == Rec::LAUNCH_GROUP) placed.push_back({r.task->name, r.task->labels.at(labels::OFFER_HOSTNAME)});
    }
    store.storeTasks(toStore);
    return placed;
  }
};

std::vector<std::pair<int64_t, int64_t>> agentPorts() {  // the port ranges of a stock DC/OS private agent with ZK, Mesos and Marathon ports cut out
  return {{1025, 2180}, {2182, 3887}, {3889, 5049}, {5052, 8079}, {8082, 8180}, {8182, 32000}};
}
mesos::Offer agentOffer(int i, double cpus, double mem, double disk, std::vector<double> mounts = {}) {
  mesos::Offer o;
  o.id = "offer-" + std::to_string(i); o.slave_id = "agent-" + std::to_string(i); o.hostname = "host-" + std::to_string(i);
  o.framework_id = "fw";
  o.resources = {ResourceTestUtils::getUnreservedCpus(cpus), ResourceTestUtils::getUnreservedMem(mem), ResourceTestUtils::getUnreservedDisk(disk)};
  mesos::Resource ports = ResourceTestUtils::getUnreservedPorts(0, 0);
  ports.value = mesos::Value::Ranges(agentPorts());
  o.resources.push_back(ports);
  for (double m : mounts) o.resources.push_back(ResourceTestUtils::getUnreservedMountVolume(m));
  return o;
}

// frameworks/cassandra/src/main/dist/svc.yml:50-93 with the universe defaults: node pod, pod-level data volume, resource
// sets server-resources (0.5 cpu, 4096 mem, ports jmx 7199 / storage 7000 / ssl 7001 / native 9042) and
// sidecar-resources (1 cpu, 1024 mem); placement [["hostname","MAX_PER","1"]]
std::shared_ptr<PodSpec> cassandraNode(VolumeSpec::Type diskType) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = "node";
  pod->count = 3;
  pod->placementRule = MarathonConstraintParser::parse("node", "[[\"hostname\", \"MAX_PER\", \"1\"]]");
  VolumeSpec v;
  v.type = diskType; v.size = 10240; v.containerPath = "container-path"; v.role = "cassandra-role"; v.principal = "cassandra-principal";
  pod->volumes.push_back(v);
  ResourceSet server = ResourceSetBuilder("cassandra-role", ANY_ROLE, "cassandra-principal").id("server-resources").cpus(0.5).memory(4096)
                           .port("jmx", 7199).port("storage", 7000).port("ssl", 7001).port("native-client", 9042).build();
  ResourceSet sidecar = ResourceSetBuilder("cassandra-role", ANY_ROLE, "cassandra-principal").id("sidecar-resources").cpus(1).memory(1024).build();
  for (const char* name : {"server"}) { TaskSpec t; t.name = name; t.resourceSet = server; pod->tasks.push_back(t); }
  for (const char* name : {"repair", "cleanup", "backup-schema", "restore-schema"}) { TaskSpec t; t.name = name; t.resourceSet = sidecar; pod->tasks.push_back(t); }
  return pod;
}

std::vector<Step> nodeSteps(const std::shared_ptr<PodSpec>& pod, std::vector<int> indexes) {
  std::vector<Step> steps;
  for (int i : indexes) {
    Step s;
    s.name = "node-" + std::to_string(i);
    s.requirement = PodInstanceRequirement::newBuilder(PodInstance{pod, i}, {"server"});
    steps.push_back(s);
  }
  return steps;
}

}  // namespace

// C2: cassandra, 3-node ring, ROOT data volume, 10 agents (4 cpu / 16384 mem / 102400 disk each)
KAT(C2_cassandra_root_volumes_serial_rollout) {
  Cluster c;
  auto pod = cassandraNode(VolumeSpec::ROOT);
  std::vector<mesos::Offer> offers;
  for (int i = 0; i < 10; ++i) offers.push_back(agentOffer(i, 4, 16384, 102400));
  // serial deploy strategy: one node per cycle, the whole cluster re-offered every cycle
  using P = std::vector<std::pair<std::string, std::string>>;
  auto s0 = nodeSteps(pod, {0});
  assertTrue((c.cycle(offers, s0) == P{{"node-0-server", "host-0"}}));
  auto s1 = nodeSteps(pod, {1});
  assertTrue((c.cycle(offers, s1) == P{{"node-1-server", "host-1"}}));   // host-0 already holds one node-.* task
  auto s2 = nodeSteps(pod, {2});
  assertTrue((c.cycle(offers, s2) == P{{"node-2-server", "host-2"}}));
  assertEquals(15u, c.store.fetchTasks().size());                         // 5 task specs per node are stored, one launched
  // node-1 replaced in place (permanent failure): its own stale tasks do not count against host-1
  PodInstanceRequirement replace = PodInstanceRequirement::newBuilder(PodInstance{pod, 1}, {"server"});
  replace.recoveryType = RecoveryType::PERMANENT;
  std::vector<Step> sr(1);
  sr[0].requirement = replace;
  std::vector<mesos::Offer> onlyOneAndZero = {offers[0], offers[1]};
  assertTrue((c.cycle(onlyOneAndZero, sr) == P{{"node-1-server", "host-1"}}));
}
// the same ring in ONE cycle (parallel strategy): counters are the cycle-start snapshot, offers are consumed whole
KAT(C2_cassandra_parallel_rollout_one_cycle) {
  Cluster c;
  auto pod = cassandraNode(VolumeSpec::ROOT);
  std::vector<mesos::Offer> offers;
  for (int i = 0; i < 10; ++i) offers.push_back(agentOffer(i, 4, 16384, 102400));
  offers[1] = agentOffer(1, 1.0, 16384, 102400);   // too little cpu: 0.1 executor + 0.5 server + 1 sidecar
  mesos::Offer busy = agentOffer(3, 4, 16384, 102400);
  busy.resources[3].value = mesos::Value::Ranges({{1025, 6999}, {7001, 32000}});   // storage port 7000 taken
  offers[3] = busy;
  auto steps = nodeSteps(pod, {0, 1, 2});
  using P = std::vector<std::pair<std::string, std::string>>;
  assertTrue((c.cycle(offers, steps) == P{{"node-0-server", "host-0"}, {"node-1-server", "host-2"}, {"node-2-server", "host-4"}}));
}
// MOUNT data volume: agents carry two 20480 MOUNT disks; the first one that fits is taken whole
KAT(C2_cassandra_mount_volumes) {
  Cluster c;
  auto pod = cassandraNode(VolumeSpec::MOUNT);
  std::vector<mesos::Offer> offers;
  offers.push_back(agentOffer(0, 4, 16384, 102400));                    // no MOUNT disk at all
  offers.push_back(agentOffer(1, 4, 16384, 102400, {5000, 8000}));      // both too small
  offers.push_back(agentOffer(2, 4, 16384, 102400, {5000, 20480}));     // second one fits
  offers.push_back(agentOffer(3, 4, 16384, 102400, {20480, 20480}));
  auto steps = nodeSteps(pod, {0, 1});
  auto recs = c.scheduler.resourceOffers(offers, steps);
  std::vector<std::string> reservedMounts;
  for (auto& r : recs)
    if (r.type == Rec::CREATE) reservedMounts.push_back(r.offerId + ":" + std::to_string((int)r.resource.value.scalar));
  assertTrue((reservedMounts == std::vector<std::string>{"offer-2:20480", "offer-3:20480"}));
  assertEquals(2, countType(recs, Rec::LAUNCH_GROUP));
}

// C3: hdfs (frameworks/hdfs/src/main/dist/svc.yml, Main.java:197-217): journal x3, name x2, data x3; 0.3 cpu / 2048 mem /
// 5000 disk ROOT; static port triples; journal and name avoid each other and themselves, data avoids itself; 1000 agents,
// every 20th of which already has one of the static ports occupied.
namespace {
std::shared_ptr<PodSpec> hdfsPod(const std::string& type, std::vector<int> ports, RulePtr rule) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = type;
  pod->placementRule = rule;
  ResourceSetBuilder b("hdfs-role", ANY_ROLE, "hdfs-principal");
  b.id(type + "-resources").cpus(0.3).memory(2048);
  for (int p : ports) b.port(type + "-" + std::to_string(p), p);
  b.addRootVolume(5000, type + "-data");
  TaskSpec t;
  t.name = "node";
  t.resourceSet = b.build();
  pod->tasks.push_back(t);
  return pod;
}
}  // namespace
KAT(C3_hdfs_colocate_avoid_on_1000_agents) {
  Cluster c;
  RulePtr journalRule = std::make_shared<AndRule>(TaskTypeRule::avoid("journal"), TaskTypeRule::avoid("name"));
  RulePtr nameRule = std::make_shared<AndRule>(TaskTypeRule::avoid("name"), TaskTypeRule::avoid("journal"));
  RulePtr dataRule = TaskTypeRule::avoid("data");
  auto journal = hdfsPod("journal", {8485, 8480}, journalRule);
  auto name = hdfsPod("name", {9001, 9002, 8019}, nameRule);
  auto data = hdfsPod("data", {9003, 9004, 9005}, dataRule);
  std::vector<mesos::Offer> offers;
  for (int i = 0; i < 1000; ++i) {
    mesos::Offer o = agentOffer(i, 8, 32768, 204800);
    if (i % 20 == 0) o.resources[3].value = mesos::Value::Ranges({{1025, 8484}, {8486, 9000}, {9002, 9002}, {9004, 32000}});  // 8485, 9001, 9003 busy
    offers.push_back(o);
  }
  using P = std::vector<std::pair<std::string, std::string>>;
  auto stepOf = [](const std::shared_ptr<PodSpec>& pod, int i) {
    Step s;
    s.requirement = PodInstanceRequirement::newBuilder(PodInstance{pod, i}, {"node"});
    return s;
  };
  // serial phases as the hdfs deploy plan runs them: journal 0..2, name 0..1, data 0..2, one step per cycle
  std::vector<std::pair<std::shared_ptr<PodSpec>, int>> order = {{journal, 0}, {journal, 1}, {journal, 2}, {name, 0}, {name, 1}, {data, 0}, {data, 1}, {data, 2}};
  P all;
  for (auto& [pod, i] : order) {
    std::vector<Step> steps = {stepOf(pod, i)};
    auto placed = c.cycle(offers, steps);
    assertEquals(1u, placed.size());
    all.push_back(placed[0]);
  }
  // host-0 lacks 8485 / 9001 / 9003, so nothing lands there; journal 1,2,3; name avoids journal hosts: 4,5; data avoids only data: 1,2,3
  assertTrue((all == P{{"journal-0-node", "host-1"}, {"journal-1-node", "host-2"}, {"journal-2-node", "host-3"}, {"name-0-node", "host-4"},
                       {"name-1-node", "host-5"}, {"data-0-node", "host-1"}, {"data-1-node", "host-2"}, {"data-2-node", "host-3"}}));
  // the whole plan in one cycle against an empty state store: nothing to avoid yet, so it is pure first fit with whole-offer removal
  Cluster fresh;
  std::vector<Step> steps;
  for (auto& [pod, i] : order) steps.push_back(stepOf(pod, i));
  auto placed = fresh.cycle(offers, steps);
  assertTrue((placed == P{{"journal-0-node", "host-1"}, {"journal-1-node", "host-2"}, {"journal-2-node", "host-3"}, {"name-0-node", "host-4"},
                          {"name-1-node", "host-5"}, {"data-0-node", "host-6"}, {"data-1-node", "host-7"}, {"data-2-node", "host-8"}}));
}
