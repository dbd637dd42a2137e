// kat_multi.cpp — the multi-service fan-out (S/scheduler/multi/MultiServiceEventClient.java:312-372) as one batched engine cycle.
// T/scheduler/multi/MultiServiceEventClientTest.java mocks the services ("consume the first offer", "consume the last offer",
// "no change"); here they are real services whose pods have those effects.  The second half checks the batched cycle against
// the reference's own algorithm run literally: one service after the other, each through its own PlanScheduler over the offers
// the earlier services left.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct ServiceFixture {
  std::string name;
  StateStore stateStore;
  OfferEvaluator evaluator;
  std::vector<Step> steps;
  explicit ServiceFixture(const std::string& n) : name(n), evaluator(*kat::library(), stateStore, n) {}
  void recordOperations(const std::vector<Rec>& recs) {
    std::vector<mesos::TaskInfo> store;
    for (auto& r : recs) if (r.type == Rec::STORE_TASK) store.push_back(*r.task);
    stateStore.storeTasks(store);
  }
};

mesos::Offer numberedOffer(int k, double cpus) {  // MultiServiceEventClientTest.getOffer(k), with something in it
  mesos::Offer o = OfferTestUtils::getCompleteOffer(ResourceTestUtils::getUnreservedCpus(cpus));
  o.id = std::to_string(k); o.slave_id = "agent-" + std::to_string(k); o.hostname = "host-" + std::to_string(k);
  return o;
}

std::vector<MultiServiceEventClient::Service> view(std::vector<std::unique_ptr<ServiceFixture>>& services) {
  std::vector<MultiServiceEventClient::Service> v;
  for (auto& s : services) v.push_back({&s->evaluator, &s->steps});
  return v;
}

// the reference, literally (MultiServiceEventClient.java:326-345 around PlanScheduler.resourceOffers)
std::vector<Rec> sequentialFanOut(const std::vector<mesos::Offer>& offers, std::vector<std::unique_ptr<ServiceFixture>>& services) {
  std::vector<Rec> all;
  std::vector<mesos::Offer> remaining = offers;
  for (auto& s : services) {
    PlanScheduler ps(s->evaluator);
    auto recs = ps.resourceOffers(remaining, s->steps);
    all.insert(all.end(), recs.begin(), recs.end());
    if (!remaining.empty() && !recs.empty()) {  // OfferUtils.filterOutAcceptedOffers
      std::set<std::string> used;
      for (auto& r : recs) used.insert(r.offerId);
      std::vector<mesos::Offer> left;
      for (auto& o : remaining) if (!used.count(o.id)) left.push_back(o);
      remaining.swap(left);
    }
  }
  return all;
}

std::string trace(const std::vector<Rec>& recs) {  // offer id and operation type of every recommendation, in order
  std::string s;
  for (auto& r : recs) {
    s += r.offerId + ":";
    s += r.type == Rec::RESERVE ? 'R' : r.type == Rec::UNRESERVE ? 'U' : r.type == Rec::CREATE ? 'C' : r.type == Rec::DESTROY ? 'D' : r.type == Rec::LAUNCH_GROUP ? 'L' : '-';
    if (r.task) s += "(" + r.task->name + ")";
    s += ' ';
  }
  return s;
}

}  // namespace

KAT(MultiServiceEventClientTest_offerNoClientsDeclineLong) {  // :151-164
  MultiServiceEventClient client(*kat::library());
  assertTrue(client.offers({}, {}).empty());
  assertTrue(client.offers({numberedOffer(1, 1), numberedOffer(2, 1), numberedOffer(3, 1)}, {}).empty());
}
KAT(MultiServiceEventClientTest_emptyOffersHitAllServices) {  // :267-308: with nothing on offer every service is still cranked
  std::vector<std::unique_ptr<ServiceFixture>> services;
  for (int k = 1; k <= 2; ++k) {
    services.push_back(std::make_unique<ServiceFixture>(std::to_string(k)));
    services.back()->steps.resize(1);
    services.back()->steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(1.0);
  }
  MultiServiceEventClient client(*kat::library());
  assertTrue(client.offers({}, view(services)).empty());
  for (auto& s : services) assertTrue(s->steps[0].isPrepared());
}
KAT(MultiServiceEventClientTest_offerPruning) {  // :310-373
  // services 1,4,7 consume the first remaining offer; 2,5,8 the last remaining one; 3,6,9 nothing.  Offer k holds k cpus on top
  // of the executor's 0.1, so "the last remaining" is "the only one left with 7 / 6 / 5 cpus".
  std::vector<std::unique_ptr<ServiceFixture>> services;
  const double wants[9] = {0.5, 7, 100, 0.5, 6, 100, 0.5, 5, 100};
  for (int k = 0; k < 9; ++k) {
    services.push_back(std::make_unique<ServiceFixture>(std::to_string(k + 1)));
    services.back()->steps.resize(1);
    services.back()->steps[0].requirement = PodInstanceRequirementTestUtils::getCpuRequirement(wants[k]);
  }
  std::vector<mesos::Offer> offers;
  for (int k = 1; k <= 7; ++k) offers.push_back(numberedOffer(k, k));
  MultiServiceEventClient client(*kat::library());
  auto recs = client.offers(offers, view(services));
  // the offer each service's recommendations name: 1 ate the first, 2 the last, 3 nothing, 4 the (new) first, ...
  const char* expected[9] = {"1", "7", "", "2", "6", "", "3", "5", ""};
  size_t at = 0;
  for (int k = 0; k < 9; ++k) {
    if (!*expected[k]) { assertTrue(services[(size_t)k]->steps[0].isPrepared()); continue; }
    assertTrue(services[(size_t)k]->steps[0].status == Step::STARTING);
    for (int q = 0; q < 6; ++q) assertEquals(std::string(expected[k]), recs[at++].offerId);   // RRRRL-
  }
  assertEquals(at, recs.size());
  for (auto& r : recs) assertFalse(r.offerId == "4");   // only the middle offer is left at the end
}

// Two services with the SAME pod type, task names and placement constraint: each counts only its own tasks.
KAT(MultiService_tasksOfOtherServicesDoNotCount) {
  auto uniqueHostPod = [] {
    auto pod = std::make_shared<PodSpec>();
    pod->type = "hello";
    pod->placementRule = MarathonConstraintParser::parse("hello", "[[\"hostname\", \"UNIQUE\"]]");
    TaskSpec t; t.name = "server"; t.resourceSet = PodInstanceRequirementTestUtils::getCpuResourceSet(1.0);
    pod->tasks = {t};
    return pod;
  };
  std::vector<std::unique_ptr<ServiceFixture>> services;
  services.push_back(std::make_unique<ServiceFixture>("alpha"));
  services.push_back(std::make_unique<ServiceFixture>("beta"));
  auto podA = uniqueHostPod(), podB = uniqueHostPod();
  std::vector<mesos::Offer> round1 = {numberedOffer(1, 4)};
  // alpha deploys hello-0 on host-1
  services[0]->steps.resize(1);
  services[0]->steps[0].requirement = PodInstanceRequirement::newBuilder(PodInstance{podA, 0}, {"server"});
  MultiServiceEventClient client(*kat::library());
  auto recs = client.offers(round1, view(services));
  assertEquals(6u, recs.size());
  services[0]->recordOperations(recs);
  // next cycle: host-1 is offered again (new offer id), then host-2.  alpha's hello-1 must avoid host-1 (its own hello-0 is
  // there); beta's hello-0 and hello-1 see no tasks of their own anywhere.
  mesos::Offer again = numberedOffer(1, 4); again.id = "1b";
  mesos::Offer again2 = numberedOffer(1, 4); again2.id = "1c";   // a second offer from the same agent
  std::vector<mesos::Offer> round2 = {again, again2, numberedOffer(2, 4), numberedOffer(3, 4)};
  services[0]->steps.assign(1, Step());
  services[0]->steps[0].requirement = PodInstanceRequirement::newBuilder(PodInstance{podA, 1}, {"server"});
  services[1]->steps.assign(2, Step());
  services[1]->steps[0].requirement = PodInstanceRequirement::newBuilder(PodInstance{podB, 0}, {"server"});
  services[1]->steps[1].requirement = PodInstanceRequirement::newBuilder(PodInstance{podB, 1}, {"server"});
  recs = client.offers(round2, view(services));
  assertEquals(18u, recs.size());
  assertEquals(std::string("2"), recs[0].offerId);    // alpha hello-1: host-1 is taken by alpha's hello-0
  assertEquals(std::string("1b"), recs[6].offerId);   // beta hello-0: alpha's task on host-1 is none of its business
  assertEquals(std::string("1c"), recs[12].offerId);  // beta hello-1: counters are the cycle-start snapshot (beta had no tasks)
}

// The batched cycle against the reference's sequential fan-out, on a mixed scenario: services with deployed tasks, placement
// rules (per-service counters), ports, volumes, a relaunch of an existing pod, steps that are not pending, offers that run out.
namespace {
struct Scenario {
  std::vector<std::unique_ptr<ServiceFixture>> services;
  std::vector<mesos::Offer> offers;
};
Scenario mixedScenario() {
  Scenario sc;
  for (const char* n : {"cassandra", "kafka", "hdfs"}) sc.services.push_back(std::make_unique<ServiceFixture>(n));
  auto pod = [](const std::string& type, double cpus, const char* constraint, std::vector<int> ports, double disk) {
    auto p = std::make_shared<PodSpec>();
    p->type = type;
    if (constraint) p->placementRule = MarathonConstraintParser::parse(type, constraint);
    TaskSpec t; t.name = "node";
    auto b = PodInstanceRequirementTestUtils::builder();
    b.cpus(cpus);
    for (int q : ports) b.port("p" + std::to_string(q), q);
    if (disk > 0) b.addRootVolume(disk, "data");
    t.resourceSet = b.build();
    p->tasks = {t};
    return p;
  };
  auto offer = [](int k, double cpus, const char* zone, std::vector<std::pair<int, int>> ports, double disk) {
    mesos::Offer o = numberedOffer(k, cpus);
    o.domain = mesos::DomainInfo{"region-1", zone};
    for (auto& pr : ports) o.resources.push_back(ResourceTestUtils::getUnreservedPorts(pr.first, pr.second));
    if (disk > 0) o.resources.push_back(ResourceTestUtils::getUnreservedDisk(disk));
    return o;
  };
  for (int k = 1; k <= 12; ++k) sc.offers.push_back(offer(k, 1.0 + (k % 4), k % 3 == 0 ? "zone-a" : k % 3 == 1 ? "zone-b" : "zone-c", {{9000 + (k % 2), 9005}, {7000, 7000 + (k % 3)}}, k % 2 ? 5000 : 0));
  auto cass = pod("node", 2.0, "[[\"hostname\", \"MAX_PER\", \"1\"]]", {9001, 0}, 1000);
  auto kafka = pod("node", 1.0, "[[\"@zone\", \"GROUP_BY\", \"3\"]]", {0, 0}, 0);   // same pod type name as cassandra's, on purpose
  auto hdfs = pod("data", 0.5, nullptr, {7001}, 0);
  // round 1 deploys a few pods of each service so that round 2 has counters to read
  auto steps = [](std::shared_ptr<PodSpec> p, int from, int to) {
    std::vector<Step> s;
    for (int i = from; i < to; ++i) { Step st; st.requirement = PodInstanceRequirement::newBuilder(PodInstance{p, i}, {"node"}); s.push_back(st); }
    return s;
  };
  sc.services[0]->steps = steps(cass, 0, 2);
  sc.services[1]->steps = steps(kafka, 0, 3);
  sc.services[2]->steps = steps(hdfs, 0, 1);
  MultiServiceEventClient client(*kat::library());
  auto recs = client.offers(sc.offers, view(sc.services));
  size_t perService[3] = {0, 0, 0};
  for (size_t s = 0; s < 3; ++s) for (auto& st : sc.services[s]->steps) perService[s] += st.status == Step::STARTING;
  assertEquals(2u, perService[0]); assertEquals(3u, perService[1]); assertEquals(1u, perService[2]);
  // record each service's stores in ITS state store (names collide across services: cassandra's and kafka's "node-0-node")
  size_t at = 0;
  for (size_t s = 0; s < 3; ++s) {
    std::vector<Rec> mine;
    for (auto& st : sc.services[s]->steps) {
      if (st.status != Step::STARTING) continue;
      do { mine.push_back(recs[at]); } while (recs[at++].type != Rec::STORE_TASK);
    }
    sc.services[s]->recordOperations(mine);
  }
  assertEquals(at, recs.size());
  // round 2: fresh offer ids from the same agents, more pods, one step already complete, one without a requirement
  for (auto& o : sc.offers) o.id += "-r2";
  sc.services[0]->steps = steps(cass, 2, 6);
  sc.services[0]->steps[1].status = Step::COMPLETE;
  sc.services[1]->steps = steps(kafka, 3, 8);
  sc.services[1]->steps.insert(sc.services[1]->steps.begin() + 2, Step());
  sc.services[2]->steps = steps(hdfs, 1, 9);
  return sc;
}
}  // namespace
KAT(MultiService_batchedCycleEqualsSequentialFanOut) {
  Scenario batched = mixedScenario(), sequential = mixedScenario();
  MultiServiceEventClient client(*kat::library());
  const std::string a = trace(client.offers(batched.offers, view(batched.services)));
  const std::string b = trace(sequentialFanOut(sequential.offers, sequential.services));
  assertEquals(b, a);
  if (std::getenv("KAT_TRACE")) std::cerr << a << "\n";
  assertTrue(a.size() > 200);   // something was placed
  for (size_t s = 0; s < batched.services.size(); ++s)
    for (size_t k = 0; k < batched.services[s]->steps.size(); ++k)
      assertTrue(batched.services[s]->steps[k].status == sequential.services[s]->steps[k].status);
}
