// kat_resident.cpp — the offers a scheduler holds between cycles as ONE resident table (sdk::ResidentOffers): what the
// reference's queue does with arriving and rescinded offers (S/framework/OfferQueue.java:46-85), what PlanScheduler does with
// accepted ones (S/scheduler/plan/PlanScheduler.java:61-66), and that sending only the changed rows (oe_update_offers) gives
// exactly what a full re-encode of the current list gives (S/framework/OfferProcessor.java:240-262 hands the evaluator that list).
// The reference has no vector for this path (its queue holds protobufs, there is nothing to send): every expectation here is the
// plain cycle over `resident.list()`, computed by the same engine / oracle through the plain encoder.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;
using Rec = OfferRecommendation;

namespace {

struct Fixture {
  StateStore stateStore;
  OfferEvaluator evaluator, plain;
  Fixture() : evaluator(*kat::library(), stateStore, TestConstants::SERVICE_NAME), plain(*kat::library(), stateStore, TestConstants::SERVICE_NAME) {}
};

mesos::Offer agentOffer(const std::string& agent, int serial, double cpus, double mem = 4096, std::vector<std::pair<int, int>> ports = {{1025, 2180}, {2182, 32000}},
                        const std::string& rack = "r1", const std::string& zone = "z1") {
  mesos::Offer o = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(cpus), ResourceTestUtils::getUnreservedMem(mem), ResourceTestUtils::getUnreservedDisk(10000)});
  mesos::Resource p = ResourceTestUtils::unreservedScalar("ports", 0);
  std::vector<std::pair<int64_t, int64_t>> rg;
  for (auto& x : ports) rg.push_back({x.first, x.second});
  p.value = mesos::Value::Ranges(rg);
  o.resources.push_back(p);
  o.id = agent + "#" + std::to_string(serial);
  o.slave_id = agent; o.hostname = "host-" + agent;
  mesos::Attribute a;
  a.name = "rack"; a.value = mesos::Value::Text(rack);
  o.attributes.push_back(a);
  o.domain = mesos::DomainInfo{"region-1", zone};
  return o;
}

// offer id of the launch (or "" when nothing was placed), per requirement
std::vector<std::string> winners(const std::vector<std::vector<Rec>>& all) {
  std::vector<std::string> out;
  for (auto& recs : all) out.push_back(recs.empty() ? std::string() : recs[0].offerId);
  return out;
}
std::string shapes(const std::vector<std::vector<Rec>>& all) {
  std::string s;
  for (auto& recs : all) {
    for (auto& r : recs) s += r.type == Rec::RESERVE ? 'R' : r.type == Rec::CREATE ? 'C' : r.type == Rec::LAUNCH_GROUP ? 'L' : r.type == Rec::STORE_TASK ? '-' : '?';
    s += '|';
  }
  return s;
}

PodInstanceRequirement portReq(double cpus, int index, std::optional<RulePtr> rule = std::nullopt) {
  ResourceSetBuilder b = PodInstanceRequirementTestUtils::builder();
  b.cpus(cpus).memory(512).port("http", 8080).port("dyn", 0);
  return PodInstanceRequirementTestUtils::getRequirement(b.build(), TestConstants::POD_TYPE, index, rule);
}

// One cycle both ways: over the resident table (delta upload) and over the plain list (fresh encoder, fresh engine).
void sameCycle(Fixture& f, ResidentOffers& resident, const std::vector<PodInstanceRequirement>& reqs) {
  const std::vector<mesos::Offer> list = resident.list();
  auto want = f.plain.evaluateCycle(reqs, list);
  auto got = f.evaluator.evaluateCycle(reqs, resident);
  assertEquals(winners(want).size(), winners(got).size());
  for (size_t j = 0; j < want.size(); ++j) assertEquals(winners(want)[j], winners(got)[j]);
  assertEquals(shapes(want), shapes(got));
  for (size_t j = 0; j < want.size(); ++j)   // the dynamic port the winner was given
    for (size_t k = 0; k < want[j].size(); ++k)
      if (want[j][k].type == Rec::RESERVE && want[j][k].resource.name == "ports")
        assertEquals(want[j][k].resource.value.ranges[0].first, got[j][k].resource.value.ranges[0].first);
}

}  // namespace

KAT(ResidentOffers_arrivingRescindedAndAcceptedOffers) {
  Fixture f;
  ResidentOffers resident;
  for (int a = 0; a < 6; ++a) resident.offer(agentOffer("agent-" + std::to_string(a), 0, 1.0 + a));
  std::vector<PodInstanceRequirement> reqs = {portReq(2.0, 0), portReq(4.0, 1), portReq(9.0, 2)};
  sameCycle(f, resident, reqs);                                     // cycle 1: the whole table is sent
  const Engine* e = f.evaluator.residentEngine(resident);
  assertEquals((int64_t)1, e->fullLoads);
  const int64_t table = e->rowsSent;                                // 6 offers and the spare rows behind them
  assertTrue(table >= 6);
  assertEquals(4u, resident.list().size());                         // two offers were accepted and left the list
  // between the cycles: agent-1 (accepted before) offers again, agent-0 offers more, agent-5 is rescinded, agent-6 is new
  resident.offer(agentOffer("agent-1", 1, 8.0));
  resident.offer(agentOffer("agent-0", 1, 3.0));
  resident.remove("agent-5#0");
  resident.offer(agentOffer("agent-6", 0, 2.5));
  sameCycle(f, resident, {portReq(2.5, 3), portReq(2.5, 4), portReq(7.0, 5), portReq(50.0, 6)});
  assertEquals((int64_t)1, e->fullLoads);                           // nothing but the changed rows crossed the boundary
  assertEquals((int64_t)1, e->deltaLoads);
  assertEquals(table + 2 /* accepted in cycle 1 */ + 4, e->rowsSent);
  // a cycle in which nothing arrived: only the rows of the offers accepted in cycle 2
  const int64_t before = e->rowsSent;
  sameCycle(f, resident, {portReq(1.0, 7)});
  assertEquals((int64_t)2, e->deltaLoads);
  assertTrue(e->rowsSent - before <= 3);
}

KAT(ResidentOffers_newDictionaryValuesAndPlacementRules) {
  Fixture f;
  ResidentOffers resident;
  for (int a = 0; a < 5; ++a) resident.offer(agentOffer("agent-" + std::to_string(a), 0, 4.0, 4096, {{1025, 32000}}, a < 3 ? "r1" : "r2", a % 2 ? "z1" : "z2"));
  RulePtr onR2 = std::make_shared<AttributeRule>(std::make_shared<ExactMatcher>("rack:r2"));
  sameCycle(f, resident, {portReq(1.0, 0, onR2), portReq(1.0, 1)});
  // a rack, a zone and a hostname the dictionaries have never seen arrive in a delta; the rule of the next cycle names them
  resident.offer(agentOffer("agent-9", 0, 4.0, 4096, {{1025, 32000}}, "r7", "z9"));
  RulePtr onR7 = std::make_shared<AttributeRule>(std::make_shared<ExactMatcher>("rack:r7"));
  RulePtr inZ9 = std::make_shared<ZoneRule>(std::make_shared<ExactMatcher>("z9"));
  sameCycle(f, resident, {portReq(1.0, 2, onR7), portReq(1.0, 3, inZ9), portReq(1.0, 4, onR2)});
  const Engine* e = f.evaluator.residentEngine(resident);
  assertEquals((int64_t)1, e->fullLoads);
  assertEquals((int64_t)1, e->deltaLoads);
}

KAT(ResidentOffers_anOfferThatOutgrowsTheTableForcesAFullReload) {
  Fixture f;
  ResidentOffers resident;
  for (int a = 0; a < 4; ++a) resident.offer(agentOffer("agent-" + std::to_string(a), 0, 2.0));          // two port ranges each
  sameCycle(f, resident, {portReq(1.0, 0)});
  resident.offer(agentOffer("agent-2", 1, 2.0, 4096, {{1025, 2000}, {2100, 3000}, {4000, 5000}, {8000, 9000}}));   // four ranges: no slot for them
  sameCycle(f, resident, {portReq(1.0, 1), portReq(1.0, 2)});
  const Engine* e = f.evaluator.residentEngine(resident);
  assertEquals((int64_t)2, e->fullLoads);
  // ... and the table is resident again afterwards
  resident.offer(agentOffer("agent-3", 1, 5.0));
  sameCycle(f, resident, {portReq(4.0, 3)});
  assertEquals((int64_t)2, e->fullLoads);
  assertEquals((int64_t)1, e->deltaLoads);
}

KAT(ResidentOffers_roundRobinAttributeAddsAColumn) {
  Fixture f;
  ResidentOffers resident;
  for (int a = 0; a < 6; ++a) resident.offer(agentOffer("agent-" + std::to_string(a), 0, 4.0, 4096, {{1025, 32000}}, "r" + std::to_string(a % 3)));
  sameCycle(f, resident, {portReq(1.0, 0)});
  // RoundRobinByAttribute over "rack" needs a round-robin key column the table was not built with
  RulePtr rr = std::make_shared<RoundRobinByAttributeRule>("rack", std::optional<int>(3), std::make_shared<RegexMatcher>(TestConstants::POD_TYPE + "-.*"));
  sameCycle(f, resident, {portReq(1.0, 1, rr), portReq(1.0, 2, rr)});
  const Engine* e = f.evaluator.residentEngine(resident);
  assertEquals((int64_t)2, e->fullLoads);
}
