// What a scheduler sees through the C++ host layer (sdk::OfferEvaluator over a resident offer table, the mirror of the reference's
// OfferEvaluator / PlanScheduler.resourceOffers): mesos::Offer objects in, OfferRecommendations out, per cycle — encoder, loads,
// device cycle and the winners' materialisation together, against the encoder alone (tools/encoder_bench.cpp) and the engine
// alone (bench.py).  Cycle 0 sends the whole table (packed wire form); the cycles after it re-offer the agents whose offers were
// accepted plus 1 % of the others and send only those rows.
//   g++ -O2 -std=c++17 -o tools/host_cycle_bench tools/host_cycle_bench.cpp -ldl
//   tools/host_cycle_bench dcos_commons_b200/liboe_b200.so [n_offers] [n_requirements] [cycles]
#include <chrono>
#include <cstdio>
#include <random>

#include "../dcos_commons_b200/host/offer_evaluator.hpp"

using namespace sdk;

static mesos::Resource scalar(const char* name, double v) {
  mesos::Resource r;
  r.name = name; r.value = mesos::Value::Scalar(v);
  return r;
}
static mesos::Offer makeOffer(int i, int serial, std::mt19937_64& rng) {
  mesos::Offer o;
  o.id = "offer-" + std::to_string(i) + "-" + std::to_string(serial);
  o.framework_id = "fw"; o.slave_id = "agent-" + std::to_string(i); o.hostname = "host-" + std::to_string(i) + ".example.com";
  o.resources = {scalar("cpus", 2 + rng() % 63), scalar("mem", 1024.0 * (4 + rng() % 509)), scalar("disk", 1024.0 * (50 + rng() % 1951))};
  if (rng() % 4 == 3) o.resources.push_back(scalar("gpus", 8));
  mesos::Resource p = scalar("ports", 0);
  std::vector<std::pair<int64_t, int64_t>> rg = {{1025, 2180}, {2182, 3887}, {3889, 5049}, {5052, 8079}, {8082, 8180}, {8182, 32000}};
  if (rng() % 100 < 3) { rg[3] = {5052, 7000}; rg.insert(rg.begin() + 4, {7002, 8079}); }
  p.value = mesos::Value::Ranges(rg);
  o.resources.push_back(p);
  mesos::Attribute rack, tier;
  rack.name = "rack"; rack.value = mesos::Value::Text("r" + std::to_string(rng() % 200));
  tier.name = "tier"; tier.value = mesos::Value::Text(rng() % 3 == 0 ? "gold" : rng() % 2 ? "silver" : "bronze");
  o.attributes = {rack, tier};
  const int zone = (int)(rng() % 48);
  o.domain = mesos::DomainInfo{"region-" + std::to_string(zone / 6), "zone-" + std::to_string(zone)};
  return o;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s liboe_b200.so [n_offers] [n_requirements] [cycles]\n", argv[0]); return 2; }
  const int N = argc > 2 ? atoi(argv[2]) : 200000, M = argc > 3 ? atoi(argv[3]) : 2000, cycles = argc > 4 ? atoi(argv[4]) : 5;
  const EngineLibrary lib = EngineLibrary::open(argv[1]);
  std::mt19937_64 rng(20240607);
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };

  ResidentOffers resident;
  {
    const auto t0 = now();
    for (int i = 0; i < N; ++i) resident.offer(makeOffer(i, 0, rng));
    printf("%d offers queued (ResidentOffers::offer): %.1f ms\n", N, ms(t0, now()));
  }
  // four pod shapes, M instances in plan order (interleaved), a zone rule on one of them
  std::vector<std::shared_ptr<PodSpec>> pods;
  for (int k = 0; k < 4; ++k) {
    auto pod = std::make_shared<PodSpec>();
    pod->type = "pod-" + std::to_string(k);
    TaskSpec t;
    t.name = "server";
    ResourceSetBuilder b("role", ANY_ROLE, "principal");
    b.id("rs").cpus(1.0 + k).memory(1024.0 * (1 + k)).port("http", 8080 + k).port("dyn", 0);
    t.resourceSet = b.build();
    pod->tasks.push_back(t);
    if (k == 3) pod->placementRule = std::make_shared<MaxPerZoneRule>(M, std::make_shared<RegexMatcher>("pod-3-.*"));
    pods.push_back(pod);
  }
  StateStore store;
  OfferEvaluator evaluator(lib, store, "svc");
  int next_index = 0;
  for (int cyc = 0; cyc < cycles; ++cyc) {
    std::vector<PodInstanceRequirement> reqs;
    for (int j = 0; j < M; ++j, ++next_index) reqs.push_back(PodInstanceRequirement::newBuilder(PodInstance{pods[(size_t)(j % 4)], next_index / 4}, {"server"}));
    const auto t0 = now();
    const auto recs = evaluator.evaluateCycle(reqs, resident);
    const auto t1 = now();
    size_t placed = 0, n_recs = 0;
    std::vector<std::string> accepted_agents;
    for (auto& r : recs) { placed += !r.empty(); n_recs += r.size(); }
    const Engine* e = evaluator.residentEngine(resident);
    printf("cycle %d: %8.3f ms for %d requirements over %zu standing offers — %zu placed, %zu recommendations; loads so far: %lld whole tables, %lld deltas, %lld rows sent\n",
           cyc, ms(t0, t1), M, resident.list().size() + placed, placed, n_recs, (long long)e->fullLoads, (long long)e->deltaLoads, (long long)e->rowsSent);
    // between the cycles: the agents whose offers were accepted offer again, and 1 % of the others change
    int serial = cyc + 1, reoffered = 0;
    for (size_t r = 0; r < resident.rows(); ++r)
      if (!resident.row(r)) { resident.offer(makeOffer((int)r, serial, rng)); ++reoffered; }
    for (int k = 0; k < N / 100; ++k) { const int i = (int)(rng() % (uint64_t)N); resident.offer(makeOffer(i, serial, rng)); }
    (void)reoffered;
  }
  return 0;
}
