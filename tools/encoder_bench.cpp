// What the host side of the boundary costs: mesos::Offer objects -> columnar offer table, one thread.
//   full    every offer of the list encoded (what a cycle costs when the table is rebuilt every time)
//   delta   1 % of the offers changed since the last cycle, encoded through sdk::ResidentOffers' dictionaries
// C4- and C5-shaped offers (synth.py's distributions: 4 scalars, six port ranges with 3 % holes, rack + tier attributes, a fault
// domain).  No engine involved: this is encoder time only.
//   g++ -O2 -std=c++17 -o tools/encoder_bench tools/encoder_bench.cpp -ldl && tools/encoder_bench [n_offers]
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
  const int N = argc > 1 ? atoi(argv[1]) : 200000;
  std::mt19937_64 rng(20240607);
  std::vector<mesos::Offer> offers;
  offers.reserve((size_t)N);
  for (int i = 0; i < N; ++i) offers.push_back(makeOffer(i, 0, rng));
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto ns_per = [&](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b, double n) {
    return std::chrono::duration<double, std::nano>(b - a).count() / n;
  };
  // one requirement so that the cycle is complete (its cost is negligible)
  auto pod = std::make_shared<PodSpec>();
  pod->type = "pod";
  TaskSpec t;
  t.name = "server";
  ResourceSetBuilder b("role", ANY_ROLE, "principal");
  b.id("rs").cpus(1.0).memory(1024).port("http", 8080).port("dyn", 0);
  t.resourceSet = b.build();
  pod->tasks.push_back(t);
  PodInstanceRequirement req = PodInstanceRequirement::newBuilder(PodInstance{pod, 0}, {"server"});
  EncoderRequirement er;
  er.requirement = &req;

  // ---- full encode, three times (the first one also pays for the first touch of the vectors)
  for (int rep = 0; rep < 3; ++rep) {
    CycleEncoder enc;
    const auto t0 = now();
    EncodedCycle c = enc.encode(offers, {}, {er});
    const auto t1 = now();
    printf("full encode   %8d offers: %7.1f ns per offer (%6.1f ms), table %.1f MB, R=%d A=%d\n", N, ns_per(t0, t1, N), ns_per(t0, t1, 1) / 1e6,
           (c.o_scalar.size() * 8 + c.o_port.size() * 4 + c.o_attr.size() * 4 + c.o_present.size() * 4 * 6) / 1e6, c.R, c.A);
  }
  // ---- resident table: first cycle full, then 1 % of the agents re-offer per cycle
  OfferDictionaries dicts;
  std::vector<const mesos::Offer*> slots;
  for (auto& o : offers) slots.push_back(&o);
  int shape[8] = {-1, -1, -1, -1, -1, -1, -1, 0};
  std::vector<int64_t> dirty;
  for (int cyc = 0; cyc < 4; ++cyc) {
    CycleEncoder enc(SchedulerConfig(), &dicts);
    CycleEncoder::ResidentRows rr;
    rr.slots = &slots; rr.dirty = &dirty;
    std::copy(shape, shape + 8, rr.shape);
    const auto t0 = now();
    EncodedCycle c = enc.encodeResident(rr, {}, {er});
    const auto t1 = now();
    const double rows = c.full_reload ? N : (double)dirty.size();
    printf("resident cycle %d: %s, %7.0f rows encoded: %7.1f ns per row, %8.3f ms for the cycle\n", cyc, c.full_reload ? "whole table" : "changed rows", rows,
           ns_per(t0, t1, std::max(rows, 1.0)), ns_per(t0, t1, 1) / 1e6);
    if (c.full_reload) { const int s2[8] = {c.S, c.K, c.R, c.D, c.A, c.E, c.Q, 0}; std::copy(s2, s2 + 8, shape); }
    dirty.clear();
    for (int k = 0; k < N / 100; ++k) {
      const int i = (int)(rng() % (uint64_t)N);
      offers[(size_t)i] = makeOffer(i, cyc + 1, rng);
      dirty.push_back(i);
    }
    std::sort(dirty.begin(), dirty.end());
    dirty.erase(std::unique(dirty.begin(), dirty.end()), dirty.end());
  }
  return 0;
}
