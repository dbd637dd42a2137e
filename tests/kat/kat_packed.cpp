// kat_packed.cpp — the compact wire form of an offer table as the host layer builds it (sdk::PackedOffers, oe_offer_packing in
// include/oe_abi.h).  The reference has no such form (its offers stay protobufs, S/offer/MesosResourcePool.java:76-147); what is
// pinned here is that packing loses nothing: every narrow column widens back to the encoder's full-width one, and a cycle over
// the packed table gives what the cycle over the plain table gives.
#include "kat.hpp"

using namespace sdk;
using namespace testutils;

namespace {

mesos::Offer packedOffer(int a, double cpus, std::vector<std::pair<int64_t, int64_t>> ports, const std::string& rack, const std::string& zone, bool gpus = false) {
  mesos::Offer o = OfferTestUtils::getOffer({ResourceTestUtils::getUnreservedCpus(cpus), ResourceTestUtils::getUnreservedMem(1024.0 * (4 + a)),
                                            ResourceTestUtils::getUnreservedDisk(10000 + a)});
  if (gpus) o.resources.push_back(ResourceTestUtils::unreservedScalar("gpus", 8));
  mesos::Resource p = ResourceTestUtils::unreservedScalar("ports", 0);
  p.value = mesos::Value::Ranges(ports);
  o.resources.push_back(p);
  o.id = "offer-" + std::to_string(a);
  o.slave_id = "agent-" + std::to_string(a); o.hostname = "host-" + std::to_string(a);
  mesos::Attribute at;
  at.name = "rack"; at.value = mesos::Value::Text(rack);
  o.attributes.push_back(at);
  if (!zone.empty()) o.domain = mesos::DomainInfo{"region-" + zone.substr(0, 1), zone};
  return o;
}

std::vector<mesos::Offer> offerList(int n, bool oddScalar = false) {
  const std::vector<std::pair<int64_t, int64_t>> stock = {{1025, 2180}, {2182, 3887}, {3889, 32000}}, holes = {{1025, 2180}, {2182, 3000}, {3100, 3887}, {3889, 32000}};
  std::vector<mesos::Offer> offers;
  for (int a = 0; a < n; ++a)
    offers.push_back(packedOffer(a, oddScalar && a == n / 2 ? 0.1 : 1.0 + a % 7, a % 9 == 4 ? holes : stock, "r" + std::to_string(a % 5),
                                 a % 11 == 3 ? std::string() : std::string(a % 2 ? "a" : "b") + "-z" + std::to_string(a % 4), a % 3 == 0));
  return offers;
}

template <class T>
int64_t widened(const void* col, int bytes, size_t k, bool absentIsAllOnes) {
  if (bytes == 4) return (int64_t) static_cast<const T*>(col)[k];
  if (bytes == 2) { const uint16_t v = static_cast<const uint16_t*>(col)[k]; return absentIsAllOnes && v == 0xffff ? -1 : v; }
  const uint8_t v = static_cast<const uint8_t*>(col)[k];
  return absentIsAllOnes && v == 0xff ? -1 : v;
}

PodInstanceRequirement req(double cpus, int index, std::optional<RulePtr> rule = std::nullopt) {
  ResourceSetBuilder b = PodInstanceRequirementTestUtils::builder();
  b.cpus(cpus).memory(512).port("http", 8080).port("dyn", 0);
  return PodInstanceRequirementTestUtils::getRequirement(b.build(), TestConstants::POD_TYPE, index, rule);
}

void roundTrip(const EncodedCycle& c, bool expectF32) {
  const PackedOffers packed(c);
  const oe_offer_table t = packed.table();
  const oe_offer_packing pk = packed.packing();
  const size_t N = (size_t)c.N;
  assertEquals(c.N, t.n_offers);
  assertEquals(expectF32 ? 1 : 0, (int)pk.scalar_is_f32);
  for (size_t k = 0; k < c.o_scalar.size(); ++k) {
    const double v = pk.scalar_is_f32 ? (double)reinterpret_cast<const float*>(t.scalar)[k] : t.scalar[k];
    assertTrue(v == c.o_scalar[k]);
  }
  for (size_t k = 0; k < c.o_present.size(); ++k) assertEquals((int64_t)c.o_present[k], widened<uint32_t>(t.scalar_present, pk.present_bytes, k, false));
  for (size_t k = 0; k < c.o_mprof.size(); ++k) assertEquals((int64_t)c.o_mprof[k], widened<int32_t>(t.mount_profile, pk.mount_profile_bytes, k, true));
  for (size_t k = 0; k < N; ++k) assertEquals((int64_t)c.o_zone[k], widened<int32_t>(t.zone_id, pk.zone_bytes, k, true));
  for (size_t k = 0; k < N; ++k) assertEquals((int64_t)c.o_region[k], widened<int32_t>(t.region_id, pk.region_bytes, k, true));
  for (size_t k = 0; k < c.o_attr.size(); ++k) assertEquals((int64_t)c.o_attr[k], widened<int32_t>(t.attr_id, pk.attr_bytes, k, true));
  for (size_t k = 0; k < c.o_rr.size(); ++k) assertEquals((int64_t)c.o_rr[k], widened<int32_t>(t.rr_key, pk.rr_bytes, k, true));
  for (size_t k = 0; k < c.o_exec.size(); ++k) assertEquals((int64_t)c.o_exec[k], widened<int32_t>(t.exec_id, pk.exec_bytes, k, true));
  for (size_t k = 0; k < N; ++k) assertEquals(c.o_agent[k], t.agent_id[k]);
  if (pk.hostname_is_agent) { assertTrue(t.hostname_id == nullptr); assertTrue(c.o_host == c.o_agent); }
  else for (size_t k = 0; k < N; ++k) assertEquals(c.o_host[k], t.hostname_id[k]);
  if (pk.alloc_role_is_const) { assertTrue(t.alloc_role == nullptr); for (size_t k = 0; k < N; ++k) assertEquals(pk.alloc_role_const, c.o_alloc[k]); }
  else for (size_t k = 0; k < N; ++k) assertEquals(c.o_alloc[k], t.alloc_role[k]);
  if (pk.port_tuple_bytes) {
    assertTrue(t.port_range == nullptr);
    assertTrue((size_t)pk.n_port_tuples * 4 <= std::max<size_t>(N, 4096));
    for (int s = 0; s < c.S; ++s)
      for (size_t i = 0; i < N; ++i) {
        const int64_t id = widened<uint32_t>(pk.port_tuple_id, pk.port_tuple_bytes, (size_t)s * N + i, false);
        assertTrue(id >= 0 && id < pk.n_port_tuples);
        for (int r = 0; r < c.R; ++r) assertEquals(c.o_port[((size_t)s * c.R + r) * N + i], pk.port_tuples[(size_t)id * c.R + r]);
      }
    for (int a = 0; a < pk.n_port_tuples; ++a)       // distinct: no list twice
      for (int b = a + 1; b < pk.n_port_tuples; ++b)
        assertFalse(std::equal(pk.port_tuples + (size_t)a * c.R, pk.port_tuples + (size_t)(a + 1) * c.R, pk.port_tuples + (size_t)b * c.R));
  } else {
    for (size_t k = 0; k < c.o_port.size(); ++k) assertEquals(c.o_port[k], t.port_range[k]);
  }
  assertTrue(packed.bytes() < c.o_scalar.size() * 8 + c.o_port.size() * 4 + N * 4 * 6);
}

EncodedCycle encodeList(const std::vector<mesos::Offer>& offers, const std::vector<PodInstanceRequirement>& reqs) {
  CycleEncoder enc;
  std::vector<EncoderRequirement> ers;
  for (auto& r : reqs) { EncoderRequirement er; er.requirement = &r; ers.push_back(er); }
  return enc.encode(offers, {}, ers);
}

}  // namespace

KAT(PackedOffers_everyColumnWidensBackToTheEncodersOwn) {
  const std::vector<PodInstanceRequirement> reqs = {req(1.0, 0), req(2.0, 1)};
  roundTrip(encodeList(offerList(40), reqs), true);
  roundTrip(encodeList(offerList(300), reqs), true);    // more zones, racks and agents than one table of 40 has
  roundTrip(encodeList(offerList(40, true), reqs), false);   // 0.1 cpus does not survive fp32: the scalars stay fp64
  roundTrip(encodeList(offerList(1), reqs), true);
}

KAT(PackedOffers_aCycleOverThePackedTableIsTheCycleOverThePlainOne) {
  StateStore stateStore;
  const std::vector<mesos::Offer> offers = offerList(120);
  RulePtr onR2 = std::make_shared<AttributeRule>(std::make_shared<ExactMatcher>("rack:r2"));
  RulePtr inZone = std::make_shared<ZoneRule>(std::make_shared<ExactMatcher>("a-z1"));
  std::vector<PodInstanceRequirement> reqs;
  for (int j = 0; j < 24; ++j) reqs.push_back(req(1.0 + j % 5, j, j % 4 == 1 ? std::optional<RulePtr>(onR2) : j % 4 == 2 ? std::optional<RulePtr>(inZone) : std::nullopt));
  OfferEvaluator packed(*kat::library(), stateStore, TestConstants::SERVICE_NAME), plain(*kat::library(), stateStore, TestConstants::SERVICE_NAME);
  plain.packedWire = false;
  const auto want = plain.evaluateCycle(reqs, offers);
  const auto got = packed.evaluateCycle(reqs, offers);
  assertEquals(want.size(), got.size());
  size_t placed = 0;
  for (size_t j = 0; j < want.size(); ++j) {
    assertEquals(want[j].size(), got[j].size());
    placed += !want[j].empty();
    for (size_t k = 0; k < want[j].size(); ++k) {
      assertEquals(want[j][k].offerId, got[j][k].offerId);
      assertEquals((int)want[j][k].type, (int)got[j][k].type);
      if (want[j][k].resource.name == "ports" && !want[j][k].resource.value.ranges.empty())
        assertEquals(want[j][k].resource.value.ranges[0].first, got[j][k].resource.value.ranges[0].first);
    }
  }
  assertTrue(placed >= 12);
  if (kat::library()->load_offers_packed && !std::getenv("OE_HOST_WIRE")) assertTrue(packed.offerBytesSent > 0);
  assertEquals((int64_t)0, plain.offerBytesSent);
}
