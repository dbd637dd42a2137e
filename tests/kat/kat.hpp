// kat.hpp — a few dozen lines of test harness for the known-answer tests that restate the reference's JUnit vectors
// (T/ = sdk/scheduler/src/test/java/com/mesosphere/sdk/) through the C++ host layer and the C ABI.
//
// The same binary runs against the CPU oracle (pinning the ORACLE to the reference's vectors: `pytest -m "not gpu"`) and
// against the CUDA engine (`pytest -m gpu`), selected by --lib / --prefix.
#pragma once
#include <functional>
#include <iostream>
#include <map>
#include <sstream>
#include <string>

#include "../../dcos_commons_b200/host/offer_evaluator.hpp"

namespace kat {

// Test-only loader: the same ABI exported under another prefix (the CPU oracle's "oeo_" entry points).  The product header knows
// nothing about prefixes; this subclass resolves the names itself.
struct OracleLibrary : sdk::EngineLibrary {
  OracleLibrary(const std::string& file, const std::string& prefix) {
    void* h = dlopen(file.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) throw std::runtime_error(std::string("cannot load ") + file + ": " + dlerror());
    path = file;
    bind([&](const char* name) {
      const std::string renamed = prefix + (std::string(name).rfind("oe_", 0) == 0 ? std::string(name).substr(3) : std::string(name));
      void* s = dlsym(h, renamed.c_str());
      if (!s) throw std::runtime_error(file + " does not export " + renamed);
      return s;
    });
  }
};

struct Failure : std::runtime_error { using std::runtime_error::runtime_error; };

inline std::map<std::string, std::function<void()>>& registry() { static std::map<std::string, std::function<void()>> r; return r; }
inline const sdk::EngineLibrary*& library() { static const sdk::EngineLibrary* l = nullptr; return l; }
struct Registrar { Registrar(const char* name, std::function<void()> fn) { registry()[name] = std::move(fn); } };

#define KAT(name)                                              \
  static void kat_##name();                                    \
  static ::kat::Registrar kat_reg_##name(#name, kat_##name);   \
  static void kat_##name()

#define KAT_FAIL(msg)                                                                       \
  do { std::ostringstream o__; o__ << __FILE__ << ":" << __LINE__ << ": " << msg; throw ::kat::Failure(o__.str()); } while (0)
#define assertTrue(x) do { if (!(x)) KAT_FAIL("assertTrue(" #x ")"); } while (0)
#define assertFalse(x) do { if ((x)) KAT_FAIL("assertFalse(" #x ")"); } while (0)
#define assertEquals(a, b)                                                                                       \
  do { auto a__ = (a); auto b__ = (b); if (!(a__ == b__)) KAT_FAIL("assertEquals(" #a ", " #b "): expected <" << a__ << "> but was <" << b__ << ">"); } while (0)
#define assertThrows(expr, Ex)                                                               \
  do { bool t__ = false; try { expr; } catch (const Ex&) { t__ = true; } if (!t__) KAT_FAIL(#expr " did not throw " #Ex); } while (0)

}  // namespace kat

// ---- the reference's test fixtures, restated (T/testutils/*.java, T/scheduler/plan/PodInstanceRequirementTestUtils.java)
namespace testutils {
using namespace sdk;

namespace TestConstants {  // T/testutils/TestConstants.java
inline const std::string SERVICE_NAME = "service-name", HOSTNAME = "test-hostname", PRINCIPAL = "test-principal", ROLE = "test-role",
                         PRE_RESERVED_ROLE = "base-role", TASK_NAME = "test-task-name", TASK_TYPE = "test-task-type", POD_TYPE = "pod-type",
                         RESOURCE_SET_ID = "test-resource-set-id", CONTAINER_PATH = "test-container-path", OFFER_ID = "test-offer-id",
                         FRAMEWORK_ID = "test-framework-id", AGENT_ID = "test-slave-id", ZONE = "zone", LOCAL_REGION = "local",
                         REMOTE_REGION = "remote", MOUNT_ROOT = "test-mount-root";
inline mesos::DomainInfo LOCAL_DOMAIN_INFO() { return mesos::DomainInfo{LOCAL_REGION, ZONE}; }
inline mesos::DomainInfo REMOTE_DOMAIN_INFO() { return mesos::DomainInfo{REMOTE_REGION, ZONE}; }
}  // namespace TestConstants

namespace ResourceTestUtils {  // T/testutils/ResourceTestUtils.java
inline mesos::Resource unreservedScalar(const std::string& name, double v, const std::string& preReservedRole = ANY_ROLE) {
  mesos::Resource r;
  r.name = name; r.value = mesos::Value::Scalar(v);
  if (preReservedRole != ANY_ROLE) {
    mesos::ReservationInfo st;
    st.type = mesos::ReservationInfo::STATIC; st.role = preReservedRole;
    r.reservations.push_back(st);
  }
  return r;
}
inline mesos::Resource getUnreservedCpus(double v, const std::string& role = ANY_ROLE) { return unreservedScalar("cpus", v, role); }
inline mesos::Resource getUnreservedMem(double v, const std::string& role = ANY_ROLE) { return unreservedScalar("mem", v, role); }
inline mesos::Resource getUnreservedDisk(double v, const std::string& role = ANY_ROLE) { return unreservedScalar("disk", v, role); }
inline mesos::Resource getUnreservedPorts(int begin, int end, const std::string& role = ANY_ROLE) {
  mesos::Resource r = unreservedScalar("ports", 0, role);
  r.value = mesos::Value::Ranges({{begin, end}});
  return r;
}
inline mesos::Resource getUnreservedMountVolume(double size, std::optional<std::string> profile = std::nullopt) {
  mesos::Resource r = unreservedScalar("disk", size);
  mesos::DiskInfo d;
  d.source = mesos::DiskInfo::MOUNT; d.profile = profile;
  r.disk = d;
  return r;
}
}  // namespace ResourceTestUtils

namespace OfferTestUtils {  // T/testutils/OfferTestUtils.java:54-117
inline mesos::Offer getEmptyOfferBuilder() {
  mesos::Offer o;
  o.id = TestConstants::OFFER_ID; o.framework_id = TestConstants::FRAMEWORK_ID; o.slave_id = TestConstants::AGENT_ID; o.hostname = TestConstants::HOSTNAME;
  return o;
}
inline mesos::Offer getOffer(const std::vector<mesos::Resource>& resources) {
  mesos::Offer o = getEmptyOfferBuilder();
  o.resources = resources;
  return o;
}
inline mesos::Offer getCompleteOffer(const std::vector<mesos::Resource>& resources, const std::string& preReservedRole = ANY_ROLE) {
  mesos::Offer o = getEmptyOfferBuilder();
  o.resources = {ResourceTestUtils::getUnreservedCpus(0.1, preReservedRole), ResourceTestUtils::getUnreservedMem(256, preReservedRole),
                 ResourceTestUtils::getUnreservedDisk(512, preReservedRole)};
  o.resources.insert(o.resources.end(), resources.begin(), resources.end());
  return o;
}
inline mesos::Offer getCompleteOffer(const mesos::Resource& r) { return getCompleteOffer(std::vector<mesos::Resource>{r}); }
}  // namespace OfferTestUtils

namespace PodTestUtils {  // T/testutils/PodTestUtils.java
inline PodInstance getPodInstance(int index) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = TestConstants::POD_TYPE;
  pod->preReservedRole = TestConstants::PRE_RESERVED_ROLE;
  TaskSpec t;
  t.name = TestConstants::TASK_NAME;
  t.resourceSet = ResourceSetBuilder(TestConstants::ROLE, TestConstants::PRE_RESERVED_ROLE, TestConstants::PRINCIPAL).id(TestConstants::RESOURCE_SET_ID).cpus(1.0).build();
  pod->tasks.push_back(t);
  return PodInstance{pod, index};
}
}  // namespace PodTestUtils

namespace PodInstanceRequirementTestUtils {  // T/scheduler/plan/PodInstanceRequirementTestUtils.java
inline PodInstanceRequirement getRequirement(const ResourceSet& rs, const std::string& type, int index, std::optional<RulePtr> rule = std::nullopt) {
  auto pod = std::make_shared<PodSpec>();
  pod->type = type;
  pod->placementRule = rule;
  TaskSpec t;
  t.name = TestConstants::TASK_NAME;
  t.resourceSet = rs;
  pod->tasks.push_back(t);
  return PodInstanceRequirement::newBuilder(PodInstance{pod, index}, {TestConstants::TASK_NAME});
}
inline PodInstanceRequirement getRequirement(const ResourceSet& rs, int index) { return getRequirement(rs, TestConstants::POD_TYPE, index); }
inline ResourceSetBuilder builder(const std::string& preReservedRole = ANY_ROLE) {
  ResourceSetBuilder b(TestConstants::ROLE, preReservedRole, TestConstants::PRINCIPAL);
  b.id(TestConstants::RESOURCE_SET_ID);
  return b;
}
inline ResourceSet getCpuResourceSet(double v, const std::string& preReservedRole = ANY_ROLE) { return builder(preReservedRole).cpus(v).build(); }
inline PodInstanceRequirement getCpuRequirement(double v, int index = 0) { return getRequirement(getCpuResourceSet(v), index); }
inline PodInstanceRequirement getCpuRequirement(double v, const std::string& preReservedRole) { return getRequirement(getCpuResourceSet(v, preReservedRole), 0); }
inline PodInstanceRequirement getRootVolumeRequirement(double cpus, double disk, int index = 0) {
  return getRequirement(builder().cpus(cpus).addRootVolume(disk, TestConstants::CONTAINER_PATH).build(), index);
}
inline PodInstanceRequirement getMountVolumeRequirement(double cpus, double disk, std::vector<std::string> profiles = {}, int index = 0) {
  return getRequirement(builder().cpus(cpus).addMountVolume(disk, TestConstants::CONTAINER_PATH, std::move(profiles)).build(), index);
}
inline PodInstanceRequirement getPortRequirement(std::vector<int> ports) {
  ResourceSetBuilder b = builder();
  for (int p : ports) b.port("test-port-" + std::to_string(p), p);
  return getRequirement(b.build(), 0);
}
}  // namespace PodInstanceRequirementTestUtils

// T/testutils/TaskTestUtils.java + TestConstants.TASK_INFO: a deployed task with just a name and an agent
inline mesos::TaskInfo getTaskInfo(const std::string& name = TestConstants::TASK_NAME) {
  mesos::TaskInfo t;
  t.name = name; t.task_id = name + "__uuid"; t.slave_id = TestConstants::AGENT_ID;
  return t;
}

inline int countType(const std::vector<OfferRecommendation>& recs, OfferRecommendation::Type t) {
  int n = 0;
  for (auto& r : recs) n += r.type == t;
  return n;
}

}  // namespace testutils
