// specification.hpp — C++ mirror of the parts of com.mesosphere.sdk.specification that the offer-evaluation path reads
// (S/ = sdk/scheduler/src/main/java/com/mesosphere/sdk/): ResourceSpec, PortSpec, VolumeSpec, ResourceSet, TaskSpec,
// PodSpec, PodInstance, and scheduler/plan/PodInstanceRequirement.
#pragma once
#include <memory>
#include <optional>
#include <string>
#include <vector>

#include "placement.hpp"

namespace sdk {

constexpr const char* ANY_ROLE = "*";  // S/offer/Constants.java

// S/specification/RangeSpec.java:12-54 — a missing or zero end means 65535
struct RangeSpec {
  int begin = 0, end = 65535;
  RangeSpec(int b, std::optional<int> e = std::nullopt) : begin(b), end((!e || *e == 0) ? 65535 : *e) {}
};

// S/specification/ResourceSpec.java:12-28 and S/specification/PortSpec.java:20-135 folded into one value type: a
// PortSpec is a ResourceSpec named "ports" whose value is the single range [port, port] (port 0 = pick dynamically).
struct ResourceSpec {
  std::string name;
  double value = 0.0;                 // scalar resources
  std::string role, preReservedRole = ANY_ROLE, principal;
  bool isPort = false;
  int port = 0;                       // PortSpec.getPort()
  std::string portName;
  std::vector<RangeSpec> ranges;      // PortSpec.getRanges(): constraint on a dynamic port
  std::vector<std::string> networkNames;

  static ResourceSpec scalar(std::string name, double value, std::string role, std::string preReservedRole, std::string principal) {
    ResourceSpec r;
    r.name = std::move(name); r.value = value; r.role = std::move(role); r.preReservedRole = std::move(preReservedRole);
    r.principal = std::move(principal);
    return r;
  }
  static ResourceSpec portSpec(std::string portName, int port, std::string role, std::string preReservedRole, std::string principal,
                               std::vector<RangeSpec> ranges = {}, std::vector<std::string> networkNames = {}) {
    ResourceSpec r;
    r.name = "ports"; r.isPort = true; r.port = port; r.portName = std::move(portName); r.role = std::move(role);
    r.preReservedRole = std::move(preReservedRole); r.principal = std::move(principal); r.ranges = std::move(ranges);
    r.networkNames = std::move(networkNames);
    return r;
  }
};

// S/specification/VolumeSpec.java:13-38 (a VolumeSpec is a "disk" ResourceSpec with a type, a path and profiles)
struct VolumeSpec {
  enum Type { ROOT, PATH, MOUNT };
  Type type = ROOT;
  double size = 0.0;
  std::string containerPath;
  std::vector<std::string> profiles;
  std::string role, preReservedRole = ANY_ROLE, principal;
};

// S/specification/ResourceSet.java:12-20
struct ResourceSet {
  std::string id;
  std::vector<ResourceSpec> resources;  // insertion order: cpus, gpus, mem, then ports (YAMLToInternalMappers.java:554-570)
  std::vector<VolumeSpec> volumes;
};

// S/specification/DefaultResourceSet.java builder
struct ResourceSetBuilder {
  ResourceSet rs;
  std::string role, preReservedRole, principal;
  ResourceSetBuilder(std::string role_, std::string preReservedRole_, std::string principal_)
      : role(std::move(role_)), preReservedRole(std::move(preReservedRole_)), principal(std::move(principal_)) {}
  ResourceSetBuilder& id(std::string v) { rs.id = std::move(v); return *this; }
  ResourceSetBuilder& scalar(const std::string& name, double v) { rs.resources.push_back(ResourceSpec::scalar(name, v, role, preReservedRole, principal)); return *this; }
  ResourceSetBuilder& cpus(double v) { return scalar("cpus", v); }
  ResourceSetBuilder& gpus(double v) { return scalar("gpus", v); }
  ResourceSetBuilder& memory(double v) { return scalar("mem", v); }
  ResourceSetBuilder& port(const std::string& name, int p, std::vector<RangeSpec> ranges = {}, std::vector<std::string> networks = {}) {
    rs.resources.push_back(ResourceSpec::portSpec(name, p, role, preReservedRole, principal, std::move(ranges), std::move(networks)));
    return *this;
  }
  ResourceSetBuilder& addVolume(VolumeSpec::Type t, double size, const std::string& path, std::vector<std::string> profiles = {}) {
    VolumeSpec v;
    v.type = t; v.size = size; v.containerPath = path; v.profiles = std::move(profiles);
    v.role = role; v.preReservedRole = preReservedRole; v.principal = principal;
    rs.volumes.push_back(v);
    return *this;
  }
  ResourceSetBuilder& addRootVolume(double size, const std::string& path) { return addVolume(VolumeSpec::ROOT, size, path); }
  ResourceSetBuilder& addMountVolume(double size, const std::string& path, std::vector<std::string> profiles = {}) {
    return addVolume(VolumeSpec::MOUNT, size, path, std::move(profiles));
  }
  ResourceSet build() const { return rs; }
};

// S/specification/TaskSpec.java (the fields the path reads)
struct TaskSpec {
  std::string name;
  ResourceSet resourceSet;
  bool usesTLS = false;  // getTransportEncryption() non-empty: adds an always-pass TLS stage (out of scope otherwise)
};

// S/specification/PodSpec.java:19-81
struct PodSpec {
  std::string type;
  int count = 1;
  std::vector<TaskSpec> tasks;
  std::optional<RulePtr> placementRule;
  std::vector<VolumeSpec> volumes;  // pod-level (executor) volumes
  std::string preReservedRole = ANY_ROLE;
};

// S/specification/PodInstance.java:8-32
struct PodInstance {
  std::shared_ptr<const PodSpec> pod;
  int index = 0;
  std::string getName() const { return pod->type + "-" + std::to_string(index); }
};
// S/offer/TaskUtils.java: task instance name = "<pod>-<index>-<task>"
inline std::string taskInstanceName(const PodInstance& pi, const std::string& taskSpecName) { return pi.getName() + "-" + taskSpecName; }

enum class RecoveryType { NONE, TRANSIENT, PERMANENT };  // S/scheduler/recovery/RecoveryType.java

// S/scheduler/plan/PodInstanceRequirement.java:47-69
struct PodInstanceRequirement {
  PodInstance podInstance;
  std::vector<std::string> tasksToLaunch;
  RecoveryType recoveryType = RecoveryType::NONE;
  static PodInstanceRequirement newBuilder(const PodInstance& pi, std::vector<std::string> tasks) {
    PodInstanceRequirement r;
    r.podInstance = pi; r.tasksToLaunch = std::move(tasks);
    return r;
  }
};

// S/scheduler/SchedulerConfig.java:160-170,358-367 — the default executor's overhead, a TreeMap (cpus, disk, mem)
struct SchedulerConfig {
  std::vector<std::pair<std::string, double>> executorResources{{"cpus", 0.1}, {"disk", 256.0}, {"mem", 32.0}};
  const std::vector<std::pair<std::string, double>>& getExecutorResources() const { return executorResources; }
};

}  // namespace sdk
