// mesos.hpp — plain C++ stand-ins for the org.apache.mesos.Protos messages that the offer-evaluation path reads.
//
// The reference consumes protobufs (org.apache.mesos:mesos:1.9.0, sdk/scheduler/build.gradle:166); protobuf is a data
// carrier only, so the host layer models just the fields the path touches.  Field names follow mesos.proto.
#pragma once
#include <cstdint>
#include <map>
#include <optional>
#include <string>
#include <utility>
#include <vector>

namespace mesos {

struct Value {
  enum Type { SCALAR, RANGES, SET, TEXT };
  Type type = SCALAR;
  double scalar = 0.0;
  std::vector<std::pair<int64_t, int64_t>> ranges;  // closed [begin, end]
  std::vector<std::string> set;
  std::string text;

  static Value Scalar(double v) { Value x; x.type = SCALAR; x.scalar = v; return x; }
  static Value Ranges(std::vector<std::pair<int64_t, int64_t>> r) { Value x; x.type = RANGES; x.ranges = std::move(r); return x; }
  static Value Set(std::vector<std::string> s) { Value x; x.type = SET; x.set = std::move(s); return x; }
  static Value Text(std::string t) { Value x; x.type = TEXT; x.text = std::move(t); return x; }
};

struct ReservationInfo {
  enum Type { STATIC, DYNAMIC };
  Type type = DYNAMIC;
  std::string role;
  std::string principal;
  std::map<std::string, std::string> labels;  // "resource_id", "framework_id", "namespace"
};

struct DiskInfo {
  enum SourceType { NONE, PATH, MOUNT };   // NONE = root disk (no Source message)
  SourceType source = NONE;
  std::optional<std::string> profile;      // Source.profile
  std::optional<std::string> persistence_id;
  std::string container_path;
};

struct Resource {
  std::string name;
  Value value;
  std::vector<ReservationInfo> reservations;       // refined-reservation stack; the last entry is the effective one
  std::optional<std::string> allocation_role;      // allocation_info.role
  std::optional<DiskInfo> disk;
};

struct Attribute {
  std::string name;
  Value value;
};

struct DomainInfo {
  std::string region, zone;  // fault_domain.region.name / fault_domain.zone.name
};

struct Offer {
  std::string id, framework_id, slave_id, hostname;
  std::vector<Resource> resources;
  std::vector<Attribute> attributes;
  std::vector<std::string> executor_ids;
  std::optional<DomainInfo> domain;
};

struct ExecutorInfo {
  std::string executor_id;
  std::vector<Resource> resources;
};

struct TaskInfo {
  std::string name, task_id, slave_id;
  std::map<std::string, std::string> labels;
  std::vector<Resource> resources;
  ExecutorInfo executor;
  std::map<std::string, int64_t> discovery_ports;  // discovery.ports: port name -> number (how a dynamic port is remembered)
};

}  // namespace mesos
