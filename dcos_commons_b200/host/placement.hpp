// placement.hpp — C++ mirror of com.mesosphere.sdk.offer.evaluate.placement (P/ = sdk/scheduler/src/main/java/com/
// mesosphere/sdk/offer/evaluate/placement/): StringMatchers, the PlacementRule tree and MarathonConstraintParser.
//
// The rules here are DESCRIPTIONS: they hold the same fields as the reference classes, print the same toString(), and
// know how to emit themselves as postfix rule nodes of include/oe_abi.h.  They never evaluate an offer on the host —
// PlacementRule::filter() in the reference becomes a bit of the feasibility matrix computed by the engine.  Only the
// StringMatchers run on the host, once per dictionary value (include/oe_abi.h `matcher_bits`).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <optional>
#include <regex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/oe_abi.h"
#include "mesos.hpp"

namespace sdk {

// ---- S/offer/taskdata/AttributeStringUtils.java ------------------------------------------------------------------------
struct AttributeStringUtils {
  static std::string join(const std::string& name, const std::string& value) { return name + ":" + value; }  // :41-43
  // :50-58 — split on the first ':'; throws when there is none
  static std::pair<std::string, std::string> split(const std::string& attribute) {
    const size_t k = attribute.find(':');
    if (k == std::string::npos) throw std::invalid_argument("Unable to split attribute into name:value elements: " + attribute);
    return {attribute.substr(0, k), attribute.substr(k + 1)};
  }
  // :28-35 — StringTokenizer on ';' (empty tokens are dropped)
  static std::vector<std::string> toStringList(const std::string& joined) {
    std::vector<std::string> out;
    size_t b = 0;
    while (b <= joined.size()) {
      size_t e = joined.find(';', b);
      if (e == std::string::npos) e = joined.size();
      if (e > b) out.push_back(joined.substr(b, e - b));
      b = e + 1;
    }
    return out;
  }
  // :107-145
  static std::string toString(const mesos::Value& v) {
    char buf[64];
    std::string out;
    switch (v.type) {
      case mesos::Value::RANGES:
        out = "[";
        for (size_t k = 0; k < v.ranges.size(); ++k) {
          std::snprintf(buf, sizeof buf, "%lld-%lld", (long long)v.ranges[k].first, (long long)v.ranges[k].second);
          out += (k ? "," : "") + std::string(buf);
        }
        return out + "]";
      case mesos::Value::SCALAR:
        std::snprintf(buf, sizeof buf, "%.3f", v.scalar);  // String.format(Locale.ROOT, "%.3f")
        return buf;
      case mesos::Value::SET:
        out = "{";
        for (size_t k = 0; k < v.set.size(); ++k) out += (k ? "," : "") + v.set[k];
        return out + "}";
      case mesos::Value::TEXT: return v.text;
    }
    return out;
  }
  static std::string toString(const mesos::Attribute& a) { return a.name + ":" + toString(a.value); }  // :81-84
  static std::string toString(const std::vector<mesos::Attribute>& attrs) {                            // :66-72
    std::string out;
    for (size_t k = 0; k < attrs.size(); ++k) out += (k ? ";" : "") + toString(attrs[k]);
    return out;
  }
};

// ---- matchers ---------------------------------------------------------------------------------------------------------
// P/StringMatcher.java:15
struct StringMatcher {
  virtual ~StringMatcher() = default;
  virtual bool matches(const std::string& value) const = 0;
  virtual std::string toString() const = 0;
};
using MatcherPtr = std::shared_ptr<const StringMatcher>;

// P/AnyMatcher.java:21
struct AnyMatcher : StringMatcher {
  static MatcherPtr create() { return std::make_shared<AnyMatcher>(); }
  bool matches(const std::string&) const override { return true; }
  std::string toString() const override { return "AnyMatcher{}"; }
};

// P/ExactMatcher.java:15-48.  Numbers are normalised through DecimalFormat("0.###") on both sides.
struct ExactMatcher : StringMatcher {
  std::string str;
  explicit ExactMatcher(const std::string& s) : str(normalise(s)) {}
  static MatcherPtr create(const std::string& s) { return std::make_shared<ExactMatcher>(s); }
  static MatcherPtr createAttribute(const std::string& name, const std::string& value) { return create(AttributeStringUtils::join(name, value)); }
  bool matches(const std::string& value) const override { return str == normalise(value); }
  std::string toString() const override { return "ExactMatcher{str='" + str + "'}"; }

  // Double.valueOf(s) succeeds -> DecimalFormat("0.###").format(d); otherwise the string itself.  Restated for the
  // inputs the path sees (decimal literals with optional exponent and an optional trailing d/D/f/F, NaN, Infinity);
  // hexadecimal floating literals are not recognised.
  static std::string normalise(const std::string& s) {
    std::string t = s;
    size_t b = 0, e = t.size();
    while (b < e && (unsigned char)t[b] <= ' ') ++b;   // Double.valueOf trims whitespace
    while (e > b && (unsigned char)t[e - 1] <= ' ') --e;
    t = t.substr(b, e - b);
    if (t.empty()) return s;
    std::string core = t;
    if (core.size() > 1 && (core.back() == 'd' || core.back() == 'D' || core.back() == 'f' || core.back() == 'F')) core.pop_back();
    std::string unsignedCore = core;
    bool neg = false;
    if (!unsignedCore.empty() && (unsignedCore[0] == '+' || unsignedCore[0] == '-')) { neg = unsignedCore[0] == '-'; unsignedCore = unsignedCore.substr(1); }
    if (unsignedCore == "NaN") return "\xEF\xBF\xBD";            // DecimalFormat prints U+FFFD for NaN
    if (unsignedCore == "Infinity") return std::string(neg ? "-" : "") + "\xE2\x88\x9E";  // U+221E
    // Java's grammar: digits [. digits] [e[+-]digits]; no hex, no "inf"/"nan" spellings
    bool digit = false;
    size_t k = 0;
    while (k < unsignedCore.size() && std::isdigit((unsigned char)unsignedCore[k])) { ++k; digit = true; }
    if (k < unsignedCore.size() && unsignedCore[k] == '.') { ++k; while (k < unsignedCore.size() && std::isdigit((unsigned char)unsignedCore[k])) { ++k; digit = true; } }
    if (!digit) return s;
    if (k < unsignedCore.size() && (unsignedCore[k] == 'e' || unsignedCore[k] == 'E')) {
      ++k;
      if (k < unsignedCore.size() && (unsignedCore[k] == '+' || unsignedCore[k] == '-')) ++k;
      bool ed = false;
      while (k < unsignedCore.size() && std::isdigit((unsigned char)unsignedCore[k])) { ++k; ed = true; }
      if (!ed) return s;
    }
    if (k != unsignedCore.size()) return s;
    const double d = std::strtod(core.c_str(), nullptr);
    char buf[400];
    std::snprintf(buf, sizeof buf, "%.3f", d);       // HALF_EVEN on the exact binary value, like DecimalFormat
    std::string out = buf;
    while (!out.empty() && out.back() == '0') out.pop_back();
    if (!out.empty() && out.back() == '.') out.pop_back();
    if (out == "-0") return "-0";
    return out;
  }
};

// P/RegexMatcher.java:14-33.  java.util.regex full-match; std::regex (ECMAScript) agrees on the subset the SDK's
// constraints use (literals, '.', classes, alternation, * + ? {m,n}); possessive quantifiers, look-behind and \p{..}
// classes are not available here.
struct RegexMatcher : StringMatcher {
  std::string pattern;
  std::regex re;
  explicit RegexMatcher(const std::string& p) : pattern(p), re(p, std::regex::ECMAScript) {}
  static MatcherPtr create(const std::string& p) { return std::make_shared<RegexMatcher>(p); }
  static MatcherPtr createAttribute(const std::string& name, const std::string& value) { return create(AttributeStringUtils::join(name, value)); }
  bool matches(const std::string& value) const override { return std::regex_match(value, re); }
  std::string toString() const override { return "RegexMatcher{pattern='" + pattern + "'}"; }
};

// ---- rule tree ---------------------------------------------------------------------------------------------------------
enum class PlacementField { HOSTNAME, ZONE, REGION, ATTRIBUTE };  // P/PlacementField.java

// What a rule needs from the encoder while it emits its postfix nodes (implemented in encoder.hpp).
struct RuleEmitter {
  virtual ~RuleEmitter() = default;
  virtual void node(int op, int a = 0, int b = 0, int c = 0) = 0;
  virtual int matcherOverAttributes(const StringMatcher& m) = 0;  // word offset into matcher_bits
  virtual int matcherOverHostnames(const StringMatcher& m) = 0;
  virtual int matcherOverZones(const StringMatcher& m) = 0;
  virtual int matcherOverRegions(const StringMatcher& m) = 0;
  virtual int counterSet(int keyspace, const StringMatcher& taskFilter, const std::string& rrAttributeName) = 0;
  virtual int typeSet(const std::string& podType) = 0;
  virtual int agentId(const std::string& slaveId) = 0;
  virtual int regionId(const std::string& region) = 0;
};
// Dictionary values a rule refers to by name must exist before the tables are laid out.
struct RuleDictionary {
  virtual ~RuleDictionary() = default;
  virtual void needAgent(const std::string&) = 0;
  virtual void needRegion(const std::string&) = 0;
  virtual void needPodType(const std::string&) = 0;
  virtual void needRoundRobinAttribute(const std::string& name) = 0;
  virtual void needTaskFilter(const StringMatcher&) = 0;
};

// P/PlacementRule.java:38-53.  filter() is what the engine computes; see the header comment.
struct PlacementRule {
  virtual ~PlacementRule() = default;
  virtual std::string toString() const = 0;
  virtual std::vector<PlacementField> getPlacementFields() const = 0;
  virtual void declare(RuleDictionary&) const {}
  virtual void emit(RuleEmitter&) const = 0;
};
using RulePtr = std::shared_ptr<const PlacementRule>;

inline std::string joinRules(const std::vector<RulePtr>& rules) {
  std::string s = "[";
  for (size_t k = 0; k < rules.size(); ++k) s += (k ? ", " : "") + rules[k]->toString();
  return s + "]";
}
inline std::vector<PlacementField> unionFields(const std::vector<RulePtr>& rules) {
  std::vector<PlacementField> out;
  for (auto& r : rules) for (auto f : r->getPlacementFields()) out.push_back(f);
  return out;
}

// P/PassthroughRule.java:25, P/InvalidPlacementRule.java:34
struct PassthroughRule : PlacementRule {
  std::string toString() const override { return "PassthroughRule{}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_TRUE); }
};
struct InvalidPlacementRule : PlacementRule {
  std::string constraints, exception;
  InvalidPlacementRule(std::string c, std::string x) : constraints(std::move(c)), exception(std::move(x)) {}
  std::string toString() const override { return "InvalidPlacementRule{constraints=" + constraints + ", exception=" + exception + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_FALSE); }
};

// P/AndRule.java:35-69 (all children evaluated; zero children fail), P/OrRule.java:36-63, P/NotRule.java:28-37
struct AndRule : PlacementRule {
  std::vector<RulePtr> rules;
  explicit AndRule(std::vector<RulePtr> r) : rules(std::move(r)) {}
  AndRule(RulePtr a, RulePtr b) : rules{std::move(a), std::move(b)} {}
  std::string toString() const override { return "AndRule{rules=" + joinRules(rules) + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return unionFields(rules); }
  void declare(RuleDictionary& d) const override { for (auto& r : rules) r->declare(d); }
  void emit(RuleEmitter& e) const override { for (auto& r : rules) r->emit(e); e.node(OE_R_AND, (int)rules.size()); }
};
struct OrRule : PlacementRule {
  std::vector<RulePtr> rules;
  explicit OrRule(std::vector<RulePtr> r) : rules(std::move(r)) {}
  std::string toString() const override { return "OrRule{rules=" + joinRules(rules) + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return unionFields(rules); }
  void declare(RuleDictionary& d) const override { for (auto& r : rules) r->declare(d); }
  void emit(RuleEmitter& e) const override { for (auto& r : rules) r->emit(e); e.node(OE_R_OR, (int)rules.size()); }
};
struct NotRule : PlacementRule {
  RulePtr rule;
  explicit NotRule(RulePtr r) : rule(std::move(r)) {}
  std::string toString() const override { return "NotRule{rule=" + rule->toString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return rule->getPlacementFields(); }
  void declare(RuleDictionary& d) const override { rule->declare(d); }
  void emit(RuleEmitter& e) const override { rule->emit(e); e.node(OE_R_NOT); }
};

// P/AgentRule.java:98-115 (+ require/avoid helpers :33-96)
struct AgentRule : PlacementRule {
  std::string agentId;
  explicit AgentRule(std::string id) : agentId(std::move(id)) {}
  std::string toString() const override { return "AgentRule{agentId=" + agentId + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {}; }
  void declare(RuleDictionary& d) const override { d.needAgent(agentId); }
  void emit(RuleEmitter& e) const override { e.node(OE_R_AGENT_EQ, e.agentId(agentId)); }
  static RulePtr require(const std::vector<std::string>& ids) {
    if (ids.size() == 1) return std::make_shared<AgentRule>(ids[0]);
    std::vector<RulePtr> rs;
    for (auto& id : ids) rs.push_back(std::make_shared<AgentRule>(id));
    return std::make_shared<OrRule>(rs);
  }
  static RulePtr avoid(const std::vector<std::string>& ids) { return std::make_shared<NotRule>(require(ids)); }
};

// P/StringMatcherRule.java:48-60 with P/AttributeRule.java:30-60, P/HostnameRule.java:26-52, P/ZoneRule.java:25-53,
// P/RegionRule.java:25-51
struct StringMatcherRule : PlacementRule {
  std::string name;
  MatcherPtr matcher;
  StringMatcherRule(std::string n, MatcherPtr m) : name(std::move(n)), matcher(std::move(m)) {}
  std::string toString() const override { return name + "{matcher=" + matcher->toString() + "}"; }
};
struct AttributeRule : StringMatcherRule {
  explicit AttributeRule(MatcherPtr m) : StringMatcherRule("AttributeRule", std::move(m)) {}
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ATTRIBUTE}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_ATTR_ANY, e.matcherOverAttributes(*matcher)); }
};
struct HostnameRule : StringMatcherRule {
  explicit HostnameRule(MatcherPtr m) : StringMatcherRule("HostnameRule", std::move(m)) {}
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::HOSTNAME}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_HOST_IN, e.matcherOverHostnames(*matcher)); }
};
struct ZoneRule : StringMatcherRule {
  explicit ZoneRule(MatcherPtr m) : StringMatcherRule("ZoneRule", std::move(m)) {}
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ZONE}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_ZONE_IN, e.matcherOverZones(*matcher)); }
};
struct RegionRule : StringMatcherRule {
  explicit RegionRule(MatcherPtr m) : StringMatcherRule("RegionRule", std::move(m)) {}
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::REGION}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_REGION_IN, e.matcherOverRegions(*matcher)); }
};

// P/RuleFactory.java + the four *RuleFactory singletons: require(matcher) / avoid(matcher) = Not(require)
template <class R>
struct RuleFactoryOf {
  static RulePtr require(MatcherPtr m) { return std::make_shared<R>(std::move(m)); }
  static RulePtr avoid(MatcherPtr m) { return std::make_shared<NotRule>(require(std::move(m))); }
};
using AttributeRuleFactory = RuleFactoryOf<AttributeRule>;
using HostnameRuleFactory = RuleFactoryOf<HostnameRule>;
using ZoneRuleFactory = RuleFactoryOf<ZoneRule>;
using RegionRuleFactory = RuleFactoryOf<RegionRule>;

// P/IsLocalRegionRule.java:24-67.  The local domain is process-wide state set at registration
// (S/framework/FrameworkScheduler.java:109).
struct IsLocalRegionRule : PlacementRule {
  static std::optional<mesos::DomainInfo>& localDomain() { static std::optional<mesos::DomainInfo> d; return d; }
  static void setLocalDomain(const mesos::DomainInfo& d) { localDomain() = d; }
  static void clearLocalDomain() { localDomain().reset(); }
  std::string toString() const override { return "IsLocalRegionRule{}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::REGION}; }
  void declare(RuleDictionary& d) const override { if (localDomain()) d.needRegion(localDomain()->region); }
  void emit(RuleEmitter& e) const override {
    if (localDomain()) e.node(OE_R_LOCAL_REGION, e.regionId(localDomain()->region), 1);
    else e.node(OE_R_LOCAL_REGION, 0, 0);
  }
};

// P/MaxPerRule.java:21-87 and its four key providers
struct MaxPerRule : PlacementRule {
  int max;
  MatcherPtr taskFilter;  // may be null, as in the reference's one-argument constructors
  MaxPerRule(int m, MatcherPtr f) : max(m), taskFilter(std::move(f)) {
    if (max < 1) throw std::invalid_argument("max must be at least 1");  // ValidationUtils.atLeastOne, MaxPerRule.java:35
  }
  const StringMatcher& filterOrThrow() const {
    // P/MaxPerRule.java:60 dereferences the filter for every task: a null filter is a NullPointerException there
    if (!taskFilter) throw std::logic_error("MaxPerRule without a task filter (NullPointerException in the reference)");
    return *taskFilter;
  }
  void declare(RuleDictionary& d) const override { if (taskFilter) d.needTaskFilter(*taskFilter); }
  std::string filterString() const { return taskFilter ? taskFilter->toString() : "null"; }
};
struct MaxPerHostnameRule : MaxPerRule {
  explicit MaxPerHostnameRule(int m, MatcherPtr f = nullptr) : MaxPerRule(m, std::move(f)) {}
  std::string toString() const override { return "MaxPerHostnameRule{max=" + std::to_string(max) + ", task-filter=" + filterString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::HOSTNAME}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_MAXPER_HOST, e.counterSet(OE_KS_HOSTNAME, filterOrThrow(), ""), max); }
};
struct MaxPerZoneRule : MaxPerRule {
  explicit MaxPerZoneRule(int m, MatcherPtr f = nullptr) : MaxPerRule(m, std::move(f)) {}
  std::string toString() const override { return "MaxPerZoneRule{max=" + std::to_string(max) + ", task-filter=" + filterString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ZONE}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_MAXPER_ZONE, e.counterSet(OE_KS_ZONE, filterOrThrow(), ""), max); }
};
struct MaxPerRegionRule : MaxPerRule {
  explicit MaxPerRegionRule(int m, MatcherPtr f = nullptr) : MaxPerRule(m, std::move(f)) {}
  std::string toString() const override { return "MaxPerRegionRule{max=" + std::to_string(max) + ", task-filter=" + filterString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::REGION}; }
  void emit(RuleEmitter& e) const override { e.node(OE_R_MAXPER_REGION, e.counterSet(OE_KS_REGION, filterOrThrow(), ""), max); }
};
// P/MaxPerAttributeRule.java:69-153: a missing task filter becomes AnyMatcher (:96-100)
struct MaxPerAttributeRule : MaxPerRule {
  MatcherPtr attributeMatcher;
  MaxPerAttributeRule(int m, MatcherPtr am, MatcherPtr f = nullptr)
      : MaxPerRule(m, f ? std::move(f) : AnyMatcher::create()), attributeMatcher(std::move(am)) {}
  std::string toString() const override {
    return "MaxPerAttributeRule{max=" + std::to_string(max) + ", matcher=" + attributeMatcher->toString() + ", task-filter=" + filterString() + "}";
  }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ATTRIBUTE}; }
  void emit(RuleEmitter& e) const override {
    e.node(OE_R_MAXPER_ATTR, e.counterSet(OE_KS_ATTR, filterOrThrow(), ""), max, e.matcherOverAttributes(*attributeMatcher));
  }
};

// P/AbstractRoundRobinRule.java:30-151 and its four key providers; a missing task filter becomes AnyMatcher (:32-35)
struct AbstractRoundRobinRule : PlacementRule {
  std::optional<int> distinctKeyCount;
  MatcherPtr taskFilter;
  AbstractRoundRobinRule(MatcherPtr f, std::optional<int> n) : distinctKeyCount(n), taskFilter(f ? std::move(f) : AnyMatcher::create()) {}
  std::string countString() const { return distinctKeyCount ? "Optional[" + std::to_string(*distinctKeyCount) + "]" : "Optional.empty"; }
  void declare(RuleDictionary& d) const override { d.needTaskFilter(*taskFilter); }
  void emitWith(RuleEmitter& e, int keyspace, const std::string& attr) const {
    e.node(OE_R_RR, e.counterSet(keyspace, *taskFilter, attr), distinctKeyCount ? *distinctKeyCount : -1);
  }
};
struct RoundRobinByHostnameRule : AbstractRoundRobinRule {
  explicit RoundRobinByHostnameRule(std::optional<int> n, MatcherPtr f = nullptr) : AbstractRoundRobinRule(std::move(f), n) {}
  std::string toString() const override { return "RoundRobinByHostnameRule{agent-count=" + countString() + ", task-filter=" + taskFilter->toString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::HOSTNAME}; }
  void emit(RuleEmitter& e) const override { emitWith(e, OE_KS_HOSTNAME, ""); }
};
struct RoundRobinByZoneRule : AbstractRoundRobinRule {
  explicit RoundRobinByZoneRule(std::optional<int> n, MatcherPtr f = nullptr) : AbstractRoundRobinRule(std::move(f), n) {}
  std::string toString() const override { return "RoundRobinByZoneRule{zone-count=" + countString() + ", task-filter=" + taskFilter->toString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ZONE}; }
  void emit(RuleEmitter& e) const override { emitWith(e, OE_KS_ZONE, ""); }
};
struct RoundRobinByRegionRule : AbstractRoundRobinRule {
  explicit RoundRobinByRegionRule(std::optional<int> n, MatcherPtr f = nullptr) : AbstractRoundRobinRule(std::move(f), n) {}
  std::string toString() const override { return "RoundRobinByRegionRule{region-count=" + countString() + ", task-filter=" + taskFilter->toString() + "}"; }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::REGION}; }
  void emit(RuleEmitter& e) const override { emitWith(e, OE_KS_REGION, ""); }
};
struct RoundRobinByAttributeRule : AbstractRoundRobinRule {
  std::string attributeName;
  RoundRobinByAttributeRule(std::string name, std::optional<int> n, MatcherPtr f = nullptr)
      : AbstractRoundRobinRule(std::move(f), n), attributeName(std::move(name)) {}
  std::string toString() const override {
    return "RoundRobinByAttributeRule{attribute=" + attributeName + ", attribute-count=" + countString() + ", task-filter=" + taskFilter->toString() + "}";
  }
  std::vector<PlacementField> getPlacementFields() const override { return {PlacementField::ATTRIBUTE}; }
  void declare(RuleDictionary& d) const override { AbstractRoundRobinRule::declare(d); d.needRoundRobinAttribute(attributeName); }
  void emit(RuleEmitter& e) const override { emitWith(e, OE_KS_RR_ATTR, attributeName); }
};

// P/TaskTypeRule.java:36-217 (type read from the task_type label, P/TaskTypeLabelConverter.java)
struct TaskTypeRule : PlacementRule {
  enum BehaviorType { COLOCATE, AVOID };
  std::string typeToFind;
  BehaviorType behavior;
  TaskTypeRule(std::string t, BehaviorType b) : typeToFind(std::move(t)), behavior(b) {}
  static RulePtr avoid(const std::string& t) { return std::make_shared<TaskTypeRule>(t, AVOID); }
  static RulePtr colocateWith(const std::string& t) { return std::make_shared<TaskTypeRule>(t, COLOCATE); }
  std::string toString() const override {
    return "TaskTypeRule{type=" + typeToFind + ", converter=TaskTypeLabelConverter{}, behavior=" + (behavior == AVOID ? "AVOID" : "COLOCATE") + "}";
  }
  std::vector<PlacementField> getPlacementFields() const override { return {}; }
  void declare(RuleDictionary& d) const override { d.needPodType(typeToFind); }
  void emit(RuleEmitter& e) const override { e.node(behavior == AVOID ? OE_R_TYPE_AVOID : OE_R_TYPE_COLOCATE, e.typeSet(typeToFind)); }
};

// ---- P/PlacementUtils.java ---------------------------------------------------------------------------------------------
struct PlacementUtils {
  static PlacementField getField(const std::string& fieldName) {  // :95-108
    if (fieldName == "hostname" || fieldName == "@hostname") return PlacementField::HOSTNAME;
    if (fieldName == "@region") return PlacementField::REGION;
    if (fieldName == "@zone") return PlacementField::ZONE;
    return PlacementField::ATTRIBUTE;
  }
  // :37-62
  static std::optional<RulePtr> getAgentPlacementRule(const std::vector<std::string>& avoidAgents, const std::vector<std::string>& collocateAgents) {
    if (!avoidAgents.empty()) {
      if (!collocateAgents.empty()) return RulePtr(std::make_shared<AndRule>(AgentRule::avoid(avoidAgents), AgentRule::require(collocateAgents)));
      return AgentRule::avoid(avoidAgents);
    }
    if (!collocateAgents.empty()) return AgentRule::require(collocateAgents);
    return std::nullopt;
  }
  // :126-141 — does the pod's rule tree read the offer's region (zone)?  SchedulerBuilder asks before it injects an
  // IsLocalRegionRule (S/scheduler/SchedulerBuilder.java: region-aware placement); takes PodSpec::placementRule.
  static bool placementRuleReferencesField(PlacementField field, const std::optional<RulePtr>& rule) {
    if (!rule || !*rule) return false;
    for (PlacementField f : (*rule)->getPlacementFields()) if (f == field) return true;
    return false;
  }
  static bool placementRuleReferencesRegion(const std::optional<RulePtr>& rule) { return placementRuleReferencesField(PlacementField::REGION, rule); }
  static bool placementRuleReferencesZone(const std::optional<RulePtr>& rule) { return placementRuleReferencesField(PlacementField::ZONE, rule); }
};

// ---- P/MarathonConstraintParser.java -----------------------------------------------------------------------------------
struct MarathonConstraintParser {
  struct ParseError : std::runtime_error { using std::runtime_error::runtime_error; };  // the reference's IOException

  // :61-87
  static RulePtr parse(const std::string& podName, const std::string& marathonConstraints) {
    if (marathonConstraints.empty() || marathonConstraints == "[]") return std::make_shared<PassthroughRule>();
    try {
      const auto rows = splitConstraints(marathonConstraints);
      MatcherPtr taskFilter = RegexMatcher::create(podName + "-.*");
      if (rows.size() == 1) return parseRow(taskFilter, rows[0]);
      std::vector<RulePtr> rowRules;
      for (auto& row : rows) rowRules.push_back(parseRow(taskFilter, row));
      return std::make_shared<AndRule>(rowRules);
    } catch (const ParseError& e) {
      return std::make_shared<InvalidPlacementRule>(marathonConstraints, e.what());
    }
  }

  // :129-170 — JSON flat list, JSON nested list, else "a:b:c,d:e"
  static std::vector<std::vector<std::string>> splitConstraints(const std::string& s) {
    std::vector<std::string> flat;
    if (parseJsonStringList(s, &flat)) return {flat};
    std::vector<std::vector<std::string>> nested;
    if (parseJsonNestedList(s, &nested)) return nested;
    std::vector<std::vector<std::string>> rows;
    for (auto& rowStr : escapedSplit(s, ',')) rows.push_back(escapedSplit(rowStr, ':'));
    return rows;
  }

  // :179-215
  static std::vector<std::string> escapedSplit(const std::string& str, char split) {
    std::vector<std::string> vals;
    std::string buf;
    bool escaped = false;
    for (char c : str) {
      if (escaped) {
        if (c == split) buf.push_back(c);
        else { buf.push_back('\\'); buf.push_back(c); }
        escaped = false;
      } else if (c == '\\') {
        escaped = true;
      } else if (c == split) {
        vals.push_back(trim(buf));
        buf.clear();
      } else {
        buf.push_back(c);
      }
    }
    if (escaped) buf.push_back('\\');
    vals.push_back(trim(buf));
    return vals;
  }

 private:
  static std::string trim(const std::string& s) {
    size_t b = 0, e = s.size();
    while (b < e && (unsigned char)s[b] <= ' ') ++b;
    while (e > b && (unsigned char)s[e - 1] <= ' ') --e;
    return s.substr(b, e - b);
  }
  static void skipWs(const std::string& s, size_t& k) { while (k < s.size() && std::isspace((unsigned char)s[k])) ++k; }
  static bool jsonString(const std::string& s, size_t& k, std::string* out) {
    if (k >= s.size() || s[k] != '"') return false;
    ++k;
    out->clear();
    while (k < s.size() && s[k] != '"') {
      if (s[k] == '\\' && k + 1 < s.size()) {
        const char c = s[k + 1];
        switch (c) {
          case 'n': out->push_back('\n'); break;
          case 't': out->push_back('\t'); break;
          case 'r': out->push_back('\r'); break;
          case 'b': out->push_back('\b'); break;
          case 'f': out->push_back('\f'); break;
          default: out->push_back(c);
        }
        k += 2;
      } else {
        out->push_back(s[k++]);
      }
    }
    if (k >= s.size()) return false;
    ++k;
    return true;
  }
  static bool jsonList(const std::string& s, size_t& k, std::vector<std::string>* out) {
    skipWs(s, k);
    if (k >= s.size() || s[k] != '[') return false;
    ++k;
    out->clear();
    skipWs(s, k);
    if (k < s.size() && s[k] == ']') { ++k; return true; }
    for (;;) {
      skipWs(s, k);
      std::string v;
      if (!jsonString(s, k, &v)) return false;
      out->push_back(v);
      skipWs(s, k);
      if (k < s.size() && s[k] == ',') { ++k; continue; }
      if (k < s.size() && s[k] == ']') { ++k; return true; }
      return false;
    }
  }
  static bool parseJsonStringList(const std::string& s, std::vector<std::string>* out) {
    size_t k = 0;
    if (!jsonList(s, k, out)) return false;
    skipWs(s, k);
    return k == s.size();  // FAIL_ON_TRAILING_TOKENS
  }
  static bool parseJsonNestedList(const std::string& s, std::vector<std::vector<std::string>>* out) {
    size_t k = 0;
    skipWs(s, k);
    if (k >= s.size() || s[k] != '[') return false;
    ++k;
    out->clear();
    skipWs(s, k);
    if (k < s.size() && s[k] == ']') { ++k; skipWs(s, k); return k == s.size(); }
    for (;;) {
      std::vector<std::string> row;
      if (!jsonList(s, k, &row)) return false;
      out->push_back(row);
      skipWs(s, k);
      if (k < s.size() && s[k] == ',') { ++k; continue; }
      if (k < s.size() && s[k] == ']') { ++k; break; }
      return false;
    }
    skipWs(s, k);
    return k == s.size();
  }

  static bool equalsIgnoreCase(const std::string& a, const char* b) {
    size_t n = 0;
    for (; b[n]; ++n) if (n >= a.size() || std::toupper((unsigned char)a[n]) != b[n]) return false;
    return n == a.size();
  }
  static std::string required(const std::string& op, const std::optional<std::string>& p) {  // :217-227
    if (!p) throw ParseError("Missing required parameter for operator '" + op + "'.");
    return *p;
  }
  static int parseInt(const std::string& op, const std::string& v) {
    // Integer.parseInt: optional sign, decimal digits only
    size_t k = (!v.empty() && (v[0] == '+' || v[0] == '-')) ? 1 : 0;
    if (k == v.size()) throw ParseError("Unable to parse max parameter as integer for '" + op + "' operation: " + v);
    for (size_t q = k; q < v.size(); ++q)
      if (!std::isdigit((unsigned char)v[q])) throw ParseError("Unable to parse max parameter as integer for '" + op + "' operation: " + v);
    return std::atoi(v.c_str());
  }

  // :99-121 and the operator classes :245-502
  static RulePtr parseRow(const MatcherPtr& taskFilter, const std::vector<std::string>& row) {
    if (row.size() < 2 || row.size() > 3) throw ParseError("Invalid number of entries in rule. Expected 2 or 3, got " + std::to_string(row.size()));
    const std::string& fieldName = row[0];
    const std::string& op = row[1];
    std::optional<std::string> parameter;
    if (row.size() >= 3) parameter = row[2];
    const PlacementField field = PlacementUtils::getField(fieldName);
    auto matcherRule = [&](MatcherPtr plain, MatcherPtr attr, bool avoid) -> RulePtr {
      switch (field) {
        case PlacementField::HOSTNAME: return avoid ? HostnameRuleFactory::avoid(plain) : HostnameRuleFactory::require(plain);
        case PlacementField::ZONE: return avoid ? ZoneRuleFactory::avoid(plain) : ZoneRuleFactory::require(plain);
        case PlacementField::REGION: return avoid ? RegionRuleFactory::avoid(plain) : RegionRuleFactory::require(plain);
        default: return avoid ? AttributeRuleFactory::avoid(attr) : AttributeRuleFactory::require(attr);
      }
    };
    auto maxPer = [&](int max) -> RulePtr {
      switch (field) {
        case PlacementField::HOSTNAME: return std::make_shared<MaxPerHostnameRule>(max, taskFilter);
        case PlacementField::ZONE: return std::make_shared<MaxPerZoneRule>(max, taskFilter);
        case PlacementField::REGION: return std::make_shared<MaxPerRegionRule>(max, taskFilter);
        default: {
          MatcherPtr m = RegexMatcher::createAttribute(fieldName, ".*");
          return std::make_shared<AndRule>(AttributeRuleFactory::require(m), std::make_shared<MaxPerAttributeRule>(max, m, taskFilter));
        }
      }
    };
    if (equalsIgnoreCase(op, "UNIQUE")) return maxPer(1);
    if (equalsIgnoreCase(op, "CLUSTER") || equalsIgnoreCase(op, "IS")) {
      const std::string p = required(op, parameter);
      return matcherRule(ExactMatcher::create(p), ExactMatcher::createAttribute(fieldName, p), false);
    }
    if (equalsIgnoreCase(op, "GROUP_BY")) {
      std::optional<int> num;
      if (parameter) num = parseInt(op, *parameter);
      switch (field) {
        case PlacementField::HOSTNAME: return std::make_shared<RoundRobinByHostnameRule>(num, taskFilter);
        case PlacementField::ZONE: return std::make_shared<RoundRobinByZoneRule>(num, taskFilter);
        case PlacementField::REGION: return std::make_shared<RoundRobinByRegionRule>(num, taskFilter);
        default: return std::make_shared<RoundRobinByAttributeRule>(fieldName, num, taskFilter);
      }
    }
    if (equalsIgnoreCase(op, "LIKE") || equalsIgnoreCase(op, "UNLIKE")) {
      const std::string p = required(op, parameter);
      return matcherRule(RegexMatcher::create(p), RegexMatcher::createAttribute(fieldName, p), equalsIgnoreCase(op, "UNLIKE"));
    }
    if (equalsIgnoreCase(op, "MAX_PER")) return maxPer(parseInt(op, required(op, parameter)));
    throw ParseError("Unsupported operator: '" + op + "' (expected one of: UNIQUE, CLUSTER, GROUP_BY, LIKE, UNLIKE, or MAX_PER)");
  }
};

}  // namespace sdk
