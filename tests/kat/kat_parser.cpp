// kat_parser.cpp — PT/MarathonConstraintParserTest.java:21-230 restated: the operator -> rule-tree goldens (compared through
// toString, exactly as the reference does) and the constraint-string splitting.
#include "kat.hpp"

using namespace sdk;
using S = std::string;
using Rows = std::vector<std::vector<std::string>>;

namespace {
const S POD_NAME = "hello";
S unescape(S s) { for (auto& c : s) if (c == '\'') c = '"'; return s; }
S parse(const S& c) { return MarathonConstraintParser::parse(POD_NAME, c)->toString(); }
void sameInAllThreeForms(const S& expected, const S& nested, const S& flat, const S& colon) {
  assertEquals(expected, parse(unescape(nested)));
  assertEquals(expected, parse(unescape(flat)));
  assertEquals(expected, parse(colon));
}
std::ostream& operator<<(std::ostream& o, const Rows& r) {
  o << "[";
  for (auto& row : r) { o << "["; for (auto& v : row) o << "'" << v << "',"; o << "]"; }
  return o << "]";
}
}  // namespace

KAT(MarathonConstraintParserTest_testSplitConstraints) {
  assertEquals((Rows{{"a", "b", "c"}}), MarathonConstraintParser::splitConstraints(unescape("['a', 'b', 'c']")));
  assertEquals((Rows{{"a", "b", "c"}}), MarathonConstraintParser::splitConstraints(unescape("[['a', 'b', 'c']]")));
  assertEquals((Rows{{"a", "b", "c"}, {"d", "e"}}), MarathonConstraintParser::splitConstraints(unescape("[['a', 'b', 'c'], ['d', 'e']]")));
  assertEquals((Rows{{"a"}, {}}), MarathonConstraintParser::splitConstraints(unescape("[['a'], []]")));
  assertEquals((Rows{{"a", "b", "c"}, {"d", "e"}, {"f", "g", ""}}), MarathonConstraintParser::splitConstraints("a:b:c,d:e,f:g:"));
  assertEquals((Rows{{"a", "b", "c"}, {"", "d", "e"}, {"", "f", ""}}), MarathonConstraintParser::splitConstraints("a:b:c,:d:e,:f:"));
  assertEquals((Rows{{"a", "b", "c"}, {"", "d", "e"}, {"", "f", ""}}), MarathonConstraintParser::splitConstraints(" a : b : c , : d : e , : f : "));
  assertEquals((Rows{{"", "", ""}, {"", ""}}), MarathonConstraintParser::splitConstraints("::,:"));
}
KAT(MarathonConstraintParserTest_testUniqueOperator) {
  sameInAllThreeForms("MaxPerHostnameRule{max=1, task-filter=RegexMatcher{pattern='hello-.*'}}", "[['hostname', 'UNIQUE']]", "['hostname', 'UNIQUE']", "hostname:UNIQUE");
  sameInAllThreeForms("AndRule{rules=[AttributeRule{matcher=RegexMatcher{pattern='rack-id:.*'}}, "
                      "MaxPerAttributeRule{max=1, matcher=RegexMatcher{pattern='rack-id:.*'}, task-filter=RegexMatcher{pattern='hello-.*'}}]}",
                      "[['rack-id', 'UNIQUE']]", "['rack-id', 'UNIQUE']", "rack-id:UNIQUE");
}
KAT(MarathonConstraintParserTest_testClusterOperator) {
  sameInAllThreeForms("AttributeRule{matcher=ExactMatcher{str='rack-id:rack-1'}}", "[['rack-id', 'CLUSTER', 'rack-1']]", "['rack-id', 'CLUSTER', 'rack-1']", "rack-id:CLUSTER:rack-1");
  sameInAllThreeForms("HostnameRule{matcher=ExactMatcher{str='a.specific.node.com'}}", "[['hostname', 'CLUSTER', 'a.specific.node.com']]",
                      "['hostname', 'CLUSTER', 'a.specific.node.com']", "hostname:CLUSTER:a.specific.node.com");
}
KAT(MarathonConstraintParserTest_testGroupByOperator) {
  sameInAllThreeForms("RoundRobinByAttributeRule{attribute=rack-id, attribute-count=Optional.empty, task-filter=RegexMatcher{pattern='hello-.*'}}",
                      "[['rack-id', 'GROUP_BY']]", "['rack-id', 'GROUP_BY']", "rack-id:GROUP_BY");
  sameInAllThreeForms("RoundRobinByAttributeRule{attribute=rack-id, attribute-count=Optional[3], task-filter=RegexMatcher{pattern='hello-.*'}}",
                      "[['rack-id', 'GROUP_BY', '3']]", "['rack-id', 'GROUP_BY', '3']", "rack-id:GROUP_BY:3");
  sameInAllThreeForms("RoundRobinByAttributeRule{attribute=zone, attribute-count=Optional[3], task-filter=RegexMatcher{pattern='hello-.*'}}",
                      "[['zone', 'GROUP_BY', '3']]", "['zone', 'GROUP_BY', '3']", "zone:GROUP_BY:3");
  sameInAllThreeForms("RoundRobinByHostnameRule{agent-count=Optional.empty, task-filter=RegexMatcher{pattern='hello-.*'}}", "[['hostname', 'GROUP_BY']]",
                      "['hostname', 'GROUP_BY']", "hostname:GROUP_BY");
  sameInAllThreeForms("RoundRobinByHostnameRule{agent-count=Optional[3], task-filter=RegexMatcher{pattern='hello-.*'}}", "[['hostname', 'GROUP_BY', '3']]",
                      "['hostname', 'GROUP_BY', '3']", "hostname:GROUP_BY:3");
}
KAT(MarathonConstraintParserTest_testLikeOperator) {
  sameInAllThreeForms("AttributeRule{matcher=RegexMatcher{pattern='rack-id:rack-[1-3]'}}", "[['rack-id', 'LIKE', 'rack-[1-3]']]", "['rack-id', 'LIKE', 'rack-[1-3]']",
                      "rack-id:LIKE:rack-[1-3]");
  sameInAllThreeForms("HostnameRule{matcher=RegexMatcher{pattern='rack-[1-3]'}}", "[['hostname', 'LIKE', 'rack-[1-3]']]", "['hostname', 'LIKE', 'rack-[1-3]']",
                      "hostname:LIKE:rack-[1-3]");
}
KAT(MarathonConstraintParserTest_testIsOperator) {
  sameInAllThreeForms("AttributeRule{matcher=ExactMatcher{str='foo:bar'}}", "[['foo', 'IS', 'bar']]", "['foo', 'IS', 'bar']", "foo:IS:bar");
  sameInAllThreeForms("RegionRule{matcher=ExactMatcher{str='bar'}}", "[['@region', 'IS', 'bar']]", "['@region', 'IS', 'bar']", "@region:IS:bar");
  sameInAllThreeForms("ZoneRule{matcher=ExactMatcher{str='bar'}}", "[['@zone', 'IS', 'bar']]", "['@zone', 'IS', 'bar']", "@zone:IS:bar");
  sameInAllThreeForms("HostnameRule{matcher=ExactMatcher{str='bar'}}", "[['@hostname', 'IS', 'bar']]", "['@hostname', 'IS', 'bar']", "@hostname:IS:bar");
}
KAT(MarathonConstraintParserTest_testUnlikeOperator) {
  sameInAllThreeForms("NotRule{rule=AttributeRule{matcher=RegexMatcher{pattern='rack-id:rack-[7-9]'}}}", "[['rack-id', 'UNLIKE', 'rack-[7-9]']]",
                      "['rack-id', 'UNLIKE', 'rack-[7-9]']", "rack-id:UNLIKE:rack-[7-9]");
  sameInAllThreeForms("NotRule{rule=HostnameRule{matcher=RegexMatcher{pattern='rack-[7-9]'}}}", "[['hostname', 'UNLIKE', 'rack-[7-9]']]",
                      "['hostname', 'UNLIKE', 'rack-[7-9]']", "hostname:UNLIKE:rack-[7-9]");
}
KAT(MarathonConstraintParserTest_testMaxPerOperator) {
  sameInAllThreeForms("AndRule{rules=[AttributeRule{matcher=RegexMatcher{pattern='rack-id:.*'}}, "
                      "MaxPerAttributeRule{max=2, matcher=RegexMatcher{pattern='rack-id:.*'}, task-filter=RegexMatcher{pattern='hello-.*'}}]}",
                      "[['rack-id', 'MAX_PER', '2']]", "['rack-id', 'MAX_PER', '2']", "rack-id:MAX_PER:2");
  sameInAllThreeForms("MaxPerHostnameRule{max=2, task-filter=RegexMatcher{pattern='hello-.*'}}", "[['hostname', 'MAX_PER', '2']]", "['hostname', 'MAX_PER', '2']",
                      "hostname:MAX_PER:2");
}
KAT(MarathonConstraintParserTest_testManyOperators) {
  const S expected =
      "AndRule{rules=[MaxPerHostnameRule{max=1, task-filter=RegexMatcher{pattern='hello-.*'}}, "
      "AttributeRule{matcher=ExactMatcher{str='rack-id:rack-1'}}, "
      "RoundRobinByAttributeRule{attribute=rack-id, attribute-count=Optional.empty, task-filter=RegexMatcher{pattern='hello-.*'}}, "
      "AttributeRule{matcher=RegexMatcher{pattern='rack-id:rack-[1-3]'}}, "
      "NotRule{rule=AttributeRule{matcher=RegexMatcher{pattern='rack-id:rack-[7-9]'}}}, "
      "AndRule{rules=[AttributeRule{matcher=RegexMatcher{pattern='rack-id:.*'}}, "
      "MaxPerAttributeRule{max=2, matcher=RegexMatcher{pattern='rack-id:.*'}, task-filter=RegexMatcher{pattern='hello-.*'}}]}]}";
  assertEquals(expected, parse(unescape("[['hostname', 'UNIQUE'], ['rack-id', 'CLUSTER', 'rack-1'], ['rack-id', 'GROUP_BY'], "
                                        "['rack-id', 'LIKE', 'rack-[1-3]'], ['rack-id', 'UNLIKE', 'rack-[7-9]'],['rack-id', 'MAX_PER', '2']]")));
  assertEquals(expected, parse("hostname:UNIQUE,rack-id:CLUSTER:rack-1,rack-id:GROUP_BY,rack-id:LIKE:rack-[1-3],rack-id:UNLIKE:rack-[7-9],rack-id:MAX_PER:2"));
}
KAT(MarathonConstraintParserTest_testEscapedCommaRegex) {
  assertEquals(S("AttributeRule{matcher=RegexMatcher{pattern='rack-id:rack-{1,3}'}}"), parse("rack-id:LIKE:rack-{1\\,3}"));
}
KAT(MarathonConstraintParserTest_emptyAndInvalid) {
  assertEquals(S("PassthroughRule{}"), parse(""));
  assertEquals(S("PassthroughRule{}"), parse("[]"));
  assertTrue(parse("hostname:FOO").rfind("InvalidPlacementRule{", 0) == 0);    // unknown operator
  assertTrue(parse("hostname").rfind("InvalidPlacementRule{", 0) == 0);        // too few entries
  assertTrue(parse("a:b:c:d").rfind("InvalidPlacementRule{", 0) == 0);         // too many
  assertTrue(parse("hostname:MAX_PER").rfind("InvalidPlacementRule{", 0) == 0);   // missing parameter
  assertTrue(parse("hostname:MAX_PER:x").rfind("InvalidPlacementRule{", 0) == 0); // not an integer
  assertEquals(S("MaxPerHostnameRule{max=1, task-filter=RegexMatcher{pattern='hello-.*'}}"), parse("hostname:unique"));  // operators ignore case
}
