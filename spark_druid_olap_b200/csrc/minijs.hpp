// A side-effect-free expression subset of JavaScript, enough for the functions the reference puts into
// JavascriptFilterSpec / JavascriptAggregationSpec when the predicate or the aggregate argument is plain string /
// arithmetic work:
//   docs/benchmark/druid/queries/q3.json, q5.json, q8.json      function(x) { return(x < '1995-03-15T08:00:00.000Z') }
//   src/test/.../FilterPlanningTest.scala:131-157               return(((s_name).toUpperCase())  ==  ("S1"));
//   src/test/resources/tpch_index_task.json.template:142-155    function(current, a, b) { return current + (a * b); }
//   jscodegen/JSCodeGenerator.scala:356-391,453-465             ((l) op (r)), Math.abs/floor/ceil/sqrt, -(x)
//   jscodegen/JSAggGenerator.scala:36-62                        current + x, Math.min(current, x), Math.max(current, x)
// Nothing here runs per row: a filter function is evaluated once per dictionary entry while the plan is compiled (the
// scan sees an id bitset, like a regex filter), and an aggregate argument is translated into a postfix program over
// metric columns that the scan kernel evaluates in registers.  Functions that need anything else (statements other
// than one `return`, `var`, `new`, Java interop such as org.joda...) are refused with SDO_UNSUPPORTED.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace sdo {
namespace minijs {

struct Value {
  enum Kind { Null, Bool, Num, Str } kind = Null;
  bool b = false;
  double n = 0;
  std::string s;
  static Value null() { return Value(); }
  static Value boolean(bool v) { Value x; x.kind = Bool; x.b = v; return x; }
  static Value number(double v) { Value x; x.kind = Num; x.n = v; return x; }
  static Value string(std::string v) { Value x; x.kind = Str; x.s = std::move(v); return x; }
};

struct Node {
  enum Kind { Literal, Param, Unary, Binary, Ternary, Method, Call } kind = Literal;
  Value lit;               // Literal
  int param = -1;          // Param: index into the function's parameter list
  std::string op;          // Unary: ! - + ; Binary: || && == != === !== < <= > >= + - * / % ; Method / Call: name
  std::vector<std::unique_ptr<Node>> kids;  // Unary 1, Binary 2, Ternary 3, Method receiver + args, Call args
  bool property = false;   // Method without call parentheses (.length)
};

struct Function {
  std::vector<std::string> params;
  std::unique_ptr<Node> body;  // the returned expression
};

// `function [name](p, ...) { return <expr>; }`; throws Unsupported outside the subset, Invalid on syntax errors
Function parse_function(const std::string& src);
// a bare expression; identifiers become parameters in order of first appearance (Druid's `expression` attribute)
Function parse_expression(const std::string& src);

Value eval(const Node& n, const std::vector<Value>& args);
bool truthy(const Value& v);
double to_number(const Value& v);
std::string to_string(const Value& v);

// ---- numeric programs for the scan kernel ---------------------------------------------------------------------------
// postfix over doubles; PUSH_PARAM's arg = parameter index, PUSH_CONST's arg = index into consts
enum NumOp : uint8_t { N_PUSH_PARAM = 0, N_PUSH_CONST = 1, N_ADD = 2, N_SUB = 3, N_MUL = 4, N_DIV = 5, N_NEG = 6, N_ABS = 7,
                       N_MIN = 8, N_MAX = 9, N_FLOOR = 10, N_CEIL = 11, N_SQRT = 12, N_MOD = 13 };
struct NumProgram {
  std::vector<uint8_t> code, arg;
  std::vector<double> consts;
  int max_depth = 0;
};
// throws Unsupported when the tree uses strings, comparisons or anything else that is not arithmetic on numbers
NumProgram compile_numeric(const Node& n);
bool uses_param(const Node& n, int param);

// What a JavascriptAggregationSpec's fnAggregate does with its first parameter (`current`):
//   return current + E            -> Sum
//   return Math.min(current, E)   -> Min        return Math.max(current, E) -> Max
// E must not mention `current`.  Returns the argument expression re-rooted (moved out of `f`).
enum class Fold { Sum, Min, Max };
std::unique_ptr<Node> split_fold(Function* f, Fold* kind);

}  // namespace minijs
}  // namespace sdo
