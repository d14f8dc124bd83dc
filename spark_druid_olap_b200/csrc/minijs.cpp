// Expression subset of JavaScript: tokenizer, recursive-descent parser, host evaluator, numeric compiler.  See minijs.hpp.
#include "minijs.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "timeutil.hpp"  // Unsupported / Invalid

namespace sdo {
namespace minijs {
namespace {

struct Token {
  enum Kind { End, Ident, Number, String, Punct } kind = End;
  std::string text;
  double num = 0;
};

struct Lexer {
  const std::string& s;
  size_t i = 0;
  explicit Lexer(const std::string& src) : s(src) {}
  [[noreturn]] void bad(const std::string& why) const { throw Invalid("javascript: " + why + " at offset " + std::to_string(i) + " of `" + s + "`"); }
  Token next() {
    for (;;) {
      while (i < s.size() && isspace((unsigned char)s[i])) i++;
      if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '/') { while (i < s.size() && s[i] != '\n') i++; continue; }
      if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '*') {
        size_t e = s.find("*/", i + 2);
        if (e == std::string::npos) bad("unterminated comment");
        i = e + 2;
        continue;
      }
      break;
    }
    Token t;
    if (i >= s.size()) return t;
    const char c = s[i];
    if (isalpha((unsigned char)c) || c == '_' || c == '$') {
      size_t b = i;
      while (i < s.size() && (isalnum((unsigned char)s[i]) || s[i] == '_' || s[i] == '$')) i++;
      t.kind = Token::Ident;
      t.text = s.substr(b, i - b);
      return t;
    }
    if (isdigit((unsigned char)c) || (c == '.' && i + 1 < s.size() && isdigit((unsigned char)s[i + 1]))) {
      const char* b = s.c_str() + i;
      char* e = nullptr;
      if (c == '0' && i + 1 < s.size() && (s[i + 1] == 'x' || s[i + 1] == 'X')) t.num = (double)strtoull(b, &e, 16);
      else t.num = strtod(b, &e);
      if (e == b) bad("bad number");
      i += (size_t)(e - b);
      t.kind = Token::Number;
      return t;
    }
    if (c == '\'' || c == '"') {
      i++;
      std::string out;
      for (;;) {
        if (i >= s.size()) bad("unterminated string");
        char d = s[i++];
        if (d == c) break;
        if (d != '\\') { out.push_back(d); continue; }
        if (i >= s.size()) bad("unterminated string");
        d = s[i++];
        switch (d) {
          case 'n': out.push_back('\n'); break;
          case 't': out.push_back('\t'); break;
          case 'r': out.push_back('\r'); break;
          case '0': out.push_back('\0'); break;
          case 'b': out.push_back('\b'); break;
          case 'f': out.push_back('\f'); break;
          case 'v': out.push_back('\v'); break;
          case 'u': {
            if (i + 4 > s.size()) bad("bad \\u escape");
            unsigned cp = (unsigned)strtoul(s.substr(i, 4).c_str(), nullptr, 16);
            i += 4;
            if (cp < 0x80) out.push_back((char)cp);
            else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 63))); }
            else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 63))); out.push_back((char)(0x80 | (cp & 63))); }
            break;
          }
          default: out.push_back(d);
        }
      }
      t.kind = Token::String;
      t.text = out;
      return t;
    }
    static const char* puncts[] = {"===", "!==", "==", "!=", "<=", ">=", "&&", "||", "(", ")", "{", "}", ",", ";", ".", "?", ":",
                                   "<", ">", "+", "-", "*", "/", "%", "!"};
    for (const char* p : puncts) {
      size_t n = strlen(p);
      if (s.compare(i, n, p) == 0) {
        i += n;
        t.kind = Token::Punct;
        t.text = p;
        return t;
      }
    }
    // =, ++, [, new ... : statements and constructs outside the subset
    throw Unsupported(std::string("javascript: character '") + c + "' is outside the supported expression subset");
  }
};

struct Parser {
  Lexer lex;
  Token tok;
  std::vector<std::string>* params;
  bool implicit_params;  // bare expressions: unknown identifiers become parameters
  Parser(const std::string& src, std::vector<std::string>* ps, bool implicit) : lex(src), params(ps), implicit_params(implicit) { advance(); }
  void advance() { tok = lex.next(); }
  bool is(const char* p) const { return tok.kind == Token::Punct && tok.text == p; }
  bool ident(const char* p) const { return tok.kind == Token::Ident && tok.text == p; }
  void expect(const char* p) {
    if (!is(p)) lex.bad(std::string("expected '") + p + "'");
    advance();
  }
  static std::unique_ptr<Node> mk(Node::Kind k) { std::unique_ptr<Node> n(new Node()); n->kind = k; return n; }
  static std::unique_ptr<Node> binary(const std::string& op, std::unique_ptr<Node> a, std::unique_ptr<Node> b) {
    auto n = mk(Node::Binary);
    n->op = op;
    n->kids.push_back(std::move(a));
    n->kids.push_back(std::move(b));
    return n;
  }

  std::unique_ptr<Node> expr() {
    auto c = logical_or();
    if (is("?")) {
      advance();
      auto a = expr();
      expect(":");
      auto b = expr();
      auto n = mk(Node::Ternary);
      n->kids.push_back(std::move(c));
      n->kids.push_back(std::move(a));
      n->kids.push_back(std::move(b));
      return n;
    }
    return c;
  }
  std::unique_ptr<Node> logical_or() {
    auto a = logical_and();
    while (is("||")) { advance(); a = binary("||", std::move(a), logical_and()); }
    return a;
  }
  std::unique_ptr<Node> logical_and() {
    auto a = equality();
    while (is("&&")) { advance(); a = binary("&&", std::move(a), equality()); }
    return a;
  }
  std::unique_ptr<Node> equality() {
    auto a = relational();
    while (is("==") || is("!=") || is("===") || is("!==")) {
      std::string op = tok.text;
      advance();
      a = binary(op, std::move(a), relational());
    }
    return a;
  }
  std::unique_ptr<Node> relational() {
    auto a = additive();
    while (is("<") || is("<=") || is(">") || is(">=")) {
      std::string op = tok.text;
      advance();
      a = binary(op, std::move(a), additive());
    }
    return a;
  }
  std::unique_ptr<Node> additive() {
    auto a = multiplicative();
    while (is("+") || is("-")) {
      std::string op = tok.text;
      advance();
      a = binary(op, std::move(a), multiplicative());
    }
    return a;
  }
  std::unique_ptr<Node> multiplicative() {
    auto a = unary();
    while (is("*") || is("/") || is("%")) {
      std::string op = tok.text;
      advance();
      a = binary(op, std::move(a), unary());
    }
    return a;
  }
  std::unique_ptr<Node> unary() {
    if (is("!") || is("-") || is("+")) {
      auto n = mk(Node::Unary);
      n->op = tok.text;
      advance();
      n->kids.push_back(unary());
      return n;
    }
    return postfix();
  }
  void arguments(Node* n) {
    expect("(");
    if (!is(")")) {
      for (;;) {
        n->kids.push_back(expr());
        if (is(",")) { advance(); continue; }
        break;
      }
    }
    expect(")");
  }
  std::unique_ptr<Node> postfix() {
    auto a = primary();
    while (is(".")) {
      advance();
      if (tok.kind != Token::Ident) lex.bad("expected a member name");
      auto n = mk(Node::Method);
      n->op = tok.text;
      advance();
      n->kids.push_back(std::move(a));
      if (is("(")) arguments(n.get());
      else n->property = true;
      a = std::move(n);
    }
    return a;
  }
  std::unique_ptr<Node> literal(Value v) {
    auto n = mk(Node::Literal);
    n->lit = std::move(v);
    return n;
  }
  std::unique_ptr<Node> primary() {
    if (tok.kind == Token::Number) { auto n = literal(Value::number(tok.num)); advance(); return n; }
    if (tok.kind == Token::String) { auto n = literal(Value::string(tok.text)); advance(); return n; }
    if (is("(")) {
      advance();
      auto e = expr();
      expect(")");
      return e;
    }
    if (tok.kind != Token::Ident) lex.bad("expected an expression");
    std::string id = tok.text;
    advance();
    if (id == "null" || id == "undefined") return literal(Value::null());
    if (id == "true") return literal(Value::boolean(true));
    if (id == "false") return literal(Value::boolean(false));
    if (id == "Infinity") return literal(Value::number(std::numeric_limits<double>::infinity()));
    if (id == "NaN") return literal(Value::number(std::numeric_limits<double>::quiet_NaN()));
    if (id == "new" || id == "var" || id == "function" || id == "typeof" || id == "if" || id == "for" || id == "while" || id == "let" || id == "const")
      throw Unsupported("javascript: '" + id + "' is outside the supported expression subset");
    auto it = std::find(params->begin(), params->end(), id);
    if (it != params->end()) {
      auto n = mk(Node::Param);
      n->param = (int)(it - params->begin());
      return n;
    }
    if (id == "Math" || id == "Number") {
      // Math.f(args), Number.POSITIVE_INFINITY, Number(x)
      if (id == "Number" && is("(")) {
        auto n = mk(Node::Call);
        n->op = "Number";
        arguments(n.get());
        return n;
      }
      expect(".");
      if (tok.kind != Token::Ident) lex.bad("expected a member name");
      std::string m = tok.text;
      advance();
      if (id == "Number") {
        if (m == "POSITIVE_INFINITY") return literal(Value::number(std::numeric_limits<double>::infinity()));
        if (m == "NEGATIVE_INFINITY") return literal(Value::number(-std::numeric_limits<double>::infinity()));
        if (m == "MAX_VALUE") return literal(Value::number(std::numeric_limits<double>::max()));
        if (m == "MIN_VALUE") return literal(Value::number(std::numeric_limits<double>::denorm_min()));
        if (m == "NaN") return literal(Value::number(std::numeric_limits<double>::quiet_NaN()));
        throw Unsupported("javascript: Number." + m);
      }
      if (!is("(")) {
        if (m == "PI") return literal(Value::number(3.141592653589793));
        if (m == "E") return literal(Value::number(2.718281828459045));
        throw Unsupported("javascript: Math." + m);
      }
      auto n = mk(Node::Call);
      n->op = "Math." + m;
      arguments(n.get());
      return n;
    }
    if (is("(")) {
      if (id != "parseFloat" && id != "parseInt" && id != "String" && id != "isNaN" && id != "Boolean")
        throw Unsupported("javascript: call of '" + id + "' is outside the supported expression subset");
      auto n = mk(Node::Call);
      n->op = id;
      arguments(n.get());
      return n;
    }
    if (implicit_params) {
      params->push_back(id);
      auto n = mk(Node::Param);
      n->param = (int)params->size() - 1;
      return n;
    }
    throw Unsupported("javascript: identifier '" + id + "' is neither a parameter nor a supported global");
  }
};

std::string number_to_string(double v) {
  if (std::isnan(v)) return "NaN";
  if (std::isinf(v)) return v > 0 ? "Infinity" : "-Infinity";
  if (v == 0) return "0";
  char buf[40];
  if (v == std::floor(v) && std::fabs(v) < 1e21) {
    snprintf(buf, sizeof buf, "%.0f", v);
    return buf;
  }
  // shortest representation that round-trips
  for (int prec = 1; prec <= 17; prec++) {
    snprintf(buf, sizeof buf, "%.*g", prec, v);
    if (strtod(buf, nullptr) == v) break;
  }
  std::string s = buf;
  // %g exponent form "1e-07" -> JS "1e-7"
  size_t e = s.find('e');
  if (e != std::string::npos) {
    std::string mant = s.substr(0, e), ex = s.substr(e + 1);
    bool neg = !ex.empty() && ex[0] == '-';
    size_t k = (ex[0] == '-' || ex[0] == '+') ? 1 : 0;
    while (k + 1 < ex.size() && ex[k] == '0') k++;
    s = mant + "e" + (neg ? "-" : "+") + ex.substr(k);
  }
  return s;
}

double string_to_number(const std::string& s) {
  size_t b = 0, e = s.size();
  while (b < e && isspace((unsigned char)s[b])) b++;
  while (e > b && isspace((unsigned char)s[e - 1])) e--;
  if (b == e) return 0.0;
  std::string t = s.substr(b, e - b);
  if (t == "Infinity" || t == "+Infinity") return std::numeric_limits<double>::infinity();
  if (t == "-Infinity") return -std::numeric_limits<double>::infinity();
  // JS accepts decimal literals and 0x integers; strtod also takes "inf"/"nan"/hex floats, which JS does not
  for (char c : t)
    if (!(isdigit((unsigned char)c) || c == '.' || c == 'e' || c == 'E' || c == '+' || c == '-' || c == 'x' || c == 'X' ||
          (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F')))
      return std::numeric_limits<double>::quiet_NaN();
  char* end = nullptr;
  double v;
  if (t.size() > 2 && t[0] == '0' && (t[1] == 'x' || t[1] == 'X')) v = (double)strtoull(t.c_str(), &end, 16);
  else {
    if (t.find_first_of("xX") != std::string::npos) return std::numeric_limits<double>::quiet_NaN();
    v = strtod(t.c_str(), &end);
  }
  if (!end || *end) return std::numeric_limits<double>::quiet_NaN();
  return v;
}

double parse_float_prefix(const std::string& s) {
  size_t b = 0;
  while (b < s.size() && isspace((unsigned char)s[b])) b++;
  const char* p = s.c_str() + b;
  // longest decimal-literal prefix
  size_t k = 0;
  if (p[k] == '+' || p[k] == '-') k++;
  if (strncmp(p + k, "Infinity", 8) == 0) return p[0] == '-' ? -std::numeric_limits<double>::infinity() : std::numeric_limits<double>::infinity();
  size_t digits = 0;
  while (isdigit((unsigned char)p[k])) { k++; digits++; }
  if (p[k] == '.') { k++; while (isdigit((unsigned char)p[k])) { k++; digits++; } }
  if (!digits) return std::numeric_limits<double>::quiet_NaN();
  if (p[k] == 'e' || p[k] == 'E') {
    size_t j = k + 1;
    if (p[j] == '+' || p[j] == '-') j++;
    if (isdigit((unsigned char)p[j])) { while (isdigit((unsigned char)p[j])) j++; k = j; }
  }
  return strtod(std::string(p, k).c_str(), nullptr);
}

bool loose_equal(const Value& a, const Value& b) {
  if (a.kind == b.kind) {
    switch (a.kind) {
      case Value::Null: return true;
      case Value::Bool: return a.b == b.b;
      case Value::Num: return a.n == b.n;
      case Value::Str: return a.s == b.s;
    }
  }
  if (a.kind == Value::Null || b.kind == Value::Null) return false;  // null equals only null / undefined
  return to_number(a) == to_number(b);
}
bool strict_equal(const Value& a, const Value& b) { return a.kind == b.kind && loose_equal(a, b); }

const std::string& need_string(const Value& v, const std::string& method) {
  if (v.kind != Value::Str) throw Invalid("javascript: TypeError: ." + method + " on a value that is not a string");
  return v.s;
}

}  // namespace

bool truthy(const Value& v) {
  switch (v.kind) {
    case Value::Null: return false;
    case Value::Bool: return v.b;
    case Value::Num: return v.n != 0 && !std::isnan(v.n);
    case Value::Str: return !v.s.empty();
  }
  return false;
}
double to_number(const Value& v) {
  switch (v.kind) {
    case Value::Null: return 0.0;
    case Value::Bool: return v.b ? 1.0 : 0.0;
    case Value::Num: return v.n;
    case Value::Str: return string_to_number(v.s);
  }
  return 0.0;
}
std::string to_string(const Value& v) {
  switch (v.kind) {
    case Value::Null: return "null";
    case Value::Bool: return v.b ? "true" : "false";
    case Value::Num: return number_to_string(v.n);
    case Value::Str: return v.s;
  }
  return "";
}

Function parse_function(const std::string& src) {
  Function f;
  Parser p(src, &f.params, false);
  if (!p.ident("function")) p.lex.bad("expected 'function'");
  p.advance();
  if (p.tok.kind == Token::Ident) p.advance();  // optional name
  p.expect("(");
  if (!p.is(")")) {
    for (;;) {
      if (p.tok.kind != Token::Ident) p.lex.bad("expected a parameter name");
      f.params.push_back(p.tok.text);
      p.advance();
      if (p.is(",")) { p.advance(); continue; }
      break;
    }
  }
  p.expect(")");
  p.expect("{");
  while (p.is(";")) p.advance();
  if (!p.ident("return")) {
    if (p.tok.kind == Token::Ident) throw Unsupported("javascript: statement '" + p.tok.text + "' (only a single return is supported)");
    p.lex.bad("expected 'return'");
  }
  p.advance();
  f.body = p.expr();
  while (p.is(";")) p.advance();
  p.expect("}");
  while (p.is(";")) p.advance();
  if (p.tok.kind != Token::End) p.lex.bad("trailing input after the function");
  return f;
}

Function parse_expression(const std::string& src) {
  Function f;
  Parser p(src, &f.params, true);
  f.body = p.expr();
  if (p.tok.kind != Token::End) p.lex.bad("trailing input after the expression");
  return f;
}

Value eval(const Node& n, const std::vector<Value>& args) {
  switch (n.kind) {
    case Node::Literal: return n.lit;
    case Node::Param: return n.param < (int)args.size() ? args[n.param] : Value::null();
    case Node::Unary: {
      Value v = eval(*n.kids[0], args);
      if (n.op == "!") return Value::boolean(!truthy(v));
      if (n.op == "-") return Value::number(-to_number(v));
      return Value::number(to_number(v));
    }
    case Node::Ternary: return truthy(eval(*n.kids[0], args)) ? eval(*n.kids[1], args) : eval(*n.kids[2], args);
    case Node::Binary: {
      if (n.op == "||") { Value a = eval(*n.kids[0], args); return truthy(a) ? a : eval(*n.kids[1], args); }
      if (n.op == "&&") { Value a = eval(*n.kids[0], args); return truthy(a) ? eval(*n.kids[1], args) : a; }
      Value a = eval(*n.kids[0], args), b = eval(*n.kids[1], args);
      if (n.op == "==") return Value::boolean(loose_equal(a, b));
      if (n.op == "!=") return Value::boolean(!loose_equal(a, b));
      if (n.op == "===") return Value::boolean(strict_equal(a, b));
      if (n.op == "!==") return Value::boolean(!strict_equal(a, b));
      if (n.op == "<" || n.op == "<=" || n.op == ">" || n.op == ">=") {
        if (a.kind == Value::Str && b.kind == Value::Str) {
          // code-unit order; equals UTF-8 byte order except between supplementary and high-BMP characters
          int c = a.s.compare(b.s);
          return Value::boolean(n.op == "<" ? c < 0 : n.op == "<=" ? c <= 0 : n.op == ">" ? c > 0 : c >= 0);
        }
        double x = to_number(a), y = to_number(b);
        return Value::boolean(n.op == "<" ? x < y : n.op == "<=" ? x <= y : n.op == ">" ? x > y : x >= y);
      }
      if (n.op == "+") {
        if (a.kind == Value::Str || b.kind == Value::Str) return Value::string(to_string(a) + to_string(b));
        return Value::number(to_number(a) + to_number(b));
      }
      double x = to_number(a), y = to_number(b);
      if (n.op == "-") return Value::number(x - y);
      if (n.op == "*") return Value::number(x * y);
      if (n.op == "/") return Value::number(x / y);
      if (n.op == "%") return Value::number(std::fmod(x, y));
      throw Unsupported("javascript: operator " + n.op);
    }
    case Node::Method: {
      Value r = eval(*n.kids[0], args);
      std::vector<Value> a;
      for (size_t i = 1; i < n.kids.size(); i++) a.push_back(eval(*n.kids[i], args));
      const std::string& m = n.op;
      if (n.property) {
        if (m == "length") return Value::number((double)need_string(r, m).size());
        throw Unsupported("javascript: property ." + m);
      }
      if (m == "toString") {
        if (r.kind == Value::Null) throw Invalid("javascript: TypeError: .toString on null");
        return Value::string(to_string(r));
      }
      const std::string& s = need_string(r, m);
      auto clampi = [&](double v, double lo, double hi) { return std::isnan(v) ? lo : std::min(std::max(std::trunc(v), lo), hi); };
      const double len = (double)s.size();
      if (m == "toUpperCase" || m == "toLowerCase") {
        std::string o = s;  // ASCII letters; other characters pass through unchanged
        for (char& c : o) c = (char)(m == "toUpperCase" ? (c >= 'a' && c <= 'z' ? c - 32 : c) : (c >= 'A' && c <= 'Z' ? c + 32 : c));
        return Value::string(o);
      }
      if (m == "concat") {
        std::string o = s;
        for (auto& v : a) o += to_string(v);
        return Value::string(o);
      }
      if (m == "trim") {
        size_t b = 0, e = s.size();
        while (b < e && isspace((unsigned char)s[b])) b++;
        while (e > b && isspace((unsigned char)s[e - 1])) e--;
        return Value::string(s.substr(b, e - b));
      }
      if (m == "substr") {
        double st = a.empty() ? 0 : to_number(a[0]);
        st = std::isnan(st) ? 0 : std::trunc(st);
        if (st < 0) st = std::max(0.0, len + st);
        st = std::min(st, len);
        double cnt = a.size() > 1 ? clampi(to_number(a[1]), 0, len - st) : len - st;
        return Value::string(s.substr((size_t)st, (size_t)cnt));
      }
      if (m == "substring" || m == "slice") {
        double b = a.empty() ? 0 : to_number(a[0]), e = a.size() > 1 && a[1].kind != Value::Null ? to_number(a[1]) : len;
        if (m == "slice") {
          b = std::isnan(b) ? 0 : std::trunc(b);
          e = std::isnan(e) ? 0 : std::trunc(e);
          if (b < 0) b = std::max(0.0, len + b);
          if (e < 0) e = std::max(0.0, len + e);
          b = std::min(b, len);
          e = std::min(e, len);
          return Value::string(e > b ? s.substr((size_t)b, (size_t)(e - b)) : std::string());
        }
        b = clampi(b, 0, len);
        e = clampi(e, 0, len);
        if (b > e) std::swap(b, e);
        return Value::string(s.substr((size_t)b, (size_t)(e - b)));
      }
      if (m == "charAt") {
        double k = a.empty() ? 0 : to_number(a[0]);
        k = std::isnan(k) ? 0 : std::trunc(k);
        return Value::string(k >= 0 && k < len ? s.substr((size_t)k, 1) : std::string());
      }
      if (m == "indexOf") {
        size_t pos = s.find(a.empty() ? std::string("undefined") : to_string(a[0]));
        return Value::number(pos == std::string::npos ? -1.0 : (double)pos);
      }
      if (m == "startsWith" || m == "endsWith" || m == "includes") {
        std::string t = a.empty() ? std::string("undefined") : to_string(a[0]);
        if (m == "includes") return Value::boolean(s.find(t) != std::string::npos);
        if (t.size() > s.size()) return Value::boolean(false);
        return Value::boolean(m == "startsWith" ? s.compare(0, t.size(), t) == 0 : s.compare(s.size() - t.size(), t.size(), t) == 0);
      }
      if (m == "replace") {
        // string pattern: first occurrence only
        if (a.size() < 2 || a[0].kind != Value::Str) throw Unsupported("javascript: .replace with a non-string pattern");
        std::string o = s;
        size_t pos = o.find(a[0].s);
        if (pos != std::string::npos) o.replace(pos, a[0].s.size(), to_string(a[1]));
        return Value::string(o);
      }
      throw Unsupported("javascript: method ." + m);
    }
    case Node::Call: {
      std::vector<Value> a;
      for (auto& k : n.kids) a.push_back(eval(*k, args));
      const std::string& f = n.op;
      auto num = [&](size_t i) { return i < a.size() ? to_number(a[i]) : std::numeric_limits<double>::quiet_NaN(); };
      if (f == "Number") return Value::number(a.empty() ? 0.0 : to_number(a[0]));
      if (f == "String") return Value::string(a.empty() ? std::string() : to_string(a[0]));
      if (f == "Boolean") return Value::boolean(!a.empty() && truthy(a[0]));
      if (f == "isNaN") return Value::boolean(std::isnan(num(0)));
      if (f == "parseFloat") return Value::number(a.empty() ? num(0) : (a[0].kind == Value::Num ? a[0].n : parse_float_prefix(to_string(a[0]))));
      if (f == "parseInt") {
        if (a.size() > 1 && to_number(a[1]) != 10 && to_number(a[1]) != 0) throw Unsupported("javascript: parseInt radix other than 10");
        double v = a.empty() ? num(0) : parse_float_prefix(to_string(a[0]));
        // parseInt stops at '.', 'e': re-parse the integer prefix
        if (!a.empty()) {
          std::string t = to_string(a[0]);
          size_t b = 0;
          while (b < t.size() && isspace((unsigned char)t[b])) b++;
          size_t k = b;
          if (k < t.size() && (t[k] == '+' || t[k] == '-')) k++;
          size_t d0 = k;
          while (k < t.size() && isdigit((unsigned char)t[k])) k++;
          v = k == d0 ? std::numeric_limits<double>::quiet_NaN() : strtod(t.substr(b, k - b).c_str(), nullptr);
        }
        return Value::number(v);
      }
      if (f == "Math.abs") return Value::number(std::fabs(num(0)));
      if (f == "Math.floor") return Value::number(std::floor(num(0)));
      if (f == "Math.ceil") return Value::number(std::ceil(num(0)));
      if (f == "Math.round") return Value::number(std::floor(num(0) + 0.5));
      if (f == "Math.sqrt") return Value::number(std::sqrt(num(0)));
      if (f == "Math.log") return Value::number(std::log(num(0)));
      if (f == "Math.exp") return Value::number(std::exp(num(0)));
      if (f == "Math.pow") return Value::number(std::pow(num(0), num(1)));
      if (f == "Math.min" || f == "Math.max") {
        double r = f == "Math.min" ? std::numeric_limits<double>::infinity() : -std::numeric_limits<double>::infinity();
        for (size_t i = 0; i < a.size(); i++) {
          double v = to_number(a[i]);
          if (std::isnan(v)) return Value::number(v);
          r = f == "Math.min" ? std::min(r, v) : std::max(r, v);
        }
        return Value::number(r);
      }
      throw Unsupported("javascript: " + f);
    }
  }
  return Value::null();
}

bool uses_param(const Node& n, int param) {
  if (n.kind == Node::Param) return n.param == param;
  for (auto& k : n.kids)
    if (uses_param(*k, param)) return true;
  return false;
}

namespace {
void emit_numeric(const Node& n, NumProgram* out, int* depth) {
  auto push = [&](uint8_t code, uint8_t arg, int delta) {
    out->code.push_back(code);
    out->arg.push_back(arg);
    *depth += delta;
    out->max_depth = std::max(out->max_depth, *depth);
  };
  switch (n.kind) {
    case Node::Literal: {
      if (n.lit.kind != Value::Num) throw Unsupported("javascript aggregate: only numbers are supported in the argument expression");
      size_t k = 0;
      for (; k < out->consts.size(); k++)
        if (memcmp(&out->consts[k], &n.lit.n, 8) == 0) break;
      if (k == out->consts.size()) out->consts.push_back(n.lit.n);
      push(N_PUSH_CONST, (uint8_t)k, 1);
      return;
    }
    case Node::Param: push(N_PUSH_PARAM, (uint8_t)n.param, 1); return;
    case Node::Unary:
      if (n.op == "!") break;
      emit_numeric(*n.kids[0], out, depth);
      if (n.op == "-") push(N_NEG, 0, 0);
      return;
    case Node::Binary: {
      uint8_t code;
      if (n.op == "+") code = N_ADD;
      else if (n.op == "-") code = N_SUB;
      else if (n.op == "*") code = N_MUL;
      else if (n.op == "/") code = N_DIV;
      else if (n.op == "%") code = N_MOD;
      else break;
      emit_numeric(*n.kids[0], out, depth);
      emit_numeric(*n.kids[1], out, depth);
      push(code, 0, -1);
      return;
    }
    case Node::Call: {
      struct { const char* name; uint8_t code; int arity; } fns[] = {
          {"Math.abs", N_ABS, 1}, {"Math.floor", N_FLOOR, 1}, {"Math.ceil", N_CEIL, 1}, {"Math.sqrt", N_SQRT, 1},
          {"Math.min", N_MIN, 2}, {"Math.max", N_MAX, 2}};
      if ((n.op == "Number" || n.op == "parseFloat") && n.kids.size() == 1) {
        emit_numeric(*n.kids[0], out, depth);  // metrics already are numbers
        return;
      }
      for (auto& f : fns)
        if (n.op == f.name && (int)n.kids.size() == f.arity) {
          for (auto& k : n.kids) emit_numeric(*k, out, depth);
          push(f.code, 0, f.arity == 2 ? -1 : 0);
          return;
        }
      break;
    }
    default: break;
  }
  throw Unsupported("javascript aggregate: the argument expression must be arithmetic (+ - * / %, Math.abs/floor/ceil/sqrt/min/max) over metrics and numbers");
}
}  // namespace

NumProgram compile_numeric(const Node& n) {
  NumProgram out;
  int depth = 0;
  emit_numeric(n, &out, &depth);
  return out;
}

std::unique_ptr<Node> split_fold(Function* f, Fold* kind) {
  Node& b = *f->body;
  if (f->params.empty()) throw Unsupported("javascript aggregate: fnAggregate has no `current` parameter");
  auto is_current = [](const Node& n) { return n.kind == Node::Param && n.param == 0; };
  std::unique_ptr<Node> arg;
  if (b.kind == Node::Binary && b.op == "+" && (is_current(*b.kids[0]) || is_current(*b.kids[1]))) {
    *kind = Fold::Sum;
    arg = std::move(b.kids[is_current(*b.kids[0]) ? 1 : 0]);
  } else if (b.kind == Node::Call && (b.op == "Math.min" || b.op == "Math.max") && b.kids.size() == 2 &&
             (is_current(*b.kids[0]) || is_current(*b.kids[1]))) {
    *kind = b.op == "Math.min" ? Fold::Min : Fold::Max;
    arg = std::move(b.kids[is_current(*b.kids[0]) ? 1 : 0]);
  } else {
    throw Unsupported("javascript aggregate: fnAggregate must be `current + e`, `Math.min(current, e)` or `Math.max(current, e)`");
  }
  if (uses_param(*arg, 0)) throw Unsupported("javascript aggregate: the argument expression refers to `current`");
  return arg;
}

}  // namespace minijs
}  // namespace sdo
