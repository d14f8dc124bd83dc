"""TEST INFRASTRUCTURE ONLY (see oracle/druid_oracle.py's header): an independent Python reading of the JavaScript
expression subset the engine accepts in JavascriptFilterSpec / JavascriptAggregationSpec.

What the subset has to cover is set by the reference's own uses:
  docs/benchmark/druid/queries/q3.json, q5.json, q8.json    function(x) { return(x < '1995-03-15T08:00:00.000Z') }
  src/test/.../FilterPlanningTest.scala:131-157             return(((s_name).toUpperCase())  ==  ("S1"));
  src/test/resources/tpch_index_task.json.template:142-155  function(current, a, b) { return current + (a * b); }
  jscodegen/JSCodeGenerator.scala:356-391,453-465           ((l) op (r)), Math.abs/floor/ceil/sqrt, -(x)
  jscodegen/JSAggGenerator.scala:36-62                      current + x, Math.min(current, x), Math.max(current, x)
The language semantics followed are ECMAScript's (ToNumber / ToString / ToBoolean, abstract equality and relational
comparison); parity unpinned: the reference has no test that asserts a value computed by one of these functions.

The evaluator works on Python scalars (None = null, bool, float, str) for filters and on numpy float64 arrays for the
argument of an aggregate."""
import math
import re

import numpy as np

_TOKEN = re.compile(r"""\s*(?:
    (?P<num>0[xX][0-9a-fA-F]+|(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?)
  | (?P<id>[A-Za-z_$][A-Za-z0-9_$]*)
  | (?P<str>'(?:[^'\\]|\\.)*'|"(?:[^"\\]|\\.)*")
  | (?P<op>===|!==|==|!=|<=|>=|&&|\|\||[(){},;.?:<>+\-*/%!])
)""", re.X)

_ESC = {"n": "\n", "t": "\t", "r": "\r", "0": "\0", "b": "\b", "f": "\f", "v": "\v"}


class Unsupported(Exception):
    pass


def _tokens(src):
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    pos, out = 0, []
    while True:
        if not src[pos:].strip():
            return out
        m = _TOKEN.match(src, pos)
        if not m:
            raise Unsupported("javascript: cannot read `%s`" % src[pos:pos + 20].strip())
        pos = m.end()
        if m.group("num") is not None:
            t = m.group("num")
            out.append(("num", float(int(t, 16)) if t[:2].lower() == "0x" else float(t)))
        elif m.group("id") is not None:
            out.append(("id", m.group("id")))
        elif m.group("str") is not None:
            body = m.group("str")[1:-1]
            body = re.sub(r"\\u([0-9a-fA-F]{4})", lambda k: chr(int(k.group(1), 16)), body)
            out.append(("str", re.sub(r"\\(.)", lambda k: _ESC.get(k.group(1), k.group(1)), body)))
        else:
            out.append(("op", m.group("op")))


class _Parser:
    def __init__(self, toks, params, implicit):
        self.t, self.i, self.params, self.implicit = toks, 0, params, implicit

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else ("end", None)

    def take(self):
        tok = self.peek()
        self.i += 1
        return tok

    def at(self, op):
        return self.peek() == ("op", op)

    def need(self, op):
        if not self.at(op):
            raise Unsupported("javascript: expected %r, found %r" % (op, self.peek()[1]))
        self.i += 1

    def binary_level(self, ops, lower):
        node = lower()
        while self.peek()[0] == "op" and self.peek()[1] in ops:
            op = self.take()[1]
            node = ("bin", op, node, lower())
        return node

    def expr(self):
        c = self.binary_level({"||"}, self.l_and)
        if self.at("?"):
            self.i += 1
            a = self.expr()
            self.need(":")
            return ("cond", c, a, self.expr())
        return c

    def l_and(self):
        return self.binary_level({"&&"}, self.l_eq)

    def l_eq(self):
        return self.binary_level({"==", "!=", "===", "!=="}, self.l_rel)

    def l_rel(self):
        return self.binary_level({"<", "<=", ">", ">="}, self.l_add)

    def l_add(self):
        return self.binary_level({"+", "-"}, self.l_mul)

    def l_mul(self):
        return self.binary_level({"*", "/", "%"}, self.l_unary)

    def l_unary(self):
        if self.peek()[0] == "op" and self.peek()[1] in "!-+":
            op = self.take()[1]
            return ("un", op, self.l_unary())
        node = self.primary()
        while self.at("."):
            self.i += 1
            kind, name = self.take()
            if kind != "id":
                raise Unsupported("javascript: member name expected")
            if self.at("("):
                node = ("method", name, node, self.args())
            else:
                node = ("prop", name, node)
        return node

    def args(self):
        self.need("(")
        out = []
        if not self.at(")"):
            out.append(self.expr())
            while self.at(","):
                self.i += 1
                out.append(self.expr())
        self.need(")")
        return out

    def primary(self):
        kind, v = self.take()
        if kind == "num":
            return ("lit", v)
        if kind == "str":
            return ("lit", v)
        if kind == "op" and v == "(":
            e = self.expr()
            self.need(")")
            return e
        if kind != "id":
            raise Unsupported("javascript: expression expected, found %r" % (v,))
        consts = {"null": None, "undefined": None, "true": True, "false": False, "Infinity": math.inf, "NaN": math.nan}
        if v in consts:
            return ("lit", consts[v])
        if v in ("new", "var", "function", "typeof", "if", "for", "while", "let", "const"):
            raise Unsupported("javascript: %s is outside the subset" % v)
        if v in self.params:
            return ("param", self.params.index(v))
        if v == "Number" and self.at("("):
            return ("call", "Number", self.args())
        if v in ("Math", "Number"):
            self.need(".")
            name = self.take()[1]
            if v == "Number":
                table = {"POSITIVE_INFINITY": math.inf, "NEGATIVE_INFINITY": -math.inf, "NaN": math.nan,
                         "MAX_VALUE": 1.7976931348623157e308, "MIN_VALUE": 5e-324}
                if name not in table:
                    raise Unsupported("javascript: Number.%s" % name)
                return ("lit", table[name])
            if not self.at("("):
                if name in ("PI", "E"):
                    return ("lit", math.pi if name == "PI" else math.e)
                raise Unsupported("javascript: Math.%s" % name)
            return ("call", "Math." + name, self.args())
        if self.at("("):
            if v not in ("parseFloat", "parseInt", "String", "isNaN", "Boolean"):
                raise Unsupported("javascript: call of %s" % v)
            return ("call", v, self.args())
        if self.implicit:
            self.params.append(v)
            return ("param", len(self.params) - 1)
        raise Unsupported("javascript: unknown identifier %s" % v)


def parse_function(src):
    """-> (params, body)"""
    toks = _tokens(src)
    p = _Parser(toks, [], False)
    if p.take() != ("id", "function"):
        raise Unsupported("javascript: 'function' expected")
    if p.peek()[0] == "id":
        p.take()
    p.need("(")
    while not p.at(")"):
        kind, name = p.take()
        if kind != "id":
            raise Unsupported("javascript: parameter name expected")
        p.params.append(name)
        if p.at(","):
            p.i += 1
    p.need(")")
    p.need("{")
    while p.at(";"):
        p.i += 1
    if p.take() != ("id", "return"):
        raise Unsupported("javascript: only a single return statement is supported")
    body = p.expr()
    while p.at(";"):
        p.i += 1
    p.need("}")
    while p.at(";"):
        p.i += 1
    if p.peek()[0] != "end":
        raise Unsupported("javascript: trailing input")
    return p.params, body


def parse_expression(src):
    p = _Parser(_tokens(src), [], True)
    body = p.expr()
    if p.peek()[0] != "end":
        raise Unsupported("expression: trailing input")
    return p.params, body


# ---- scalar semantics ----------------------------------------------------------------------------------------------
def truthy(v):
    if v is None:
        return False
    if isinstance(v, bool):
        return v
    if isinstance(v, float):
        return v != 0 and not math.isnan(v)
    return len(v) > 0


_NUM_LITERAL = re.compile(r"[+-]?(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?$")


def to_number(v):
    if v is None:
        return 0.0
    if isinstance(v, bool):
        return 1.0 if v else 0.0
    if isinstance(v, float):
        return v
    s = v.strip()
    if s == "":
        return 0.0
    if s in ("Infinity", "+Infinity"):
        return math.inf
    if s == "-Infinity":
        return -math.inf
    if re.match(r"0[xX][0-9a-fA-F]+$", s):
        return float(int(s, 16))
    return float(s) if _NUM_LITERAL.match(s) else math.nan


def to_string(v):
    if v is None:
        return "null"
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, str):
        return v
    if math.isnan(v):
        return "NaN"
    if math.isinf(v):
        return "Infinity" if v > 0 else "-Infinity"
    if v == int(v) and abs(v) < 1e21:
        return str(int(v))
    r = repr(v)  # shortest round-trip digits, like Number::toString
    if "e" in r:
        mant, ex = r.split("e")
        r = mant + "e" + ("-" if ex[0] == "-" else "+") + ex.lstrip("+-").lstrip("0")
    return r


def _loose_eq(a, b):
    if a is None or b is None:
        return a is None and b is None
    if type(a) is type(b):
        return a == b
    return to_number(a) == to_number(b)


def _float_prefix(s):
    m = re.match(r"\s*([+-]?(?:Infinity|(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?))", s)
    if not m:
        return math.nan
    t = m.group(1)
    return float(t.replace("Infinity", "inf"))


def _int_index(v, default=0.0):
    x = to_number(v) if v is not None else default
    return 0.0 if math.isnan(x) else float(math.trunc(x)) if not math.isinf(x) else x


def evaluate(node, args):
    k = node[0]
    if k == "lit":
        return node[1]
    if k == "param":
        return args[node[1]] if node[1] < len(args) else None
    if k == "un":
        v = evaluate(node[2], args)
        if node[1] == "!":
            return not truthy(v)
        return -to_number(v) if node[1] == "-" else to_number(v)
    if k == "cond":
        return evaluate(node[2], args) if truthy(evaluate(node[1], args)) else evaluate(node[3], args)
    if k == "bin":
        op = node[1]
        a = evaluate(node[2], args)
        if op == "||":
            return a if truthy(a) else evaluate(node[3], args)
        if op == "&&":
            return evaluate(node[3], args) if truthy(a) else a
        b = evaluate(node[3], args)
        if op == "==":
            return _loose_eq(a, b)
        if op == "!=":
            return not _loose_eq(a, b)
        if op in ("===", "!=="):
            same = type(a) is type(b) and (a == b)
            return same if op == "===" else not same
        if op in ("<", "<=", ">", ">="):
            if isinstance(a, str) and isinstance(b, str):
                x, y = a.encode("utf-16-be", "surrogatepass"), b.encode("utf-16-be", "surrogatepass")  # code-unit order
            else:
                x, y = to_number(a), to_number(b)
            return {"<": x < y, "<=": x <= y, ">": x > y, ">=": x >= y}[op]
        if op == "+":
            if isinstance(a, str) or isinstance(b, str):
                return to_string(a) + to_string(b)
            return to_number(a) + to_number(b)
        x, y = to_number(a), to_number(b)
        if op == "-":
            return x - y
        if op == "*":
            return x * y
        if op == "/":
            if y == 0:
                return math.nan if x == 0 or math.isnan(x) else math.copysign(math.inf, x) * math.copysign(1.0, y)
            return x / y
        if op == "%":
            return math.fmod(x, y) if y != 0 and not math.isinf(x) else math.nan
        raise Unsupported(op)
    if k == "prop":
        recv = evaluate(node[2], args)
        if node[1] == "length" and isinstance(recv, str):
            return float(len(recv.encode("utf-8")))  # the engine's strings are UTF-8 bytes
        raise Unsupported("javascript: property .%s" % node[1])
    if k == "method":
        name, recv = node[1], evaluate(node[2], args)
        a = [evaluate(x, args) for x in node[3]]
        if name == "toString":
            if recv is None:
                raise TypeError("toString on null")
            return to_string(recv)
        if not isinstance(recv, str):
            raise TypeError(".%s on a value that is not a string" % name)
        s = recv
        if name == "toUpperCase":
            return "".join(c.upper() if "a" <= c <= "z" else c for c in s)
        if name == "toLowerCase":
            return "".join(c.lower() if "A" <= c <= "Z" else c for c in s)
        if name == "concat":
            return s + "".join(to_string(x) for x in a)
        if name == "trim":
            return s.strip(" \t\n\r\f\v")
        n = float(len(s))
        if name == "substr":
            st = _int_index(a[0]) if a else 0.0
            if st < 0:
                st = max(0.0, n + st)
            st = min(st, n)
            cnt = min(max(_int_index(a[1]), 0.0), n - st) if len(a) > 1 else n - st
            return s[int(st):int(st + cnt)]
        if name == "substring":
            b = min(max(_int_index(a[0]) if a else 0.0, 0.0), n)
            e = min(max(_int_index(a[1]), 0.0), n) if len(a) > 1 and a[1] is not None else n
            b, e = min(b, e), max(b, e)
            return s[int(b):int(e)]
        if name == "slice":
            b = _int_index(a[0]) if a else 0.0
            e = _int_index(a[1]) if len(a) > 1 and a[1] is not None else n
            b = max(0.0, n + b) if b < 0 else min(b, n)
            e = max(0.0, n + e) if e < 0 else min(e, n)
            return s[int(b):int(e)] if e > b else ""
        if name == "charAt":
            i = _int_index(a[0]) if a else 0.0
            return s[int(i)] if 0 <= i < n else ""
        if name == "indexOf":
            return float(s.find(to_string(a[0]) if a else "undefined"))
        if name in ("startsWith", "endsWith", "includes"):
            t = to_string(a[0]) if a else "undefined"
            return s.startswith(t) if name == "startsWith" else s.endswith(t) if name == "endsWith" else (t in s)
        if name == "replace":
            if len(a) < 2 or not isinstance(a[0], str):
                raise Unsupported("replace with a non-string pattern")
            return s.replace(a[0], to_string(a[1]), 1)
        raise Unsupported("javascript: method .%s" % name)
    if k == "call":
        name = node[1]
        a = [evaluate(x, args) for x in node[2]]
        num = lambda i: to_number(a[i]) if i < len(a) else math.nan  # noqa: E731
        if name == "Number":
            return to_number(a[0]) if a else 0.0
        if name == "String":
            return to_string(a[0]) if a else ""
        if name == "Boolean":
            return bool(a) and truthy(a[0])
        if name == "isNaN":
            return math.isnan(num(0))
        if name == "parseFloat":
            if not a:
                return math.nan
            return a[0] if isinstance(a[0], float) else _float_prefix(to_string(a[0]))
        if name == "parseInt":
            m = re.match(r"\s*([+-]?\d+)", to_string(a[0])) if a else None
            return float(m.group(1)) if m else math.nan
        fn1 = {"Math.abs": abs, "Math.floor": math.floor, "Math.ceil": math.ceil, "Math.sqrt": math.sqrt, "Math.log": math.log,
               "Math.exp": math.exp, "Math.round": lambda x: math.floor(x + 0.5)}
        if name in fn1:
            x = num(0)
            if math.isnan(x) or math.isinf(x):
                return x if name != "Math.abs" else abs(x)
            return float(fn1[name](x))
        if name == "Math.pow":
            return float(num(0) ** num(1))
        if name in ("Math.min", "Math.max"):
            xs = [to_number(x) for x in a]
            if any(math.isnan(x) for x in xs):
                return math.nan
            return float(min(xs)) if name == "Math.min" else float(max(xs))
        raise Unsupported("javascript: %s" % name)
    raise Unsupported(k)


# ---- numeric expressions over columns --------------------------------------------------------------------------------
def evaluate_numeric(node, columns):
    """columns: list of float64 arrays, one per parameter.  IEEE double arithmetic element by element."""
    k = node[0]
    if k == "lit":
        if isinstance(node[1], bool) or not isinstance(node[1], float):
            raise Unsupported("numeric expression: literal %r" % (node[1],))
        return np.float64(node[1])
    if k == "param":
        return columns[node[1]]
    if k == "un" and node[1] in "+-":
        v = evaluate_numeric(node[2], columns)
        return -v if node[1] == "-" else v
    if k == "bin" and node[1] in "+-*/%":
        a, b = evaluate_numeric(node[2], columns), evaluate_numeric(node[3], columns)
        with np.errstate(all="ignore"):
            return {"+": np.add, "-": np.subtract, "*": np.multiply, "/": np.divide, "%": np.fmod}[node[1]](a, b)
    if k == "call":
        a = [evaluate_numeric(x, columns) for x in node[2]]
        one = {"Math.abs": np.abs, "Math.floor": np.floor, "Math.ceil": np.ceil, "Math.sqrt": np.sqrt, "Number": lambda x: x, "parseFloat": lambda x: x}
        if node[1] in one and len(a) == 1:
            return one[node[1]](a[0])
        if node[1] in ("Math.min", "Math.max") and len(a) == 2:
            return (np.minimum if node[1] == "Math.min" else np.maximum)(a[0], a[1])
    raise Unsupported("numeric expression: %r" % (node[:2],))


def split_fold(params, body):
    """fnAggregate's body -> ('sum' | 'min' | 'max', argument expression); parameter 0 is `current`."""
    cur = ("param", 0)
    if body[0] == "bin" and body[1] == "+" and cur in (body[2], body[3]):
        return "sum", body[3] if body[2] == cur else body[2]
    if body[0] == "call" and body[1] in ("Math.min", "Math.max") and len(body[2]) == 2 and cur in body[2]:
        return body[1][5:], body[2][1] if body[2][0] == cur else body[2][0]
    raise Unsupported("fnAggregate must fold with +, Math.min or Math.max")
