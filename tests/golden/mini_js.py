"""A small tree-walking interpreter for the subset of JavaScript / TypeScript that the reference's
audio front-end is written in.  TEST INFRASTRUCTURE ONLY (used by make_ref_exec.py and tests/).

Why it exists: no JavaScript engine is installed in the build image or on the GPU boxes, so the
reference's own text (apps/web/src/App.tsx) could never be *executed*; every golden vector was
the output of a restatement.  This interpreter runs the reference's source text itself -- sliced
out of App.tsx at run time, never committed -- so that the committed fixtures are outputs of the
reference's code rather than of somebody's reading of it.  It is not Node: the language semantics
below are this file's implementation of ECMA-262, kept deliberately small and tested on their own
(tests/test_ref_exec.py) so that each rule the audio path leans on is visible:

  * Number = IEEE-754 binary64 (Python float); + - * / never raise (x/0 -> +-Infinity or NaN)
  * relational operators on numbers with NaN -> false; `undefined` converts to NaN
  * Math.floor / Math.sqrt / Math.max / Math.min with their NaN, Infinity and -0 rules
  * typed arrays: Float32Array stores round to nearest binary32 (ties to even), Int16Array stores
    apply ToInt16 (NaN / +-Infinity -> 0, truncate toward zero, modulo 2^16, two's complement);
    reads outside the array give `undefined`; `new X(len)` applies ToIndex (RangeError)
  * TypeScript type annotations are parsed and dropped (parameters, return types, declarations)

Grammar covered: function declarations, arrow functions, const / let, for, if / else, return,
blocks, expression statements; expressions with assignment (= += -= *= /=), ?:, || &&, equality,
relational, + - * / %, unary - + !, postfix / prefix ++ --, member access, indexing, calls, `new`,
object literals, number (decimal, hex) / string / boolean / null literals.
"""
from __future__ import annotations

import math
import re
import struct

import numpy as np


class JSUndefined:
    _inst = None

    def __new__(cls):
        if cls._inst is None:
            cls._inst = super().__new__(cls)
        return cls._inst

    def __repr__(self):
        return "undefined"

    def __bool__(self):
        return False


undefined = JSUndefined()


class JSRangeError(Exception):
    pass


class JSTypeError(Exception):
    pass


class JSSyntaxError(Exception):
    pass


# ------------------------------------------------------------------------------------------------
# values
# ------------------------------------------------------------------------------------------------
def to_number(v) -> float:
    if v is undefined:
        return math.nan
    if v is None:
        return 0.0
    if isinstance(v, bool):
        return 1.0 if v else 0.0
    if isinstance(v, (int, float)):
        return float(v)
    raise JSTypeError(f"ToNumber of {type(v).__name__} is outside the supported subset")


def to_boolean(v) -> bool:
    if v is undefined or v is None:
        return False
    if isinstance(v, bool):
        return v
    if isinstance(v, (int, float)):
        return not (v == 0 or v != v)
    if isinstance(v, str):
        return len(v) > 0
    return True          # objects


def to_int16(v: float) -> int:
    """ECMA-262 7.1.9 ToInt16."""
    v = to_number(v)
    if v != v or math.isinf(v):
        return 0
    i = math.trunc(v) % 65536
    return i - 65536 if i >= 32768 else i


def to_float32(v: float) -> float:
    """Nearest binary32 (ties to even) as a Python float; overflow -> +-Infinity (what a Float32Array store does)."""
    v = to_number(v)
    if v != v:
        return math.nan
    try:
        return struct.unpack("<f", struct.pack("<f", v))[0]
    except OverflowError:
        return math.copysign(math.inf, v)


def to_index(v) -> int:
    """ECMA-262 7.1.22 ToIndex (typed-array lengths)."""
    n = to_number(v)
    if n != n:
        return 0
    if math.isinf(n):
        raise JSRangeError("Invalid typed array length")
    n = math.trunc(n)
    if n < 0 or n > 2 ** 53 - 1:
        raise JSRangeError("Invalid typed array length")
    return int(n)


class ArrayBuffer:
    def __init__(self, view: np.ndarray):
        self.view = view

    def tobytes(self) -> bytes:
        return self.view.tobytes()


class TypedArray:
    dtype = None

    def __init__(self, arg=0.0):
        if isinstance(arg, TypedArray):
            self.a = np.array(arg.a, dtype=self.dtype)
        elif isinstance(arg, np.ndarray):
            self.a = np.array(arg, dtype=self.dtype)
        else:
            n = to_index(arg)
            if n > 1 << 31:
                raise JSRangeError("Array buffer allocation failed")
            self.a = np.zeros(n, dtype=self.dtype)

    # property access from the interpreter ---------------------------------------------------
    def js_get(self, key):
        if key == "length":
            return float(self.a.size)
        if key == "buffer":
            return ArrayBuffer(self.a)
        if key == "set":
            return self._set
        idx = _canonical_index(key)
        if idx is None:
            return undefined
        if 0 <= idx < self.a.size:
            return float(self.a[idx])
        return undefined

    def js_set(self, key, value):
        idx = _canonical_index(key)
        if idx is None:
            raise JSTypeError(f"typed-array property {key!r} is outside the supported subset")
        conv = self.convert(value)                 # the conversion happens even when the index is outside
        if 0 <= idx < self.a.size:
            self.a[idx] = conv

    def _set(self, src, offset=0.0):
        """%TypedArray%.prototype.set(source, offset)."""
        off = to_index(offset)
        if not isinstance(src, TypedArray):
            raise JSTypeError("TypedArray.set: only typed-array sources are supported")
        if off + src.a.size > self.a.size:
            raise JSRangeError("offset is out of bounds")
        if type(src) is type(self):
            self.a[off:off + src.a.size] = src.a
        else:
            for k in range(src.a.size):
                self.a[off + k] = self.convert(float(src.a[k]))
        return undefined


def _canonical_index(key):
    if isinstance(key, bool):
        return None
    if isinstance(key, (int, float)):
        if key != key or math.isinf(key) or key != math.floor(key):
            return None
        if key == 0 and math.copysign(1.0, key) < 0:
            return None                            # "-0" is not an integer index: reads undefined, writes dropped
        return int(key)
    return None


class Float32Array(TypedArray):
    dtype = np.float32

    @staticmethod
    def convert(v):
        return np.float32(to_float32(v))


class Int16Array(TypedArray):
    dtype = np.int16

    @staticmethod
    def convert(v):
        return np.int16(to_int16(v))


class JSObject:
    """A plain object: property bag.  Host objects handed to the interpreter are these too."""

    def __init__(self, **props):
        self.props = dict(props)

    def js_get(self, key):
        return self.props.get(_prop_key(key), undefined)

    def js_set(self, key, value):
        self.props[_prop_key(key)] = value


def _prop_key(key):
    if isinstance(key, float) and key == math.floor(key) and not math.isinf(key):
        return str(int(key))
    return key if isinstance(key, str) else str(key)


def js_floor(x):
    x = to_number(x)
    if x != x or math.isinf(x) or x == 0:
        return x                                   # NaN, +-Infinity and +-0 are returned as they are
    return float(math.floor(x))


def js_sqrt(x):
    x = to_number(x)
    if x != x or x < 0:
        return math.nan
    return math.sqrt(x) if not math.isinf(x) else x


def js_max(*args):
    """Math.max: NaN if any argument is NaN (all are converted first); +0 beats -0; no arguments -> -Infinity."""
    vals = [to_number(a) for a in args]
    best = -math.inf
    nan = False
    for v in vals:
        if v != v:
            nan = True
        elif v > best or (v == 0 and best == 0 and math.copysign(1.0, v) > 0):
            best = v
    return math.nan if nan else best


def js_min(*args):
    vals = [to_number(a) for a in args]
    best = math.inf
    nan = False
    for v in vals:
        if v != v:
            nan = True
        elif v < best or (v == 0 and best == 0 and math.copysign(1.0, v) < 0):
            best = v
    return math.nan if nan else best


def js_abs(x):
    return abs(to_number(x))


def make_globals() -> dict:
    """The standard-library objects the audio path touches."""
    return {
        "Math": JSObject(floor=js_floor, sqrt=js_sqrt, max=js_max, min=js_min, abs=js_abs),
        "Float32Array": Float32Array,
        "Int16Array": Int16Array,
        "undefined": undefined,
        "NaN": math.nan,
        "Infinity": math.inf,
    }


# ------------------------------------------------------------------------------------------------
# tokens
# ------------------------------------------------------------------------------------------------
_TOKEN = re.compile(r"""
    (?P<ws>\s+|//[^\n]*|/\*.*?\*/)
  | (?P<num>0[xX][0-9a-fA-F]+|(?:\d+\.\d*|\.\d+|\d+)(?:[eE][+-]?\d+)?)
  | (?P<id>[A-Za-z_$][A-Za-z0-9_$]*)
  | (?P<str>"(?:[^"\\]|\\.)*"|'(?:[^'\\]|\\.)*')
  | (?P<op>===|!==|=>|\+\+|--|\+=|-=|\*=|/=|&&|\|\||\?\.|==|!=|<=|>=|[-+*/%<>=!?:;,.(){}\[\]|&])
""", re.X | re.S)


def tokenize(src: str):
    out, pos = [], 0
    while pos < len(src):
        m = _TOKEN.match(src, pos)
        if not m:
            raise JSSyntaxError(f"cannot tokenize at {src[pos:pos + 30]!r}")
        pos = m.end()
        kind = m.lastgroup
        if kind == "ws":
            continue
        out.append((kind, m.group(kind)))
    out.append(("eof", ""))
    return out


# ------------------------------------------------------------------------------------------------
# parser (AST = nested tuples)
# ------------------------------------------------------------------------------------------------
class Parser:
    def __init__(self, src: str):
        self.t = tokenize(src)
        self.i = 0

    # -- helpers --
    def peek(self, k=0):
        return self.t[min(self.i + k, len(self.t) - 1)]

    def at(self, text, k=0):
        kind, s = self.peek(k)
        return kind in ("op", "id") and s == text

    def eat(self, text):
        if not self.at(text):
            raise JSSyntaxError(f"expected {text!r}, found {self.peek()[1]!r} (token {self.i})")
        self.i += 1

    def maybe(self, text):
        if self.at(text):
            self.i += 1
            return True
        return False

    def ident(self):
        kind, s = self.peek()
        if kind != "id":
            raise JSSyntaxError(f"expected an identifier, found {s!r}")
        self.i += 1
        return s

    def skip_type(self, stops):
        """Skips a TypeScript type up to (not including) one of `stops` at nesting depth 0."""
        depth = 0
        while True:
            kind, s = self.peek()
            if kind == "eof":
                raise JSSyntaxError("unterminated type annotation")
            if depth == 0 and kind == "op" and s in stops:
                return
            if s in ("<", "(", "[", "{"):
                depth += 1
            elif s in (">", ")", "]", "}"):
                depth -= 1
            self.i += 1

    # -- statements --
    def program(self):
        body = []
        while self.peek()[0] != "eof":
            body.append(self.statement())
        return ("block", body)

    def block(self):
        self.eat("{")
        body = []
        while not self.at("}"):
            body.append(self.statement())
        self.eat("}")
        return ("block", body)

    def statement(self):
        if self.at("{"):
            return self.block()
        if self.at("function"):
            return self.function_decl()
        if self.at("async") and self.at("function", 1):
            raise JSSyntaxError("async functions are outside the supported subset")
        if self.at("const") or self.at("let") or self.at("var"):
            d = self.var_decl()
            self.maybe(";")
            return d
        if self.at("if"):
            self.i += 1
            self.eat("(")
            c = self.expression()
            self.eat(")")
            a = self.statement()
            b = None
            if self.maybe("else"):
                b = self.statement()
            return ("if", c, a, b)
        if self.at("for"):
            self.i += 1
            self.eat("(")
            init = None
            if not self.at(";"):
                init = self.var_decl() if (self.at("let") or self.at("const") or self.at("var")) else ("expr", self.expression())
            self.eat(";")
            cond = None if self.at(";") else self.expression()
            self.eat(";")
            upd = None if self.at(")") else self.expression()
            self.eat(")")
            return ("for", init, cond, upd, self.statement())
        if self.at("return"):
            self.i += 1
            e = None
            if not self.at(";") and not self.at("}"):
                e = self.expression()
            self.maybe(";")
            return ("return", e)
        if self.at(";"):
            self.i += 1
            return ("block", [])
        e = self.expression()
        self.maybe(";")
        return ("expr", e)

    def var_decl(self):
        kind = self.peek()[1]
        self.i += 1
        decls = []
        while True:
            name = self.ident()
            if self.maybe(":"):
                self.skip_type(("=", ";", ","))
            init = None
            if self.maybe("="):
                init = self.assignment()
            decls.append((name, init))
            if not self.maybe(","):
                break
        return ("var", kind, decls)

    def params(self):
        self.eat("(")
        names = []
        while not self.at(")"):
            names.append(self.ident())
            self.maybe("?")
            if self.maybe(":"):
                self.skip_type((",", ")"))
            self.maybe(",")
        self.eat(")")
        return names

    def function_decl(self):
        self.eat("function")
        name = self.ident()
        names = self.params()
        if self.maybe(":"):
            self.skip_type(("{",))
        return ("funcdecl", name, names, self.block())

    # -- expressions --
    def expression(self):
        e = self.assignment()
        while self.maybe(","):
            e = ("comma", e, self.assignment())
        return e

    def is_arrow(self):
        """At '(' : is this the parameter list of an arrow function?"""
        if self.peek()[0] == "id" and self.at("=>", 1):
            return True
        if not self.at("("):
            return False
        depth, k = 0, 0
        while True:
            kind, s = self.peek(k)
            if kind == "eof":
                return False
            if s in ("(", "[", "{", "<") and kind == "op":
                depth += 1
            elif s in (")", "]", "}", ">") and kind == "op":
                depth -= 1
                if depth == 0:
                    return self.at("=>", k + 1) or self.at(":", k + 1) and self._arrow_after_type(k + 2)
            k += 1

    def _arrow_after_type(self, k):
        while True:
            kind, s = self.peek(k)
            if kind == "eof" or s in (";", ")"):
                return False
            if s == "=>":
                return True
            k += 1

    def arrow(self):
        if self.peek()[0] == "id":
            names = [self.ident()]
        else:
            names = self.params()
            if self.maybe(":"):
                self.skip_type(("=>",))
        self.eat("=>")
        if self.at("{"):
            return ("func", names, self.block(), False)
        return ("func", names, self.assignment(), True)

    def assignment(self):
        if self.is_arrow():
            return self.arrow()
        left = self.conditional()
        for op in ("=", "+=", "-=", "*=", "/="):
            if self.at(op):
                if left[0] not in ("name", "member", "index"):
                    raise JSSyntaxError("invalid assignment target")
                self.i += 1
                return ("assign", op, left, self.assignment())
        return left

    def conditional(self):
        c = self.binary(0)
        if self.maybe("?"):
            a = self.assignment()
            self.eat(":")
            b = self.assignment()
            return ("cond", c, a, b)
        return c

    LEVELS = [("||",), ("&&",), ("===", "!==", "==", "!="), ("<", ">", "<=", ">="), ("+", "-"), ("*", "/", "%")]

    def binary(self, level):
        if level == len(self.LEVELS):
            return self.unary()
        left = self.binary(level + 1)
        while True:
            kind, s = self.peek()
            if kind == "op" and s in self.LEVELS[level]:
                self.i += 1
                right = self.binary(level + 1)
                left = ("logic" if s in ("||", "&&") else "bin", s, left, right)
            else:
                return left

    def unary(self):
        kind, s = self.peek()
        if kind == "op" and s in ("-", "+", "!"):
            self.i += 1
            return ("unary", s, self.unary())
        if kind == "op" and s in ("++", "--"):
            self.i += 1
            return ("update", s, True, self.unary())
        e = self.postfix()
        if self.at("++") or self.at("--"):
            op = self.peek()[1]
            self.i += 1
            return ("update", op, False, e)
        return e

    def postfix(self):
        e = self.primary()
        while True:
            if self.maybe("."):
                e = ("member", e, self.ident())
            elif self.at("["):
                self.i += 1
                k = self.expression()
                self.eat("]")
                e = ("index", e, k)
            elif self.at("("):
                e = ("call", e, self.arguments())
            else:
                return e

    def arguments(self):
        self.eat("(")
        args = []
        while not self.at(")"):
            args.append(self.assignment())
            self.maybe(",")
        self.eat(")")
        return args

    def primary(self):
        kind, s = self.peek()
        if kind == "num":
            self.i += 1
            return ("num", float(int(s, 16)) if s[:2] in ("0x", "0X") else float(s))
        if kind == "str":
            self.i += 1
            return ("str", bytes(s[1:-1], "utf-8").decode("unicode_escape"))
        if kind == "id":
            if s == "new":
                self.i += 1
                callee = self.primary()
                while self.maybe("."):
                    callee = ("member", callee, self.ident())
                args = self.arguments() if self.at("(") else []
                return ("new", callee, args)
            if s in ("true", "false"):
                self.i += 1
                return ("lit", s == "true")
            if s == "null":
                self.i += 1
                return ("lit", None)
            self.i += 1
            return ("name", s)
        if self.at("("):
            self.i += 1
            e = self.expression()
            self.eat(")")
            return e
        if self.at("{"):
            self.i += 1
            props = []
            while not self.at("}"):
                k = self.ident() if self.peek()[0] == "id" else self.primary()[1]
                self.eat(":")
                props.append((k, self.assignment()))
                self.maybe(",")
            self.eat("}")
            return ("object", props)
        raise JSSyntaxError(f"unexpected token {s!r} (token {self.i})")


# ------------------------------------------------------------------------------------------------
# evaluator
# ------------------------------------------------------------------------------------------------
class _Return(Exception):
    def __init__(self, value):
        self.value = value


class Scope:
    def __init__(self, parent=None):
        self.vars = {}
        self.consts = set()
        self.parent = parent

    def lookup(self, name):
        s = self
        while s is not None:
            if name in s.vars:
                return s
            s = s.parent
        return None


class Function:
    def __init__(self, interp, params, body, is_expr, scope, name="<arrow>"):
        self.interp, self.params, self.body, self.is_expr, self.scope, self.name = interp, params, body, is_expr, scope, name

    def __call__(self, *args):
        sc = Scope(self.scope)
        for k, p in enumerate(self.params):
            sc.vars[p] = args[k] if k < len(args) else undefined
        if self.is_expr:
            return self.interp.ev(self.body, sc)
        try:
            self.interp.exec(self.body, sc)
        except _Return as r:
            return r.value
        return undefined


class Interpreter:
    def __init__(self, env: dict | None = None):
        self.globals = Scope()
        self.globals.vars.update(make_globals())
        if env:
            self.globals.vars.update(env)
        self.steps = 0

    def run(self, src: str):
        self.exec(Parser(src).program(), self.globals, new_scope=False)

    def get(self, name):
        return self.globals.vars[name]

    # -- statements --
    def exec(self, node, sc, new_scope=True):
        kind = node[0]
        if kind == "block":
            inner = Scope(sc) if new_scope else sc
            for st in node[1]:                                  # function declarations hoist
                if st[0] == "funcdecl":
                    inner.vars[st[1]] = Function(self, st[2], st[3], False, inner, st[1])
            for st in node[1]:
                if st[0] != "funcdecl":
                    self.exec(st, inner)
        elif kind == "expr":
            self.ev(node[1], sc)
        elif kind == "var":
            for name, init in node[2]:
                sc.vars[name] = undefined if init is None else self.ev(init, sc)
                if node[1] == "const":
                    sc.consts.add(name)
        elif kind == "if":
            if to_boolean(self.ev(node[1], sc)):
                self.exec(node[2], sc)
            elif node[3] is not None:
                self.exec(node[3], sc)
        elif kind == "for":
            # `let` in a for head gets a fresh binding per iteration in ECMAScript; no closure in the subset
            # captures the loop variable, so one scope for the whole loop is indistinguishable here
            loop = Scope(sc)
            if node[1] is not None:
                self.exec(node[1], loop)
            while node[2] is None or to_boolean(self.ev(node[2], loop)):
                self.exec(node[4], loop)
                if node[3] is not None:
                    self.ev(node[3], loop)
        elif kind == "return":
            raise _Return(undefined if node[1] is None else self.ev(node[1], sc))
        elif kind == "funcdecl":
            sc.vars[node[1]] = Function(self, node[2], node[3], False, sc, node[1])
        else:
            raise JSSyntaxError(f"statement {kind} not supported")

    # -- expressions --
    def ev(self, n, sc):
        self.steps += 1
        k = n[0]
        if k == "num" or k == "str" or k == "lit":
            return n[1]
        if k == "name":
            s = sc.lookup(n[1])
            if s is None:
                raise JSTypeError(f"ReferenceError: {n[1]} is not defined")
            return s.vars[n[1]]
        if k == "bin":
            return self.binop(n[1], self.ev(n[2], sc), self.ev(n[3], sc))
        if k == "index":
            return self.get_prop(self.ev(n[1], sc), self.ev(n[2], sc))
        if k == "member":
            return self.get_prop(self.ev(n[1], sc), n[2])
        if k == "assign":
            return self.assign(n, sc)
        if k == "update":
            old = to_number(self.ev(n[3], sc))
            new = old + 1.0 if n[1] == "++" else old - 1.0
            self.store(n[3], new, sc)
            return new if n[2] else old
        if k == "cond":
            return self.ev(n[2], sc) if to_boolean(self.ev(n[1], sc)) else self.ev(n[3], sc)
        if k == "logic":
            left = self.ev(n[2], sc)
            if n[1] == "&&":
                return self.ev(n[3], sc) if to_boolean(left) else left
            return left if to_boolean(left) else self.ev(n[3], sc)
        if k == "unary":
            v = self.ev(n[2], sc)
            if n[1] == "!":
                return not to_boolean(v)
            return -to_number(v) if n[1] == "-" else to_number(v)
        if k == "call":
            callee = n[1]
            if callee[0] == "member":
                this = self.ev(callee[1], sc)
                fn = self.get_prop(this, callee[2])
            elif callee[0] == "index":
                this = self.ev(callee[1], sc)
                fn = self.get_prop(this, self.ev(callee[2], sc))
            else:
                fn = self.ev(callee, sc)
            if not callable(fn):
                raise JSTypeError(f"TypeError: {callee} is not a function")
            return fn(*[self.ev(a, sc) for a in n[2]])
        if k == "new":
            ctor = self.ev(n[1], sc)
            if not isinstance(ctor, type):
                raise JSTypeError("`new` on a non-constructor")
            return ctor(*[self.ev(a, sc) for a in n[2]])
        if k == "func":
            return Function(self, n[1], n[2], n[3], sc)
        if k == "object":
            o = JSObject()
            for key, val in n[1]:
                o.js_set(key, self.ev(val, sc))
            return o
        if k == "comma":
            self.ev(n[1], sc)
            return self.ev(n[2], sc)
        raise JSSyntaxError(f"expression {k} not supported")

    def get_prop(self, obj, key):
        if obj is undefined or obj is None:
            raise JSTypeError(f"TypeError: cannot read properties of {obj!r} (reading {key!r})")
        if hasattr(obj, "js_get"):
            return obj.js_get(key)
        if isinstance(obj, type) and isinstance(key, str) and hasattr(obj, key):
            return getattr(obj, key)
        raise JSTypeError(f"property access on {type(obj).__name__} is outside the supported subset")

    def store(self, target, value, sc):
        if target[0] == "name":
            s = sc.lookup(target[1])
            if s is None:
                raise JSTypeError(f"ReferenceError: {target[1]} is not defined")
            if target[1] in s.consts:
                raise JSTypeError(f"TypeError: assignment to constant variable {target[1]}")
            s.vars[target[1]] = value
        elif target[0] == "member":
            self.ev(target[1], sc).js_set(target[2], value)
        elif target[0] == "index":
            self.ev(target[1], sc).js_set(self.ev(target[2], sc), value)
        else:
            raise JSSyntaxError("invalid assignment target")

    def assign(self, n, sc):
        op, target = n[1], n[2]
        if op == "=":
            v = self.ev(n[3], sc)
        else:
            v = self.binop(op[0], self.ev(target, sc), self.ev(n[3], sc))
        self.store(target, v, sc)
        return v

    @staticmethod
    def binop(op, a, b):
        if op in ("===", "!=="):
            if isinstance(a, (int, float)) and not isinstance(a, bool) and isinstance(b, (int, float)) and not isinstance(b, bool):
                eq = float(a) == float(b)
            elif type(a) is not type(b):
                eq = False
            else:
                eq = a == b if isinstance(a, (str, bool)) else a is b
            return eq if op == "===" else not eq
        if op in ("==", "!="):
            if (a is None or a is undefined) and (b is None or b is undefined):
                eq = True
            elif a is None or a is undefined or b is None or b is undefined:
                eq = False
            else:
                eq = Interpreter.binop("===", a, b)
            return eq if op == "==" else not eq
        if op == "+" and (isinstance(a, str) or isinstance(b, str)):
            raise JSTypeError("string concatenation is outside the supported subset")
        x, y = to_number(a), to_number(b)
        if op == "+":
            return x + y
        if op == "-":
            return x - y
        if op == "*":
            return x * y
        if op == "/":
            if y == 0:
                if x != x or x == 0:
                    return math.nan
                return math.copysign(math.inf, x) * math.copysign(1.0, y)
            try:
                return x / y
            except OverflowError:
                return math.copysign(math.inf, x) * math.copysign(1.0, y)
        if op == "%":
            if x != x or y != y or math.isinf(x) or y == 0:
                return math.nan
            if math.isinf(y):
                return x
            return math.fmod(x, y)
        if op == "<":
            return x < y
        if op == ">":
            return x > y
        if op == "<=":
            return x <= y
        if op == ">=":
            return x >= y
        raise JSSyntaxError(f"operator {op} not supported")
