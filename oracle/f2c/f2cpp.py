#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- mechanical Fortran -> C++ translation of the UNMODIFIED reference source.

    python oracle/f2c/f2cpp.py /root/reference/src/tracer/MOM_tracer_advect.F90 \
        advect_tracer advect_x advect_y  > oracle/_ref/advect_f2c.inc

There is no Fortran compiler in this image, so the reference cannot be compiled as it is.  This script
reads the reference file where it lies (nothing of it is copied into the repository: the output goes to
the git-ignored oracle/_ref/) and translates the named subroutines statement by statement into C++ that
is compiled against oracle/f2c/f2c_runtime.h -- the C++ counterpart of the stand-in modules of
oracle/fortran_stub_harness (types with the members the path reads, Fortran-bounds arrays, and the
intrinsics with gfortran's lowering).  The translation is purely syntactic:

  * every expression is parsed with Fortran's precedence and associativity and emitted FULLY
    parenthesised, so the C++ compiler has no freedom the Fortran compiler did not have
    (g++ -O2 -ffp-contract=off: no contraction, no reassociation);
  * identifiers are case-folded (Fortran is case-insensitive: `I` and `i` are one variable);
  * do/if/elseif/else/exit/cycle/return/call map one to one; array references keep their
    Fortran bounds (Arr<T,N>::operator()); whole-array assignments become f_assign;
  * SZI_(G) etc. expand as src/framework/MOM_memory_macros.h:163-175 defines them for dynamic memory.

It understands the subset of Fortran 90 that MOM_tracer_advect.F90 (and MOM_tracer_hor_diff.F90's
arithmetic) uses and stops with an error on anything else rather than guessing.
"""
from __future__ import annotations

import re
import sys

# src/framework/MOM_memory_macros.h:163-175 (dynamic memory)
SZ_MACROS = {
    "szi_": "{0}%isd:{0}%ied", "szj_": "{0}%jsd:{0}%jed", "szk_": "{0}%ke", "szk0_": "0:{0}%ke",
    "szib_": "{0}%isdb:{0}%iedb", "szjb_": "{0}%jsdb:{0}%jedb",
}
INTRINSICS = {"max": "f_max", "min": "f_min", "abs": "f_abs", "sign": "f_sign", "mod": "f_mod", "int": "f_int",
              "ceiling": "f_ceiling", "tiny": "f_tiny", "present": "f_present", "associated": "f_associated",
              "allocated": "f_allocated", "sqrt": "f_sqrt", "epsilon": "f_epsilon", "real": "f_real"}
# names whose Fortran declaration is `pointer` to a derived type: member access is ->
POINTER_NAMES = {"cs", "reg", "obc", "segment"}      # local / dummy pointers to derived types
POINTER_MEMBERS = {"tr_reg"}                          # pointer-valued components
CXX_RESERVED = {"int", "double", "float", "long", "short", "char", "bool", "new", "delete", "this", "class", "struct",
                "union", "enum", "template", "typename", "namespace", "using", "public", "private", "protected",
                "virtual", "operator", "and", "or", "not", "xor", "auto", "register", "static", "const", "volatile",
                "signed", "unsigned", "void", "default", "switch", "case", "break", "continue", "goto", "try",
                "catch", "throw", "true", "false", "nullptr", "inline", "friend", "export", "extern", "typedef"}


class F2CError(Exception):
    pass


# ----------------------------------------------------------------------------------------------------
# source -> statements
# ----------------------------------------------------------------------------------------------------
def strip_comment(line: str) -> str:
    out, q = [], None
    for ch in line:
        if q:
            out.append(ch)
            if ch == q:
                q = None
        elif ch in "'\"":
            q = ch
            out.append(ch)
        elif ch == "!":
            break
        else:
            out.append(ch)
    return "".join(out).rstrip()


def split_semicolons(s: str):
    parts, cur, q, depth = [], [], None, 0
    for ch in s:
        if q:
            cur.append(ch)
            if ch == q:
                q = None
        elif ch in "'\"":
            q = ch
            cur.append(ch)
        elif ch == ";":
            parts.append("".join(cur))
            cur = []
        else:
            cur.append(ch)
    parts.append("".join(cur))
    return [p.strip() for p in parts if p.strip()]


def statements(text: str):
    """Yield (first_line_number, statement) with comments removed, continuations joined, ';' split,
    everything outside string literals lower-cased."""
    lines = text.split("\n")
    cur, start = "", 0
    for n, raw in enumerate(lines, 1):
        if raw.lstrip().startswith("#"):
            continue
        l = strip_comment(raw)
        if not l.strip():
            continue
        s = l.strip()
        if cur:
            if s.startswith("&"):
                s = s[1:].lstrip()
        else:
            start = n
        if s.endswith("&"):
            cur += s[:-1].rstrip() + " "
            continue
        cur += s
        for st in split_semicolons(cur):
            yield start, lower_outside_strings(st)
        cur = ""


def lower_outside_strings(s: str) -> str:
    out, q = [], None
    for ch in s:
        if q:
            out.append(ch)
            if ch == q:
                q = None
        elif ch in "'\"":
            q = ch
            out.append(ch)
        else:
            out.append(ch.lower())
    return "".join(out)


# ----------------------------------------------------------------------------------------------------
# expressions
# ----------------------------------------------------------------------------------------------------
TOKEN_RE = re.compile(r"""
    (?P<ws>\s+)
  | (?P<dotop>\.(?:and|or|not|true|false|eqv|neqv|eq|ne|lt|le|gt|ge)\.)
  | (?P<num>(?:\d+\.\d*|\.\d+|\d+)(?:[ed][+-]?\d+)?(?:_\w+)?)
  | (?P<name>[a-z_]\w*)
  | (?P<str>'[^']*'|"[^"]*")
  | (?P<op>\*\*|//|==|/=|<=|>=|=>|[-+*/(),%<>=:])
""", re.X)


def tokenize(s: str):
    toks, i = [], 0
    while i < len(s):
        # a number followed by a dotted operator: "1.and." must not swallow the dot
        m = TOKEN_RE.match(s, i)
        if not m:
            raise F2CError("cannot tokenize %r at %r" % (s, s[i:i + 20]))
        kind = m.lastgroup
        txt = m.group()
        if kind == "num" and txt.endswith(".") and re.match(r"\.(?:and|or|not|eq|ne|lt|le|gt|ge|eqv|neqv)\.", s[m.end() - 1:]):
            txt = txt[:-1]
            i += len(txt)
            toks.append(("num", txt))
            continue
        i = m.end()
        if kind == "ws":
            continue
        toks.append((kind, txt))
    return toks


class Node:
    pass


class Num(Node):
    def __init__(self, t): self.t = t


class Str(Node):
    def __init__(self, t): self.t = t


class Logical(Node):
    def __init__(self, v): self.v = v


class Section(Node):      # a(:) or a(lo:hi) argument
    def __init__(self, lo, hi): self.lo, self.hi = lo, hi


class Keyword(Node):      # name=value argument
    def __init__(self, name, val): self.name, self.val = name, val


class Ref(Node):          # a%b(i)%c ...: parts = [(name, args or None), ...]
    def __init__(self, parts): self.parts = parts


class Un(Node):
    def __init__(self, op, a): self.op, self.a = op, a


class Bin(Node):
    def __init__(self, op, a, b): self.op, self.a, self.b = op, a, b


class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def accept(self, txt):
        if self.peek()[1] == txt:
            self.i += 1
            return True
        return False

    def expect(self, txt):
        if not self.accept(txt):
            raise F2CError("expected %r, got %r in %r" % (txt, self.peek(), self.t))

    # precedence (lowest first): .eqv./.neqv. < .or. < .and. < .not. < relational < +,- (binary and unary) < *,/ < **
    def expr(self):
        return self.p_or()

    def p_or(self):
        a = self.p_and()
        while self.peek()[1] == ".or.":
            self.next()
            a = Bin("||", a, self.p_and())
        return a

    def p_and(self):
        a = self.p_not()
        while self.peek()[1] == ".and.":
            self.next()
            a = Bin("&&", a, self.p_not())
        return a

    def p_not(self):
        if self.peek()[1] == ".not.":
            self.next()
            return Un("!", self.p_not())
        return self.p_rel()

    REL = {"==": "==", "/=": "!=", "<": "<", "<=": "<=", ">": ">", ">=": ">=", ".eq.": "==", ".ne.": "!=", ".lt.": "<",
           ".le.": "<=", ".gt.": ">", ".ge.": ">="}

    def p_rel(self):
        a = self.p_add()
        if self.peek()[1] in self.REL:
            op = self.REL[self.next()[1]]
            a = Bin(op, a, self.p_add())
        return a

    def p_add(self):
        # a leading sign applies to the first TERM: -a*b = -(a*b), -a + b = (-a) + b
        if self.peek()[1] in ("-", "+"):
            op = self.next()[1]
            a = self.p_mul()
            a = Un(op, a) if op == "-" else a
        else:
            a = self.p_mul()
        while self.peek()[1] in ("+", "-", "//"):
            op = self.next()[1]
            a = Bin(op, a, self.p_mul())
        return a

    def p_mul(self):
        a = self.p_pow()
        while self.peek()[1] in ("*", "/"):
            op = self.next()[1]
            a = Bin(op, a, self.p_pow())
        return a

    def p_pow(self):
        a = self.p_primary()
        if self.peek()[1] == "**":
            self.next()
            # right associative; the exponent may carry a sign
            if self.peek()[1] in ("-", "+"):
                op = self.next()[1]
                b = self.p_pow()
                b = Un("-", b) if op == "-" else b
            else:
                b = self.p_pow()
            return Bin("**", a, b)
        return a

    def p_primary(self):
        kind, txt = self.peek()
        if kind == "num":
            self.next()
            return Num(txt)
        if kind == "str":
            self.next()
            return Str(txt)
        if kind == "dotop" and txt in (".true.", ".false."):
            self.next()
            return Logical(txt == ".true.")
        if txt == "(":
            self.next()
            e = self.expr()
            self.expect(")")
            return Un("()", e)
        if kind == "name":
            return self.p_ref()
        raise F2CError("unexpected token %r in %r" % ((kind, txt), self.t))

    def p_ref(self):
        parts = []
        while True:
            kind, name = self.next()
            if kind != "name":
                raise F2CError("expected a name, got %r" % ((kind, name),))
            args = None
            if self.peek()[1] == "(":
                self.next()
                args = []
                if not self.accept(")"):
                    while True:
                        args.append(self.p_arg())
                        if self.accept(")"):
                            break
                        self.expect(",")
            parts.append((name, args))
            if not self.accept("%"):
                break
        return Ref(parts)

    def p_arg(self):
        # keyword argument
        if self.peek()[0] == "name" and self.i + 1 < len(self.t) and self.t[self.i + 1][1] == "=":
            name = self.next()[1]
            self.next()
            return Keyword(name, self.expr())
        if self.peek()[1] == ":":
            self.next()
            hi = None if self.peek()[1] in (",", ")") else self.expr()
            return Section(None, hi)
        e = self.expr()
        if self.accept(":"):
            hi = None if self.peek()[1] in (",", ")") else self.expr()
            return Section(e, hi)
        return e


def parse_expr(s: str) -> Node:
    p = Parser(tokenize(s))
    e = p.expr()
    if p.i != len(p.t):
        raise F2CError("trailing tokens in %r: %r" % (s, p.t[p.i:]))
    return e


def cname(n: str) -> str:
    return n + "_" if n in CXX_RESERVED else n


class Emitter:
    """AST -> C++ text.  `funcs` = names of the translated subroutines (calls to them are plain calls)."""

    def __init__(self, arrays):
        self.arrays = arrays      # names declared as arrays in the current subroutine

    def num(self, t: str) -> str:
        t = re.sub(r"_\w+$", "", t)
        if re.fullmatch(r"\d+", t):
            return t
        t = t.replace("d", "e")
        if t.endswith("."):
            t += "0"
        t = re.sub(r"\.e", ".0e", t)
        if t.startswith("."):
            t = "0" + t
        return t

    def e(self, n: Node) -> str:
        if isinstance(n, Num):
            return self.num(n.t)
        if isinstance(n, Str):
            return '"' + n.t[1:-1].replace("\\", "\\\\").replace('"', '\\"') + '"'
        if isinstance(n, Logical):
            return "true" if n.v else "false"
        if isinstance(n, Un):
            if n.op == "()":
                return "(" + self.e(n.a) + ")"
            return "(" + n.op + self.e(n.a) + ")"
        if isinstance(n, Bin):
            if n.op == "**":
                return "f_pow(" + self.e(n.a) + ", " + self.e(n.b) + ")"
            if n.op == "//":
                return "f_concat(" + self.e(n.a) + ", " + self.e(n.b) + ")"
            return "(" + self.e(n.a) + " " + n.op + " " + self.e(n.b) + ")"
        if isinstance(n, Ref):
            return self.ref(n)
        if isinstance(n, Keyword):
            return self.e(n.val)
        if isinstance(n, Section):
            raise F2CError("array section in an expression context")
        raise F2CError("unknown node %r" % n)

    def whole(self, args) -> bool:
        return args is not None and len(args) > 0 and all(isinstance(a, Section) and a.lo is None and a.hi is None for a in args)

    def ref(self, r: Ref) -> str:
        out = ""
        # ANY(array-section <cmp> scalar), e.g. any(frac_under_100m(:,:,k) > 0): a predicate over the section's elements
        if len(r.parts) == 1 and r.parts[0][0] == "any" and r.parts[0][1] is not None and len(r.parts[0][1]) == 1 \
                and isinstance(r.parts[0][1][0], Bin) and isinstance(r.parts[0][1][0].a, Ref):
            cmp_ = r.parts[0][1][0]
            arr = cmp_.a
            aname, aargs = arr.parts[-1]
            if aargs and any(isinstance(a, Section) for a in aargs):
                for a in aargs:
                    if isinstance(a, Section) and (a.lo is not None or a.hi is not None):
                        raise F2CError("ANY: only full ':' sections are supported")
                base = Ref(arr.parts[:-1] + [(aname, None)])
                fixed = ", ".join("0" if isinstance(a, Section) else self.e(a) for a in aargs)
                full = ", ".join("true" if isinstance(a, Section) else "false" for a in aargs)
                return "f_any_section(%s, {%s}, {%s}, [&](double x__) { return x__ %s %s; })" % (
                    self.ref(base), fixed, full, cmp_.op, self.e(cmp_.b))
        for k, (name, args) in enumerate(r.parts):
            first = (k == 0)
            if first and name in INTRINSICS and args is not None and len(r.parts) == 1:
                return INTRINSICS[name] + "(" + ", ".join(self.e(a) for a in args) + ")"
            txt = cname(name)
            if args is not None and not self.whole(args):
                txt += "(" + ", ".join(self.e(a) for a in args) + ")"
            if first:
                out = txt
            else:
                out += txt
            if k + 1 < len(r.parts):
                out += "->" if (name in POINTER_NAMES if first else name in POINTER_MEMBERS) else "."
        return out

    def is_whole_array_target(self, n: Node) -> bool:
        if not isinstance(n, Ref):
            return False
        name, args = n.parts[-1]
        if self.whole(args):
            return True
        return len(n.parts) == 1 and args is None and name in self.arrays


# ----------------------------------------------------------------------------------------------------
# declarations
# ----------------------------------------------------------------------------------------------------
def split_top(s: str, sep: str = ","):
    parts, cur, depth, q = [], [], 0, None
    for ch in s:
        if q:
            cur.append(ch)
            if ch == q:
                q = None
            continue
        if ch in "'\"":
            q = ch
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == sep and depth == 0:
            parts.append("".join(cur).strip())
            cur = []
        else:
            cur.append(ch)
    parts.append("".join(cur).strip())
    return parts


def expand_sz(s: str) -> str:
    def rep(m):
        return SZ_MACROS[m.group(1)].format(m.group(2))
    return re.sub(r"\b(szi_|szj_|szk_|szk0_|szib_|szjb_)\((\w+)\)", rep, s)


class Decl:
    def __init__(self, name, base, dims, attrs, init):
        self.name, self.base, self.dims, self.attrs, self.init = name, base, dims, attrs, init
    # base: "double" | "int" | "lbool" | ("type", tname);  dims: list of (lo, hi) strings or None


DECL_RE = re.compile(r"^(real|integer|logical|character|type\s*\(\s*\w+\s*\))(?![a-z0-9_])(.*)$")


def parse_decl(st: str):
    m = DECL_RE.match(st)
    if not m:
        return None
    head, rest = m.group(1), m.group(2)
    if "::" in rest:
        attrs_s, names_s = rest.split("::", 1)
    else:
        attrs_s, names_s = "", rest
    if head.startswith("character"):
        return []
    attrs, dims = {}, None
    for a in split_top(attrs_s.strip().lstrip(",")):
        a = a.strip()
        if not a:
            continue
        if a.startswith("dimension"):
            dims = split_top(expand_sz(a[a.index("(") + 1:a.rindex(")")]))
        elif a.startswith("intent"):
            attrs["intent"] = re.sub(r"\s", "", a[a.index("(") + 1:a.rindex(")")])
        elif a.startswith("(") and head.startswith("character"):
            pass
        else:
            attrs[a] = True
    if head == "real":
        base = "double"
    elif head == "integer":
        base = "int"
    elif head == "logical":
        base = "lbool"
    else:
        base = ("type", re.sub(r"\s", "", head)[5:-1])
    out = []
    for item in split_top(expand_sz(names_s)):
        init = None
        if "=>" in item:
            item, init = [x.strip() for x in item.split("=>", 1)]
        elif "=" in item:
            item, init = [x.strip() for x in item.split("=", 1)]
        mm = re.match(r"^(\w+)\s*(?:\((.*)\))?$", item)
        if not mm:
            raise F2CError("cannot parse declaration item %r" % item)
        d = split_top(mm.group(2)) if mm.group(2) else dims
        out.append(Decl(mm.group(1), base, d, dict(attrs), init))
    return out


# ----------------------------------------------------------------------------------------------------
# subroutines
# ----------------------------------------------------------------------------------------------------
class Sub:
    def __init__(self, name, args, line):
        self.name, self.args, self.line = name, args, line
        self.decls = {}
        self.body = []      # (line, statement)


def collect_subs(text: str, wanted):
    subs, cur = {}, None
    for ln, st in statements(text):
        m = re.match(r"^subroutine\s+(\w+)\s*(?:\((.*)\))?$", st)
        if m:
            cur = Sub(m.group(1), [a.strip() for a in m.group(2).split(",")] if m.group(2) else [], ln)
            continue
        if re.match(r"^end\s*subroutine", st):
            if cur and cur.name in wanted:
                subs[cur.name] = cur
            cur = None
            continue
        if cur is None or cur.name not in wanted:
            continue
        d = parse_decl(st) if not cur.body else None
        if d is not None:
            for x in d:
                cur.decls[x.name] = x
        else:
            cur.body.append((ln, st))
    return subs


def ctype(base) -> str:
    return base if isinstance(base, str) else base[1]


def dim_bounds(em: Emitter, d: str):
    if ":" in d:
        lo, hi = d.split(":", 1)
        return em.e(parse_expr(lo)), em.e(parse_expr(hi))
    return "1", em.e(parse_expr(d))


def signature(sub: Sub) -> str:
    ps = []
    for a in sub.args:
        d = sub.decls.get(a)
        if d is None:
            raise F2CError("%s: dummy argument %s is not declared" % (sub.name, a))
        t = ctype(d.base)
        if d.dims:
            ps.append("Arr<%s,%d> %s__" % (t, len(d.dims), cname(a)))
        elif isinstance(d.base, tuple):
            ps.append("%s *%s" % (t, cname(a)) if d.attrs.get("pointer") else "%s &%s" % (t, cname(a)))
        elif d.attrs.get("optional"):
            ps.append("Opt<%s> %s" % (t, cname(a)))
        elif d.attrs.get("intent") in ("out", "inout"):
            ps.append("%s &%s" % (t, cname(a)))
        else:
            ps.append("%s %s" % (t, cname(a)))
    return "void %s(%s)" % (sub.name, ", ".join(ps))


def translate_range(text: str, lo: int, hi: int, arrays, src_name: str) -> str:
    """Translate the statements whose first line lies in [lo, hi] (a run of balanced blocks inside a subroutine) into a
    bare C++ statement list: the including wrapper declares every variable the statements use (same names, case-folded).
    For code that lives in the middle of a subroutine too entangled to translate whole (tracer_hordiff)."""
    body = [(ln, st) for ln, st in statements(text) if lo <= ln <= hi]
    sub = Sub("range_%d_%d" % (lo, hi), [], lo)
    sub.body = body
    return translate_body(sub, set(arrays), src_name, header=False)


def translate_sub(sub: Sub, all_subs, src_name: str) -> str:
    arrays = {n for n, d in sub.decls.items() if d.dims}
    return translate_body(sub, arrays, src_name, header=True)


def translate_body(sub: Sub, arrays, src_name: str, header: bool) -> str:
    em = Emitter(arrays)
    out = ["// %s:%d  subroutine %s" % (src_name, sub.line, sub.name), signature(sub) + " {"] if header else \
          ["// %s:%d-%d (statement range)" % (src_name, sub.body[0][0], sub.body[-1][0])]
    ind = 1

    def put(s, ln=None):
        out.append("  " * ind + s + ("   // :%d" % ln if ln else ""))

    if not header:
        sub = Sub(sub.name, [], sub.line) if False else sub

    # scalars first (array bounds may use them only after the executable part sets them -- Fortran automatic arrays
    # take their bounds from dummy arguments, which are all set at entry)
    for name, d in (sub.decls.items() if header else []):
        if name in sub.args:
            continue
        t = ctype(d.base)
        if d.dims:
            continue
        if isinstance(d.base, tuple):
            put("%s *%s = nullptr;" % (t, cname(name)) if d.attrs.get("pointer") else "%s %s;" % (t, cname(name)))
        else:
            put("%s %s = f_uninit<%s>();" % (t, cname(name), t))
    for name, d in (sub.decls.items() if header else []):
        if not d.dims:
            continue
        t = ctype(d.base)
        b = ", ".join("{%s, %s}" % dim_bounds(em, x) for x in d.dims)
        if name in sub.args:
            put("Arr<%s,%d> %s = %s__.rebound({%s});" % (t, len(d.dims), cname(name), cname(name), b))
        else:
            put("Arr<%s,%d> %s = Arr<%s,%d>::automatic({%s});" % (t, len(d.dims), cname(name), t, len(d.dims), b))

    def stmt(ln, st):
        nonlocal ind
        # block structure
        m = re.match(r"^do\s+(\w+)\s*=\s*(.*)$", st)
        if m:
            var = cname(m.group(1))
            parts = split_top(m.group(2))
            lo, hi = em.e(parse_expr(parts[0])), em.e(parse_expr(parts[1]))
            if len(parts) == 3:
                step = em.e(parse_expr(parts[2]))
                put("for (int lo__ = %s, hi__ = %s, st__ = %s, n__ = f_trips(lo__, hi__, st__), c__ = (%s = lo__, 0); c__ < n__; c__++, %s += st__) {"
                    % (lo, hi, step, var, var), ln)
            else:
                put("for (int hi__ = (%s = %s, %s); %s <= hi__; %s++) {" % (var, lo, hi, var, var), ln)
            ind += 1
            return
        if re.match(r"^end\s*do$", st):
            ind -= 1
            put("}")
            return
        if st.startswith("if") and re.match(r"^if\s*\(", st):
            close = match_paren(st, st.index("("))
            cond, rest = st[st.index("(") + 1:close], st[close + 1:].strip()
            c = em.e(parse_expr(cond))
            if rest == "then":
                put("if (%s) {" % c, ln)
                ind += 1
            else:
                put("if (%s) {" % c, ln)
                ind += 1
                stmt(None, rest)
                ind -= 1
                put("}")
            return
        m = re.match(r"^else\s*if\s*\(", st)
        if m:
            close = match_paren(st, st.index("("))
            c = em.e(parse_expr(st[st.index("(") + 1:close]))
            if st[close + 1:].strip() != "then":
                raise F2CError("line %s: malformed elseif %r" % (ln, st))
            ind -= 1
            put("} else if (%s) {" % c, ln)
            ind += 1
            return
        if st == "else":
            ind -= 1
            put("} else {", ln)
            ind += 1
            return
        if re.match(r"^end\s*if$", st):
            ind -= 1
            put("}")
            return
        if st == "return":
            put("return;", ln)
            return
        if st == "exit":
            put("break;", ln)
            return
        if st == "cycle":
            put("continue;", ln)
            return
        if re.match(r"^write\s*\(", st):
            # formatted output into a message string / unit: no effect on any arithmetic
            put("/* %s */;" % st.replace("*/", "* /"), ln)
            return
        m = re.match(r"^call\s+(\w+)\s*(\(.*\))?$", st)
        if m:
            name = m.group(1)
            args = []      # (keyword or None, expression)
            if m.group(2):
                inner = m.group(2).strip()[1:-1]
                for piece in split_top(inner):
                    piece = piece.strip()
                    k = re.match(r"^(\w+)\s*=(?!=)\s*(.*)$", piece)
                    if k:
                        args.append((k.group(1), parse_expr(k.group(2))))
                    else:
                        args.append((None, parse_expr(piece)))
            if name == "mom_error":
                put("f_mom_error(%s);" % ", ".join(em.e(a) for _, a in args), ln)
                return
            a_txt = []
            for kw, a in args:
                if isinstance(a, Ref) and em.whole(a.parts[-1][1]):
                    t = em.ref(a)
                else:
                    t = em.e(a)
                # keyword argument name=expr: a tagged value the callee's C++ overload sorts into its optional slot
                a_txt.append("kw_%s(%s)" % (cname(kw), t) if kw else t)
            put("%s(%s);" % (cname(name), ", ".join(a_txt)), ln)
            return
        # pointer assignment
        if "=>" in st:
            lhs, rhs = [x.strip() for x in st.split("=>", 1)]
            put("%s = &(%s);" % (em.e(parse_expr(lhs)), em.e(parse_expr(rhs))), ln)
            return
        # assignment: first top-level '=' that is not part of ==, /=, <=, >=
        eq = find_assign(st)
        if eq < 0:
            raise F2CError("line %s: cannot classify statement %r" % (ln, st))
        lhs, rhs = parse_expr(st[:eq]), parse_expr(st[eq + 1:])
        if em.is_whole_array_target(lhs) and isinstance(rhs, Ref) and rhs.parts[-1][1] and not em.whole(rhs.parts[-1][1]) \
                and sum(isinstance(a, Section) for a in rhs.parts[-1][1]) == 1:
            # rank-1 section of a higher-rank array, e.g. a(:) = b(:,k): element loop over the one sectioned dimension
            args = rhs.parts[-1][1]
            d = [isinstance(a, Section) for a in args].index(True)
            if args[d].lo is not None or args[d].hi is not None:
                raise F2CError("line %s: only full ':' sections are supported" % ln)
            base = Ref(rhs.parts[:-1] + [(rhs.parts[-1][0], None)])
            L, B = em.ref(lhs), em.ref(base)
            idx = ["(s__ - %s.lo[0] + %s.lo[%d])" % (L, B, d) if k == d else em.e(a) for k, a in enumerate(args)]
            put("for (int s__ = %s.lo[0]; s__ <= %s.hi[0]; s__++) %s(s__) = %s(%s);" % (L, L, L, B, ", ".join(idx)), ln)
        elif em.is_whole_array_target(lhs):
            r = em.ref(rhs) if isinstance(rhs, Ref) and em.is_whole_array_target(rhs) else em.e(rhs)
            put("f_assign(%s, %s);" % (em.ref(lhs), r), ln)
        else:
            put("%s = %s;" % (em.e(lhs), em.e(rhs)), ln)

    for ln, st in sub.body:
        stmt(ln, st)
    if ind != 1:
        raise F2CError("%s: unbalanced blocks (indent %d at end)" % (sub.name, ind))
    if header:
        out.append("}")
    return "\n".join(out)


def match_paren(s: str, i: int) -> int:
    depth, q = 0, None
    for k in range(i, len(s)):
        ch = s[k]
        if q:
            if ch == q:
                q = None
            continue
        if ch in "'\"":
            q = ch
        elif ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
            if depth == 0:
                return k
    raise F2CError("unbalanced parentheses in %r" % s)


def find_assign(s: str) -> int:
    depth, q = 0, None
    for k, ch in enumerate(s):
        if q:
            if ch == q:
                q = None
            continue
        if ch in "'\"":
            q = ch
        elif ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        elif ch == "=" and depth == 0:
            prev, nxt = s[k - 1] if k else "", s[k + 1] if k + 1 < len(s) else ""
            if prev in "=/<>" or nxt in "=>":
                continue
            return k
    return -1


def main(argv):
    if len(argv) < 3:
        print(__doc__)
        return 2
    path, wanted = argv[1], [a.lower() for a in argv[2:]]
    text = open(path).read()
    if wanted and wanted[0] == "--range":      # f2cpp.py FILE --range LO HI [array names...]
        lo, hi = int(wanted[1]), int(wanted[2])
        print("// GENERATED by oracle/f2c/f2cpp.py from %s lines %d-%d -- do not edit, do not commit." % (path, lo, hi))
        print(translate_range(text, lo, hi, wanted[3:], path.split("/")[-1]))
        return 0
    subs = collect_subs(text, set(wanted))
    missing = [w for w in wanted if w not in subs]
    if missing:
        raise F2CError("subroutines not found: %s" % missing)
    print("// GENERATED by oracle/f2c/f2cpp.py from %s -- do not edit, do not commit." % path)
    for w in wanted:      # forward declarations
        print(signature(subs[w]) + ";")
    for w in wanted:
        print()
        print(translate_sub(subs[w], subs, path.split("/")[-1]))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
