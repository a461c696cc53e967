"""Mechanical execution of the reference's own Fortran statements (TEST INFRASTRUCTURE).

No Fortran compiler exists in this image, so the arithmetic subroutines of the path cannot be
compiled -- but they are written in a small, regular subset of Fortran 90 (do loops, block and
one-line ifs, assignments to scalars / array elements, calls, intrinsics sqrt exp max min abs sign),
and in that subset an array reference `a(i,j)` is also a valid Python call.  This module reads a
subroutine straight out of the reference checkout, applies the C preprocessor conditionals, joins
continuation lines, rewrites the control statements line by line into Python and executes the result
with IEEE doubles (Python floats: same operations, same left-to-right order, no contraction).
Nothing is transcribed by hand: the statements that run are the reference's text.

It is used ONLY by tests/golden/make_reference_vectors.py, in the build container (the reference
does not travel to the GPU box); the vectors it produces are committed under tests/golden/.

Not supported (not needed by the subroutines used): where/forall, array-valued expressions other
than `a(:,:) = scalar`, derived types, I/O, goto, select case, character data.
"""
from __future__ import annotations

import math
import re

import numpy as np


# ---------------------------------------------------------------------------------------------
# Fortran arrays: explicit lower bounds, element read by call, element write by item assignment
# ---------------------------------------------------------------------------------------------
class FArr:
    def __init__(self, bounds, dtype=float, fill=0):
        """bounds: [(lo, hi), ...] in Fortran order (first index fastest)."""
        self.lo = tuple(int(b[0]) for b in bounds)
        shape = tuple(int(b[1]) - int(b[0]) + 1 for b in bounds)
        self.a = np.full(shape, fill, dtype=dtype)

    @classmethod
    def from_numpy(cls, arr, los=None):
        """arr indexed [j-1, i-1] (C order, as the repo's (ny_block, nx_block) planes) or 1-D / 3-D
        [k-1, j-1, i-1]; becomes a Fortran array (i, j[, k]) with lower bounds `los` (default 1)."""
        f = cls.__new__(cls)
        f.a = np.array(np.transpose(arr))            # own copy, index order (i, j, k)
        f.lo = tuple(los) if los is not None else (1,) * f.a.ndim
        return f

    def to_numpy(self):
        return np.ascontiguousarray(np.transpose(self.a))

    def _ix(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        return tuple(int(i) - l for i, l in zip(idx, self.lo))

    def __call__(self, *idx):
        ix = self._ix(idx)
        if any(i < 0 for i in ix):
            raise IndexError(f"index {idx} below lower bounds {self.lo}")
        v = self.a[ix]
        return v.item()

    def __setitem__(self, idx, val):
        ix = self._ix(idx)
        if any(i < 0 for i in ix):
            raise IndexError(f"index {idx} below lower bounds {self.lo}")
        self.a[ix] = val

    def fill(self, val):
        self.a[...] = val


def _sign(a, b):
    return math.copysign(abs(a), b)


def _real(x, *_, **__):
    return float(x)


INTRINSICS = {"sqrt": math.sqrt, "exp": math.exp, "max": max, "min": min, "abs": abs, "sign": _sign,
              "real": _real, "present": lambda x: x is not None, "int": int, "mod": lambda a, b: math.fmod(a, b)}


# ---------------------------------------------------------------------------------------------
# source preparation
# ---------------------------------------------------------------------------------------------
def preprocess(text: str, defined=()) -> list:
    """C preprocessor conditionals (#ifdef / #ifndef / #if [!]defined(X) [&& ||] / #else / #endif),
    comments, continuation lines.  Returns logical lines, lower-cased."""
    defined = {d.lower() for d in defined}
    out, stack = [], []          # stack of (parent_active, this_branch_active, any_branch_taken)

    def active():
        return all(s[1] for s in stack) if stack else True

    def evaluate(cond):
        c = cond.lower()
        c = re.sub(r"defined\s*\(?\s*(\w+)\s*\)?", lambda m: str(m.group(1) in defined), c)
        c = c.replace("&&", " and ").replace("||", " or ").replace("!", " not ")
        c = re.sub(r"\b([a-z_]\w*)\b", lambda m: m.group(1) if m.group(1) in ("and", "or", "not", "true", "false")
                   else str(m.group(1) in defined), c)
        c = c.replace("true", "True").replace("false", "False")
        return bool(eval(c))

    for raw in text.splitlines():
        s = raw.strip()
        if s.startswith("#"):
            d = s[1:].strip()
            if d.startswith("ifdef"):
                stack.append([active(), d.split()[1].lower() in defined, False])
            elif d.startswith("ifndef"):
                stack.append([active(), d.split()[1].lower() not in defined, False])
            elif d.startswith("if"):
                stack.append([active(), evaluate(d[2:]), False])
            elif d.startswith("elif"):
                taken = stack[-1][1] or stack[-1][2]
                stack[-1][2] = taken
                stack[-1][1] = (not taken) and evaluate(d[4:])
            elif d.startswith("else"):
                taken = stack[-1][1] or stack[-1][2]
                stack[-1][2] = taken
                stack[-1][1] = not taken
            elif d.startswith("endif"):
                stack.pop()
            continue
        if not active():
            continue
        out.append(raw)

    # comments (no character data with '!' in the routines used), blank lines, continuations
    logical, cur = [], ""
    for raw in out:
        line = raw.split("!")[0].rstrip()
        if not line.strip():
            continue
        line = line.strip()
        if line.startswith("&"):
            line = line[1:].strip()
        if line.endswith("&"):
            cur += line[:-1] + " "
            continue
        cur += line
        logical.append(re.sub(r"\s+", " ", cur).strip().lower())
        cur = ""
    return logical


def find_subroutine(lines: list, name: str) -> list:
    name = name.lower()
    start = next(i for i, l in enumerate(lines) if re.match(rf"subroutine {name}\b", l))
    end = next(i for i in range(start, len(lines)) if re.match(rf"end subroutine( {name})?\b", lines[i]))
    return lines[start:end]


_DECL = re.compile(r"^(integer|real|logical|character|type\s*\()")


def split_top(s: str, sep=",") -> list:
    """Split at separators that are not inside parentheses."""
    parts, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == sep and depth == 0:
            parts.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        parts.append(cur.strip())
    return parts


def parse_parameters(lines: list, ns: dict) -> dict:
    """Evaluate `..., parameter :: a = expr, b = expr` declarations in order into ns."""
    for l in lines:
        if "parameter" in l and "::" in l and _DECL.match(l):
            for item in split_top(l.split("::", 1)[1]):
                if "=" in item:
                    k, v = item.split("=", 1)
                    ns[k.strip()] = eval(expr(v), dict(INTRINSICS), ns)
    return ns


# ---------------------------------------------------------------------------------------------
# expression / statement rewriting
# ---------------------------------------------------------------------------------------------
_POWOP = re.compile(r"\*\*\s*([23])\b")


def _ipow(x, n):
    r = x
    for _ in range(n - 1):
        r = r * x
    return r


INTRINSICS["_ipow"] = _ipow


def _rewrite_pow(e: str) -> str:
    """operand**N (N = 2, 3; operand = name, name(...), (...)) -> _ipow(operand, N)."""
    while True:
        m = _POWOP.search(e)
        if not m:
            return e
        end = m.start()
        while end > 0 and e[end - 1] == " ":
            end -= 1
        i = end
        if e[i - 1] == ")":
            depth = 0
            while True:
                i -= 1
                if e[i] == ")":
                    depth += 1
                elif e[i] == "(":
                    depth -= 1
                    if depth == 0:
                        break
        while i > 0 and (e[i - 1].isalnum() or e[i - 1] in "_."):
            i -= 1
        e = e[:i] + "_ipow(" + e[i:end] + ", " + m.group(1) + ")" + e[m.end():]


def expr(e: str) -> str:
    e = re.sub(r"([\d.])_(dbl_kind|int_kind|real_kind|r8)\b", r"\1", e)   # 1.0_dbl_kind, 2106._dbl_kind
    e = re.sub(r"(\d\.?\d*)d([+-]?\d+)", r"\1e\2", e)                     # 1.0d0
    e = e.replace("/=", "!=")
    e = re.sub(r"\s*%\s*", ".", e)                                        # this_block%ilo
    for f, p in ((".and.", " and "), (".or.", " or "), (".not.", " not "), (".true.", " True "),
                 (".false.", " False "), (".eq.", "=="), (".ne.", "!="), (".gt.", ">"), (".lt.", "<"),
                 (".ge.", ">="), (".le.", "<=")):
        e = e.replace(f, p)
    e = re.sub(r",\s*kind\s*=\s*\w+", "", e)                              # real(x, kind=dbl_kind)
    # integer powers as repeated products, the way compilers expand them (x**2 -> x*x, x**3 -> (x*x)*x);
    # Python's float ** goes through libm pow(), which is not the same operation
    e = _rewrite_pow(e)
    return e.strip()


def _matching_paren(s: str, start: int) -> int:
    depth = 0
    for i in range(start, len(s)):
        if s[i] == "(":
            depth += 1
        elif s[i] == ")":
            depth -= 1
            if depth == 0:
                return i
    raise ValueError(s)


def _assignment(stmt: str) -> str:
    # find the '=' of the assignment: first '=' at depth 0 that is not part of ==, /=, <=, >=
    depth = 0
    for i, ch in enumerate(stmt):
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        elif ch == "=" and depth == 0 and stmt[i + 1:i + 2] != "=" and stmt[i - 1] not in "=/<>":
            lhs, rhs = stmt[:i].strip(), stmt[i + 1:].strip()
            m = re.match(r"^([a-z_]\w*)\s*\((.*)\)$", lhs)
            if m:
                idx = m.group(2)
                if re.fullmatch(r"[\s:,]*", idx):                          # a(:,:) = scalar
                    return f"{m.group(1)}.fill({expr(rhs)})"
                return f"{m.group(1)}[{expr(idx)}] = {expr(rhs)}"
            return f"{lhs} = {expr(rhs)}"
    raise ValueError("not an assignment: " + stmt)


def _simple(stmt: str, callable_names) -> str:
    if stmt.startswith("call "):
        m = re.match(r"call ([a-z_]\w*)\s*(\(.*\))?$", stmt)
        if m.group(1) not in callable_names:
            return "pass"                                                  # timers, abort_ice, ...
        return f"{m.group(1)}{expr(m.group(2) or '()')}"
    if stmt == "return":
        return "return"
    if stmt in ("continue",) or stmt.startswith(("write", "print")):
        return "pass"
    return _assignment(stmt)


def translate(sub_lines: list, callable_names=()) -> tuple:
    """(python source of a function body, dummy argument names, local array declarations)."""
    header = sub_lines[0]
    m = re.match(r"subroutine ([a-z_]\w*)\s*\((.*)\)", header)
    name, args = m.group(1), [a.strip() for a in m.group(2).split(",")]
    decl_arrays = {}            # name -> list of "lo:hi" strings
    optional = set()
    body, ind = [], 1
    for l in sub_lines[1:]:
        if l.startswith(("use ", "implicit ", "include ", "save", "external ")):
            continue
        if _DECL.match(l) and "::" in l:
            attrs, names = l.split("::", 1)
            dm = re.search(r"dimension\s*\((.*?)\)\s*(,|$)", attrs)
            common = None
            if dm:
                k = attrs.index("dimension")
                close = _matching_paren(attrs, attrs.index("(", k))
                common = split_top(attrs[attrs.index("(", k) + 1:close])
            for item in split_top(names):
                im = re.match(r"^([a-z_]\w*)\s*(\((.*)\))?", item)
                nm = im.group(1)
                dims = split_top(im.group(3)) if im.group(3) else common
                if dims:
                    decl_arrays[nm] = (dims, "logical" if attrs.startswith("logical") else
                                       "integer" if attrs.startswith("integer") else "real")
                if "optional" in attrs:
                    optional.add(nm)
            continue
        pad = "    " * ind
        if re.match(r"^do [a-z_]\w*\s*=", l):
            mm = re.match(r"^do ([a-z_]\w*)\s*=\s*(.*)$", l)
            parts = split_top(mm.group(2))
            lo, hi = expr(parts[0]), expr(parts[1])
            step = expr(parts[2]) if len(parts) > 2 else "1"
            body.append(f"{pad}for {mm.group(1)} in _frange({lo}, {hi}, {step}):")
            ind += 1
        elif l in ("enddo", "end do", "endif", "end if"):
            ind -= 1
        elif re.match(r"^(else\s*if|elseif)\s*\(", l):
            close = _matching_paren(l, l.index("("))
            body.append("    " * (ind - 1) + f"elif {expr(l[l.index('(') + 1:close])}:")
        elif l == "else":
            body.append("    " * (ind - 1) + "else:")
        elif l.startswith("if"):
            close = _matching_paren(l, l.index("("))
            cond, rest = expr(l[l.index("(") + 1:close]), l[close + 1:].strip()
            if rest == "then":
                body.append(f"{pad}if {cond}:")
                ind += 1
            else:
                body.append(f"{pad}if {cond}: {_simple(rest, callable_names)}")
        else:
            body.append(pad + _simple(l, callable_names))
    locals_ = {k: v for k, v in decl_arrays.items() if k not in args}
    return name, args, optional, locals_, body


def _frange(lo, hi, step=1):
    return range(int(lo), int(hi) + (1 if step > 0 else -1), int(step))


class Module:
    """Subroutines of one or more reference files, executable; `ns` holds module-level names."""

    def __init__(self, ns=None):
        self.ns = dict(INTRINSICS)
        self.ns["_frange"] = _frange
        self.ns["FArr"] = FArr
        if ns:
            self.ns.update(ns)
        self.source = {}

    def load(self, lines: list, names, callable_names=()):
        for n in names:
            name, args, optional, locals_, body = translate(find_subroutine(lines, n), tuple(callable_names) + tuple(names))
            sig = ", ".join(a + ("=None" if a in optional else "") for a in args)
            pre = []
            for k, (dims, typ) in locals_.items():
                b = ", ".join("(" + (d.split(":")[0] if ":" in d else "1") + ", " + d.split(":")[-1] + ")" for d in
                              (expr(x) for x in dims))
                dt = {"logical": "bool", "integer": "int", "real": "float"}[typ]
                pre.append(f"    {k} = FArr([{b}], dtype={dt})")
            src = f"def {name}({sig}):\n" + "\n".join(pre + body) + "\n    return locals()\n"
            self.source[name] = src
            exec(compile(src, f"<reference:{name}>", "exec"), self.ns)
        return self

    def call(self, name, **kw):
        return self.ns[name.lower()](**{k.lower(): v for k, v in kw.items()})
