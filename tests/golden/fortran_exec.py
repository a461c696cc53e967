"""Mechanical execution of the reference's own Fortran statements (TEST INFRASTRUCTURE).

No Fortran compiler exists in this image, so the subroutines of the path cannot be compiled -- but
they are written in a small, regular subset of Fortran 90, and in that subset an array reference
`a(i,j)` is also a valid Python call.  This module reads a subroutine or function straight out of
the reference checkout, applies the C preprocessor conditionals, joins continuation lines, rewrites
the statements line by line into Python and executes the result with IEEE doubles (Python floats:
same operations, same left-to-right order, no contraction; integer `/` truncates as in Fortran).
Nothing is transcribed by hand: the statements that run are the reference's text.

Subset: declarations (explicit-shape, deferred-shape, pointer, optional, intent, local parameters),
do / do while / named do, block and one-line if, select case (values, ranges, strings, default),
assignments to scalars, elements, sections and whole arrays, pointer assignment, derived-type
components (`halo%srcLocalAddr(1,n)`, `all_blocks(n)%ilo`), allocate / deallocate / allocated,
array constructors, calls with positional and keyword arguments whose non-intent(in) scalars are
copied back, functions with a result variable, module variables (globals of the Module), the
intrinsics sqrt exp max min abs sign mod real int size count any maxval minval huge present trim.
A statement outside the subset (I/O aside, which is skipped) raises if execution reaches it
(`UNTRANSLATED` lists them); calls to routines that were not loaded (timers, MPI_*) are no-ops.

It is used ONLY by tests/golden/make_reference_vectors.py and tests/golden/make_halo_vectors.py, in
the build container (the reference does not travel to the GPU box); the vectors they produce are
committed under tests/golden/.
"""
from __future__ import annotations

import ast
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
        if len(idx) != self.a.ndim:
            raise IndexError(f"rank mismatch: {len(idx)} subscripts for rank {self.a.ndim}")
        out = []
        for i, l in zip(idx, self.lo):
            if isinstance(i, slice):
                out.append(slice(None if i.start is None else int(i.start) - l,
                                 None if i.stop is None else int(i.stop) - l + 1))
            else:
                k = int(i) - l
                if k < 0:
                    raise IndexError(f"index {idx} below lower bounds {self.lo}")
                out.append(k)
        return tuple(out)

    def __call__(self, *idx):
        ix = self._ix(idx)
        if any(isinstance(i, slice) for i in ix):          # array section: a view, lower bounds 1
            f = FArr.__new__(FArr)
            f.a = self.a[ix]
            f.lo = (1,) * f.a.ndim
            return f
        v = self.a[ix]
        return v.item() if isinstance(v, np.generic) else v

    def __setitem__(self, idx, val):
        ix = self._ix(idx)
        self.a[ix] = val.a if isinstance(val, FArr) else val

    def fill(self, val):
        self.a[...] = val.a if isinstance(val, FArr) else val

    # whole-array comparisons (count(a /= 0), any(abs(a) /= n)): plain numpy logical arrays
    def __ne__(self, o):
        return self.a != (o.a if isinstance(o, FArr) else o)

    def __eq__(self, o):
        return self.a == (o.a if isinstance(o, FArr) else o)

    def __gt__(self, o):
        return self.a > (o.a if isinstance(o, FArr) else o)

    def __lt__(self, o):
        return self.a < (o.a if isinstance(o, FArr) else o)

    __hash__ = object.__hash__


class Obj:
    """A derived-type variable: components are attributes."""

    def __repr__(self):
        return "Obj(" + ", ".join(f"{k}={v!r}" for k, v in vars(self).items() if not isinstance(v, FArr)) + ")"


class _Null:
    """A disassociated pointer (present as an actual argument, unlike None = absent optional)."""

    def __repr__(self):
        return "NULL"


NULL = _Null()


def _sign(a, b):
    return math.copysign(abs(a), b)


def _real(x, *_, **__):
    return float(x)


def _mod(a, b):
    r = math.fmod(a, b)
    return int(r) if isinstance(a, int) and isinstance(b, int) else r


def _div(a, b):
    """Fortran `/`: truncating for two integers, IEEE division otherwise."""
    if isinstance(a, int) and isinstance(b, int) and not isinstance(a, bool):
        q = abs(a) // abs(b)
        return q if (a >= 0) == (b >= 0) else -q
    return a / b


def _size(x, dim=None):
    return x.a.shape[dim - 1] if dim else x.a.size


def _aset(cur, val):
    """`name = value` where name may be a whole array (filled in place) or a scalar (rebound)."""
    if isinstance(cur, FArr):
        cur.fill(val)
        return cur
    return val


def _setcomp(obj, comp, val):
    """`obj%comp = value`: fills an array component, rebinds anything else."""
    cur = getattr(obj, comp, None)
    if isinstance(cur, FArr) and not isinstance(val, FArr):
        cur.fill(val)
    else:
        setattr(obj, comp, val)


def _out(f, r, key, old):
    """Value of a non-intent(in) dummy after a call, for writing back into the actual argument."""
    names = getattr(f, "__fargs__", None)
    if names is None or not isinstance(r, dict):
        return old
    if isinstance(key, int):
        if key >= len(names):
            return old
        key = names[key]
    return r[key] if key in getattr(f, "__fout__", ()) and key in r else old


def _alloc(bounds, kind=None):
    a = FArr(bounds, dtype=object)
    if kind == "derived":
        for k in np.ndindex(a.a.shape):
            a.a[k] = Obj()
    return a


def _abort(msg=""):
    raise RuntimeError("abort_ice: " + str(msg))


def fstub(args, out=()):
    """Mark a Python stand-in for a Fortran subroutine: dummy names in order, non-intent(in) ones."""
    def deco(fn):
        fn.__fargs__, fn.__fout__ = list(args), set(out)
        return fn
    return deco


def _abs(x):
    return np.abs(x.a) if isinstance(x, FArr) else abs(x)


INTRINSICS = {"sqrt": math.sqrt, "exp": math.exp, "max": max, "min": min, "abs": _abs, "sign": _sign,
              "any": lambda x: bool(np.any(x)), "huge": lambda x: 2147483647 if isinstance(x, int) else 1.7976931348623157e308,
              "_vec": lambda v: np.array(v, dtype=object), "count": lambda x: int(np.count_nonzero(x)),
              "maxval": lambda x: max(x.a.ravel().tolist()) if _size(x) else -2147483647,
              "minval": lambda x: min(x.a.ravel().tolist()) if _size(x) else 2147483647,
              "real": _real, "present": lambda x: x is not None, "int": int, "mod": _mod, "_div": _div,
              "size": _size, "allocated": lambda x: isinstance(x, FArr), "associated": lambda x: isinstance(x, FArr),
              "trim": lambda s: s.rstrip(), "_aset": _aset, "_setcomp": _setcomp, "_out": _out, "_alloc": _alloc,
              "_all": slice(None), "Obj": Obj, "NULL": NULL, "abort_ice": _abort}


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
    """Logical lines of `subroutine name` / `function name`, header included, `end` line excluded."""
    name = name.lower()
    head = re.compile(rf"^(recursive )?(subroutine|function) {name}\b")
    start = next(i for i, l in enumerate(lines) if head.match(l))
    end = next(i for i in range(start + 1, len(lines)) if re.match(r"^end ?(subroutine|function)\b", lines[i]))
    return lines[start:end]


_DECL = re.compile(r"^(integer|real|double precision|logical|character|type\s*\()")


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
    e = e.replace("(/", "_vec([").replace("/)", "])")                      # array constructor
    e = e.replace("/=", "!=")
    e = re.sub(r"\s*%\s*", ".", e)                                        # this_block%ilo
    for f, p in ((".and.", " and "), (".or.", " or "), (".not.", " not "), (".true.", " True "),
                 (".false.", " False "), (".eq.", "=="), (".ne.", "!="), (".gt.", ">"), (".lt.", "<"),
                 (".ge.", ">="), (".le.", "<=")):
        e = e.replace(f, p)
    e = re.sub(r",\s*kind\s*=\s*\w+", "", e)                              # real(x, kind=dbl_kind)
    e = re.sub(r"(?<=[(,])\s*:\s*(?=[,)])", "_all", e)                    # a(:,n): whole extent
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


_NAME = re.compile(r"^[a-z_]\w*$")


def _top_level(s: str, ch: str) -> int:
    """Index of the LAST occurrence of ch outside parentheses, or -1."""
    depth, found = 0, -1
    for i, c in enumerate(s):
        if c == "(":
            depth += 1
        elif c == ")":
            depth -= 1
        elif c == ch and depth == 0:
            found = i
    return found


def _split_assignment(stmt: str):
    """(lhs, op, rhs) with op '=' or '=>'; None if stmt is not an assignment."""
    depth, quote = 0, None
    for i, ch in enumerate(stmt):
        if quote:
            if ch == quote:
                quote = None
        elif ch in "'\"":
            quote = ch
        elif ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        elif ch == "=" and depth == 0 and stmt[i + 1:i + 2] != "=" and stmt[i - 1] not in "=/<>":
            if stmt[i + 1:i + 2] == ">":
                return stmt[:i].strip(), "=>", stmt[i + 2:].strip()
            return stmt[:i].strip(), "=", stmt[i + 1:].strip()
    return None


class _Ctx:
    """What the translator knows about the routine being translated."""

    def __init__(self, callable_names, derived_arrays):
        self.callable = set(callable_names)
        self.derived = set(derived_arrays)
        self.scalars = set()      # declared here, scalar (dummies included)
        self.arrays = set()       # declared here, array (explicit or deferred shape)
        self.assigned = set()     # simple names rebound by a statement
        self.result = None        # function result variable

    def declared(self, name):
        return name in self.scalars or name in self.arrays


def _assignment(stmt: str, ctx: _Ctx) -> str:
    parts = _split_assignment(stmt)
    if parts is None:
        raise ValueError("not an assignment: " + stmt)
    lhs, op, rhs = parts
    value = expr(rhs)
    k = _top_level(lhs, "%")
    if k >= 0:                                                             # component of a derived type
        base, comp = expr(lhs[:k]), lhs[k + 1:].strip()
        m = re.match(r"^([a-z_]\w*)\s*\((.*)\)$", comp)
        if m:
            return f"{base}.{m.group(1)}[{expr(m.group(2))}] = {value}"
        if op == "=>":
            return f"{base}.{comp} = {value}"
        return f"_setcomp({base}, {comp!r}, {value})"
    m = re.match(r"^([a-z_]\w*)\s*\((.*)\)$", lhs)
    if m:
        idx = m.group(2)
        if re.fullmatch(r"[\s:,]*", idx):                                  # a(:,:) = scalar or array
            return f"{m.group(1)}.fill({value})"
        return f"{m.group(1)}[{expr(idx)}] = {value}"
    ctx.assigned.add(lhs)
    if op == "=>" or lhs in ctx.scalars:
        return f"{lhs} = {value}"
    if lhs in ctx.arrays:
        return f"{lhs} = _aset({lhs}, {value})"
    return f"{lhs} = _aset(globals().get({lhs!r}), {value})"               # module variable


def _call(stmt: str, ctx: _Ctx) -> list:
    m = re.match(r"call ([a-z_]\w*)\s*(\((.*)\))?$", stmt)
    name, args = m.group(1), split_top(m.group(3) or "")
    if name == "abort_ice":
        return [f"abort_ice({m.group(3) or ''})"]
    if name not in ctx.callable:
        return ["pass"]                                                    # timers, diagnostics
    actual, back = [], []
    for pos, a in enumerate(args):
        km = re.match(r"^([a-z_]\w*)\s*=(?!=)\s*(.*)$", a)
        key, val = (repr(km.group(1)), km.group(2).strip()) if km else (str(pos), a)
        actual.append((km.group(1) + "=" if km else "") + expr(val))
        if _NAME.match(val) and ctx.declared(val):                         # a variable that can come back
            back.append(f"{val} = _out({name}, _r, {key}, {val})")
            ctx.assigned.add(val)
            continue
        em = re.match(r"^([a-z_]\w*)\s*\((.*)\)$", val)                    # an array element: icellt(iblk)
        if em and em.group(1) in ctx.arrays and ":" not in em.group(2):
            back.append(f"{em.group(1)}[{expr(em.group(2))}] = _out({name}, _r, {key}, {expr(val)})")
    return [f"_r = {name}({', '.join(actual)})"] + back


def _simple(stmt: str, ctx: _Ctx) -> list:
    if stmt.startswith("call "):
        return _call(stmt, ctx)
    if stmt == "return":
        return ["return " + (ctx.result or "locals()")]
    if stmt == "exit":
        return ["break"]
    if stmt == "cycle":
        return ["continue"]
    if stmt == "continue" or stmt.startswith(("write", "print", "nullify", "format")):
        return ["pass"]
    m = re.match(r"^allocate\s*\((.*)\)$", stmt)
    if m:
        out = []
        for item in split_top(m.group(1)):
            if re.match(r"^stat\s*=", item):
                out.append(f"{item.split('=')[1].strip()} = 0")
                continue
            open_ = item.rindex("(", 0, _matching_open(item))
            target, dims = item[:open_].strip(), split_top(item[open_ + 1:-1])
            bounds = ", ".join("(" + (expr(d.split(":")[0]) if ":" in d else "1") + ", " + expr(d.split(":")[-1]) + ")"
                               for d in dims)
            kind = "'derived'" if target in ctx.derived else "None"
            if "%" in target:
                out.append(f"{expr(target)} = _alloc([{bounds}], {kind})")
            else:
                ctx.assigned.add(target)
                out.append(f"{target} = _alloc([{bounds}], {kind})")
        return out
    m = re.match(r"^deallocate\s*\((.*)\)$", stmt)
    if m:
        out = []
        for item in split_top(m.group(1)):
            if re.match(r"^stat\s*=", item):
                out.append(f"{item.split('=')[1].strip()} = 0")
            elif "%" in item:
                out.append(f"{expr(item)} = None")
            else:
                ctx.assigned.add(item)
                out.append(f"{item} = None")
        return out
    return [_assignment(stmt, ctx)]


def _matching_open(item: str) -> int:
    """Index just past the '(' that matches the final ')' of item (item ends with ')')."""
    depth = 0
    for i in range(len(item) - 1, -1, -1):
        if item[i] == ")":
            depth += 1
        elif item[i] == "(":
            depth -= 1
            if depth == 0:
                return i + 1
    raise ValueError(item)


_PYKEYWORD = re.compile(r"(?<![.\w])(is|as|from|lambda|pass|del|def|class|with|yield|try|raise|global|import|assert|"
                        r"except|finally|nonlocal|break|for|len)(?![.\w])")
UNTRANSLATED = []      # (routine, statement, why): statements that raise if execution reaches them

_TYPE_INIT = (("integer", "0"), ("real", "0.0"), ("double", "0.0"), ("logical", "False"), ("character", "''"),
              ("type", "Obj()"))


def translate(sub_lines: list, callable_names=(), derived_arrays=()):
    """One subroutine / function of the reference -> the source of a Python function.

    Returns (name, python source).  Subroutines return locals() (the caller copies the non-intent(in)
    scalars back); functions return their result variable."""
    ctx = _Ctx(callable_names, derived_arrays)
    m = re.match(r"^(?:recursive )?(subroutine|function) ([a-z_]\w*)\s*(?:\((.*?)\))?\s*(?:result\s*\(\s*(\w+)\s*\))?$",
                 sub_lines[0])
    kind, name = m.group(1), m.group(2)
    args = [a.strip() for a in (m.group(3) or "").split(",") if a.strip()]
    if kind == "function":
        ctx.result = m.group(4) or name
    optional, not_in, init = set(), set(), []
    body, ind, selects = [], 1, []

    def emit(stmts, level=None):
        pad = "    " * (ind if level is None else level)
        body.extend(pad + s for s in stmts)

    for l in sub_lines[1:]:
        l = _PYKEYWORD.sub(r"\1_", l)                                       # `is`, `lambda`, ... as variable names
        if l.startswith(("use ", "implicit ", "include ", "save", "external ", "intrinsic ")) or l == "contains":
            continue
        if _DECL.match(l) and "::" in l:
            attrs, names = l.split("::", 1)
            common = None
            if "dimension" in attrs:
                k = attrs.index("dimension")
                close = _matching_paren(attrs, attrs.index("(", k))
                common = split_top(attrs[attrs.index("(", k) + 1:close])
            typ = next(v for key, v in _TYPE_INIT if attrs.startswith(key))
            for item in split_top(names):
                im = re.match(r"^([a-z_]\w*)\s*(\((.*)\))?\s*(=.*)?$", item)
                nm = im.group(1)
                dims = split_top(im.group(3)) if im.group(3) else common
                if "optional" in attrs:
                    optional.add(nm)
                if "intent(in)" not in attrs.replace(" ", ""):
                    not_in.add(nm)
                if "parameter" in attrs:                                   # local named constant
                    ctx.scalars.add(nm)
                    init.append(f"{nm} = {expr(im.group(4)[1:])}")
                    continue
                if dims:
                    ctx.arrays.add(nm)
                    if nm in args:
                        continue
                    if any(d.strip() == ":" for d in dims):                # deferred shape
                        init.append(f"{nm} = {'NULL' if 'pointer' in attrs else 'None'}")
                    else:
                        b = ", ".join("(" + (expr(d.split(":")[0]) if ":" in d else "1") + ", " +
                                      expr(d.split(":")[-1]) + ")" for d in dims)
                        dt = {"False": "bool", "0": "int", "0.0": "float"}.get(typ, "object")
                        init.append(f"{nm} = FArr([{b}], dtype={dt})")
                else:
                    ctx.scalars.add(nm)
                    if nm not in args:
                        init.append(f"{nm} = {'NULL' if 'pointer' in attrs else typ}")
            continue
        l = re.sub(r"^[a-z_]\w*\s*:\s*(?=do\b)", "", l)                     # construct names
        if re.match(r"^do [a-z_]\w*\s*=", l):
            mm = re.match(r"^do ([a-z_]\w*)\s*=\s*(.*)$", l)
            parts = split_top(mm.group(2))
            step = expr(parts[2]) if len(parts) > 2 else "1"
            emit([f"for {mm.group(1)} in _frange({expr(parts[0])}, {expr(parts[1])}, {step}):"])
            ind += 1
            emit(["pass"])
        elif re.match(r"^do while\s*\(", l):
            emit([f"while {expr(l[l.index('(') + 1:_matching_paren(l, l.index('('))])}:"])
            ind += 1
            emit(["pass"])
        elif re.match(r"^end ?(do|if)\b", l):
            ind -= 1
        elif re.match(r"^select case\s*\(", l):
            depth = len(selects)
            emit([f"_sel{depth} = {expr(l[l.index('(') + 1:_matching_paren(l, l.index('('))])}"])
            selects.append(True)                                           # True: no case emitted yet
            ind += 1
        elif l.startswith("case"):
            depth = len(selects) - 1
            if l.replace(" ", "") == "casedefault":
                emit(["else:" if not selects[depth] else "if True:"], ind - 1)
            else:
                tests = []
                for item in split_top(l[l.index("(") + 1:_matching_paren(l, l.index("("))]):
                    if _top_level(item, ":") >= 0:
                        lo, hi = item.split(":")
                        tests.append(f"({expr(lo)} <= _sel{depth} <= {expr(hi)})")
                    else:
                        tests.append(f"_sel{depth} == {expr(item)}")
                emit([("if " if selects[depth] else "elif ") + " or ".join(tests) + ":"], ind - 1)
            selects[depth] = False
            emit(["pass"])
        elif re.match(r"^end ?select\b", l):
            selects.pop()
            ind -= 1
        elif re.match(r"^(else\s*if|elseif)\s*\(", l):
            close = _matching_paren(l, l.index("("))
            emit([f"elif {expr(l[l.index('(') + 1:close])}:"], ind - 1)
            emit(["pass"])
        elif l == "else":
            emit(["else:"], ind - 1)
            emit(["pass"])
        elif re.match(r"^if\s*\(", l):
            close = _matching_paren(l, l.index("("))
            cond, rest = expr(l[l.index("(") + 1:close]), l[close + 1:].strip()
            emit([f"if {cond}:"])
            if rest == "then":
                ind += 1
                emit(["pass"])
            else:
                emit(_simple(rest, ctx), ind + 1)
        else:
            try:
                stmts = _simple(l, ctx)
                for s_ in stmts:
                    ast.parse(s_)
                emit(stmts)
            except Exception as e:                                         # outside the subset: fails if reached
                UNTRANSLATED.append((name, l, repr(e)))
                emit([f"raise NotImplementedError({l!r})"])
    glob = sorted(n for n in ctx.assigned if not ctx.declared(n) and n != ctx.result and _NAME.match(n))
    if ctx.result and not ctx.declared(ctx.result):
        init.insert(0, f"{ctx.result} = 0")
    sig = ", ".join(a + ("=None" if a in optional else "") for a in args)
    head = [f"def {name}({sig}):"] + (["    global " + ", ".join(glob)] if glob else []) + ["    " + s for s in init]
    tail = ["    return " + (ctx.result or "locals()")]
    return name, args, not_in, "\n".join(head + body + tail) + "\n"


def _frange(lo, hi, step=1):
    return range(int(lo), int(hi) + (1 if step > 0 else -1), int(step))


class _FortranDivision(ast.NodeTransformer):
    """`/` between two integers truncates in Fortran: route every division through _div()."""

    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            return ast.copy_location(ast.Call(func=ast.Name(id="_div", ctx=ast.Load()), args=[node.left, node.right],
                                              keywords=[]), node)
        return node


class Module:
    """Subroutines of one or more reference files, executable; `ns` holds module-level names."""

    def __init__(self, ns=None):
        self.ns = dict(INTRINSICS)
        self.ns["_frange"] = _frange
        self.ns["FArr"] = FArr
        if ns:
            self.ns.update(ns)
        self.source = {}

    def load(self, lines: list, names, callable_names=(), derived_arrays=()):
        known = set(callable_names) | set(n.lower() for n in names) | \
            {k for k, v in self.ns.items() if hasattr(v, "__fargs__")}
        for n in names:
            name, args, not_in, src = translate(find_subroutine(lines, n), known, derived_arrays)
            self.source[name] = src
            tree = ast.fix_missing_locations(_FortranDivision().visit(ast.parse(src)))
            exec(compile(tree, f"<reference:{name}>", "exec"), self.ns)
            self.ns[name].__fargs__, self.ns[name].__fout__ = args, not_in
        return self

    def call(self, name, **kw):
        return self.ns[name.lower()](**{k.lower(): v for k, v in kw.items()})
