"""Host execution of the REAL device sources (TEST INFRASTRUCTURE -- see tests/emu/cuda_emu.h).

`build()` copies cice5_b200/csrc/*.cu{,h} into tests/emu/_build/src with three mechanical rewrites and
compiles them with g++ against tests/emu/cuda_emu.{h,cpp}:

  * `#include <cuda_runtime.h>` / `<cuda.h>`            -> `#include "cuda_emu.h"`
  * `kernel<<<grid, block, smem, stream>>>(args...)`    -> `emu::launch(grid, block, smem, stream, kernel, args...)`
  * `asm volatile("<ptx>" : outs : ins : clobbers);`    -> the host stand-in of that instruction: the cross-rank
    flags (st.release.sys / ld.acquire.sys / globaltimer) and what the TMA kernel is made of (mbarrier init /
    arrive / expect_tx / try_wait, cp.async.bulk.tensor 2-D and 3-D with zero fill outside the tensor,
    bar.sync id, fences); `emu_ptx_unavailable()` for anything else
  * `extern __shared__ T x[];` -> the block's dynamic buffer; `__shared__ T x[N];` -> per-block storage

Nothing else is touched: prep, halo, fold, strength, finish, both subcycle kernels (the TMA ring with its
producer warp, work queue and item list, and the register sweep), the program builder, the CUDA-graph path
and the multi-rank flag barrier all run from the product's own text.  The
result, tests/emu/_build/libevp_emu.so, exports the product's C ABI; `patched()` swaps it into
cice5_b200.evp for the duration of a test.  The product never loads it: cice5_b200 has no reference to
this directory, and a missing CUDA library is still a hard error there.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "cice5_b200", "csrc")
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(BUILD, "libevp_emu.so")

_LAUNCH = re.compile(r"([A-Za-z_]\w*)\s*<<<(.*?)>>>\s*\(", re.S)


def _balanced(s: str, i: int) -> int:
    """Index of the ')' that closes the '(' at s[i]."""
    depth, quote = 0, None
    for k in range(i, len(s)):
        ch = s[k]
        if quote:
            if ch == "\\":
                continue
            if ch == quote and s[k - 1] != "\\":
                quote = None
        elif ch == '"':
            quote = ch
        elif ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
            if depth == 0:
                return k
    raise ValueError("unbalanced parentheses")


def _operands(section: str) -> list:
    """C expressions of `"constraint"(expr), ...`."""
    out, i = [], 0
    while True:
        m = re.compile(r'"[^"]*"\s*\(').search(section, i)
        if not m:
            return out
        close = _balanced(section, m.end() - 1)
        out.append(section[m.end():close].strip())
        i = close + 1


def _rewrite_asm(text: str) -> str:
    out, i = [], 0
    pat = re.compile(r"asm\s+volatile\s*\(")
    while True:
        m = pat.search(text, i)
        if not m:
            out.append(text[i:])
            return "".join(out)
        close = _balanced(text, m.end() - 1)
        body = text[m.end():close]
        end = text.index(";", close) + 1
        # split template : outputs : inputs : clobbers at top-level colons outside string literals
        parts, depth, quote, cur = [], 0, None, ""
        k = 0
        while k < len(body):
            ch = body[k]
            if quote:
                cur += ch
                if ch == quote and body[k - 1] != "\\":
                    quote = None
            elif ch == '"':
                quote = ch
                cur += ch
            elif ch == "(":
                depth += 1
                cur += ch
            elif ch == ")":
                depth -= 1
                cur += ch
            elif ch == ":" and depth == 0:
                if body[k + 1:k + 2] == ":":            # '::' = empty section
                    parts.append(cur)
                    parts.append("")
                    cur = ""
                    k += 2
                    continue
                parts.append(cur)
                cur = ""
            else:
                cur += ch
            k += 1
        parts.append(cur)
        template = " ".join(re.findall(r'"((?:[^"\\]|\\.)*)"', parts[0]))
        outs = _operands(parts[1]) if len(parts) > 1 else []
        ins = _operands(parts[2]) if len(parts) > 2 else []
        if "st.release.sys.global.u64" in template or "st.release.gpu.global.u64" in template:
            rep = f"emu_st_release_u64((void *)({ins[0]}), ({ins[1]}));"
        elif "ld.acquire.sys.global.u64" in template or "ld.acquire.gpu.global.u64" in template:
            rep = f"{outs[0]} = emu_ld_acquire_u64((const void *)({ins[0]}));"
        elif "globaltimer" in template:
            rep = f"{outs[0]} = emu::now_ns();"
        elif "mbarrier.init" in template:
            rep = f"emu_mbar_init({ins[0]}, {ins[1]});"
        elif "mbarrier.arrive.expect_tx" in template:
            rep = f"emu_mbar_expect_tx({ins[0]}, {ins[1]});"
        elif "mbarrier.arrive" in template:
            rep = f"emu_mbar_arrive({ins[0]});"
        elif "mbarrier.try_wait.parity" in template:
            rep = f"{outs[0]} = emu_mbar_try_wait({ins[0]}, {ins[1]});"
        elif "cp.async.bulk.tensor.3d" in template:
            rep = f"emu_tma_load({ins[0]}, {ins[1]}, {ins[2]}, 3, {ins[3]}, {ins[4]}, {ins[5]});"
        elif "cp.async.bulk.tensor.2d" in template:
            rep = f"emu_tma_load({ins[0]}, {ins[1]}, {ins[2]}, 2, {ins[3]}, {ins[4]}, 0);"
        elif template.startswith("bar.arrive"):
            rep = f"emu_bar_arrive({template.split()[1].rstrip(',')}, {ins[0]});"
        elif template.startswith("bar.sync"):
            rep = f"emu_bar_sync({template.split()[1].rstrip(',')}, {ins[0]});"
        elif template.startswith("prefetch."):
            rep = f"(void)({ins[0]});"                  # a hint: no architectural effect
        elif template.startswith("fence."):
            rep = "std::atomic_thread_fence(std::memory_order_seq_cst);"
        else:
            first = template.split(";")[0].strip().replace('"', "'")[:60]
            rep = "".join(f"{o} = {{}}; " for o in outs if re.fullmatch(r"[A-Za-z_]\w*", o)) + \
                f'emu_ptx_unavailable("{first}");'
        out.append(text[i:m.start()] + rep)
        i = end


_SHARED = re.compile(r"^([ \t]*)__shared__\s+(?:__align__\(\d+\)\s+)?([A-Za-z_][\w ]*?)\s+([A-Za-z_]\w*\s*(?:\[[^;]*)?);",
                     re.M)
_shared_ids = [0]


def _shared_decl(m) -> str:
    """`__shared__ T a[N], b;` -> references into per-block storage (several blocks are resident at a time)."""
    indent, typ, decls = m.group(1), m.group(2).strip(), m.group(3)
    out = []
    for d in [x.strip() for x in re.split(r",(?![^\[]*\])", decls)]:
        dm = re.match(r"^([A-Za-z_]\w*)\s*((?:\[[^\]]*\])*)$", d)
        name, dims = dm.group(1), dm.group(2)
        _shared_ids[0] += 1
        out.append(f"{indent}auto &{name} = *reinterpret_cast<{typ} (*){dims}>(emu::block_static({_shared_ids[0]}, "
                   f"sizeof({typ}{dims})));")
    return "\n".join(out)


def transform(text: str) -> str:
    text = re.sub(r"#include\s*<cuda_runtime\.h>|#include\s*<cuda\.h>", '#include "cuda_emu.h"', text)
    text = _LAUNCH.sub(lambda m: f"emu::launch({m.group(2)}, {m.group(1)}, ", text)
    text = _rewrite_asm(text)
    # dynamic shared memory: the block's buffer; static shared variables: per host thread (the blocks of a
    # launch run one after the other on their stream's thread, and shared memory starts undefined anyway)
    text = re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?([\w ]+?)\s*(\w+)\s*\[\s*\]\s*;",
                  r"\1 *\2 = (\1 *)emu::dynamic_smem();", text)
    text = _SHARED.sub(_shared_decl, text)
    assert "__shared__" not in text
    return text


def build(force: bool = False, defines=()) -> str:
    """defines: extra -D flags (kernel variants, e.g. ("-DEVP_TMA_MERGE_UV=1",)): a separate library each."""
    tag = "".join(re.sub(r"\W+", "_", d) for d in defines)
    lib = LIB if not defines else LIB.replace(".so", f"{tag}.so")
    return _build(lib, force, tuple(defines))


def _build(LIB: str, force: bool, defines: tuple) -> str:
    srcdir = os.path.join(BUILD, "src")
    os.makedirs(srcdir, exist_ok=True)
    sources = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")))
    plain = sorted(f for f in os.listdir(CSRC) if f.endswith(".h"))      # plain C headers: included as they are
    deps = [os.path.join(CSRC, f) for f in sources + plain] + [os.path.join(HERE, f) for f in ("cuda_emu.h", "cuda_emu.cpp", "emu.py")]
    deps.append(os.path.join(ROOT, "include", "evp_b200.h"))
    if not force and os.path.exists(LIB) and all(os.path.getmtime(d) <= os.path.getmtime(LIB) for d in deps):
        return LIB
    for f in sources:
        out = transform(open(os.path.join(CSRC, f)).read())
        out = out.replace('"../../include/evp_b200.h"', f'"{os.path.join(ROOT, "include", "evp_b200.h")}"')
        with open(os.path.join(srcdir, f + (".cpp" if f.endswith(".cu") else "")), "w") as fh:
            fh.write(out)
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-pthread",
           "-DEVP_STRICT=1", *defines, "-w", "-I", HERE, "-I", srcdir, "-I", CSRC, "-o", LIB,
           os.path.join(srcdir, "evp_abi.cu.cpp"), os.path.join(HERE, "cuda_emu.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + " ".join(cmd) + "\n" + (r.stdout + r.stderr)[-6000:])
    return LIB


@contextlib.contextmanager
def patched(kernel: str = "tma", defines=()):
    """cice5_b200.evp bound to the host build of its own sources inside the `with` block;
    kernel = "tma" (the product's default subcycle kernel) or "sweep"."""
    from cice5_b200 import evp
    lib_path = build(defines=defines)
    saved_libs, saved_env = dict(evp._libs), os.environ.get("EVP_B200_KERNEL")
    os.environ["EVP_B200_KERNEL"] = kernel
    lib = C.CDLL(lib_path)
    lib.evp_gpu_strerror.restype = C.c_char_p
    lib.evp_gpu_last_cuda_error.restype = C.c_char_p
    evp._libs.clear()
    evp._libs.update({True: lib, False: lib})
    try:
        yield evp
    finally:
        evp._libs.clear()
        evp._libs.update(saved_libs)
        if saved_env is None:
            os.environ.pop("EVP_B200_KERNEL", None)
        else:
            os.environ["EVP_B200_KERNEL"] = saved_env
