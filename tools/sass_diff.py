"""Does the kernel measured in profiles/ still ship?  Builds the strict product library at an older
commit (git archive into a scratch directory) and at the working tree, and compares, per kernel,
registers / stack (cuobjdump -res-usage) and the SASS opcode sequence (cuobjdump -sass, operands
and predicates dropped): the longest run of identical opcodes is the hot loop of the subcycle
kernels, and it must be the same one that was timed.

    python tools/sass_diff.py fb2be07            # the commit the round-1 profiles were taken on

No GPU needed (nvcc cross-compiles sm_100a).  Template arguments added since (REVISED = false)
are matched by the first template argument only.
"""
from __future__ import annotations

import difflib
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
        "-shared", "-fmad=false", "-DEVP_STRICT=1"]


def build(src_root: str, out: str) -> None:
    subprocess.run([*NVCC, "-o", out, os.path.join(src_root, "cice5_b200/csrc/evp_abi.cu")], check=True,
                   capture_output=True)


def kernels(lib: str) -> dict[str, dict]:
    res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True, check=True).stdout
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    out: dict[str, dict] = {}
    for m in re.finditer(r"Function (\S+):\n\s*(REG:\d+ STACK:\d+)", res):
        out[m.group(1)] = {"res": m.group(2), "ops": []}
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = out.setdefault(m.group(1), {"res": "?", "ops": []})
        elif cur is not None and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            ins = re.sub(r"/\*[0-9a-f]+\*/", "", line).strip()
            cur["ops"].append(re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].rstrip(";"))
    return out


def key(mangled: str) -> str:
    name = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
    name = re.sub(r"\(.*", "", name)
    m = re.match(r"(void )?(evp::\w+)(<(\w+)(, false)?>)?$", name)      # REVISED = false is the old kernel
    if m:
        return m.group(2) + (f"<{m.group(4)}>" if m.group(4) else "")
    return name


def main() -> None:
    old_rev = sys.argv[1] if len(sys.argv) > 1 else "fb2be07"
    with tempfile.TemporaryDirectory() as tmp:
        tar = subprocess.run(["git", "-C", ROOT, "archive", old_rev, "cice5_b200/csrc", "include"],
                             capture_output=True, check=True).stdout
        subprocess.run(["tar", "-x", "-C", tmp], input=tar, check=True)
        build(tmp, os.path.join(tmp, "old.so"))
        build(ROOT, os.path.join(tmp, "new.so"))
        old = {key(k): v for k, v in kernels(os.path.join(tmp, "old.so")).items()}
        new = {key(k): v for k, v in kernels(os.path.join(tmp, "new.so")).items()}
    print(f"| kernel | {old_rev}: regs/stack, SASS | working tree | longest identical opcode run |")
    print("|---|---|---|---|")
    for k in sorted(new):
        if k not in old:
            continue
        a, b = old[k], new[k]
        sm = difflib.SequenceMatcher(None, a["ops"], b["ops"], autojunk=False)
        run = max((blk.size for blk in sm.get_matching_blocks()), default=0)
        same = "identical" if a["ops"] == b["ops"] else f"{run} of {len(b['ops'])}"
        print(f"| `{k}` | {a['res']}, {len(a['ops'])} | {b['res']}, {len(b['ops'])} | {same} |")


if __name__ == "__main__":
    main()
