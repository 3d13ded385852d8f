#!/usr/bin/env python3
"""Regenerate docs/sass/: one SASS listing per kernel of hpnn_b200/lib/libhpnn_core.so plus an index
with the instruction mix.  Usage: python tools/dump_sass.py  (after python -m hpnn_b200.build)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "hpnn_b200", "lib", "libhpnn_core.so")
OUT = os.path.join(ROOT, "docs", "sass")
NOTABLE = ["DFMA", "DADD", "DMUL", "DMMA", "LDG", "STG", "LDGSTS", "LDS", "STS", "MEMBAR", "CCTL", "BAR", "SHFL", "MUFU",
           "ERRBAR", "UTCHMMA", "UTMALDG", "LDTM", "SYNCS", "UTCBAR", "RED", "ATOM"]


def demangle(name):
    try:
        return subprocess.run(["c++filt", name], capture_output=True, text=True, check=True).stdout.strip()
    except Exception:
        return name


def short(dem):
    m = re.match(r"(?:void )?(?:[A-Za-z0-9_]+::)*([A-Za-z0-9_]+)(<[^>]*>)?", dem)
    base = m.group(1)
    targs = m.group(2) or ""
    targs = re.sub(r"\(bool\)|\(int\)", "", targs)
    targs = re.sub(r"[<> ]", "", targs).replace(",", "_").replace("true", "1").replace("false", "0")
    return base + ("_" + targs if targs else "")


def main():
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass"):
            os.remove(os.path.join(OUT, f))
    parts = re.split(r"^\s*Function : ", text, flags=re.M)[1:]
    rows = []
    for part in parts:
        mangled, body = part.split("\n", 1)
        name = short(demangle(mangled.strip()))
        lines = []
        for ln in body.split("\n"):  # keep address + instruction, drop the encodings
            m = re.match(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?;)\s*/\*", ln)
            if m:
                lines.append(f"/*{m.group(1)}*/ {m.group(2)}")
            elif ln.strip().startswith(".") and "/*" not in ln:
                lines.append(ln.strip())  # labels / header flags
        with open(os.path.join(OUT, name + ".sass"), "w") as fh:
            fh.write("Function : " + demangle(mangled.strip()) + "\n" + "\n".join(lines) + "\n")
        ops = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", body, flags=re.M)
        cnt = collections.Counter(o.split(".")[0] for o in ops)
        mix = ", ".join(f"{k} {cnt[k]}" for k in NOTABLE if cnt.get(k))
        rows.append((name, len(ops), mix))
    rows.sort()
    with open(os.path.join(OUT, "README.md"), "w") as fh:
        fh.write("# SASS listings (sm_100a)\n\nOne file per kernel, written by `python tools/dump_sass.py` from "
                 "`cuobjdump -sass hpnn_b200/lib/libhpnn_core.so`.\nInstruction mix = mnemonic counts in the listing "
                 "(static, not dynamic).\n\n| kernel | instructions | notable mnemonics |\n|---|---|---|\n")
        for name, n, mix in rows:
            fh.write(f"| `{name}` | {n} | {mix} |\n")
        fh.write("\n`UTCHMMA` = `tcgen05.mma`, `LDTM` = `tcgen05.ld`, `UTMALDG` = TMA (`cp.async.bulk.tensor`), "
                 "`SYNCS` = mbarrier ops, `UTCBAR` = `tcgen05.commit`, `DMMA` = FP64 `mma.sync.m8n8k4`, "
                 "`LDGSTS` = `cp.async`.\n")
    print(f"{len(rows)} kernels -> {OUT}")


if __name__ == "__main__":
    sys.exit(main())
