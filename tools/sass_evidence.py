#!/usr/bin/env python3
"""Static SASS / resource table of every tensor-core, TMA and fp64 kernel of libnmf_b200.so
(-> profiles/rNN_sass_evidence.md).  CPU only: cuobjdump on the in-tree library."""
import collections
import re
import subprocess
import sys

LIB = "nmf_b200/libnmf_b200.so"
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMAPF", "ELECT", "DFMA"]


def demangle(sym):
    out = subprocess.run(["c++filt", sym], capture_output=True, text=True).stdout.strip()
    return re.sub(r"\(.*", "", out).replace("void nmfb::", "").replace("nmfb::", "")


def main(out_path):
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, counts = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = demangle(m.group(1))
            counts[cur] = collections.Counter()
        elif cur:
            for k in KEYS:
                if re.search(r"\b" + k + r"\b", line):
                    counts[cur][k] += 1
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    regs, fn = {}, None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            fn = demangle(m.group(1))
        m = re.search(r"REG:(\d+) STACK:(\d+)", line)
        if m and fn:
            regs[fn] = m.groups()
    rows = []
    for k, c in counts.items():
        if any(c[x] for x in ("UTCHMMA", "UTMALDG", "LDTM")) or any(t in k for t in ("admm_apply", "chol", "update", "kl_vaux", "tikhonov", "ingest_data", "absmax")):
            r = regs.get(k, ("?", "?"))
            rows.append(f"| `{k}` | {r[0]} | {r[1]} | " + " | ".join(str(c[x]) for x in KEYS) + " |")
    with open(out_path, "w") as f:
        f.write("# SASS evidence (cuobjdump -sass / -res-usage of nmf_b200/libnmf_b200.so, sm_100a, CUDA 12.9)\n\n"
                "Static instruction counts per kernel instantiation (not executed counts): `UTCHMMA` = `tcgen05.mma`,\n"
                "`UTCBAR` = `tcgen05.commit`, `LDTM` = `tcgen05.ld`, `UTMALDG` = TMA tensor load (`cp.async.bulk.tensor`),\n"
                "`UTMAPF` = TMA prefetch to L2, `ELECT` = `elect.sync`, `DFMA` = fp64 FMA.  Made by `tools/sass_evidence.py`.\n\n"
                "| kernel | regs | stack B | " + " | ".join(KEYS) + " |\n|---|---|---|" + "---|" * len(KEYS) + "\n"
                + "\n".join(rows) + "\n")
    print(len(rows), "kernels ->", out_path)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "profiles/sass_evidence.md")
