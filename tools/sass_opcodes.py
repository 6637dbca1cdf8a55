"""Per-kernel SASS opcode summary of libdaft_b200.so (evidence that the hot kernels are tcgen05 / TMEM / bulk-copy code).

    python tools/sass_opcodes.py > profiles/r02_sass_opcodes.txt

Counts, per kernel of the shipped library (cuobjdump -sass): UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / .st),
UBLKCP (cp.async.bulk, 1-D TMA engine copies), UTMALDG (tensor-map TMA loads), UTCBAR (tcgen05.commit), SYNCS (mbarrier),
total instructions, and local-memory instructions (LDL / STL: spills and printf argument buffers)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ubisoft-laforge-daft-exprt_b200", "libdaft_b200.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTCBAR", "SYNCS", "LDL", "STL", "MUFU"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = {}, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = dict.fromkeys(OPS, 0)
            kernels[cur]["total"] = 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            kernels[cur]["total"] += 1
            op = m.group(1).split(".")[0]
            if op in kernels[cur]:
                kernels[cur][op] += 1
    demangle = subprocess.run(["c++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    print(f"# {os.path.relpath(LIB, ROOT)}: {len(kernels)} kernels (cuobjdump -sass, sm_100a)")
    print(f"# {'kernel':70s} {'total':>7s} " + " ".join(f"{o:>7s}" for o in OPS))
    tot = dict.fromkeys(OPS + ["total"], 0)
    for (name, c), dm in sorted(zip(kernels.items(), demangle), key=lambda kv: -kv[0][1]["UTCHMMA"] * 100000 - kv[0][1]["total"]):
        short = re.sub(r"\(.*", "", dm).replace("dx::", "")[:70]
        print(f"  {short:70s} {c['total']:7d} " + " ".join(f"{c[o]:7d}" for o in OPS))
        for k in tot:
            tot[k] += c[k]
    print(f"  {'ALL':70s} {tot['total']:7d} " + " ".join(f"{tot[o]:7d}" for o in OPS))


if __name__ == "__main__":
    sys.exit(main())
