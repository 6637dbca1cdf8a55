"""Registers, stack (spills + local arrays) and static shared memory of every kernel of libdaft_b200.so.

    python tools/res_usage.py > profiles/r02_res_usage.txt
"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ubisoft-laforge-daft-exprt_b200", "libdaft_b200.so")


def main():
    out = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, name = [], None
    for line in out.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and name:
            dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            dem = re.sub(r"\(.*", "", dem).replace("void ", "").replace("dx::", "")
            rows.append((dem, int(m.group(1)), int(m.group(2)), int(m.group(3))))
            name = None
    print("# cuobjdump -res-usage of libdaft_b200.so (sm_100a): registers per thread, stack bytes (spills + local arrays), static shared bytes")
    print(f"{'kernel':60s} {'regs':>5s} {'stack':>6s} {'smem':>6s}")
    for r in sorted(rows, key=lambda x: (-x[2], -x[1])):
        print(f"{r[0]:60s} {r[1]:5d} {r[2]:6d} {r[3]:6d}")


if __name__ == "__main__":
    main()
