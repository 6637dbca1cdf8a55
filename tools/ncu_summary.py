"""Summarise ncu outputs brought back in gpurun_out/ into small text files for profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
    python tools/ncu_summary.py full gpurun_out/prof.ncu-rep      > profiles/rNN_<kernel>_full.txt
    python tools/ncu_summary.py traffic gpurun_out/prof.ncu-rep profiles/rNN_<kernel>_full.txt   # updates profiles/ncu_traffic.json
    (full / traffic take an optional last argument: keep only launches whose kernel name contains it)
"""
import csv
import subprocess
import sys
from collections import defaultdict

KEEP = (
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.max.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max", "sm__cycles_active.avg",
    "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active",
)


def launches(path):
    rows = list(csv.reader(open(path, errors="replace")))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, data = rows[h], rows[h + 1:]
    ik, iv, iu, ig = (hdr.index(n) for n in ("Kernel Name", "Metric Value", "Metric Unit", "Grid Size"))
    agg = defaultdict(lambda: [0, 0.0])
    for r in data:
        if len(r) <= iv:
            continue
        v = float(r[iv].replace(",", ""))
        v = v / 1000 if r[iu] == "ns" else (v * 1000 if r[iu] == "ms" else v)
        name = r[ik].split("(")[0].replace("void ", "")
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)")
    print(f"# {sum(v[0] for v in agg.values())} launches, {tot:.1f} us total")
    print(f"{'kernel':48s} {'n':>4s} {'us':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:48s} {v[0]:4d} {v[1]:10.1f} {100 * v[1] / tot:6.1f}%")


def _rows(path, only=None):
    """Rows of the raw page; `only`: keep launches whose kernel name contains it (a report may hold several kernels);
    launches without a duration (a capture that was cut short) are dropped."""
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik, it = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
    data = [r for r in data if (only is None or only in r[ik]) and r[it] not in ("", "nan", "-nan")]
    return hdr, units, data


def full(path, only=None):
    hdr, units, data = _rows(path, only)
    ik = hdr.index("Kernel Name")
    print(f"# ncu --set full --clock-control none: {len(data)} captured launches of {data[0][ik]}")
    for i, h in enumerate(hdr):
        if h in KEEP:
            print(f"{h:75s} {units[i]:14s} " + "  ".join(r[i] for r in data))


def traffic(path, source, only=None):
    """DRAM bytes per launch (read + write, averaged over the captured launches) → profiles/ncu_traffic.json,
    which bench.py reports as roofline.traffic."""
    import json
    import os
    hdr, units, data = _rows(path, only)
    ik, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = sum(float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]] for r in data)
    name = data[0][ik].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
    cur = json.load(open(out)) if os.path.exists(out) else {}
    cur[name] = {"dram_bytes_per_launch": tot / len(data), "captured_launches": len(data), "source": source}
    json.dump(cur, open(out, "w"), indent=1, sort_keys=True)
    print(name, cur[name])


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
