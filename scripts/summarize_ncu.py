"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/."""
import csv
import subprocess
import sys
from collections import defaultdict


def launch_list(path, out):
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        if len(r) <= iv:
            continue
        v = float(r[iv].replace(",", ""))
        v = v / 1000 if r[iu] == "ns" else (v * 1000 if r[iu] == "ms" else v)
        name = r[ik].split("(")[0]
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(out, "w") as f:
        f.write("# per-kernel device time from `ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares)\n")
        f.write("# source: %s\n%-62s %6s %12s %10s %7s\n" % (path, "kernel", "n", "total_us", "avg_us", "share"))
        for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("%-62s %6d %12.1f %10.2f %6.1f%%\n" % (k[:62], v[0], v[1], v[1] / v[0], 100 * v[1] / tot))


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__cluster_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "smsp__cycles_active.avg", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def full(rep, out):
    txt = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    with open(out, "w") as f:
        f.write("# selected metrics from `ncu --set full --clock-control none --import-source on` (%s)\n" % rep)
        for r in rows[2:]:
            f.write("\n== %s  (launch id %s)\n" % (r[ik].split("(")[0], r[0]))
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write("  %-66s %16s %s\n" % (k, r[i], units[i]))


def traffic(rep, out, pattern="track_persistent"):
    """mean dram__bytes_read + dram__bytes_write per launch of the kernels matching `pattern` -> small JSON read by bench.py"""
    import json
    txt = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ik, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    vals = [float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]] for r in rows[2:] if pattern in r[ik]]
    json.dump({"kernel": pattern, "launches": len(vals), "dram_bytes_per_launch": sum(vals) / len(vals),
               "source": "ncu --set full --clock-control none, %s, dram__bytes_read.sum + dram__bytes_write.sum, cold L2" % rep}, open(out, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "list":
        launch_list(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3])
