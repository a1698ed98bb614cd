"""Per-kernel roofline table from an `ncu --set full` report: DRAM and L2 throughput of every captured launch against the
measured peaks (HBM: MEASURED_PEAKS.json; L2: scripts/micro/l2_bw.cu on the same pool)."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L2_PEAK = 18200.0        # GB/s, scripts/micro/l2_bw.cu (32-64 MB read loop, 16-byte loads, all SMs)


def val(row, hdr, units, key):
    i = hdr.index(key)
    v = float(row[i].replace(",", ""))
    u = units[i]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}
    return v * scale.get(u, 1)


def main(rep, out):
    hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    txt = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    with open(out, "w") as f:
        f.write("# per-launch DRAM / L2 throughput from %s (cold L2: ncu flushes caches between kernels)\n" % os.path.basename(rep))
        f.write("# peaks: HBM %.0f GB/s (MEASURED_PEAKS.json), L2 %.0f GB/s (scripts/micro/l2_bw.cu)\n" % (hbm, L2_PEAK))
        f.write("%-34s %9s %10s %9s %7s %10s %9s %7s %6s\n" % ("kernel", "time_us", "dram_MB", "dram_GB/s", "ofHBM", "l2_MB", "l2_GB/s", "ofL2", "warps%"))
        for r in rows[2:]:
            t = val(r, hdr, units, "gpu__time_duration.sum")
            d = val(r, hdr, units, "dram__bytes_read.sum") + val(r, hdr, units, "dram__bytes_write.sum")
            l2 = 32.0 * float(r[hdr.index("lts__t_sectors.sum")].replace(",", ""))       # 32-byte sectors through the L2 tags
            wa = float(r[hdr.index("sm__warps_active.avg.pct_of_peak_sustained_active")])
            name = r[ik].split("(")[0].replace("void ", "")
            f.write("%-34s %9.2f %10.3f %9.1f %6.1f%% %10.3f %9.1f %6.1f%% %6.1f\n" %
                    (name[:34], t * 1e6, d / 1e6, d / t / 1e9, 100 * d / t / 1e9 / hbm, l2 / 1e6, l2 / t / 1e9, 100 * l2 / t / 1e9 / L2_PEAK, wa))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
