"""One line per kernel launch from an .ncu-rep: time, registers, occupancy, DRAM bytes, busiest pipe.

usage: python scripts/ncu_table.py REPORT.ncu-rep
"""
import csv
import io
import subprocess
import sys


def main():
    text = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    hdr = rows[0]

    def g(r, name, default=""):
        return r[hdr.index(name)] if name in hdr else default

    # the capture may start mid-overlay: keep one whole overlay = the 11 launches from the first
    # k1_count on (k1_count, k1_spine, k2_place, histogram, 4 sort passes, p1, k4, k5)
    body = rows[2:]
    for i, r in enumerate(body):
        if "k1_count" in g(r, "Kernel Name"):
            body = body[i:i + 11]
            break
    rows = rows[:2] + body
    if len(sys.argv) > 3 and sys.argv[2] == "--traffic":
        # {"workload": ..., "kernels": {short name: mean dram bytes (read + write) per launch}}
        import json
        import re

        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        acc = {}
        for r in rows[2:]:
            name = re.sub(r"<.*", "", g(r, "Kernel Name").split("(")[0].replace("void ", "")).replace("_kernel", "")
            tot = 0.0
            for col in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(g(r, col)) * unit[rows[1][hdr.index(col)]]
            acc.setdefault(name, []).append(tot)
        out = {"workload": sys.argv[3], "source": sys.argv[1].split("/")[-1],
               "kernels": {k: round(sum(v) / len(v)) for k, v in acc.items()}}
        print(json.dumps(out, indent=1))
        return
    pipes = [h for h in hdr if h.startswith("sm__inst_executed_pipe_") and h.endswith(".avg.pct_of_peak_sustained_active")]
    print("kernel | us | regs | warps% | dram rd MB | dram wr MB | dram% | L2% | issue% | top pipe | warp inst M")
    for r in rows[2:]:
        name = g(r, "Kernel Name").split("(")[0].replace("void ", "")[:34]
        best = max(pipes, key=lambda p: float(r[hdr.index(p)] or 0)) if pipes else ""
        bestv = float(r[hdr.index(best)]) if best else 0
        print(" | ".join([
            name,
            f"{float(g(r, 'gpu__time_duration.sum')):.1f}",
            g(r, "launch__registers_per_thread"),
            f"{float(g(r, 'sm__warps_active.avg.pct_of_peak_sustained_active')):.0f}",
            f"{float(g(r, 'dram__bytes_read.sum')):.0f}",
            f"{float(g(r, 'dram__bytes_write.sum')):.0f}",
            f"{float(g(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')):.0f}",
            f"{float(g(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed')):.0f}",
            f"{float(g(r, 'sm__issue_active.avg.pct_of_peak_sustained_elapsed')):.0f}",
            best.replace("sm__inst_executed_pipe_", "").split(".")[0] + f" {bestv:.0f}%",
            f"{float(g(r, 'smsp__inst_executed.sum')) / 1e6:.0f}",
        ]))


if __name__ == "__main__":
    main()
