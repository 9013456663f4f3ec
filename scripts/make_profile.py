"""Turns the scratch output of scripts/evidence_run.sh TAG (gpurun_out/) into the tracked evidence
under profiles/: bench lines, launch list, raw metrics, traffic.json and the ncu summary.
usage: python scripts/make_profile.py v8 "build description" """
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
COLS = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum"]
PER_OVERLAY = {"rs_onesweep": 4}
ORDER = ["rs_onesweep", "k5_emit", "k4_resolve", "k2_place", "k1_count", "p1_sites", "rs_global_hist", "k1_spine"]


def sh(*cmd):
    return subprocess.run(list(cmd), capture_output=True, text=True, cwd=ROOT).stdout


def main():
    tag, build = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    rep = os.path.join(G, f"prof_{tag}.ncu-rep")
    # round-1 tags were v1..v8 (files r01_vN_*); later rounds name the round in the tag itself (r02, r02b, ...)
    pre = "" if tag.startswith("r") else "r01_"
    rnd = tag[1:3].lstrip("0") if tag.startswith("r") else "1"
    for w in range(1, 6):
        shutil.copy(os.path.join(G, f"{tag}_bench_config{w}.json"), os.path.join(P, f"{pre}{tag}_bench_config{w}.json"))
    for a, b in ((f"{tag}_bench_config5_reference_arm.json", f"{pre}{tag}_bench_config5_reference_arm.json"),
                 (f"{tag}_launches.csv", f"{pre}{tag}_launches.csv"),
                 (f"{tag}_file_pipeline_config5.json", f"{pre}{tag}_file_pipeline_config5.json")):
        if os.path.exists(os.path.join(G, a)):
            shutil.copy(os.path.join(G, a), os.path.join(P, b))
    # raw metrics of one overlay
    rows = list(csv.reader(io.StringIO(sh("ncu", "-i", rep, "--page", "raw", "--csv"))))
    hdr = rows[0]
    idx = [hdr.index(c) for c in COLS]
    body = rows[2:]
    for i, r in enumerate(body):
        if "k1_count" in r[hdr.index("Kernel Name")]:
            body = body[i:i + 11]
            break
    with open(os.path.join(P, f"{pre}{tag}_ncu_raw.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(COLS)
        w.writerow([rows[1][i] for i in idx])
        for r in body:
            w.writerow([r[i] for i in idx])
    open(os.path.join(P, "traffic.json"), "w").write(sh("python", "scripts/ncu_table.py", rep, "--traffic", "config5"))
    tbl = sh("python", "scripts/ncu_table.py", rep).strip().split("\n")
    # shares from the launch list
    lrows = list(csv.reader(open(os.path.join(P, f"{pre}{tag}_launches.csv"))))
    start = next(i for i, r in enumerate(lrows) if r and r[0] == "ID")
    lh = lrows[start]
    ni, vi = lh.index("Kernel Name"), lh.index("Metric Value")
    acc = collections.defaultdict(list)
    for r in lrows[start + 2:]:
        if len(r) > vi:
            name = r[ni].split("(")[0].replace("void ", "").split("<")[0].replace("mspb::", "").replace("_kernel", "")
            acc[name].append(float(r[vi].replace(",", "")) / 1e3)
    ncu = {k: sum(acc[k]) / len(acc[k]) for k in ORDER}
    d = json.load(open(os.path.join(P, f"{pre}{tag}_bench_config5.json")))
    live = {k: v["ms_total"] / v["launches"] * 1e3 for k, v in d["kernels"].items()}
    nt = sum(ncu[k] * PER_OVERLAY.get(k, 1) for k in ORDER)
    lt = sum(live[k] * PER_OVERLAY.get(k, 1) for k in ORDER)

    def lines(k, n=6):
        out = sh("python", "scripts/ncu_lines.py", rep, k, str(n))
        i = out.index("-- by samples")
        return out[i + len("-- by samples\n"):].rstrip()

    rf = d["roofline"]
    traffic = json.load(open(os.path.join(P, "traffic.json")))["kernels"]
    md = [f"# Round {rnd}, pipeline {tag}: `ncu --set full` of one overlay (config 5, one B200)\n",
          "Command (after the same command exited 0 without ncu; `scripts/evidence_run.sh`):\n",
          '`ncu --set full --clock-control none --import-source on -k regex:"k5_emit|k4_resolve|rs_onesweep|k1_count|'
          f'k1_spine|p1_sites|k2_place|rs_global" -s 24 -c 24 -o prof_{tag} python bench.py --steps 1 --warmup 3 '
          '--no-cpu-baseline`\n',
          "The table is one whole overlay: the 11 launches from the first `k1_count` of the capture on "
          f"(`scripts/ncu_table.py`).  Build: {build}.  Times under ncu are cold-cache and serialised; the live "
          f"CUDA-event times are in `{pre}{tag}_bench_config5.json` (`kernels`).  Columns: duration, registers, achieved "
          "occupancy, DRAM read / write per launch, DRAM and L2 throughput in % of peak, issue-slot utilisation, "
          "busiest pipe, executed warp instructions.\n",
          "| " + tbl[0] + " |", "|" + "---|" * len(tbl[0].split(" | "))]
    md += ["| " + r + " |" for r in tbl[1:]]
    md += ["\n## Share of the step: ncu launch list against the live event times\n",
           f"`{pre}{tag}_launches.csv` is the `--metrics gpu__time_duration.sum --clock-control none` pass of "
           "`bench.py --steps 2 --warmup 3` (warm-up, timed, kernel-timing and end-to-end overlays, plus the "
           "`load_tables` index builds, whose sort passes share kernel names with the overlay's and pull the ncu "
           "averages of `rs_onesweep` / `rs_global_hist` slightly off).  Share of the overlay kernels' summed time, "
           f"beside the live CUDA-event share from `{pre}{tag}_bench_config5.json`:\n",
           "| kernel | launches per overlay | ncu avg us | ncu share | live avg us | live share |",
           "|---|---|---|---|---|---|"]
    for k in ORDER:
        n = PER_OVERLAY.get(k, 1)
        md.append(f"| {k} | {n} | {ncu[k]:.1f} | {ncu[k] * n / nt:.3f} | {live[k]:.1f} | {live[k] * n / lt:.3f} |")
    md.append(f"\nDominant kernel, live: `{rf['kernel']}` {rf['avg_launch_ms']:.3f} ms per launch, "
              f"{rf['algorithmic_bytes_per_launch'] / 1e9:.3f} GB algorithmic = {rf['achieved']:.0f} GB/s = "
              f"{rf['frac']:.3f} of the measured copy rate ({rf['peak']:.0f} GB/s); DRAM traffic under ncu "
              f"{traffic.get(rf['kernel'], 0) / 1e9:.2f} GB per launch (`traffic.json`).\n")
    md.append("## Where the warps wait (stall samples by source line, `scripts/ncu_lines.py`)\n")
    for k in ("k5_emit", "k4_resolve", "k2_place", "k1_count", "rs_onesweep", "p1_sites"):
        md.append(f"**{k}**\n\n```\n{lines(k)}\n```\n")
    notes = os.path.join(P, f"{pre}{tag}_reading.md")
    if os.path.exists(notes):
        md.append(open(notes).read())
    open(os.path.join(P, f"{pre}{tag}_ncu_full.md"), "w").write("\n".join(md))
    print("wrote", f"profiles/{pre}{tag}_ncu_full.md")


if __name__ == "__main__":
    main()
