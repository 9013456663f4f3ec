"""Per-source-line summary of one kernel from an .ncu-rep captured with --import-source on.

usage: python scripts/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [TOP] [LAUNCH_INDEX]
Prints the source lines carrying the most executed warp instructions and stall samples.
"""
import csv
import io
import subprocess
import sys


def main():
    rep, regex = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    text = subprocess.run(
        ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
         "regex:" + regex], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    lines = []
    path, hdr, launch, seen_funcs = None, None, -1, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            path = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            if path is not None and (not seen_funcs or path in seen_funcs and seen_funcs[0] == path):
                launch += 1
                seen_funcs = [path]
            elif path not in seen_funcs:
                seen_funcs.append(path)
            continue
        if hdr is None or r[0] == "" or launch != which:
            continue
        try:
            ins = int(r[hdr.index("Instructions Executed")])
            smp = int(r[hdr.index("# Samples")])
        except ValueError:
            continue
        lines.append((path, int(r[0]), r[1].strip(), ins, smp))
    tot_i = sum(x[3] for x in lines) or 1
    tot_s = sum(x[4] for x in lines) or 1
    print(f"launch {which}: {tot_i} warp instructions, {tot_s} samples")
    print("-- by instructions")
    for p, ln, src, ins, smp in sorted(lines, key=lambda x: -x[3])[:top]:
        print(f"{100 * ins / tot_i:5.1f}% ins {100 * smp / tot_s:5.1f}% smp  {p}:{ln}  {src[:90]}")
    print("-- by samples")
    for p, ln, src, ins, smp in sorted(lines, key=lambda x: -x[4])[:top]:
        print(f"{100 * ins / tot_i:5.1f}% ins {100 * smp / tot_s:5.1f}% smp  {p}:{ln}  {src[:90]}")


if __name__ == "__main__":
    main()
