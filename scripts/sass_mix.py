"""Instruction mix of the library's kernels from `cuobjdump -sass` (static counts per kernel): which load / store
forms, barriers, votes, atomics and async-copy instructions the sm_100a code actually contains.
usage: python scripts/sass_mix.py [msprime_b200/_mspb.so] > profiles/rNN_sass_mix.md"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "msprime_b200/_mspb.so"
text = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
kern = None
mix = collections.OrderedDict()
for line in text.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name).replace("mspb::", "").replace("(anonymous namespace)::", "")
        kern = name
        mix[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        mix[kern]["total"] += 1
        base = op.split(".")[0]
        key = base
        if base in ("LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG", "RED"):
            width = "128" if ".128" in op else "64" if ".64" in op else "8" if ".U8" in op or ".S8" in op else \
                "16" if ".U16" in op or ".S16" in op else "32"
            key = f"{base}.{width}"
        mix[kern][key] += 1
show = ["total", "LDG.128", "LDG.64", "LDG.32", "LDG.8", "STG.128", "STG.64", "STG.32", "STG.8", "LDS.128", "LDS.64", "LDS.32",
        "STS.128", "STS.64", "STS.32", "STS.8", "ATOMS.32", "ATOMG.32", "ATOMG.64", "RED.32", "RED.64", "BAR", "VOTE",
        "SHFL", "LDGSTS", "UBLKCP", "UTMALDG", "SYNCS"]
want = ("k1_count", "k1_spine", "k2_place", "rs_global_hist", "rs_onesweep_kernel<unsigned int, 7", "p1_sites_kernel<0, false",
        "k4_resolve", "k5_emit_kernel<1, false, true", "k0_link", "sf_store_steps", "sf_below", "sf_afs_kernel<true",
        "site_rows_noalloc")
print("| kernel | " + " | ".join(show) + " |")
print("|---|" + "---|" * len(show))
for k, c in mix.items():
    if any(w in k for w in want):
        print(f"| `{k[:60]}` | " + " | ".join(str(c.get(s, 0)) for s in show) + " |")
