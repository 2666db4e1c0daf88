#!/usr/bin/env python
"""profiles/rNN_ncu_full_summary.md from one round's captures:

  python tools/ncu_summary.py TAG RAW.csv LAUNCHES.csv PLAIN.json "COMMAND" > profiles/TAG_ncu_full_summary.md

RAW.csv      `ncu -i TAG_full.ncu-rep --page raw --csv`
LAUNCHES.csv `ncu --metrics gpu__time_duration.sum --clock-control none --csv` launch list
PLAIN.json   bench line of the same command run without ncu (its `kernels` are the CUDA-event shares)
"""
import collections
import csv
import json
import subprocess
import sys

tag, raw, launches, plain, cmd = sys.argv[1:6]
dig = subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "ncu_digest.py"), raw],
                     capture_output=True, text=True, check=True).stdout.splitlines()
head, body = dig[:2], dig[2:]
pick = {}
for line in body:
    name = line.split("|")[1].strip()
    if name == "k_leaf_refit" and float(line.split("|")[2]) < 10:      # the reset-only call
        continue
    pick.setdefault(name, line)
order = ["void k_action<1>", "k_broadphase", "k_substep", "k_narrowphase", "k_solve", "k_leaf_refit", "k_observations",
         "void k_rays<1>", "void k_rewards<1>"]
order = [n for n in order if n in pick]
out = [f"# {tag} — `ncu --set full --clock-control none` of the stage kernels (N=16000 worlds, mid-episode step)\n",
       f"Command (after the same command had exited 0 without ncu): `{cmd}`; digest: `ncu -i {tag}_full.ncu-rep --page raw "
       "--csv` → `tools/ncu_digest.py` → this file (`tools/ncu_summary.py`).\n",
       "The step graph runs the shard as 2 world slices on parallel branches, so every captured launch covers HALF the shard "
       "(8064 / 7936 worlds) and ncu times it alone, cold; `profiles/traffic.json` holds the mean DRAM bytes per launch x2 "
       "(a whole-shard launch).  One launch per kernel (k_leaf_refit: a whole-shard-slice call, not the reset-only one):\n"]
out += head + [pick[n] for n in order]
out.append("\nAchieved HBM bandwidth per kernel (SURVEY.md §8d per-kernel evidence), half-shard launches timed alone, "
           "against the measured 6551 GB/s:\n")
out.append("| kernel | dur us | DRAM rd+wr MB | DRAM GB/s | % of 6551 | occupancy % | regs | IPC | lanes/inst |")
out.append("|---|---|---|---|---|---|---|---|---|")
for n in order:
    c = [x.strip() for x in pick[n].split("|")]
    dur, rd, wr = float(c[2]), float(c[3]), float(c[4])
    gbs = (rd + wr) * 1e6 / (dur * 1e-6) / 1e9
    out.append(f"| {n} | {dur:.1f} | {rd + wr:.1f} | {gbs:.0f} | {100 * gbs / 6551:.1f} | {c[5]} | {c[6]} | {c[11]} | {c[10]} |")

ev = {k["kernel"]: k for k in json.loads(open(plain).read().strip().splitlines()[-1])["kernels"]}
tot, cnt = collections.defaultdict(float), collections.Counter()
for row in csv.DictReader([l for l in open(launches) if not l.startswith("==")]):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    n = row["Kernel Name"].split("(")[0].replace("hs::", "").replace("void ", "").split("<")[0]
    if n.startswith("at::"):
        continue
    v, u = float(row["Metric Value"].replace(",", "")), row["Metric Unit"]
    tot[n] += v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    cnt[n] += 1
s = sum(tot.values())
out += ["\n## Launch list vs CUDA-event shares\n",
        f"`profiles/{tag}_launches_ncu.csv` (half-shard launches, cold and serialised, so only the SHARES compare) against the "
        "`kernels` of the same command run without ncu (whole-shard launches of one un-graphed step, CUDA events).\n",
        "| kernel | ncu launches | ncu avg us (half shard) | ncu share % | event avg ms (whole shard) | event share % |",
        "|---|---|---|---|---|---|"]
for n in sorted(tot, key=lambda n: -tot[n]):
    e = ev.get(n)
    right = f"{e['avg_ms']:.4f} x{e['launches_per_step']:.0f} | {100 * e['share']:.1f}" if e else "— | —"
    out.append(f"| {n} | {cnt[n]} | {tot[n] / cnt[n]:.1f} | {100 * tot[n] / s:.1f} | {right} |")
print("\n".join(out))
