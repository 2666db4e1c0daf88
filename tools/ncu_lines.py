#!/usr/bin/env python
"""Per-source-line instruction / stall attribution of one profiled kernel.

Joins the SASS page of an `ncu --set full --import-source on` report (instructions executed and
stall samples per SASS instruction) with `nvdisasm -gi` line info of the same cubin (which carries
the whole inline chain), so a line of the kernel body is charged with everything inlined under it.

  python tools/ncu_lines.py REPORT.ncu-rep OBJECT.o KERNEL_SUBSTR [--skip N] [--depth D] [--top T]

--depth 0 charges the outermost frame (the line in the kernel body), 1 the frame below it, ...;
--inner charges the innermost frame instead.
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import tempfile


def sass_lineinfo(obj, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
    cubins = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")]
    out = subprocess.run(["nvdisasm", "-gi", "-c"] + cubins, capture_output=True, text=True).stdout
    # "7k_solve" must not match "17k_solve_positions" or "14k_solve_global": anchor both ends
    pat = re.compile(r"(?<![0-9])" + re.escape(kernel) + (r"(?=[EI])" if kernel[-1].isalnum() and "IL" not in kernel else ""))
    frames, cur, infn, table = [], [], False, {}
    for line in out.splitlines():
        if line.startswith(".text."):
            infn = pat.search(line) is not None
            continue
        if line.lstrip().startswith("//---") and infn and ".text." in line and not pat.search(line):
            infn = False
        if not infn:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', line)
        if m:
            cur.append((os.path.basename(m.group(1)), int(m.group(2))))
            if not m.group(3):
                frames, cur = cur, []       # chain complete: innermost ... outermost
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*);", line)
        if m:
            table[int(m.group(1), 16)] = (list(frames), m.group(2).strip())
    return table


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("obj")
    ap.add_argument("kernel", help="kernel name as ncu shows it (regex), e.g. k_solve")
    ap.add_argument("--mangled", default=None, help="substring of the mangled name in the cubin (default: <len><kernel>)")
    ap.add_argument("--skip", type=int, default=0, help="launches of this kernel to skip in the report")
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--inner", action="store_true")
    ap.add_argument("--top", type=int, default=40)
    a = ap.parse_args()
    table = sass_lineinfo(a.obj, a.mangled or f"{len(a.kernel)}{a.kernel}")
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "-k", "regex:^" + a.kernel + "$",
                          "--launch-skip", str(a.skip), "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    H = {n: i for i, n in enumerate(rows[hdr])}
    base = None
    inst = collections.Counter(); thr = collections.Counter(); stall = collections.Counter(); longsb = collections.Counter()
    tot_i = tot_s = 0
    for r in rows[hdr + 1:]:
        if r and r[0] in ("Address", "Kernel Name"):
            break                           # next launch's section
        if len(r) < len(H):
            continue
        addr = int(r[0], 16)
        base = addr if base is None else base
        off = addr - base
        fr, _ = table.get(off, ([], ""))
        if not fr:
            key = ("?", 0)
        elif a.inner:
            key = fr[0]
        else:
            key = fr[max(0, len(fr) - 1 - a.depth)]
        n = int(r[H["Instructions Executed"]]); t = int(r[H["Thread Instructions Executed"]])
        s = int(r[H["Warp Stall Sampling (All Samples)"]])
        inst[key] += n; thr[key] += t; stall[key] += s
        if "stall_long_sb" in H:
            longsb[key] += int(r[H["stall_long_sb"]])
        tot_i += n; tot_s += s
    print(f"kernel {a.kernel}: {tot_i} warp-instructions, {tot_s} stall samples")
    print(f"{'file:line':32s} {'winst':>10s} {'%':>6s} {'lanes':>6s} {'samples':>8s} {'%':>6s} {'long_sb':>8s}")
    for key, n in inst.most_common(a.top):
        print(f"{key[0] + ':' + str(key[1]):32s} {n:10d} {100.0 * n / tot_i:6.2f} {thr[key] / max(n, 1):6.1f} "
              f"{stall[key]:8d} {100.0 * stall[key] / max(tot_s, 1):6.2f} {longsb[key]:8d}")


if __name__ == "__main__":
    main()
