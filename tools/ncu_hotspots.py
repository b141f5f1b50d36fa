#!/usr/bin/env python
"""Aggregate an ncu SASS source page by CUDA source line (needs -lineinfo at compile time).

  ncu -i prof.ncu-rep --page source --csv > src.csv
  python tools/ncu_hotspots.py src.csv laser_slam_b200/libzhsm.so <kernel-name-substring> [top]

ncu's CSV source page is SASS-only; nvdisasm -g gives the file:line of every SASS instruction of the same
cubin.  Instructions are matched by text, in order, inside the function whose name contains the substring
(and the __noinline__ device functions it calls, which ncu lists after the kernel)."""
import csv
import collections
import os
import re
import subprocess
import sys
import tempfile


def disasm_lines(so, kernel_sub):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    out = []  # (func, instr_text, file, line)
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, text=True).stdout
        if kernel_sub not in txt:
            continue
        func, cur = None, ("?", 0)
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                func = m.group(1)
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m and func:
                out.append((func, int(m.group(1), 16), m.group(2).strip(), cur[0], cur[1]))
    return out


def main():
    src_csv, so, ksub = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    # locate the header row (starts with "Address")
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    which = int(os.environ.get("NCU_LAUNCH", "0"))  # a report with several launches repeats the header per launch
    heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    hi = heads[min(which, len(heads) - 1)]
    end = heads[heads.index(hi) + 1] if heads.index(hi) + 1 < len(heads) else len(rows)
    body = [r for r in rows[hi + 1:end] if len(r) == len(hdr) and r[0] != "Address"]
    dis = disasm_lines(so, ksub)
    # group disassembly per function, index by offset
    per_func = collections.defaultdict(dict)
    for func, off, text, f, l in dis:
        per_func[func][off] = (text, f, l)
    # split ncu rows into address-contiguous runs (one per function) and match each run to a function by its
    # first instructions' text
    runs, cur = [], []
    prev = None
    for r in body:
        a = int(r[col["Address"]], 16)
        if prev is not None and a != prev + 16:
            runs.append(cur)
            cur = []
        cur.append((a, r))
        prev = a
    if cur:
        runs.append(cur)
    agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
    total_samples = 0
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for run in runs:
        base = run[0][0]
        n = len(run)
        cand = [fn for fn, d in per_func.items() if len(d) >= n - 8 and len(d) <= n + 8]
        best = None
        for fn in cand:
            d = per_func[fn]
            ok = sum(1 for k in range(min(n, 40)) if d.get(16 * k, ("",))[0].split()[0:1] == run[k][1][col["Source"]].split()[0:1])
            if best is None or ok > best[0]:
                best = (ok, fn)
        fn = best[1] if best else None
        for a, r in run:
            off = a - base
            info = per_func.get(fn, {}).get(off)
            key = (info[1], info[2]) if info else ("?", 0)
            s = int(r[col["# Samples"]] or 0)
            ie = int(r[col["Instructions Executed"]] or 0)
            te = int(r[col["Thread Instructions Executed"]] or 0)
            agg[key][0] += s
            agg[key][1] += ie
            agg[key][2] += te
            for h in stall_cols:
                v = int(r[col[h]] or 0)
                if v:
                    agg[key][3][h] += v
            total_samples += s
    tot_ie = sum(v[1] for v in agg.values())
    print(f"total samples {total_samples}, warp instructions {tot_ie}")
    print(f"{'file:line':32s} {'samples%':>8s} {'inst%':>7s} {'thr/inst':>8s}  top stalls")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        st = ", ".join(f"{k[6:]}={c * 100 // max(1, v[0])}%" for k, c in v[3].most_common(3))
        print(f"{key[0] + ':' + str(key[1]):32s} {100.0 * v[0] / max(1, total_samples):8.2f} {100.0 * v[1] / max(1, tot_ie):7.2f} "
              f"{v[2] / max(1, v[1]):8.1f}  {st}")


if __name__ == "__main__":
    main()
