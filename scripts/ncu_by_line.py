#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` SASS dump by CUDA source line.

ncu's CSV export of the source page carries per-SASS-instruction counters but
not the source line; `nvdisasm -g` of the same cubin (built with -lineinfo)
does.  Both list the kernel's instructions in address order, so they are
aligned by position (and checked by opcode).

usage: ncu_by_line.py <ncu_source.csv> <cubin> <mangled-kernel-substring> [top N] [helper-lines-below]
"""
import collections
import csv
import re
import subprocess
import sys


HELPER_BELOW = 0   # lines below this (small inlined helpers at the top of the file) are attributed to their caller


def sass_lines(cubin, kern):
    out = subprocess.run(["nvdisasm", "-gi", cubin], capture_output=True, text=True).stdout
    ins, cur, on, chain, fresh = [], None, False, [], True
    for ln in out.splitlines():
        if ln.startswith("//--------------------- .text."):
            on = kern in ln
            continue
        if ln.startswith("//--------------------- ") and on and ".text." not in ln:
            on = False
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if m:
            # a run of //## lines is one inline chain, innermost frame first
            if fresh:
                chain, fresh = [], False
            if not chain:
                chain.append((m.group(1).split("/")[-1], int(m.group(2))))
            if m.group(3):
                chain.append((m.group(3).split("/")[-1], int(m.group(4))))
            cur = chain[0]
            for fr in chain:
                cur = fr
                if fr[0].endswith(".cu") and fr[1] >= HELPER_BELOW:
                    break
            continue
        fresh = True
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip(), cur))
    return ins


def main():
    src_csv, cubin, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    global HELPER_BELOW
    HELPER_BELOW = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    rows = list(csv.reader(open(src_csv)))
    hdr = rows[1]
    body = []
    for r in rows[2:]:          # first table only (ncu repeats a kernel that matches twice)
        if r and r[0] == "Kernel Name":
            break
        body.append(r)
    ins = sass_lines(cubin, kern)
    assert len(ins) == len(body), (len(ins), len(body))
    col = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed", "L1 Wavefronts Shared",
                                     "L1 Wavefronts Shared Ideal", "L2 Theoretical Sectors Global", "stall_short_sb",
                                     "stall_barrier", "stall_long_sb", "stall_wait", "stall_mio", "stall_lg",
                                     "L2 Theoretical Sectors Local")}
    agg = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    for (addr, op, loc), r in zip(ins, body):
        a = r[col["Source"]].split()
        b = op.split()
        # opcode check (predicates may lead)
        oa = [x for x in a if not x.startswith("@")][0]
        ob = [x for x in b if not x.startswith("@")][0]
        assert oa.split(".")[0] == ob.split(".")[0], (hex(addr), a, b)
        for n, i in col.items():
            if n == "Source":
                continue
            try:
                v = float(r[i])
            except ValueError:
                v = 0
            agg[loc][n] += v
            tot[n] += v
        agg[loc]["ninst"] += 1
    print("totals:", {k: int(v) for k, v in tot.items()})
    for key in ("L1 Wavefronts Shared", "# Samples", "Instructions Executed"):
        print(f"\n== top lines by {key}")
        for loc, c in sorted(agg.items(), key=lambda kv: -kv[1][key])[:top]:
            print(f"{loc[0] if loc else '?'}:{loc[1] if loc else 0:5d}  {key}={int(c[key]):>12d} ({100 * c[key] / max(tot[key], 1):5.1f}%)"
                  f"  wf={int(c['L1 Wavefronts Shared']):>11d} ideal={int(c['L1 Wavefronts Shared Ideal']):>11d}"
                  f"  inst={int(c['Instructions Executed']):>10d} smp={int(c['# Samples']):>6d}"
                  f"  ssb={int(c['stall_short_sb'])} bar={int(c['stall_barrier'])} lsb={int(c['stall_long_sb'])} wait={int(c['stall_wait'])} mio={int(c['stall_mio'])} lg={int(c['stall_lg'])}")


if __name__ == "__main__":
    main()
