#!/usr/bin/env python
"""Aggregates the source page of an ncu report (ncu -i X.ncu-rep --page source --csv --print-source cuda,sass) by CUDA source line:
warp-stall samples, instructions executed and shared-memory wavefronts per line, per kernel instance.
usage: ncu_source_hot.py <source.csv> [kernel substring] [top N]"""
import csv, sys, collections

def main():
    path = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    rows = list(csv.reader(open(path)))
    i = 0; seen = collections.Counter()
    while i < len(rows):
        if rows[i] and rows[i][0] == "File Path":
            fn = rows[i + 1][1]; hdr = rows[i + 2]; i += 3
            body = []
            while i < len(rows) and not (rows[i] and rows[i][0] == "File Path"):
                body.append(rows[i]); i += 1
            if want not in fn: continue
            seen[fn] += 1
            if "Warp Stall Sampling (All Samples)" not in hdr: continue
            cs = hdr.index("Warp Stall Sampling (All Samples)"); ci = hdr.index("Instructions Executed"); cw = hdr.index("L1 Wavefronts Shared")
            # the cuda view rows have a "Line No"; the sass rows have an address: use the cuda view (first block of the section)
            agg = {}
            for r in body:
                if len(r) <= cw or not r[0].strip().isdigit(): continue
                try: s = float(r[cs] or 0); n = float(r[ci] or 0); w = float(r[cw] or 0)
                except ValueError: continue
                if s or n: agg[int(r[0])] = (s, n, w, r[1].strip()[:110])
            ts = sum(v[0] for v in agg.values()) or 1; tn = sum(v[1] for v in agg.values()) or 1; tw = sum(v[2] for v in agg.values()) or 1
            print(f"==== {fn} #{seen[fn]}: samples {ts:.0f}, instructions {tn:.0f}, shared wavefronts {tw:.0f}")
            for ln, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
                print(f"  line {ln:4d} samples {100 * v[0] / ts:5.1f} %  inst {100 * v[1] / tn:5.1f} %  smem wf {100 * v[2] / tw:5.1f} %  | {v[3]}")
        else:
            i += 1

if __name__ == "__main__":
    main()
