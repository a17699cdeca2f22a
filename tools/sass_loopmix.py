#!/usr/bin/env python
"""Instruction mix of the innermost loops of one kernel in librtdose.so (static, from cuobjdump -sass): for the superposition
kernels these are the per-radius column-offset loops of ks2_tile.  usage: sass_loopmix.py <mangled kernel name> [lib]"""
import collections, re, subprocess, sys

def main():
    fun = sys.argv[1]; lib = sys.argv[2] if len(sys.argv) > 2 else "raytracedicom_b200/librtdose.so"
    txt = subprocess.run(["cuobjdump", "-sass", "-fun", fun, lib], capture_output=True, text=True).stdout
    ins = [(int(m.group(1), 16), m.group(2).strip()) for m in (re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", l) for l in txt.splitlines()) if m]
    idx = {a: i for i, (a, _) in enumerate(ins)}
    print(f"{fun}: {len(ins)} instructions, {16 * len(ins)} bytes of code")
    loops = []
    for a, t in ins:
        m = re.search(r"0x([0-9a-f]+)", t) if "BRA" in t else None
        if m and int(m.group(1), 16) < a and int(m.group(1), 16) in idx:
            loops.append((idx[a] - idx[int(m.group(1), 16)] + 1, int(m.group(1), 16), a))
    print("instr  bytes FFMA2  FFMA FADD2 FADD FMUL  LDS  STS  MOV other   (innermost loops with >= 8 FFMA + FFMA2)")
    for n, tgt, a in sorted(loops):
        ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in ins[idx[tgt]:idx[a] + 1])
        if ops["FFMA"] + ops["FFMA2"] >= 8 and n < 700:
            named = sum(ops[k] for k in ("FFMA2", "FFMA", "FADD2", "FADD", "FMUL", "LDS", "STS", "MOV"))
            print(f"{n:5d} {16 * n:6d} {ops['FFMA2']:5d} {ops['FFMA']:5d} {ops['FADD2']:5d} {ops['FADD']:4d} {ops['FMUL']:4d} {ops['LDS']:4d} {ops['STS']:4d} {ops['MOV']:4d} {n - named:5d}")

if __name__ == "__main__":
    main()
