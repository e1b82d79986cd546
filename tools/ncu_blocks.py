#!/usr/bin/env python
"""Condense an `ncu --page source --csv --print-source sass` export into basic blocks:
   per block: #instr, warp-instructions executed, avg active threads, stall samples, memory ops.
   usage: ncu_blocks.py prof_src.csv [kernel-index]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
kern = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = dict(name=r[1], rows=[], hdr=None); kern.append(cur)
    elif r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) >= len(cur["hdr"]) - 2:
        cur["rows"].append(r)
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
k = kern[which]
H = k["hdr"]
ci = {n: i for i, n in enumerate(H)}
iS, iX, iT, iSm = ci["Source"], ci["Instructions Executed"], ci["Thread Instructions Executed"], ci["# Samples"]
iL = ci.get("stall_long_sb"); 
print(k["name"][:100], "instrs", len(k["rows"]))
tot_x = sum(int(r[iX]) for r in k["rows"]); tot_t = sum(int(r[iT]) for r in k["rows"]); tot_s = sum(int(r[iSm]) for r in k["rows"])
print("warp-inst %d thread-inst %d avg %.2f samples %d" % (tot_x, tot_t, tot_t / max(1, tot_x), tot_s))
blocks = []; b = None
for n, r in enumerate(k["rows"]):
    x = int(r[iX]); t = int(r[iT])
    op = r[iS].split()
    key = (x, t)
    isbr = any(o.startswith(("BRA", "BSSY", "BSYNC", "EXIT", "RET", "CALL", "WARPSYNC", "BRX")) for o in op[:3])
    if b is None or b["key"] != key:
        b = dict(key=key, start=n, n=0, x=x, t=t, s=0, lsb=0, mem=[]); blocks.append(b)
    b["n"] += 1; b["s"] += int(r[iSm]); b["lsb"] += int(r[iL]) if iL else 0
    src = r[iS].strip()
    if any(m in src for m in ("LDG", "STG", "LDL", "STL", "ATOM", "RED", "LDS", "STS", "LDC", "SHFL", "VOTE", "MUFU", "BRA", "BSSY", "BSYNC", "CALL", "DMUL", "DADD", "F2F")):
        b["mem"].append(src.split(";")[0][:46])
    if isbr: b = None
print("%5s %4s %10s %6s %7s %7s  ops" % ("pc", "n", "warp-inst", "thr", "samp%", "lsb%"))
for b in blocks:
    if b["x"] * b["n"] < tot_x * 0.002 and b["s"] < tot_s * 0.004: continue
    print("%5d %4d %10d %6.1f %7.2f %7.2f  %s" % (b["start"], b["n"], b["x"], b["t"] / max(1, b["x"]), 100.0 * b["s"] / max(1, tot_s), 100.0 * b["lsb"] / max(1, tot_s), " | ".join(b["mem"][:6])))
