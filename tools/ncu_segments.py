"""Summarise an `ncu --page source --csv` SASS dump per code segment (segments are cut at BAR.SYNC
and at large backward branches), to see which pipeline stage of a fused kernel stalls on what.
usage: python tools/ncu_segments.py src.csv [min_samples]"""
import csv
import re
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_")]
ins = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    def num(h):
        try:
            return float(r[col[h]])
        except ValueError:
            return 0.0
    ins.append(dict(addr=int(r[0], 16), sass=r[col["Source"]].strip(), samples=num("# Samples"),
                    execd=num("Instructions Executed"), wf=num("L1 Wavefronts Shared"),
                    wf_ideal=num("L1 Wavefronts Shared Ideal"),
                    stalls={h: num(h) for h in stall_cols}))
tot_s = sum(i["samples"] for i in ins)
tot_e = sum(i["execd"] for i in ins)
print(f"{len(ins)} SASS instructions, {tot_e:.0f} executed warp-instr, {tot_s:.0f} samples")
segs, cur = [], []
for i in ins:
    cur.append(i)
    if re.search(r"\bBAR\.SYNC|\bBAR\.ARV|\bBAR\.RED", i["sass"]):
        segs.append(cur)
        cur = []
if cur:
    segs.append(cur)
mins = float(sys.argv[2]) if len(sys.argv) > 2 else 0
for k, s in enumerate(segs):
    smp = sum(i["samples"] for i in s)
    if smp < mins:
        continue
    ex = sum(i["execd"] for i in s)
    ops = Counter()
    for i in s:
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", i["sass"])
        ops[m.group(2) if m else "?"] += i["execd"]
    st = Counter()
    for i in s:
        for h, v in i["stalls"].items():
            st[h] += v
    wf = sum(i["wf"] for i in s)
    wfi = sum(i["wf_ideal"] for i in s)
    top = ", ".join(f"{h[6:]} {v / max(smp, 1) * 100:.0f}%" for h, v in st.most_common(6))
    topo = ", ".join(f"{o} {v / max(ex, 1) * 100:.0f}%" for o, v in ops.most_common(8))
    print(f"seg {k:2d} @{s[0]['addr'] & 0xfffff:05x} n={len(s):5d} exec {ex / tot_e * 100:5.1f}% samples {smp / tot_s * 100:5.1f}% "
          f"smem wf {wf:.3g} (ideal {wfi:.3g})\n      stalls: {top}\n      ops: {topo}")
