"""Shared-memory bank-conflict model for the PFA section layout (64-bit accesses: a half-warp of 16
lanes is conflict-free iff its 16 element indices are distinct mod 16).  Prints the worst
conflict degree per stage for a candidate lane mapping.  usage: python tools/bank_check.py [P]"""
import sys
from collections import Counter

P = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N, M, HP, SQ = 15 * P, 15 * P // 2, P // 2, 15 * P + 1


def elem(q, xs):
    return q * SQ + xs


def degree(addrs):
    """addrs: list of 32 (or fewer) element indices or None (inactive) -> total wavefronts"""
    wf = 0
    for h in (0, 1):
        a = [x for x in addrs[16 * h:16 * h + 16] if x is not None]
        if not a:
            continue
        c = Counter()
        seen = set()
        for x in a:
            if x in seen:
                continue  # broadcast
            seen.add(x)
            c[x % 16] += 1
        wf += max(c.values())
    return wf


def ka_of(i):  # unit-local column permutation (P=16: pairs (ka, ka+8) adjacent)
    if P == 16:
        return ((i & 1) << 3) | (i >> 1)
    if P == 8:
        return ((i & 1) << 2) | (i >> 1)
    return i


def unit_cols(u):
    if u < 7 * P:
        kb, ka = 1 + u // P, ka_of(u % P)
    elif u < M - 1:
        kb, ka = 0, u - 7 * P + 1
    else:
        return 0, HP * 15, True
    return ka * 15 + kb, ((P - ka) % P) * 15 + (15 - kb) % 15, False


def report(name, instrs):
    worst = max(instrs) if instrs else 0
    ideal = sum(1 for _ in instrs) * 2
    print(f"{name:28s} instr {len(instrs):5d} wavefronts {sum(instrs):6d} (ideal {ideal:6d}) worst/instr {worst}")


# FX15 store: lanes (qi, g), fixed (j, kb): xs = (4g+j)*15+kb ; P=16: lane = qi*4+g (8 row pairs/warp)
G4 = P // 4
rows_per_warp = 32 // G4 if P == 16 else (15 if P == 8 else 30)
ins = []
for j in range(4):
    for kb in range(15):
        ad = [None] * 32
        for l in range(rows_per_warp * G4):
            qi, g = l // G4, l % G4
            ad[l] = elem(qi, (4 * g + j) * 15 + kb)
        ins.append(degree(ad))
report("FX15 store", ins)
# FX16: items (qi, bx) 15 lanes per row pair per half warp; fixed ax
ins = []
for ax in range(P):
    ad = [None] * 32
    for h in (0, 1):
        for b in range(15):
            ad[16 * h + b] = elem(h, ax * 15 + b)
    ins.append(degree(ad))
report("FX16/IX16 ld/st", ins)
# IX15: items (qi, ka): half-warp = ka 0..P-1 (P=16); fixed kb
ins = []
for kb in range(15):
    ad = [None] * 32
    for l in range(32):
        qi, ka = l // P, l % P
        ad[l] = elem(qi, ka * 15 + kb)
    ins.append(degree(ad))
report("IX15 ld/st", ins)
# FY16: items (unit, by): half-warp = one unit, lanes by 0..14 ; fixed ayp, side
ins = []
for u0 in range(0, M, 2):
    for ayp in range(HP):
        for side in (0, 1):
            ad = [None] * 32
            for h in (0, 1):
                u = u0 + h
                if u >= M:
                    continue
                kc, mc, sp = unit_cols(u)
                for by in range(15):
                    ad[16 * h + by] = elem(ayp * 15 + by, mc if side else kc)
            ins.append(degree(ad))
report("FY16/IY16 ld/st", ins)
# FY15: items (unit_local, j): item = ul*HP + j, 32 per round ; fixed (half, by)
ins = []
units_per_round = 32 // HP
for u0 in range(0, M - 1, units_per_round):
    for half in (0, 1):
        for by in range(15):
            ad = [None] * 32
            for l in range(32):
                u, j = u0 + l // HP, l % HP
                if u >= M - 1:
                    continue
                kc, mc, sp = unit_cols(u)
                va = j if half == 0 else (HP if j == 0 else P - j)
                ad[l] = elem((va % HP) * 15 + by, kc if va < HP else mc)
            ins.append(degree(ad))
report("FY15/IY15 ld/st", ins)
