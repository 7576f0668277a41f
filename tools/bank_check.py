"""Shared-memory bank-conflict model for the PFA section layout (64-bit accesses: a half-warp of 16
lanes is conflict-free iff its 16 element indices are distinct mod 16).  Prints the worst
conflict degree per stage for a candidate lane mapping.  usage: python tools/bank_check.py [P]"""
import sys
from collections import Counter

P = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N, M, HP, SQ = 15 * P, 15 * P // 2, P // 2, 15 * P + 1


def slot(a, b):
    return a * 15 + b if P == 16 else b * P + a


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
    return i


def unit_cols(u):
    if u < 7 * P:
        kb, ka = 1 + u // P, ka_of(u % P)
    elif u < M - 1:
        kb, ka = 0, u - 7 * P + 1
    else:
        return slot(0, 0), slot(HP, 0), True
    return slot(ka, kb), slot((P - ka) % P, (15 - kb) % 15), False


RPT = 8 if P == 16 else (15 if P == 8 else 30)
TASKS = M // RPT
G4 = P // 4


def x16_item(r, lane):
    if P == 16:
        qi, b = 2 * r + (lane >> 4), lane & 15
        return (qi, b) if b < 15 else None
    if P == 8:
        qi, b = lane & 15, 2 * r + (lane >> 4)
        return (qi, b) if qi < 15 and b < 15 else None
    return (lane, r) if lane < 30 else None


def x15_item(r, lane):
    if P == 16:
        it = 32 * r + lane
        return it >> 4, it & 15
    if P == 8:
        qi, ka = lane & 15, 2 * r + (lane >> 4)
        return (qi, ka) if qi < 15 else None
    return (lane, r) if lane < 30 else None


def y15_item(r, lane, nu):
    if P == 16:
        it = 32 * r + lane
        ul, j = it >> 3, it & 7
    elif P == 8:
        ul, j = lane & 15, 2 * r + (lane >> 4)
    else:
        ul, j = 16 * r + (lane & 15), lane >> 4
    return (ul, j) if ul < nu else None


def y15_rounds(nu):
    return (nu * HP + 31) // 32 if P == 16 else (2 if P == 8 else (nu + 15) // 16)


def fx15_lane(lane):
    qi, g = (lane & 15, lane >> 4) if P == 8 else (lane // G4, lane % G4)
    return (qi, g) if qi < RPT else None


def report(name, instrs):
    worst = max(instrs) if instrs else 0
    ideal = sum(i for _, i in instrs)
    got = sum(w for w, _ in instrs)
    print(f"{name:28s} instr {len(instrs):5d} wavefronts {got:6d} (ideal {ideal:6d}) worst/instr {max(w for w, _ in instrs)}")


def run(name, gen):
    """gen yields lists of 32 element indices (None = inactive lane), one per warp instruction"""
    ins = []
    for ad in gen:
        act = [x for x in ad if x is not None]
        if not act:
            continue
        ideal = (1 if any(x is not None for x in ad[:16]) else 0) + (1 if any(x is not None for x in ad[16:]) else 0)
        ins.append((degree(ad), ideal))
    report(name, ins)


def gen_fx15():
    for j in range(4):
        for kb in range(15):
            ad = [None] * 32
            for l in range(32):
                it = fx15_lane(l)
                if it:
                    ad[l] = elem(it[0], slot(4 * it[1] + j, kb))
            yield ad


def gen_x16():
    rounds = RPT // 2 if P == 16 else (8 if P == 8 else 15)
    for r in range(rounds):
        for a in range(P):
            ad = [None] * 32
            for l in range(32):
                it = x16_item(r, l)
                if it:
                    ad[l] = elem(it[0], slot(a, it[1]))
            yield ad


def gen_x15():
    for r in range(4):
        for b in range(15):
            ad = [None] * 32
            for l in range(32):
                it = x15_item(r, l)
                if it:
                    ad[l] = elem(it[0], slot(it[1], b))
            yield ad


def gen_y16():
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
                yield ad


def gen_y15():
    for task in range(TASKS):
        u0 = task * RPT
        nu = RPT - 1 if task == TASKS - 1 else RPT
        for r in range(y15_rounds(nu)):
            for half in (0, 1):
                for by in range(15):
                    ad = [None] * 32
                    for l in range(32):
                        it = y15_item(r, l, nu)
                        if not it:
                            continue
                        kc, mc, sp = unit_cols(u0 + it[0])
                        j = it[1]
                        va = j if half == 0 else (HP if j == 0 else P - j)
                        ad[l] = elem((va % HP) * 15 + by, kc if va < HP else mc)
                    yield ad


run("FX15 store", gen_fx15())
run("FX16/IX16 ld/st (per task)", gen_x16())
run("IX15 ld/st (per task)", gen_x15())
run("FY16/IY16 ld/st", gen_y16())
run("FY15/IY15 ld/st", gen_y15())

if len(sys.argv) > 2:  # detail: conflicted FY15 instructions
    k = 0
    for task in range(TASKS):
        nu = RPT - 1 if task == TASKS - 1 else RPT
        for r in range(y15_rounds(nu)):
            for half in (0, 1):
                for by in range(15):
                    ad = [None] * 32
                    for l in range(32):
                        it = y15_item(r, l, nu)
                        if it:
                            kc, mc, sp = unit_cols(task * RPT + it[0])
                            va = it[1] if half == 0 else (HP if it[1] == 0 else P - it[1])
                            ad[l] = elem((va % HP) * 15 + by, kc if va < HP else mc)
                    d = degree(ad)
                    if d > 2 and by == 0:
                        print("task", task, "round", r, "half", half, "deg", d,
                              [None if x is None else x % 16 for x in ad])
