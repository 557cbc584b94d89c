#!/usr/bin/env python
"""Shared-memory bank-conflict model of the FFT tile engine (csrc/fft.cu): replays the address pattern of every
shared-memory access of a tile transform and counts wavefronts (64-bit accesses: two half-warp phases, 16 bank pairs).

    python tools/bank_conflicts.py            # current layout, all production tile shapes
"""
import sys
import numpy as np

THREADS = 256


def radices(L):
    if L == 1024:
        return [32, 32]
    if L == 512:
        return [32, 16]
    out, rem = [], L
    if rem % 3 == 0:
        out.append(3); rem //= 3
    while rem > 1:
        r = 16 if rem >= 16 else rem
        out.append(r); rem //= r
    return out


def phys_default(pos):
    return pos + (pos >> 4) + (pos >> 8)


def pitch_default(L, rows, phys):
    want = 1 if rows >= 16 else 16 // rows
    base = phys(L)
    return base + ((want - base % 16) + 16) % 16


def wavefronts(addrs):
    """addrs: float2-unit addresses of the 32 lanes (or None for inactive). 64-bit access = 2 half-warp phases."""
    total = 0
    for half in (addrs[:16], addrs[16:]):
        banks = {}
        for a in half:
            if a is None:
                continue
            banks.setdefault(a % 16, set()).add(a)
        total += max((len(v) for v in banks.values()), default=0)
    return total


def stage_patterns(L, rows, phys, pitch):
    rad = radices(L)
    res = []
    n = L
    for s, R in enumerate(rad):
        S = n // R
        BF = L // R
        items = rows * BF
        last = s == len(rad) - 1
        for map_row_fast, label in ((True, "row-fast"), (False, "bf-fast")):
            tot = ideal = 0
            for it in range(items // THREADS):
                for w in range(THREADS // 32):
                    for j in range(R):
                        addrs = []
                        for lane in range(32):
                            item = w * 32 + lane + it * THREADS
                            if map_row_fast:
                                row, bf = item % rows, item // rows
                            else:
                                row, bf = item // BF, item % BF
                            blk, p = bf // S, bf % S
                            addrs.append(row * pitch + phys(blk * n + p + j * S))
                        tot += wavefronts(addrs)
                        ideal += 2
            res.append((s, R, S, label, tot / ideal))
        n //= R
    return res


def pos_of_freq(L, k):
    rad = radices(L)
    pos, n = 0, L
    for R in rad:
        S = n // R
        q = (k // (L // n)) % R
        pos += q * S
        n //= R
    return pos


def untangle_pattern(L, rows, phys, pitch):
    T2 = rows // 2
    tot = ideal = 0
    for base in range(0, T2 * L, THREADS):
        for w in range(THREADS // 32):
            a1, a2 = [], []
            for lane in range(32):
                item = base + w * 32 + lane
                r, k1 = item % T2, item // T2
                k1m = L - 1 - k1
                a1.append(r * pitch + phys(pos_of_freq(L, k1)))
                a2.append((r + T2) * pitch + phys(pos_of_freq(L, k1m)))
            tot += wavefronts(a1) + wavefronts(a2)
            ideal += 4
    return tot / ideal


def report(L, rows, phys=phys_default, pitch=None, quiet=False):
    pitch = pitch_default(L, rows, phys) if pitch is None else pitch
    pats = stage_patterns(L, rows, phys, pitch)
    un = untangle_pattern(L, rows, phys, pitch)
    if not quiet:
        print(f"L={L} rows={rows} pitch={pitch} radices={radices(L)}")
        for s, R, S, label, ratio in pats:
            print(f"   stage {s} R={R:2d} S={S:4d} {label:8s} wavefronts/ideal = {ratio:.3f}")
        print(f"   untangle reads (pass 2)          wavefronts/ideal = {un:.3f}")
    return pats, un


if __name__ == "__main__":
    shapes = [(2048, 4), (1024, 8), (512, 16), (4096, 2), (3072, 4)] if len(sys.argv) < 2 else [tuple(map(int, a.split("x"))) for a in sys.argv[1:]]
    for L, rows in shapes:
        report(L, rows)


def search(L, rows, which):
    """brute-force the row pitch (mod 32) and a few padding rules; `which` = 'pass1' or 'pass2'"""
    best = []
    rules = {
        "p4+p8": lambda pos: pos + (pos >> 4) + (pos >> 8),
        "p4": lambda pos: pos + (pos >> 4),
        "p4+p8+p12": lambda pos: pos + (pos >> 4) + (pos >> 8) + (pos >> 12),
        "p5": lambda pos: pos + (pos >> 5),
        "p3+p7": lambda pos: pos + (pos >> 3) + (pos >> 7),
        "p4+2p8": lambda pos: pos + (pos >> 4) + 2 * (pos >> 8),
        "p4+p7": lambda pos: pos + (pos >> 4) + (pos >> 7),
        "p4+p6": lambda pos: pos + (pos >> 4) + (pos >> 6),
    }
    nst = len(radices(L))
    for name, phys in rules.items():
        base = phys(L)
        for extra in range(0, 32):
            pitch = base + extra
            pats = stage_patterns(L, rows, phys, pitch)
            cost = 0.0
            for s, R, S, label, ratio in pats:
                if which == "pass1":
                    used = (label == "row-fast") if s in (0, nst - 1) else (label == "bf-fast")
                    weight = 1 if s in (0, nst - 1) else 2
                else:
                    used = label == "bf-fast"
                    weight = 1 if s == 0 else 2
                if used:
                    cost += weight * ratio
            if which == "pass2":
                cost += untangle_pattern(L, rows, phys, pitch)
            best.append((cost, name, extra, pitch))
    best.sort()
    return best[:6]
