#!/usr/bin/env python
"""Layout search for a pass-1 FFT tile (2048 rows x 4 columns) whose global traffic goes through TMA tensor copies:
the tile lands DENSE in shared memory ([pos][col], 32 bytes per row), stage 0 reads it from there, the stages exchange
through the PADDED layout ([col][phys(pos)]), and the last stage writes the result dense again ([k2][col]) for the TMA
store.  Counts shared-memory wavefronts of every access pattern for a set of padding rules / pitches / lane mappings.

    python tools/pass1_tma_layout.py
"""
import itertools

L, ROWS, NT = 2048, 4, 512
RAD = [16, 16, 8]


def wavefronts(addrs):
    total = 0
    for half in (addrs[:16], addrs[16:]):
        banks = {}
        for a in half:
            banks.setdefault(a % 16, set()).add(a)
        total += max(len(v) for v in banks.values())
    return total


def freq_of_pos(pos):
    k, n = 0, L
    for R in RAD:
        S = n // R
        q = (pos // S) % R
        k += q * (L // n)
        n //= R
    return k


def rot(bf):
    return (bf & ~15) | ((bf & 1) << 3) | ((bf >> 1) & 7)


def transpose16(x):  # 16 x 16 index transpose of an 8-bit butterfly number
    return 16 * (x % 16) + x // 16


def skewed_transpose(x):
    """like transpose16, with the low nibble of the butterfly number rotated by 2*t(a), t = (0, 1, 1, 2)[a % 4]: four consecutive
    lanes then differ in (bf/2 + bf/32) mod 4 as well as in bf/16 mod 4, which is what a pitch of 4 (mod 16) needs in BOTH the padded
    read and the dense write of the last stage"""
    a, hi = x % 16, x // 16
    return 16 * a + (hi + 2 * (0, 1, 1, 2)[a % 4]) % 16


LANE_MAPS = {
    "row-fast": lambda item, BF: (item % ROWS, item // ROWS),
    "row-fast+rot": lambda item, BF: (item % ROWS, rot(item // ROWS)),
    "row-fast+T": lambda item, BF: (item % ROWS, transpose16(item // ROWS) if BF == 256 else item // ROWS),
    "row-fast+T2": lambda item, BF: (item % ROWS, skewed_transpose(item // ROWS) if BF == 256 else item // ROWS),
    "bf-fast": lambda item, BF: (item // BF, item % BF),
}


def stage_cost(s, lane_map, addr_of):
    """average wavefronts / ideal over all warp-level accesses of stage s with the given lane map and address function"""
    n = L
    for R in RAD[:s]:
        n //= R
    R = RAD[s]
    S = n // R
    BF = L // R
    tot = ideal = 0
    for it in range(ROWS * BF // NT):
        for w in range(NT // 32):
            for j in range(R):
                addrs = []
                for lane in range(32):
                    row, bf = LANE_MAPS[lane_map](w * 32 + lane + it * NT, BF)
                    blk, p = bf // S, bf % S
                    addrs.append(addr_of(row, blk * n + p, j, S))
                tot += wavefronts(addrs)
                ideal += 2
    return tot / ideal


def main():
    rules = {
        "p4": lambda pos: pos + (pos >> 4),
        "p4+p8": lambda pos: pos + (pos >> 4) + (pos >> 8),
        "p4+p7": lambda pos: pos + (pos >> 4) + (pos >> 7),
        "p4+p7+p11": lambda pos: pos + (pos >> 4) + (pos >> 7) + (pos >> 11),
        "p3+p7": lambda pos: pos + (pos >> 3) + (pos >> 7),
        "p4+p6": lambda pos: pos + (pos >> 4) + (pos >> 6),
        "p4+p5+p7": lambda pos: pos + (pos >> 4) + (pos >> 5) + (pos >> 7),
        "p5+p7": lambda pos: pos + (pos >> 5) + (pos >> 7),
        "p4+2p7": lambda pos: pos + (pos >> 4) + 2 * (pos >> 7),
    }
    dense_in = lambda row, base, j, S: (base + j * S) * ROWS + row                       # [pos][col]
    dense_out = lambda row, base, j, S: (freq_of_pos(base) + j * (L // RAD[-1])) * ROWS + row  # [k2][col], q = j
    results = []
    for (name, phys), extra in itertools.product(rules.items(), range(0, 20)):
        pitch = phys(L) + extra
        padded = lambda row, base, j, S: row * pitch + phys(base + j * S)
        for m0, m2 in itertools.product(("row-fast", "row-fast+rot"), ("row-fast", "row-fast+T", "row-fast+T2")):
            c = {
                "s0 dense LDS": stage_cost(0, m0, dense_in),
                "s0 padded STS": stage_cost(0, m0, padded),
                "s1 LDS": stage_cost(1, "bf-fast", padded),
                "s1 STS": stage_cost(1, "bf-fast", padded),
                "s2 padded LDS": stage_cost(2, m2, padded),
                "s2 dense STS": stage_cost(2, m2, dense_out),
            }
            results.append((sum(c.values()), name, extra, pitch * ROWS * 8, m0, m2, c))
    results.sort(key=lambda r: r[0])
    for tot, name, extra, nbytes, m0, m2, c in results[:8]:
        print(f"{tot:6.3f}  {name:10s} extra={extra:2d} bytes={nbytes}  stage0={m0:13s} last={m2:11s} " +
              " ".join(f"{k}={v:.2f}" for k, v in c.items()))


if __name__ == "__main__":
    main()


def one(rule_name, extra, m0, m2):
    rules = {"p4": lambda pos: pos + (pos >> 4), "p4+p8": lambda pos: pos + (pos >> 4) + (pos >> 8)}
    phys = rules[rule_name]
    pitch = phys(L) + extra
    padded = lambda row, base, j, S: row * pitch + phys(base + j * S)
    dense_in = lambda row, base, j, S: (base + j * S) * ROWS + row
    dense_out = lambda row, base, j, S: (freq_of_pos(base) + j * (L // RAD[-1])) * ROWS + row
    return {"s0 dense LDS": stage_cost(0, m0, dense_in), "s0 padded STS": stage_cost(0, m0, padded),
            "s1": stage_cost(1, "bf-fast", padded), "s2 padded LDS": stage_cost(2, m2, padded),
            "s2 dense STS": stage_cost(2, m2, dense_out)}
