"""Host-side model of the work partition of the load-balanced peel kernel (k_peel_dynamic in
evox_b200/csrc/nds.cu): the same integer arithmetic, replayed in Python for random fronts.  It checks
what the kernel relies on: CTA ranges tile the cost line, owner() inverts them, every group of every
pair is gathered exactly once, every pair with members is committed exactly once, the number of CTAs
that share a pair is what the last-arriver count expects, and the parking slots (pair + CTA) of one
front never collide.  No GPU; the constants are read from the source so the model cannot drift."""
import re
from pathlib import Path

import numpy as np
import pytest

SRC = (Path(__file__).resolve().parent.parent / "evox_b200" / "csrc" / "nds.cu").read_text()


def _const(name):
    m = re.search(rf"#define {name} (\d+)", SRC)
    assert m, name
    return int(m.group(1))


WPG, CW, CC = _const("EVX_WPG"), _const("EVX_CW"), _const("EVX_CC")


def owner(u, T, G):                       # CTA whose range [b*T//G, (b+1)*T//G) holds cost unit u
    return ((u + 1) * G + T - 1) // T - 1


def simulate(rows_per_pair, G):
    """rows_per_pair[lp] = front members pair lp has to fetch (pre[2Q+2] in the kernel)."""
    nq = len(rows_per_pair)
    groups = [(r + 31) >> 5 for r in rows_per_pair]
    cost = [(g + (CW if g <= WPG else CC)) if r else 0 for g, r in zip(groups, rows_per_pair)]
    S = np.concatenate([[0], np.cumsum(cost)]).astype(np.int64)
    T = int(S[-1])
    gathered = [np.zeros(g, int) for g in groups]
    commits = np.zeros(nq, int)
    arrivals = np.zeros(nq, int)
    expected = np.zeros(nq, int)
    slots = set()
    for B in range(G):
        lo, hi = B * T // G, (B + 1) * T // G
        if lo >= hi:
            continue
        # first pair with S[lp+1] > lo (the kernel's binary search)
        Qa = int(np.searchsorted(S, lo, side="right") - 1)
        assert S[Qa] <= lo < S[Qa + 1] or S[Qa] == S[Qa + 1]
        warp_pairs = []
        lp = Qa
        while lp < nq and S[lp] < hi:
            tot = int(S[lp + 1] - S[lp])
            if 0 < tot <= WPG + CW:
                gq, g0 = tot - CW, int(S[lp]) + CW
                if g0 >= lo and g0 + gq <= hi:
                    warp_pairs.append(lp)
                    gathered[lp][:gq] += 1
                    commits[lp] += 1
            lp += 1
        lp = Qa
        while lp < nq and S[lp] < hi:
            tot = int(S[lp + 1] - S[lp])
            if tot > 0:
                fixed = CW if tot <= WPG + CW else CC
                gq, g0 = tot - fixed, int(S[lp]) + fixed
                assert gq == groups[lp]
                ga, gb = max(lo, g0) - g0, min(hi, g0 + gq) - g0
                if ga < gb and not (gq <= WPG and ga == 0 and gb == gq):
                    assert lp not in warp_pairs
                    bf, bl = owner(g0, T, G), owner(g0 + gq - 1, T, G)
                    parts = bl - bf + 1 if T >= G else gq
                    assert bf <= B <= bl
                    gathered[lp][ga:gb] += 1
                    if parts == 1:
                        commits[lp] += 1
                    else:
                        assert (lp + B) not in slots
                        slots.add(lp + B)
                        arrivals[lp] += 1
                        expected[lp] = parts
                        others = [bf + i if T >= G else owner(g0 + i, T, G) for i in range(parts)]
                        assert B in others and len(set(others)) == parts
            lp += 1
    for lp in range(nq):
        if groups[lp]:
            assert (gathered[lp] == 1).all(), (lp, gathered[lp])
            if expected[lp]:
                assert arrivals[lp] == expected[lp] and commits[lp] == 0     # the last arriver commits
            else:
                assert commits[lp] == 1
        else:
            assert commits[lp] == 0 and arrivals[lp] == 0
    if T:
        for u in np.random.default_rng(T).integers(0, T, 50):
            b = owner(int(u), T, G)
            assert b * T // G <= u < (b + 1) * T // G
    return T


@pytest.mark.parametrize("seed", range(12))
def test_partition_random_fronts(seed):
    rng = np.random.default_rng(seed)
    nq = int(rng.integers(1, 600))
    G = int(rng.choice([32, 80, 148, 296]))
    nf = int(rng.integers(1, 200_000))
    # members with a smaller row index than the pair's end: monotone in the pair index, like pre[2Q+2]
    cuts = np.sort(rng.integers(0, nf + 1, nq))
    kind = seed % 4
    if kind == 1:
        cuts[: nq // 2] = 0                                     # late front: low pairs have nothing to fetch
    elif kind == 2:
        cuts = np.minimum(cuts, 40)                              # tiny front: fewer cost units than CTAs
    elif kind == 3:
        cuts[:] = nf                                             # early front: every pair sees everything
    simulate([int(c) for c in cuts], G)


def test_partition_edge_cases():
    assert simulate([0, 0, 0], 296) == 0
    simulate([1], 32)
    simulate([33], 32)
    simulate([32 * 5000], 296)                                   # one huge pair shared by every CTA
    simulate([1] * 500, 296)                                     # only single-group pairs
    simulate([32 * WPG, 32 * WPG + 1] * 50, 148)                 # both sides of the warp-pair threshold
