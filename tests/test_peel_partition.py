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


# ---- k_peel_single and k_small_sort_unique: the index arithmetic of the two single-CTA kernels -------------
def _cconst(name):
    m = re.search(rf"constexpr int {name} = (\d+)", SRC)
    assert m, name
    return int(m.group(1))


@pytest.mark.parametrize("nq", [1, 2, 3, 4])
@pytest.mark.parametrize("rows", [0, 1, 7, 8, 9, 127, 128, 1016, 1017, 2033, 4065, 8192])
def test_single_cta_peel_covers_every_list_position_once(nq, rows):
    """k_peel_single: warps of a pair = row shares x 8 column shares, lanes = 8 row slots x 4 pieces; a lane of
    row share rs and slot s takes list positions base + rs*8 + s + stride*u (u < 127 per pass).  Every position
    below `rows` must be taken by exactly one (row share, slot, pass, u) - for each of the 8 x 4 column pieces."""
    lane_rows = _cconst("PS_LANE_ROWS")
    W = 32 if nq <= 1 else 16 if nq == 2 else 8
    RS = W // 8
    stride = 8 * RS
    nf = max(rows, 1)
    passes = (nf + stride * lane_rows - 1) // (stride * lane_rows)
    seen = np.zeros(rows, int)
    for p in range(passes):
        base = p * stride * lane_rows
        if rows <= base:
            continue
        most = min(lane_rows, (rows - base + stride - 1) // stride)
        for rs in range(RS):
            for slot in range(8):
                first = base + rs * 8 + slot
                for u in range(most):
                    pos = first + u * stride
                    if pos < rows:
                        seen[pos] += 1
    assert (seen == 1).all()
    # words of the super-row: column share cs (8 words) x piece q (2 words) tile the 64 words
    words = sorted(cs * 8 + 2 * q + h for cs in range(8) for q in range(4) for h in range(2))
    assert words == list(range(64))


@pytest.mark.parametrize("n", [1, 2, 3, 100, 1025, 4096, 5000, 8192])
def test_single_cta_bitonic_network_sorts(n):
    """k_small_sort_unique: the compare-exchange schedule (q-th index with bit j clear, direction by bit k) and
    the sentinel rule (index >= n sorts after every row), replayed on random keys with duplicates."""
    rng = np.random.default_rng(n)
    keys = rng.integers(0, max(2, n // 3), n)                       # many equal keys: the index decides
    P = 1
    while P < n:
        P <<= 1
    idx = np.arange(P)

    def before(a, b):
        if a >= n or b >= n:
            return a < b
        if keys[a] != keys[b]:
            return keys[a] < keys[b]
        return a < b

    k = 2
    while k <= P:
        j = k >> 1
        while j > 0:
            q = np.arange(P >> 1)
            i = ((q & ~(j - 1)) << 1) | (q & (j - 1))
            assert ((i & j) == 0).all() and len(np.unique(i)) == len(i)
            a, b = idx[i].copy(), idx[i | j].copy()
            up = (i & k) == 0
            if n <= 1100:                                             # scalar compare for the small cases
                swap = np.array([before(int(bb), int(aa)) == bool(uu) for aa, bb, uu in zip(a, b, up)])
            else:
                ka = np.where(a < n, keys[np.minimum(a, n - 1)], np.iinfo(np.int64).max)
                kb = np.where(b < n, keys[np.minimum(b, n - 1)], np.iinfo(np.int64).max)
                b_before_a = (kb < ka) | ((kb == ka) & (b < a))
                swap = b_before_a == up
            idx[i] = np.where(swap, b, a)
            idx[i | j] = np.where(swap, a, b)
            j >>= 1
        k <<= 1
    want = np.concatenate([np.argsort(keys, kind="stable"), np.arange(n, P)])
    assert (idx == want).all()
