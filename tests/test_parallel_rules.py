"""The PARALLEL reformulations the CUDA kernels rely on, pinned on the CPU against the oracle
(and, where oracle/_ref is built, against the unmodified reference):

  * list update (k_classify): inside a maximal run of consecutive leavers of a cell's bin-time
    list, run positions 0, 2, 4, ... are moved and 1, 3, 5, ... stay as strays
    (serial/parsim.cpp:234-259, the erase-then-increment loop :251-253);
    the post-move list is (bin-time list minus the moved, order kept) ++ (arrivals ordered by
    source cell, then id);
  * collisions (k_collide_* + k_resolve): with first(j) = the first entry k of j's list that is
    close to j, a collision is counted for every unordered pair {j, k} with first(j) = k and
    first(k) = j, and everybody with any close partner dies -- equal to the greedy sweep over the
    unordered_set (serial/parsim.cpp:388-420);
  * the squared-distance threshold: d2 < thr2 with thr2 the smallest double whose sqrt is
    >= EPSILON decides exactly like sqrt(d2) < EPSILON (parsim_rank.cu collision_threshold2).

These are plain numpy / Python loops over small cases; no GPU, no product code."""
import math

import numpy as np
import pytest

import oracle_lib as ol

EPSILON = 0.005

# (seed, side, ncside, n_part, steps): strays and collisions both occur in the first three, the
# fourth is a negative seed (normal cloud: very full centre cells), the fifth has ncside <= 2
CASES = [(1, 10, 100, 100000, 4), (5, 1, 20, 20000, 4), (2, 4, 16, 30000, 3), (-10, 3, 3, 100, 10),
         (3, 1, 2, 3000, 3), (9, 2, 1, 500, 3)]


def _lists(sim):
    start, ids = sim.lists()
    return [ids[start[c]:start[c + 1]] for c in range(sim.ncells)]


def _predict_list_update(bin_lists, alive, newcell):
    """run-parity rule -> post-move lists"""
    ncells = len(bin_lists)
    keep = [[] for _ in range(ncells)]
    arrivals = [[] for _ in range(ncells)]
    strays = 0
    for c, lst in enumerate(bin_lists):
        prev_removed = False
        for i in lst:
            mover = bool(alive[i]) and newcell[i] != c
            removed = mover and not prev_removed
            if removed:
                arrivals[newcell[i]].append((c, int(i)))
            else:
                keep[c].append(int(i))
                strays += mover
            prev_removed = removed
    out = []
    for c in range(ncells):
        out.append(np.array(keep[c] + [i for _, i in sorted(arrivals[c])], dtype=np.int64))
    return out, strays


def _first_close(lst, x, y, alive):
    """first[j] = position in lst of the first entry close to entry j (-1: none)"""
    n = len(lst)
    first = np.full(n, -1, dtype=np.int64)
    if n < 2:
        return first
    px, py, al = x[lst], y[lst], alive[lst].astype(bool)
    dx = px[:, None] - px[None, :]
    dy = py[:, None] - py[None, :]
    close = np.sqrt(dx * dx + dy * dy) < EPSILON  # pow(d, 2) == d * d exactly
    close &= al[:, None] & al[None, :]
    np.fill_diagonal(close, False)
    has = close.any(axis=1)
    first[has] = close.argmax(axis=1)[has]
    return first


def _predict_collisions(lists, x, y, alive):
    count, dead = 0, []
    for lst in lists:
        first = _first_close(lst, x, y, alive)
        for j, k in enumerate(first):
            if k >= 0:
                dead.append(int(lst[j]))
                if k > j and first[k] == j:
                    count += 1
    return count, np.array(sorted(dead), dtype=np.int64)


@pytest.mark.parametrize("case", CASES, ids=lambda c: " ".join(map(str, c)))
def test_list_update_and_collision_rules_match_the_oracle(case):
    seed, side, nc, n, steps = case
    sim = ol.oracle_sim(seed, side, nc, n)
    total_strays = total_col = 0
    for _ in range(steps):
        sim.phase(0)
        bin_lists = _lists(sim)
        for c, lst in enumerate(bin_lists):  # bin-time lists are id-ordered (north_star (a))
            assert np.all(np.diff(lst) > 0), c
        sim.phase(1)
        sim.phase(2)
        p = sim.particles()
        predicted, strays = _predict_list_update(bin_lists, p["alive"], p["cell_index"])
        total_strays += strays
        sim.phase(3)
        got = _lists(sim)
        for c in range(sim.ncells):
            assert np.array_equal(predicted[c], got[c]), (c, predicted[c], got[c])
        before = sim.collisions
        count, dead = _predict_collisions(got, p["x"], p["y"], p["alive"])
        sim.phase(4)
        assert sim.collisions - before == count
        after = sim.particles()["alive"]
        assert np.array_equal(np.flatnonzero((p["alive"] == 1) & (after == 0)), dead)
        total_col += count
    if n >= 20000 and nc > 2:
        assert total_col > 0 and total_strays > 0, "the case no longer exercises the rules"


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("case", [(1, 10, 100, 50000, 3), (-10, 3, 3, 100, 10)], ids=lambda c: " ".join(map(str, c)))
def test_rules_match_the_unmodified_reference(case):
    seed, side, nc, n, steps = case
    sim = ol.ref_sim(seed, side, nc, n)
    for _ in range(steps):
        sim.phase(0)
        bin_lists = _lists(sim)
        sim.phase(1)
        sim.phase(2)
        p = sim.particles()
        predicted, _ = _predict_list_update(bin_lists, p["alive"], p["cell_index"])
        sim.phase(3)
        got = _lists(sim)
        assert all(np.array_equal(a, b) for a, b in zip(predicted, got))
        before = sim.collisions
        count, dead = _predict_collisions(got, p["x"], p["y"], p["alive"])
        sim.phase(4)
        assert sim.collisions - before == count
        after = sim.particles()["alive"]
        assert np.array_equal(np.flatnonzero((p["alive"] == 1) & (after == 0)), dead)


def test_squared_distance_threshold_decides_like_the_sqrt_form():
    thr2 = EPSILON * EPSILON
    while math.sqrt(thr2) >= EPSILON:
        thr2 = math.nextafter(thr2, 0.0)
    while math.sqrt(thr2) < EPSILON:
        thr2 = math.nextafter(thr2, math.inf)
    # the definition: smallest double whose square root reaches EPSILON
    assert math.sqrt(thr2) >= EPSILON > math.sqrt(math.nextafter(thr2, 0.0))
    # every double in a window of +-2000 ulp around it, and random ones further out
    v = math.nextafter(thr2, 0.0)
    lo = hi = thr2
    for _ in range(2000):
        lo, hi = math.nextafter(lo, 0.0), math.nextafter(hi, math.inf)
        for d2 in (lo, hi):
            assert (math.sqrt(d2) < EPSILON) == (d2 < thr2)
    assert (math.sqrt(v) < EPSILON) and not (math.sqrt(thr2) < EPSILON)
    rng = np.random.default_rng(0)
    d2 = thr2 * (1.0 + (rng.random(200000) - 0.5) * 1e-9)
    assert np.array_equal(np.sqrt(d2) < EPSILON, d2 < thr2)
