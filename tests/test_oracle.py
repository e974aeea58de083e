"""Pins the CPU oracle (oracle/parsim_oracle.c): reference known answers, the reference's own
golden vectors, bit-for-bit phase agreement with the UNMODIFIED reference compiled from
/root/reference (when oracle/_ref was built), and the committed runs of the reference binary."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
VEC = json.load(open(os.path.join(HERE, "golden", "reference_vectors.json")))
RUNS = [json.loads(l) for l in open(os.path.join(HERE, "golden", "reference_runs.jsonl"))]


def test_rng_known_answers():
    k = VEC["kat"]
    assert ol.oracle_draws(1, 5).tolist() == k["draws_seed_1"]
    assert ol.oracle_draws(-10, 3).tolist() == k["draws_seed_m10"]
    x, y, vx, vy, m = ol.oracle_init(1, 2, 3, 10)
    assert [x[0], y[0], vx[0], vy[0], m[0]] == k["config1_p0_init"]


def test_config1_phase_known_answers():
    k = VEC["kat"]
    s = ol.oracle_sim(1, 2, 3, 10)
    s.phase(0)
    start, ids = s.lists()
    assert ids[start[0]:start[1]].tolist() == k["config1_cell0"]["ids"]
    mx, my, m = s.cells()
    assert (m[0], mx[0], my[0]) == (k["config1_cell0"]["m"], k["config1_cell0"]["mx"], k["config1_cell0"]["my"])
    assert ids[start[2]:start[3]].tolist() == [0, 1]
    assert [int(start[c + 1] - start[c]) for c in (3, 5, 8)] == [0, 0, 0]
    s.phase(1)
    p = s.particles()
    assert p["fx"][0] == -0.00049625823906854043 and p["fy"][0] == 0.00063400835856591516
    for ph in (2, 3, 4):
        s.phase(ph)
    p = s.particles()
    assert [p["x"][0], p["y"][0], p["vx"][0], p["vy"][0]] == k["config1_p0_after_1_step"]


@pytest.mark.parametrize("v", [v for v in VEC["run_tests"] + VEC["new_tests"] if v["fast"]],
                         ids=lambda v: " ".join(map(str, v["args"])))
def test_reference_golden_vectors(v):
    seed, side, nc, n, steps = v["args"]
    s = ol.oracle_sim(seed, side, nc, n)
    s.run(steps)
    p = s.particles()
    assert abs(p["x"][0] - v["x0"]) <= 1e-3 + 5e-4  # printed with 3 decimals
    assert abs(p["y"][0] - v["y0"]) <= 1e-3 + 5e-4
    assert s.collisions == v["collisions"]


@pytest.mark.parametrize("r", [r for r in RUNS if r["seconds"] < 3.0], ids=lambda r: " ".join(r["args"]))
def test_committed_reference_runs(r):
    seed, side, nc, n, steps = int(r["args"][0]), float(r["args"][1]), int(r["args"][2]), int(r["args"][3]), int(r["args"][4])
    s = ol.oracle_sim(seed, side, nc, n)
    s.run(steps)
    p = s.particles()
    assert f"{p['x'][0]:.3f}" == r["x0"] and f"{p['y'][0]:.3f}" == r["y0"]
    assert s.collisions == r["collisions"]


OMP_RUNS = [json.loads(l) for l in open(os.path.join(HERE, "golden", "omp_runs.jsonl"))]


@pytest.mark.parametrize("r", [r for r in OMP_RUNS if r["seconds"] < 3.0], ids=lambda r: " ".join(r["args"]))
def test_omp_semantics_match_the_one_thread_openmp_binary(r):
    """the oracle's `semantics = 1` (no strays, openMP/parsim-omp.cpp:248-278) is
    pinned by the unmodified OpenMP binary run with one thread (tests/golden/make_omp_goldens.py)."""
    seed, side, nc, n, steps = int(r["args"][0]), float(r["args"][1]), int(r["args"][2]), int(r["args"][3]), int(r["args"][4])
    s = ol.oracle_sim(seed, side, nc, n, semantics=1)
    s.run(steps)
    p = s.particles()
    assert f"{p['x'][0]:.3f}" == r["x0"] and f"{p['y'][0]:.3f}" == r["y0"]
    assert s.collisions == r["collisions"]


def test_omp_semantics_differ_from_serial_where_strays_exist():
    a, b = ol.oracle_sim(1, 10, 100, 100000), ol.oracle_sim(1, 10, 100, 100000, semantics=1)
    a.run(5); b.run(5)
    assert (a.collisions, b.collisions) == (5226, 5225)


PHASE_CASES = [(1, 2, 3, 10, 3), (-10, 3, 3, 100, 10), (1, 10, 100, 100000, 3), (5, 1, 20, 20000, 3),
               (5893, 0.05, 3, 10, 10), (-7, 10, 50, 20000, 4), (3, 1, 1, 50, 3), (3, 1, 2, 200, 3)]


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("case", PHASE_CASES, ids=lambda c: " ".join(map(str, c)))
def test_bitwise_vs_unmodified_reference(case):
    seed, side, nc, n, steps = case
    o, r = ol.oracle_sim(seed, side, nc, n), ol.ref_sim(seed, side, nc, n)
    for t in range(steps):
        for ph in range(5):
            o.phase(ph)
            r.phase(ph)
            po, pr = o.particles(), r.particles()
            for k in ("x", "y", "vx", "vy", "m", "fx", "fy", "alive"):
                assert np.array_equal(po[k], pr[k]), f"step {t} phase {ph} field {k}"
            if ph in (0, 3, 4):
                so, io = o.lists()
                sr, ir = r.lists()
                assert np.array_equal(so, sr) and np.array_equal(io, ir), f"step {t} phase {ph} lists"
            if ph == 0:
                for a, b in zip(o.cells(), r.cells()):
                    assert np.array_equal(a, b)
            assert o.collisions == r.collisions
