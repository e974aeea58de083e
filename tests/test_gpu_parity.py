"""Parity of the CUDA path (through the C ABI) against the oracle, stage by stage and end to end.
Bar: collision counts, cell lists, alive flags and cell indices exact; positions/velocities
within 1e-9 relative to the box side per stage (the force sum is reordered and uses rsqrt, so
it differs from the reference at the ulp level -- SURVEY 0 fact 5 -- while the integration,
wrap and binning arithmetic is bit-faithful), and the reference's own 1e-3 on particle 0."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from conftest import ROOT
from particlesimulation_b200 import ParticleSimulation, init_particles

pytestmark = pytest.mark.gpu

VEC = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")))
RUNS = [json.loads(l) for l in open(os.path.join(ROOT, "tests", "golden", "reference_runs.jsonl"))]


def oracle_lists_alive(o):
    """per cell: list of alive ids in list order"""
    start, ids = o.lists()
    alive = o.particles()["alive"]
    return [[int(i) for i in ids[start[c]:start[c + 1]] if alive[i]] for c in range(o.ncells)]


def gpu_lists(d, ncells):
    """per cell: alive ids ordered by the list key the collision kernel uses"""
    out = [[] for _ in range(ncells)]
    idx = np.nonzero(d["alive"])[0]
    order = idx[np.lexsort((d["list_rank"][idx], d["list_cell"][idx]))]
    for i in order:
        out[d["list_cell"][i]].append(int(i))
    return out


STAGED = [(1, 2, 3, 10, 3), (1, 1, 5, 100, 2), (-10, 3, 3, 100, 10), (5893, 0.05, 3, 10, 10),
          (8555, 0.05, 3, 10, 10), (1, 10, 100, 100000, 4), (5, 1, 20, 20000, 4), (-7, 10, 50, 20000, 5),
          (3, 1, 1, 50, 3), (3, 1, 2, 200, 3), (9, 0.3, 4, 3000, 6)]


@pytest.mark.parametrize("case", STAGED, ids=lambda c: " ".join(map(str, c)))
def test_stage_by_stage_against_oracle(case):
    seed, side, nc, n, steps = case
    o = ol.oracle_sim(seed, side, nc, n)
    g = ParticleSimulation(seed, side, nc, n)
    tol = 1e-9 * side
    for t in range(steps):
        tag = f"step {t}"
        # stage 0 <-> updateCOM
        o.phase(0)
        g.debug_stage(0)
        d = g.dump()
        po = o.particles()
        assert np.array_equal(d["alive"], po["alive"]), tag
        al = po["alive"] == 1
        assert np.array_equal(d["cell"][al], po["cell_index"][al]), tag
        mx, my, m = g.dump_cells()
        omx, omy, om = o.cells()
        assert np.allclose(m, om, rtol=1e-12, atol=0), tag
        nz = om > 0
        assert np.allclose(mx[nz], omx[nz], rtol=0, atol=tol) and np.allclose(my[nz], omy[nz], rtol=0, atol=tol), tag
        # stage 1 <-> updateForces + applyForce
        o.phase(1)
        o.phase(2)
        g.debug_stage(1)
        d = g.dump()
        po = o.particles()
        for k in ("x", "y", "vx", "vy"):
            assert np.allclose(d[k], po[k], rtol=0, atol=tol), f"{tag} {k} max {np.abs(d[k]-po[k]).max()}"
        assert np.array_equal(d["cell"][al], po["cell_index"][al]), tag
        # stage 2 <-> updateCellParticles
        o.phase(3)
        g.debug_stage(2)
        d = g.dump()
        assert gpu_lists(d, o.ncells) == oracle_lists_alive(o), tag
        # stage 3 <-> checkCollisions
        o.phase(4)
        g.debug_stage(3)
        d = g.dump()
        po = o.particles()
        assert g.result()[2] == o.collisions, tag
        assert np.array_equal(d["alive"], po["alive"]), tag
        assert np.array_equal(d["m"] == 0, po["m"] == 0), tag
        for k in ("x", "y", "vx", "vy"):
            assert np.allclose(d[k], po[k], rtol=0, atol=tol), f"{tag} {k}"
    g.close()


@pytest.mark.parametrize("v", [v for v in VEC["run_tests"] + VEC["new_tests"] if v["fast"]],
                         ids=lambda v: " ".join(map(str, v["args"])))
def test_reference_golden_vectors(v):
    seed, side, nc, n, steps = v["args"]
    g = ParticleSimulation(seed, side, nc, n)
    g.simulate(steps)
    x0, y0, col = g.result()
    assert abs(x0 - v["x0"]) <= 1e-3 + 5e-4 and abs(y0 - v["y0"]) <= 1e-3 + 5e-4
    assert col == v["collisions"]
    g.close()


@pytest.mark.slow
@pytest.mark.parametrize("v", [v for v in VEC["run_tests"] + VEC["new_tests"] if not v["fast"]],
                         ids=lambda v: " ".join(map(str, v["args"])))
def test_reference_golden_vectors_long(v):
    test_reference_golden_vectors(v)


@pytest.mark.parametrize("r", RUNS, ids=lambda r: " ".join(r["args"]))
def test_committed_reference_runs(r):
    """Dense-regime pins produced by the unmodified serial binary (tests/golden/make_goldens.py)."""
    seed, side, nc, n, steps = int(r["args"][0]), float(r["args"][1]), int(r["args"][2]), int(r["args"][3]), int(r["args"][4])
    g = ParticleSimulation(seed, side, nc, n)
    g.simulate(steps)
    x0, y0, col = g.result()
    assert col == r["collisions"]
    assert abs(x0 - float(r["x0"])) <= 1e-3 and abs(y0 - float(r["y0"])) <= 1e-3
    g.close()


OMP_RUNS = [json.loads(l) for l in open(os.path.join(ROOT, "tests", "golden", "omp_runs.jsonl"))]


@pytest.mark.parametrize("r", OMP_RUNS, ids=lambda r: "omp " + " ".join(r["args"]))
def test_omp_semantics_match_the_one_thread_openmp_binary(r):
    """Option semantics = 1: every leaver is moved, no strays -- the OpenMP
    variant's list update; counts pinned by parsim-omp with one thread."""
    seed, side, nc, n, steps = int(r["args"][0]), float(r["args"][1]), int(r["args"][2]), int(r["args"][3]), int(r["args"][4])
    g = ParticleSimulation(seed, side, nc, n, options={"semantics": 1})
    g.simulate(steps)
    x0, y0, col = g.result()
    assert col == r["collisions"]
    assert abs(x0 - float(r["x0"])) <= 1e-3 and abs(y0 - float(r["y0"])) <= 1e-3
    assert g.stats()["strays"] == 0
    g.close()


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 4), (5, 1, 20, 20000, 4), (-10, 3, 3, 100, 10)],
                         ids=lambda c: "omp " + " ".join(map(str, c)))
def test_omp_semantics_lists_match_oracle(case):
    seed, side, nc, n, steps = case
    o = ol.oracle_sim(seed, side, nc, n, semantics=1)
    g = ParticleSimulation(seed, side, nc, n, options={"semantics": 1})
    for t in range(steps):
        for ph in range(4):
            o.phase(ph)
        for st in range(3):
            g.debug_stage(st)
        assert gpu_lists(g.dump(), o.ncells) == oracle_lists_alive(o), f"step {t}"
        o.phase(4)
        g.debug_stage(3)
        assert g.result()[2] == o.collisions, f"step {t}"
        assert np.array_equal(g.dump()["alive"], o.particles()["alive"]), f"step {t}"
    g.close()


def test_cli_semantics_switch():
    cli = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    args = "1 10 100 100000 5".split()
    p = subprocess.run([cli] + args, capture_output=True, text=True, env=dict(os.environ, PARSIM_SEMANTICS="omp"))
    assert p.returncode == 0 and p.stdout == "7.872 0.263\n5225\n", p.stdout + p.stderr
    p = subprocess.run([cli] + args, capture_output=True, text=True, env=dict(os.environ, PARSIM_SEMANTICS="mpi"))
    assert p.returncode == 1 and p.stderr.startswith("Error: ")


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 6), (5, 1, 20, 20000, 6), (-7, 10, 50, 20000, 6), (5, 1, 4, 60000, 5),
                                  (9, 0.3, 4, 3000, 6)], ids=lambda c: " ".join(map(str, c)))
def test_block_per_cell_merge_matches_lane_groups(case, monkeypatch):
    """k_move_big (one block per huge cell: sorted arrivals, binary search, histogram) must build the
    same (cell, id) order as the lane-group merge; forced on ordinary cells by a tiny threshold."""
    seed, side, nc, n, steps = case
    ref = ParticleSimulation(seed, side, nc, n)
    monkeypatch.setenv("PARSIM_MOVE_BIG_MIN", "6")
    big = ParticleSimulation(seed, side, nc, n)
    monkeypatch.delenv("PARSIM_MOVE_BIG_MIN")
    for t in range(steps):
        ref.simulate(1)
        big.simulate(1)
        assert ref.result() == big.result(), f"step {t}"
        a, b = ref.dump(), big.dump()
        for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
            assert np.array_equal(a[k], b[k]), (t, k)
        ia, ib = ref.download(), big.download()  # device order = (cell, id) order
        for u, v in zip(ia, ib):
            assert np.array_equal(u, v), f"layout differs at step {t}"
    ref.close(); big.close()


def test_run_is_deterministic_and_matches_staged():
    args = (1, 10, 100, 200000)
    a = ParticleSimulation(*args); a.simulate(5)
    b = ParticleSimulation(*args); b.simulate(2); b.simulate(3)
    c = ParticleSimulation(*args)
    for _ in range(5):
        for s in range(4):
            c.debug_stage(s)
    da, db, dc = a.dump(), b.dump(), c.dump()
    for k in ("x", "y", "vx", "vy", "m", "alive"):
        assert np.array_equal(da[k], db[k]) and np.array_equal(da[k], dc[k]), k
    assert a.result() == b.result() == c.result()


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 5), (5, 1, 20, 20000, 4), (5, 1, 20, 200000, 3),
                                  (5, 1, 20, 3000000, 2), (-7, 10, 50, 200000, 6), (9, 0.3, 4, 3000, 6),
                                  (1, 5000, 20, 300000, 3), (3, 100, 1000, 800000, 5)],
                         ids=lambda c: " ".join(map(str, c)))
def test_collision_search_paths_agree(case, monkeypatch):
    """The sub-grid collision search (warp per cell, block per big cell, all-pairs fallback for
    lists beyond the staging area) must give exactly the result of the per-particle all-pairs sweep."""
    seed, side, nc, n, steps = case
    res = {}
    for mode in ("pairs", "grid"):
        monkeypatch.setenv("PARSIM_COLLIDE", mode)
        g = ParticleSimulation(seed, side, nc, n)
        g.simulate(steps)
        res[mode] = (g.result(), g.dump())
        g.close()
    assert res["pairs"][0] == res["grid"][0]
    for k in ("x", "y", "vx", "vy", "m", "alive"):
        assert np.array_equal(res["pairs"][1][k], res["grid"][1][k]), k


FORCE_VARIANTS = [0, 1, 2, 3, 4, 5, 6, 10, 11]  # = what rank_create accepts (csrc/parsim_rank.cu)


def test_unknown_force_variant_is_rejected(monkeypatch):
    from particlesimulation_b200 import ParsimError
    for v in ("7", "9", "12", "-1"):
        monkeypatch.setenv("PARSIM_FORCE_VARIANT", v)
        with pytest.raises(ParsimError):
            ParticleSimulation(1, 2, 3, 10)


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 4), (3, 100, 1000, 800000, 3), (5, 1, 20, 200000, 2),
                                  (-7, 10, 50, 200000, 4), (9, 0.3, 4, 3000, 4), (2, 7, 3, 2, 2)],
                         ids=lambda c: " ".join(map(str, c)))
def test_force_kernel_variants_agree(case, monkeypatch):
    """Every instantiation of the force + integration kernel that PARSIM_FORCE_VARIANT accepts --
    block per tile with or without the staged neighbourhood, deep unroll, 64 / 256-thread blocks --
    sums in the same order, so the runs must agree bit for bit with
    the default (which picks its instantiation from the occupancy)."""
    seed, side, nc, n, steps = case
    res = {}
    variants = FORCE_VARIANTS
    for variant in variants:
        monkeypatch.setenv("PARSIM_FORCE_VARIANT", str(variant))
        g = ParticleSimulation(seed, side, nc, n)
        g.simulate(steps)
        res[variant] = (g.result(), g.dump())
        g.close()
    for variant in variants[1:]:
        assert res[variant][0] == res[0][0], variant
        for k in ("x", "y", "vx", "vy", "m", "alive"):
            assert np.array_equal(res[variant][1][k], res[0][1][k]), (variant, k)


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 6), (5, 1, 20, 200000, 4), (-7, 10, 50, 20000, 6), (3, 1, 2, 200, 5),
                                  (1, 2, 3, 10, 4)], ids=lambda c: " ".join(map(str, c)))
def test_deferred_steps_match_synchronised_steps(case, monkeypatch):
    """parsim_run on a single strip enqueues its steps without a host round trip (the kernels read
    the particle count from device memory, grids are sized for the host's older count, the error
    flag is read at the end); PARSIM_SYNC_STEPS=1 is the step-by-step path that multi-strip runs
    and instrumented runs use.  Same bits, also when the run is cut into several parsim_run calls."""
    seed, side, nc, n, steps = case
    res = {}
    # step by step / deferred, launch by launch / deferred, replayed as CUDA graphs (what problems of
    # this size get by default) -- the latter twice over, so that the second run replays cached graphs
    for mode, env in (("sync", {"PARSIM_SYNC_STEPS": "1"}), ("eager", {"PARSIM_SYNC_STEPS": "0", "PARSIM_GRAPHS": "0"}),
                      ("graphs", {"PARSIM_SYNC_STEPS": "0", "PARSIM_GRAPHS": "1"})):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = ParticleSimulation(seed, side, nc, n, options={"keep_initial": 1})
        for rep in range(2 if mode == "graphs" else 1):
            if rep:
                g.reset()
            g.simulate(steps)
            g.simulate(1)
            g.simulate(2)
        res[mode] = (g.result(), g.dump())
        g.close()
    for mode in ("eager", "graphs"):
        assert res[mode][0] == res["sync"][0], mode
        for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
            assert np.array_equal(res[mode][1][k], res["sync"][1][k]), (mode, k)


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 6), (5, 1, 20, 20000, 5), (7, 5, 200, 40000, 5), (-7, 10, 50, 20000, 6),
                                  (9, 0.3, 4, 3000, 6), (3, 1, 2, 200, 5), (2, 4, 16, 4000, 5)], ids=lambda c: " ".join(map(str, c)))
def test_centres_of_mass_left_by_the_merge_match_k_com(case, monkeypatch):
    """The end-of-step merge leaves the NEXT step's centres of mass (k_move<.., COM>, summed in
    k_com's order from the list it has just written), and stage 0 then skips k_com.
    PARSIM_FUSE_COM=0 always runs k_com: same bits in every cell at every step, same run, same
    pair statistics -- stage by stage, through parsim_run (graphs), and with instrumentation on."""
    seed, side, nc, n, steps = case
    res = {}
    for fuse in ("0", "1"):
        monkeypatch.setenv("PARSIM_FUSE_COM", fuse)
        g = ParticleSimulation(seed, side, nc, n, options={"keep_initial": 1})
        cells = []
        for t in range(steps):
            g.debug_stage(0)
            cells.append(g.dump_cells())
            for st in (1, 2, 3):
                g.debug_stage(st)
        staged = (g.result(), g.dump())
        g.reset()
        g.simulate(steps)
        g.simulate(2)
        run = (g.result(), g.dump())
        g.reset()
        g.set_option("stats", 1)
        g.simulate(steps)
        st = g.stats()
        res[fuse] = (cells, staged, run, g.result(), st["force_pairs"], st["kernels"]["com"]["launches"])
        g.close()
    a, b = res["0"], res["1"]
    for t in range(steps):
        for q in range(3):
            assert np.array_equal(a[0][t][q], b[0][t][q]), (t, q)
    for which in (1, 2):
        assert a[which][0] == b[which][0], which
        for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
            assert np.array_equal(a[which][1][k], b[which][1][k]), (which, k)
    assert a[3] == b[3] and a[4] == b[4]
    assert a[5] == steps  # one k_com per step without the fusion


@pytest.mark.parametrize("case", [(1, 10, 100, 100000, 4, 20000), (5, 1, 20, 200000, 3, 30000), (-7, 10, 50, 20000, 5, 3000),
                                  (3, 1, 2, 200, 4, 64), (1, 2, 3, 10, 3, 64), (7, 30, 300, 2000000, 2, 1 << 19)],
                         ids=lambda c: " ".join(map(str, c)))
def test_upload_pieces_match_full_sort(case, monkeypatch):
    """Host-buffer uploads are sorted piece by piece while the next piece is still on its way over
    PCIe, and the first timestep only merges the pieces (csrc/parsim_step.inl sort_piece /
    merge_pieces).  The layout must be the one the full step-0 sort of the records gives
    (PARSIM_PIECE_SORT=0, and what parsim_reset and the device-side init use): same cell lists,
    same bits after the run -- with several pieces, with one piece, and after a reset."""
    seed, side, nc, n, steps, piece = case
    monkeypatch.setenv("PARSIM_STAGE_PIECE", str(piece))
    res = {}
    # "pieces": the default -- with more than one piece the velocity arrays travel LAST and the merge,
    # the centres of mass and the force sums of the first timestep run under their copy
    # (rank_upload_chunk / upload_prestep); "early": velocities with every piece (PARSIM_VEL_LATE=0)
    for mode in ("pieces", "early", "full"):
        monkeypatch.setenv("PARSIM_PIECE_SORT", "0" if mode == "full" else "1")
        monkeypatch.setenv("PARSIM_VEL_LATE", "0" if mode == "early" else "1")
        g = ParticleSimulation(seed, side, nc, n, options={"keep_initial": 1})
        g.debug_stage(0)
        first = g.dump()
        for st in (1, 2, 3):
            g.debug_stage(st)
        g.simulate(steps)
        res[mode] = (first, g.result(), g.dump())
        if mode == "pieces":  # the reset path restarts from the records, not from the pieces
            g.reset()
            g.simulate(steps + 1)
            res["reset"] = (first, g.result(), g.dump())
        g.close()
        if mode != "full":  # straight from the upload into parsim_run (graph-replayed steps on small problems), twice
            g = ParticleSimulation(seed, side, nc, n)
            # what was uploaded comes back (download and dump right after the upload: velocities included)
            arrs0 = init_particles(seed, side, nc, n)
            got = g.download()
            order = np.argsort(got[0], kind="stable")
            assert np.array_equal(got[0][order], np.arange(n, dtype=np.uint32)), mode
            for a0, a1 in zip(arrs0, got[1:]):
                assert np.array_equal(np.asarray(a0), a1[order]), mode
            d0 = g.dump()
            for k, a0 in zip(("x", "y", "vx", "vy", "m"), arrs0):
                assert np.array_equal(d0[k], np.asarray(a0)), (mode, k)
            g.simulate(steps + 1)
            res[mode + " run"] = (first, g.result(), g.dump())
            g.upload(*init_particles(seed, side, nc, n))
            g.simulate(1)
            g.simulate(steps)
            res[mode + " rerun"] = (first, g.result(), g.dump())
            g.close()
    for mode in ("pieces", "early", "reset", "pieces run", "pieces rerun", "early run", "early rerun"):
        assert res[mode][1] == res["full"][1], mode
        for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
            assert np.array_equal(res[mode][2][k], res["full"][2][k]), (mode, k)
    for mode in ("pieces", "early"):
        for k in ("x", "y", "vx", "vy", "m", "cell", "list_rank"):
            assert np.array_equal(res[mode][0][k], res["full"][0][k]), (mode, k)


def test_whole_strip_upload_on_a_single_rank_context():
    """parsim_upload_strip (the one-call, collective form of the explicit-id upload) on a context with
    one rank is the plain begin / chunk_ids / end sequence; the multi-rank form runs under torchrun in
    tests/test_gpu_multi.py::test_bench_under_torchrun (bench.py's end-to-end leg uses it and checks the
    result against the device-resident run)."""
    seed, side, nc, n, steps = 4, 3.0, 30, 50000, 4
    arrs = [np.ascontiguousarray(a) for a in init_particles(seed, side, nc, n)]
    a = ParticleSimulation(seed, side, nc, n, arrays=arrs)
    a.simulate(steps)
    b = ParticleSimulation(seed, side, nc, n, upload=False)
    ids = np.arange(n, dtype=np.uint32)
    b.upload_whole_strip_ptrs(n, ids.ctypes.data, [x.ctypes.data for x in arrs])
    b.simulate(steps)
    assert a.result() == b.result()
    da, db = a.dump(), b.dump()
    for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
        assert np.array_equal(da[k], db[k]), k
    a.close()
    b.close()


def test_close_encounters_throw_particles_far():
    """Two particles a few 1e-6 apart fly tens of cells in one step; the result must still be the
    reference's (alive flags and collision count exact, positions to 1e-9 of the box side)."""
    side, nc, arrs = ol.far_jump_case()
    n = len(arrs[0])
    o = ol.oracle_sim_from(side, nc, *arrs)
    g = ParticleSimulation(0, side, nc, n, arrays=arrs)
    x_before = arrs[0].copy()
    for t in range(4):
        o.step()
        g.simulate(1)
        d, p = g.dump(), o.particles()
        assert g.result()[2] == o.collisions, t
        assert np.array_equal(d["alive"], p["alive"]), t
        for k in ("x", "y", "vx", "vy"):
            assert np.allclose(d[k], p[k], rtol=1e-12, atol=1e-9 * side), (t, k, np.abs(d[k] - p[k]).max())
        if t == 0:  # the case does what it is meant to: long jumps happened
            jump = np.abs(d["x"] - x_before)
            jump = np.minimum(jump, side - jump)
            assert (jump > 3 * side / nc).sum() >= 20
    g.close()


def test_upload_arrays_path_and_errors():
    arrs = init_particles(1, 2, 3, 10)
    g = ParticleSimulation(0, 2, 3, 10, arrays=arrs)
    g.simulate(1)
    x0, y0, col = g.result()
    assert [x0, y0] == pytest.approx(VEC["kat"]["config1_p0_after_1_step"][:2], abs=1e-12) and col == 0
    with pytest.raises(Exception):
        g.debug_stage(2)  # out of order
    g.close()


@pytest.mark.parametrize("host_init", [False, True], ids=["device init (default)", "PARSIM_DEVICE_INIT=0"])
def test_cli_matches_reference_output(host_init):
    """Non-negative seeds are generated on the device by default; PARSIM_DEVICE_INIT=0 streams the
    host generator through parsim_upload_chunk.  Same output either way (the reference's)."""
    cli = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    env = dict(os.environ)
    if host_init:
        env["PARSIM_DEVICE_INIT"] = "0"
    else:
        env.pop("PARSIM_DEVICE_INIT", None)
    for args, exp in [("1 2 3 10 1", "1.570 0.056\n0\n"), ("1 10 100 100000 5", "7.872 0.263\n5226\n")]:
        p = subprocess.run([cli] + args.split(), capture_output=True, text=True, env=env)
        assert p.returncode == 0, p.stderr
        assert p.stdout == exp
        assert p.stderr.strip().endswith("s")


@pytest.mark.parametrize("args", ["1 2 3 10 1", "1 10 100 1000000 20", "-10 3 3 100 4", "5 1 20 20000 5"])
def test_patched_reference_matches_reference(args):
    """INTEGRATION.md section 2, compiled: the reference's own serial program with simulate()
    replaced by the C-ABI calls (oracle/Makefile -> oracle/_ref/parsim-patched; constructor,
    init_particles, print_result and main are the reference's) prints what the unmodified
    reference prints.  BASELINE configs[0] and configs[1] and two of the reference's vectors."""
    patched = os.path.join(ROOT, "oracle", "_ref", "parsim-patched")
    ref = os.path.join(ROOT, "oracle", "_ref", "parsim")
    if not (os.path.exists(patched) and os.path.exists(ref)):
        pytest.skip("oracle/_ref not built (reference sources absent at build time)")
    a = subprocess.run([patched] + args.split(), capture_output=True, text=True, timeout=300)
    b = subprocess.run([ref] + args.split(), capture_output=True, text=True, timeout=300)
    assert a.returncode == 0, a.stderr
    assert b.returncode == 0, b.stderr
    assert a.stdout == b.stdout
    assert a.stderr.strip().endswith("s")


@pytest.mark.parametrize("args", ["1 2 3 10 3", "5 1 20 2000 3", "-10 3 3 100 4"])
def test_cli_trace_matches_oracle(args, tmp_path):
    """PARSIM_TRACE writes the dump the reference keeps commented out in simulate()
    (serial/parsim.cpp:424-431,:442-446); the same text generated from the oracle's state must
    match line for line (3 decimals; dead particles print mass 0 and their frozen state)."""
    seed, side, nc, n, steps = args.split()
    seed, side, nc, n, steps = int(seed), float(side), int(nc), int(n), int(steps)
    cli = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    path = tmp_path / "trace.txt"
    plain = subprocess.run([cli] + args.split(), capture_output=True, text=True)
    p = subprocess.run([cli] + args.split(), capture_output=True, text=True,
                       env=dict(os.environ, PARSIM_TRACE=str(path)))
    assert p.returncode == 0, p.stderr
    assert p.stdout == plain.stdout  # the trace does not disturb the two result lines
    o = ol.oracle_sim(seed, side, nc, n)

    def particles():
        st = o.particles()
        return ["Particle %d: mass=%.3f x=%.3f y=%.3f vx=%.3f vy=%.3f" % (j, st["m"][j] if st["alive"][j] else 0.0,
                st["x"][j], st["y"][j], st["vx"][j], st["vy"][j]) for j in range(n)]

    want = particles() + ["Cell %d x: 0.000 y: 0.000 m: 0.000" % j for j in range(nc * nc)]
    for t in range(steps):
        o.step()
        want += ["t=%d" % t] + particles()
    o.close()
    got = path.read_text().splitlines()
    assert len(got) == len(want)
    bad = [(a, b) for a, b in zip(got, want) if a != b]
    assert not bad, bad[:3]


@pytest.mark.parametrize("args", [(1, 2, 3, 10), (7, 100, 1000, 300000), (0, 5, 7, 1000), (123456, 10, 100, 1 << 20)])
def test_device_init_is_bit_identical_to_host_init(args):
    """parsim_init_device (xorshift32 jump-ahead) == init_particles (serial/parsim.cpp:220-232)."""
    g = ParticleSimulation(*args, device_init=True)
    ids, x, y, vx, vy, m = g.download()
    hx, hy, hvx, hvy, hm = init_particles(*args)
    order = np.argsort(ids)
    assert np.array_equal(ids[order], np.arange(args[3], dtype=np.uint32))
    for dev, host in ((x, hx), (y, hy), (vx, hvx), (vy, hvy), (m, hm)):
        assert np.array_equal(dev[order], host)
    g.close()


def test_device_init_rejects_negative_seed():
    with pytest.raises(Exception):
        ParticleSimulation(-3, 2, 3, 10, device_init=True)


def test_reset_and_id_upload_reproduce_the_run():
    args = (1, 10, 100, 100000)
    a = ParticleSimulation(*args, options={"keep_initial": 1})
    a.simulate(5)
    first = a.result()
    a.reset()
    a.simulate(5)
    assert a.result() == first and first[2] == 5226
    # same particles handed over in a shuffled order with explicit ids
    x, y, vx, vy, m = init_particles(*args)
    perm = np.random.default_rng(0).permutation(args[3])
    arrs = [np.ascontiguousarray(v[perm]) for v in (x, y, vx, vy, m)]
    ids = np.ascontiguousarray(perm.astype(np.uint32))
    b = ParticleSimulation(*args, upload=False)
    b.upload_strip_ptrs(args[3], ids.ctypes.data, [v.ctypes.data for v in arrs])
    b.simulate(5)
    assert b.result() == first
    a.close(); b.close()


def test_upload_rejects_out_of_range_ids():
    n = 1000
    x, y, vx, vy, m = init_particles(1, 10, 100, n)
    ids = np.arange(n, dtype=np.uint32)
    ids[5] = 5 * n
    g = ParticleSimulation(1, 10, 100, n, upload=False)
    with pytest.raises(Exception):
        g.upload_strip_ptrs(n, ids.ctypes.data, [v.ctypes.data for v in (x, y, vx, vy, m)])
        g.simulate(1)
    g.close()


@pytest.mark.parametrize("args", [(1, 1, 1, 1, 3), (2, 5, 1, 2, 4), (3, 1, 2, 1, 2), (4, 7, 3, 2, 0), (5, 0.02, 2, 40, 3),
                                  (6, 1, 30, 3, 5), (-2, 3, 5, 1, 2)], ids=lambda a: " ".join(map(str, a)))
def test_degenerate_sizes_match_oracle(args):
    """One particle, one cell, more cells than particles, zero steps, a box a few EPSILON wide
    (everything collides): same answers as the oracle, bit-exact counts and alive flags."""
    seed, side, nc, n, steps = args
    g = ParticleSimulation(seed, side, nc, n)
    o = ol.oracle_sim(seed, side, nc, n)
    g.simulate(steps)
    o.run(steps)
    d, p = g.dump(), o.particles()
    assert g.result()[2] == o.collisions
    assert np.array_equal(d["alive"], p["alive"])
    for k in ("x", "y", "vx", "vy"):
        assert np.allclose(d[k], p[k], rtol=0, atol=1e-9 * side), k
    x0, y0, _ = g.result()
    assert abs(x0 - p["x"][0]) <= 1e-9 * side and abs(y0 - p["y"][0]) <= 1e-9 * side
    g.close(); o.close()


def test_full_size_properties():
    """Size-independent properties at BASELINE config 3's full size (1e8 particles), where the
    oracle cannot follow: the (cell, id) layout invariant after the run, alive + dead = n, every
    collision kills at least two, a second run from the resident state repeats the first bit for
    bit, and the count equals the serial reference's committed golden."""
    n, steps = 100_000_000, 10
    g = ParticleSimulation(7, 100.0, 1000, n, device_init=True, options={"keep_initial": 1, "stats": 1})
    g.simulate(steps)
    x0, y0, col = g.result()
    golden = [r for r in RUNS if r["args"] == ["7", "100", "1000", "100000000", "10"]][0]
    assert col == golden["collisions"] and f"{x0:.3f}" == golden["x0"] and f"{y0:.3f}" == golden["y0"]
    ids, x, y, vx, vy, m = g.download()  # device order
    deaths = g.stats()["deaths"]
    assert len(ids) + deaths == n and deaths >= 2 * col
    cw = 100.0 / 1000
    cell = np.clip((y / cw).astype(np.int64), 0, 999) * 1000 + np.clip((x / cw).astype(np.int64), 0, 999)
    key = cell * (1 << 32) + ids.astype(np.int64)
    assert np.all(np.diff(key) > 0)  # sorted by (cell, id), no duplicates
    assert np.all(m > 0) and np.all((x >= 0) & (x < 100.0) & (y >= 0) & (y < 100.0))
    del x, y, vx, vy, m, cell, key
    g.reset()
    g.simulate(steps)
    assert g.result() == (x0, y0, col)
    ids2 = g.download()[0]
    assert np.array_equal(ids, ids2)
    g.close()


def test_stats_counters():
    g = ParticleSimulation(1, 10, 100, 100000, options={"stats": 1})
    g.simulate(3)
    s = g.stats()
    assert s["steps"] == 3 and s["particle_slots"] == 300000
    assert s["kernels"]["force_integrate"]["launches"] == 3 and s["kernels"]["force_integrate"]["ms"] > 0
    assert s["movers"] > 0 and s["deaths"] > 0 and s["force_pairs"] > 0 and s["collision_pairs"] > 0
    assert g.fp64_peak_tflops() > 5
    g.close()
