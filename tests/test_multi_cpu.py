"""Host side of the N>1 path on CPU: strip arithmetic (against the C ABI and the oracle's
binning), and a world_size-2 `gloo` run of the launcher plumbing that bench.py uses under
torch.distributed.run (rendezvous on 127.0.0.1, id broadcast, strip selection, max-over-ranks
timing, the reference arm's rank gating).  No CUDA anywhere in this file."""
import ctypes as C
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from particlesimulation_b200 import _lib, strips
from particlesimulation_b200.simulation import init_particles

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(nproc, script_args, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port())] + script_args
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("nc,nranks", [(3, 1), (3, 2), (3, 3), (10, 4), (1000, 8), (3536, 8), (7, 3), (10000, 8)])
def test_strip_rows_partition_the_grid(nc, nranks):
    lib = _lib.load()
    nxt = 0
    for r in range(nranks):
        r0, rows = strips.strip_rows(nc, nranks, r)
        a, b = C.c_long(), C.c_long()
        assert lib.parsim_strip_rows(nc, nranks, r, C.byref(a), C.byref(b)) == 0
        assert (a.value, b.value) == (r0, rows)  # the library's arithmetic == the launcher's
        assert r0 == nxt and rows >= 1
        nxt = r0 + rows
    assert nxt == nc
    owner = strips.strip_of_row(np.arange(nc), nc, nranks)
    for r in range(nranks):
        r0, rows = strips.strip_rows(nc, nranks, r)
        assert np.all(owner[r0:r0 + rows] == r)
    sizes = np.bincount(owner, minlength=nranks)
    assert sizes.max() - sizes.min() <= 1  # equal-row strips


def test_strip_rows_rejects_bad_arguments():
    lib = _lib.load()
    a, b = C.c_long(), C.c_long()
    assert lib.parsim_strip_rows(4, 8, 0, C.byref(a), C.byref(b)) != 0  # more strips than rows
    assert lib.parsim_strip_rows(4, 2, 2, C.byref(a), C.byref(b)) != 0
    with pytest.raises(ValueError):
        strips.strip_rows(4, 8, 0)
    with pytest.raises(ValueError):
        strips.strip_rows(4, 2, -1)


@pytest.mark.parametrize("nc,nranks", [(10, 3), (100, 8), (7, 7), (50, 2), (1000, 8), (5, 4), (3536, 8)])
def test_balanced_rows_c_and_python_agree(nc, nranks):
    """Equal-particle strips: the C ABI's boundaries == the launcher's, they tile the
    grid with >= 1 row each, and each boundary sits where the prefix sum crosses k/R of the total."""
    lib = _lib.load()
    rng = np.random.default_rng(nc * 31 + nranks)
    x = np.arange(nc)
    shapes = [rng.integers(0, 100, nc), np.zeros(nc), rng.integers(0, 3, nc) * rng.integers(0, 1000, nc),
              np.round(1e4 * np.exp(-0.5 * ((x - nc / 2) / (0.15 * nc)) ** 2)), np.r_[np.zeros(nc - 1), 1000.0]]
    for cnt in shapes:
        cnt = np.ascontiguousarray(cnt, dtype=np.int64)
        py = strips.balanced_rows(cnt, nranks)
        out = (C.c_long * (nranks + 1))()
        assert lib.parsim_balance_rows(nc, nranks, cnt.ctypes.data_as(C.POINTER(C.c_longlong)), out) == 0
        assert list(out) == py
        assert py[0] == 0 and py[-1] == nc and all(b > a for a, b in zip(py, py[1:]))
        total, prefix = int(cnt.sum()), np.concatenate([[0], np.cumsum(cnt)])
        for k in range(1, nranks):
            want = -(-total * k // nranks)
            free = py[k - 1] + 1 < py[k] < nc - (nranks - k)  # not pinned by the one-row minimum
            if free:
                assert prefix[py[k]] >= want and prefix[py[k] - 1] < want


def test_balanced_rows_even_out_a_normal_cloud():
    """Negative seeds draw a normal cloud (sigma = 0.15 side): equal rows give the middle strips
    several times the load of the outer ones, equal-particle strips stay within one row's worth."""
    seed, side, nc, n, R = -4, 10.0, 200, 100000, 8
    x, y, vx, vy, m = init_particles(seed, side, nc, n)
    cnt = strips.row_counts(y, side, nc)
    acc = np.zeros(nc, dtype=np.int64)
    lib = _lib.load()
    assert lib.parsim_host_row_counts(side, nc, n, y.ctypes.data_as(C.POINTER(C.c_double)),
                                      acc.ctypes.data_as(C.POINTER(C.c_longlong))) == 0
    assert np.array_equal(acc, cnt) and cnt.sum() == n
    eq = np.bincount(strips.strip_of_particles(y, side, nc, R), minlength=R)
    first = strips.balanced_rows(cnt, R)
    bal = np.bincount(strips.strip_of_particles(y, side, nc, R, first), minlength=R)
    assert eq.max() > 3 * eq.min()
    assert bal.max() - bal.min() <= 2 * cnt.max()
    parts = [strips.select_strip(r, R, side, nc, x, y, vx, vy, m, first) for r in range(R)]
    assert [len(p[0]) for p in parts] == bal.tolist()


def test_cell_row_matches_oracle_binning():
    """The launcher's strip choice uses the same row as the oracle's (= the reference's) binning."""
    import oracle_lib

    seed, side, nc, n = 5, 7.5, 37, 20000
    x, y, vx, vy, m = init_particles(seed, side, nc, n)
    o = oracle_lib.oracle_sim(seed, side, nc, n)
    try:
        o.phase(0)  # updateCOM bins every particle (serial/parsim.cpp:265-285)
        start, ids = o.lists()
    finally:
        o.close()
    cells = np.empty(n, dtype=np.int64)
    cells[ids] = np.repeat(np.arange(nc * nc), np.diff(start))  # cell = col + row * nc
    assert np.array_equal(cells // nc, strips.cell_row(y, side, nc))
    for nranks in (2, 3, 8):
        owner = strips.strip_of_particles(y, side, nc, nranks)
        parts = [strips.select_strip(r, nranks, side, nc, x, y, vx, vy, m) for r in range(nranks)]
        assert sum(len(p[0]) for p in parts) == n
        allids = np.concatenate([p[0] for p in parts])
        assert np.array_equal(np.sort(allids), np.arange(n, dtype=np.uint32))
        for r, p in enumerate(parts):
            assert np.all(np.diff(p[0].astype(np.int64)) > 0)  # id-ordered slots: no id passes in the step-0 sort
            assert np.all(owner[p[0]] == r)
            assert np.array_equal(p[2], y[p[0]])


def test_gloo_world2_launcher_plumbing(tmp_path):
    seed, side, nc, n = 3, 10.0, 25, 30000
    p = _torchrun(2, [os.path.join("tests", "_gloo_worker.py"), str(tmp_path), str(seed), str(side), str(nc), str(n)])
    assert p.returncode == 0, p.stderr[-2000:]
    res = [json.load(open(tmp_path / f"rank{r}.json")) for r in range(2)]
    assert [r["rank"] for r in res] == [0, 1] and all(r["world"] == 2 for r in res)
    assert res[0]["blob"] == res[1]["blob"] == bytes((7 * i + 3) % 256 for i in range(128)).hex()
    assert res[0]["counts"] == res[1]["counts"] == [res[0]["count"], res[1]["count"]]
    assert res[0]["count"] + res[1]["count"] == n
    assert all(r["ids_ascending"] and r["in_strip"] for r in res)
    assert res[0]["first_row"] == 0 and res[0]["rows"] == 12 and res[1]["first_row"] == 12 and res[1]["rows"] == 13
    assert res[0]["id_sum"] == res[1]["id_sum"] == n * (n - 1) / 2  # every id on exactly one strip
    assert abs(res[0]["mass_sum"] - res[0]["total_mass"]) <= 1e-9 * res[0]["total_mass"]
    # timing is the max over ranks on both ranks (rank 1 slept twice as long)
    assert res[0]["slowest"] == res[1]["slowest"] >= res[1]["mine"] - 1e-9
    assert res[0]["slowest"] > res[0]["mine"]


def test_reference_arm_under_torchrun_prints_one_line():
    """`bench.py --impl reference --gpus 2` launched like the driver does: rank 0 alone runs the
    reference CPU binary and prints ONE JSON line, rank 1 exits 0 silently."""
    if not (os.path.exists(os.path.join(ROOT, "oracle", "_ref", "parsim")) or
            os.path.exists(os.path.join(ROOT, "oracle", "parsim_oracle"))):
        pytest.skip("oracle not built")
    p = _torchrun(2, ["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                      "--workload", "config2"], timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["unit"] == "particle-steps/s"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] in ("reference", "port")
    assert d["value"] > 0
