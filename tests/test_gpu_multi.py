"""Row-strip multi-GPU path (halos + migrants): N strips must reproduce the single-GPU run
bit for bit -- same collision count, same alive set, identical positions -- because every
reduction has a fixed order that does not depend on the decomposition.

Transport: with at least N GPUs the strips talk over NCCL (one GPU each); with fewer -- the
driver's 1-GPU box -- the library oversubscribes the GPUs and the strips exchange by device
copies (`LocalBus`, csrc/parsim_rank.cuh), the reference's own way of testing its distributed
variant on one machine (mpi/run_tests.sh:9-17: local oversubscribed ranks).  Only the
transport differs: pack / unpack kernels, order rules and results are the same.  On a
multi-GPU box both transports run (`transport` parameter)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from particlesimulation_b200 import ParticleSimulation

pytestmark = pytest.mark.gpu


def ngpu():
    import torch
    return torch.cuda.device_count()


def pick_transport(transport, n_gpus, monkeypatch):
    """"auto" = what parsim_create picks (NCCL with enough GPUs, else device copies); "local" forces
    the device-copy transport and only adds something where "auto" is NCCL."""
    if ngpu() < 1:
        pytest.skip("needs a GPU")
    if transport == "local":
        if ngpu() < n_gpus:
            pytest.skip("auto is already the device-copy transport here")
        monkeypatch.setenv("PARSIM_TRANSPORT", "local")
    else:
        monkeypatch.delenv("PARSIM_TRANSPORT", raising=False)


TRANSPORTS = ["auto", "local"]


CASES = [(1, 10, 100, 100000, 5), (5, 1, 20, 20000, 4), (-7, 10, 50, 20000, 5), (3, 1, 2, 200, 3),
         (9, 0.3, 4, 3000, 6), (1, 2, 3, 10, 3), (7, 30, 300, 2000000, 4)]


@pytest.mark.parametrize("transport", TRANSPORTS)
@pytest.mark.parametrize("n_gpus", [2, 3, 4, 8])
@pytest.mark.parametrize("case", CASES, ids=lambda c: " ".join(map(str, c)))
def test_strips_match_single_gpu(case, n_gpus, transport, monkeypatch):
    seed, side, nc, n, steps = case
    pick_transport(transport, n_gpus, monkeypatch)
    if n_gpus > nc:
        pytest.skip("more strips than rows")
    monkeypatch.setenv("PARSIM_CAPACITY_PCT", "400")  # tiny / skewed cases: generous strip capacity
    one = ParticleSimulation(seed, side, nc, n)
    many = ParticleSimulation(seed, side, nc, n, n_gpus=n_gpus)
    for t in range(steps):
        one.simulate(1)
        many.simulate(1)
        assert one.result() == many.result(), f"step {t}"
    a, b = one.dump(), many.dump()
    assert b["owned"].all()
    for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
        assert np.array_equal(a[k], b[k]), k
    one.close(); many.close()


@pytest.mark.parametrize("transport", TRANSPORTS)
@pytest.mark.parametrize("n_gpus", [2, 4])
def test_exchange_overflow_round(n_gpus, transport, monkeypatch):
    """More records than the eager slots of the one-round exchange: the second round carries the
    rest (neighbour messages and the far all-gather), same bits as one GPU."""
    import oracle_lib as ol
    pick_transport(transport, n_gpus, monkeypatch)
    monkeypatch.setenv("PARSIM_CAPACITY_PCT", "400")
    monkeypatch.setenv("PARSIM_EAGER_RECS", "3")
    for case in ("dense", "far"):
        if case == "dense":
            seed, side, nc, n = 1, 10, 100, 100000
            arrs = None
        else:
            side, nc, arrs = ol.far_jump_case()
            seed, n = 0, len(arrs[0])
        one = ParticleSimulation(seed, side, nc, n, arrays=arrs)
        many = ParticleSimulation(seed, side, nc, n, arrays=arrs, n_gpus=n_gpus)
        for t in range(4):
            one.simulate(1)
            many.simulate(1)
            assert one.result() == many.result(), f"{case} step {t}"
        a, b = one.dump(), many.dump()
        for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
            assert np.array_equal(a[k], b[k]), (case, k)
        assert many.stats()["migrants"] > 3 * 4
        one.close(); many.close()


@pytest.mark.parametrize("transport", TRANSPORTS)
@pytest.mark.parametrize("n_gpus", [2, 3, 4, 8])
def test_far_jumps_cross_strips(n_gpus, transport, monkeypatch):
    """Close encounters throw particles many rows away (any strip, not just the neighbour): the
    far-record exchange must keep N strips bitwise equal to one GPU."""
    import oracle_lib as ol
    pick_transport(transport, n_gpus, monkeypatch)
    monkeypatch.setenv("PARSIM_CAPACITY_PCT", "400")
    side, nc, arrs = ol.far_jump_case()
    n = len(arrs[0])
    one = ParticleSimulation(0, side, nc, n, arrays=arrs)
    many = ParticleSimulation(0, side, nc, n, arrays=arrs, n_gpus=n_gpus)
    for t in range(4):
        one.simulate(1)
        many.simulate(1)
        assert one.result() == many.result(), f"step {t}"
    a, b = one.dump(), many.dump()
    assert b["owned"].all()
    for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
        assert np.array_equal(a[k], b[k]), k
    assert many.stats()["migrants"] > 0
    one.close(); many.close()


@pytest.mark.parametrize("transport", TRANSPORTS)
@pytest.mark.parametrize("n_gpus", [2, 3, 4, 8])
@pytest.mark.parametrize("case", [(-7, 10, 50, 20000, 5), (-3, 4, 40, 60000, 4), (5, 1, 20, 20000, 3)],
                         ids=lambda c: " ".join(map(str, c)))
def test_equal_particle_strips_match_single_gpu(case, n_gpus, transport, monkeypatch):
    """strips cut by particle count (clustered, negative-seed inputs) instead of by
    row count -- same bits as one GPU, default strip capacity, and an even load."""
    from particlesimulation_b200 import strips
    from particlesimulation_b200.simulation import init_particles
    seed, side, nc, n, steps = case
    pick_transport(transport, n_gpus, monkeypatch)
    monkeypatch.delenv("PARSIM_CAPACITY_PCT", raising=False)  # balanced strips must fit the default capacity
    arrs = init_particles(seed, side, nc, n)
    first = strips.balanced_rows(strips.row_counts(arrs[1], side, nc), n_gpus)
    owner = strips.strip_of_particles(arrs[1], side, nc, n_gpus, first)
    load = np.bincount(owner, minlength=n_gpus)
    rows_max = int(strips.row_counts(arrs[1], side, nc).max())
    assert load.max() - load.min() <= 2 * rows_max  # within one row of particles of each other
    one = ParticleSimulation(seed, side, nc, n, arrays=arrs)
    many = ParticleSimulation(seed, side, nc, n, arrays=arrs, n_gpus=n_gpus, first_rows="balanced")
    assert many.first_rows == first
    for t in range(steps):
        one.simulate(1)
        many.simulate(1)
        assert one.result() == many.result(), f"step {t}"
    a, b = one.dump(), many.dump()
    assert b["owned"].all()
    for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
        assert np.array_equal(a[k], b[k]), k
    one.close(); many.close()


def test_group_upload_paths_agree(monkeypatch):
    """In-process groups copy every piece host->device ONCE and all strips ingest from that copy
    (PARSIM_SHARED_UPLOAD, the default where the GPUs can reach each other); the per-strip copy is
    the fallback.  Same records either way."""
    if ngpu() < 1:
        pytest.skip("needs a GPU")
    monkeypatch.setenv("PARSIM_CAPACITY_PCT", "400")
    res = []
    for shared in ("1", "0"):
        monkeypatch.setenv("PARSIM_SHARED_UPLOAD", shared)
        g = ParticleSimulation(-7, 10, 50, 20000, n_gpus=3)
        g.simulate(3)
        res.append((g.result(), g.dump()))
        g.close()
    assert res[0][0] == res[1][0]
    for k in ("x", "y", "vx", "vy", "m", "alive", "cell"):
        assert np.array_equal(res[0][1][k], res[1][1][k]), k


@pytest.mark.timeout(180)
@pytest.mark.parametrize("transport", TRANSPORTS)
def test_a_failing_strip_takes_the_others_with_it(transport, monkeypatch):
    """A strip whose migrant buffer overflows (forced: one record of capacity) reports it through
    the exchanged headers, so EVERY strip leaves the step with an error -- nobody is left waiting in
    a collective for a strip that gave up (the reference's MPI variant simply aborts the job)."""
    from particlesimulation_b200.simulation import ParsimError
    pick_transport(transport, 2, monkeypatch)
    monkeypatch.setenv("PARSIM_CAPACITY_PCT", "400")
    monkeypatch.setenv("PARSIM_MIG_CAP", "1")
    g = ParticleSimulation(1, 10, 100, 100000, n_gpus=2)
    with pytest.raises(ParsimError, match="error code 3|overflow"):
        g.simulate(3)
    g.close()


def test_cli_balanced_strips_match_single_gpu():
    if ngpu() < 1:
        pytest.skip("needs a GPU")
    exe = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    args = ["-7", "10", "50", "200000", "10"]  # golden: 5.957 6.317 / 46580 (tests/golden/reference_runs.jsonl)
    one = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
    env = dict(os.environ, PARSIM_NGPU="4", PARSIM_BALANCE="1")  # fewer than 4 GPUs: oversubscribed
    env.pop("PARSIM_CAPACITY_PCT", None)
    env.pop("PARSIM_TRANSPORT", None)
    many = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300, env=env)
    assert one.returncode == 0 and many.returncode == 0, one.stderr + many.stderr
    assert one.stdout == many.stdout
    assert one.stdout.split() == ["5.957", "6.317", "46580"]


def test_bad_strip_rows_are_rejected():
    if ngpu() < 1:
        pytest.skip("needs a GPU")
    from particlesimulation_b200.simulation import ParsimError
    for rows in ([0, 5, 5, 10][:3], [0, 0, 10], [1, 5, 10], [0, 5, 9]):
        with pytest.raises((ParsimError, ValueError)):
            ParticleSimulation(1, 1.0, 10, 100, n_gpus=2, first_rows=rows)


def _torchrun_upload_worker(tmp_path, mode, case, port):
    seed, side, nc, n, steps = case
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "_nccl_upload_worker.py"), str(tmp_path), mode,
           str(seed), str(side), str(nc), str(n), str(steps)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    outs = []
    for r in range(2):
        f = os.path.join(str(tmp_path), f"rank{r}.json")
        assert os.path.exists(f), (p.returncode, p.stderr[-3000:])
        outs.append(json.load(open(f)))
    return outs


@pytest.mark.parametrize("case", [(1, 10.0, 100, 200000, 5), (5, 1.0, 20, 50000, 4)], ids=lambda c: " ".join(map(str, c)))
def test_collective_strip_upload_matches_one_gpu(case, tmp_path):
    """parsim_upload_strip, one process per GPU over NCCL: every rank hands over its strip in one
    collective call (ids, x, y, m first; binning, centres of mass, ghost rows and the first force sums
    under the copy of vx, vy), twice in a row on the same contexts; same result as one GPU, bit for bit."""
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    seed, side, nc, n, steps = case
    one = ParticleSimulation(seed, side, nc, n)
    one.simulate(steps)
    want = list(one.result())
    one.close()
    outs = _torchrun_upload_worker(tmp_path, "ok", case, 29551)
    assert sum(o["count"] for o in outs) == n
    for o in outs:
        assert "error" not in o, o
        assert o["result"] == want, (o, want)


def test_collective_strip_upload_fails_on_every_rank_together(tmp_path):
    """Rank 1 hands in a NULL id array: its own call fails, and rank 0 -- whose upload was fine --
    must come back with an error too instead of waiting for rank 1 in the ghost-row exchange."""
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    outs = _torchrun_upload_worker(tmp_path, "fail", (1, 10.0, 100, 200000, 3), 29552)
    assert "NULL array" in outs[1].get("error", ""), outs
    assert "strip 1" in outs[0].get("error", ""), outs


@pytest.mark.parametrize("n_gpus", [2, 4, 8])
def test_bench_under_torchrun(n_gpus):
    if ngpu() < n_gpus:
        pytest.skip(f"needs {n_gpus} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n_gpus}",
           "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.join(ROOT, "bench.py"),
           "--gpus", str(n_gpus), "--steps", "1", "--warmup", "1", "--scale", "0.02"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["n_gpus"] == n_gpus and d["value"] > 0 and d["e2e"]["value"] > 0
    # the same problem on one GPU gives the same collision count
    one = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1",
                          "--scale", str(0.02 * n_gpus), "--no-cpu-baseline"], capture_output=True, text=True, timeout=600)
    assert one.returncode == 0, one.stderr[-3000:]
    d1 = json.loads([l for l in one.stdout.splitlines() if l.startswith("{")][-1])
    if d1["config"]["workload"].split("n_part=")[1] == d["config"]["workload"].split("n_part=")[1] and \
            d1["config"]["workload"].split("ncside=")[1].split()[0] == d["config"]["workload"].split("ncside=")[1].split()[0]:
        assert d1["result"]["collisions"] == d["result"]["collisions"]
