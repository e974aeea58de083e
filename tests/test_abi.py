"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/parsim_b200.h declares, restates init_particles exactly, and fails loudly (no CPU
fallback) when there is no CUDA device."""
import ctypes as C
import json
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from particlesimulation_b200 import _lib, init_particles, ParticleSimulation, ParsimError
from particlesimulation_b200.simulation import init_particles_py
import oracle_lib as ol

VEC = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "parsim_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(parsim_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/parsim_b200.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature"
    assert lib.parsim_abi_version() == 1


def test_kernel_names():
    lib = _lib.load()
    names = []
    k = 0
    while lib.parsim_kernel_name(k):
        names.append(lib.parsim_kernel_name(k).decode())
        k += 1
    assert "force_integrate" in names and "collide_find" in names and k <= _lib.NKERNELS


def test_host_init_known_answers():
    x, y, vx, vy, m = init_particles(1, 2, 3, 10)
    assert [x[0], y[0], vx[0], vy[0], m[0]] == VEC["kat"]["config1_p0_init"]


@pytest.mark.parametrize("args", [(1, 2, 3, 10), (-10, 3, 3, 100), (7, 100, 1000, 5000), (-50, 10000, 200, 3000)])
def test_host_init_matches_oracle_and_python(args):
    a = init_particles(*args)
    for u, v, w in zip(a, ol.oracle_init(*args), init_particles_py(*args)):
        assert np.array_equal(u, v) and np.array_equal(u, w)


def test_host_init_chunked_matches_whole():
    lib = _lib.load()
    whole = init_particles(3, 1000, 100, 1000)
    part = [np.empty(300) for _ in range(5)]
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    assert lib.parsim_host_init_particles(3, 1000.0, 100, 1000, 500, 300, *[dp(a) for a in part]) == 0
    for u, v in zip(whole, part):
        assert np.array_equal(u[500:800], v)


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(ParsimError):
        ParticleSimulation(1, 2, 3, 10)
    cli = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    if os.path.exists(cli):
        p = subprocess.run([cli, "1", "2", "3", "10", "1"], capture_output=True, text=True)
        assert p.returncode == 1 and p.stderr.startswith("Error: ") and p.stdout == ""


def test_integration_patch_is_the_compiled_one():
    """INTEGRATION.md section 2 shows the source-level patch; oracle/Makefile compiles exactly that
    text (oracle/simulate_b200.inc) into the reference's serial program.  No GPU here: the patched
    program must then fail the reference's way ("Error: ...", exit 1) -- no CPU fallback."""
    inc = open(os.path.join(ROOT, "oracle", "simulate_b200.inc")).read()
    assert inc in open(os.path.join(ROOT, "INTEGRATION.md")).read()
    exe = os.path.join(ROOT, "oracle", "_ref", "parsim-patched")
    if not os.path.exists(exe):
        pytest.skip("reference sources not present: oracle/_ref/parsim-patched not built")
    if not has_gpu():
        p = subprocess.run([exe, "1", "2", "3", "10", "1"], capture_output=True, text=True)
        assert p.returncode == 1 and p.stderr.startswith("Error: ") and p.stdout == ""
    p = subprocess.run([exe, "1", "2", "3"], capture_output=True, text=True)
    assert p.returncode == 1 and p.stderr.startswith("Error: Usage:")


def test_cli_usage_error():
    cli = os.path.join(ROOT, "particlesimulation_b200", "parsim")
    if not os.path.exists(cli):
        pytest.skip("CLI not built")
    p = subprocess.run([cli, "1", "2", "3"], capture_output=True, text=True)
    assert p.returncode == 1
    assert p.stderr.strip() == f"Error: Usage: {cli} <seed> <side_length> <grid_size> <n_particles> <n_timesteps>"
