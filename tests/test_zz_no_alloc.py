"""parsim_cell_com_static (include/parsim_b200.h): the allocation-free variant of the
centre-of-mass phase, checked through the C ABI against the existing entry points
(parsim_upload + parsim_debug_stage(0) + parsim_dump_cells) on two calls in a row.

This file is named so that pytest collects it last.
Tolerances are those of test_gpu_parity.py's stage-0 check (mass rtol 1e-12, centre 1e-9 * side);
the two paths sum every cell in the same order, so the results are in fact identical."""
import ctypes as C

import numpy as np
import pytest

from particlesimulation_b200 import ParticleSimulation, _lib, init_particles

MAX_PARTICLES = 262144  # PARSIM_STATIC_MAX_PARTICLES
MAX_CELLS = 262144      # PARSIM_STATIC_MAX_CELLS

# seed, side, ncside, n_part   (mean occupancy 10 / 50 / 50 / 187 / 1 / 1: every lane width of k_com)
CASES = [(1, 10.0, 100, 100000), (5, 1.0, 20, 20000), (3, 1.0, 1, 50), (9, 0.3, 4, 3000),
         (7, 5.0, 200, 40000), (2, 100.0, 512, MAX_PARTICLES)]


def context_cells(seed, side, ncside, n, arrays):
    g = ParticleSimulation(seed, side, ncside, n, arrays=arrays)
    g.debug_stage(0)
    out = g.dump_cells()
    g.close()
    return out


def static_cells(lib, torch, side, ncside, n, dev_in, stream=None):
    ncells = ncside * ncside
    outs = [torch.full((ncells,), float("nan"), dtype=torch.float64, device="cuda") for _ in range(3)]
    torch.cuda.synchronize()
    rc = lib.parsim_cell_com_static(side, ncside, n, *[t.data_ptr() for t in dev_in],
                                    *[t.data_ptr() for t in outs], stream)
    assert rc == 0, lib.parsim_last_error(None).decode()
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in outs]


def check(got, want, side, tag):
    (mx, my, m), (wx, wy, wm) = got, want
    assert np.allclose(m, wm, rtol=1e-12, atol=0), tag
    assert np.array_equal(m > 0, wm > 0), tag
    tol = 1e-9 * side
    assert np.allclose(mx, wx, rtol=0, atol=tol) and np.allclose(my, wy, rtol=0, atol=tol), tag
    assert np.array_equal(m, wm) and np.array_equal(mx, wx) and np.array_equal(my, wy), tag + " (bitwise)"


@pytest.mark.gpu
def test_static_scratch_variant_matches_context_path_twice_in_a_row():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    for case_a, case_b in zip(CASES, CASES[1:] + CASES[:1]):
        # two calls in a row on different problems: the second sees the first one's scratch
        for seed, side, ncside, n in (case_a, case_b):
            arrays = init_particles(seed, side, ncside, n)
            want = context_cells(seed, side, ncside, n, arrays)
            dev_in = [torch.from_numpy(np.ascontiguousarray(arrays[k])).cuda() for k in (0, 1, 4)]
            for rep in range(2):  # and twice on the same problem
                got = static_cells(lib, torch, side, ncside, n, dev_in)
                check(got, want, side, f"{seed} {side} {ncside} {n} call {rep}")


@pytest.mark.gpu
def test_static_scratch_variant_on_a_side_stream_allocates_nothing():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    seed, side, ncside, n = 11, 2.0, 64, 50000
    arrays = init_particles(seed, side, ncside, n)
    want = context_cells(seed, side, ncside, n, arrays)
    dev_in = [torch.from_numpy(np.ascontiguousarray(arrays[k])).cuda() for k in (0, 1, 4)]
    st = torch.cuda.Stream()
    check(static_cells(lib, torch, side, ncside, n, dev_in, st.cuda_stream), want, side, "first")
    free_before = torch.cuda.mem_get_info()[0]
    outs = [torch.empty(ncside * ncside, dtype=torch.float64, device="cuda") for _ in range(3)]
    free_mid = torch.cuda.mem_get_info()[0]
    for _ in range(2):
        rc = lib.parsim_cell_com_static(side, ncside, n, *[t.data_ptr() for t in dev_in],
                                        *[t.data_ptr() for t in outs], st.cuda_stream)
        assert rc == 0, lib.parsim_last_error(None).decode()
    st.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_mid <= free_before
    check([t.cpu().numpy() for t in outs], want, side, "side stream")
    # an empty set: every cell is 0, 0, 0 whatever the previous call left behind
    rc = lib.parsim_cell_com_static(side, ncside, 0, None, None, None, *[t.data_ptr() for t in outs], st.cuda_stream)
    assert rc == 0, lib.parsim_last_error(None).decode()
    st.synchronize()
    assert all(float(t.abs().max()) == 0.0 for t in outs)


def test_static_scratch_variant_rejects_sizes_beyond_its_maximum():
    """Argument checks come before any CUDA call, so this runs without a GPU."""
    lib = _lib.load()
    p = C.c_void_p(256)  # never dereferenced
    for side, ncside, n, word in ((1.0, 4, MAX_PARTICLES + 1, "PARSIM_STATIC_MAX_PARTICLES"),
                                  (1.0, 513, 10, "PARSIM_STATIC_MAX_CELLS"), (1.0, 0, 10, "ncside"),
                                  (1.0, 4, -1, "n_part"), (0.0, 4, 10, "side")):
        assert lib.parsim_cell_com_static(side, ncside, n, p, p, p, p, p, p, None) == 1
        assert word in lib.parsim_last_error(None).decode()
    assert lib.parsim_cell_com_static(1.0, 4, 10, p, p, p, None, p, p, None) == 1
    assert "NULL" in lib.parsim_last_error(None).decode()
    assert lib.parsim_cell_com_static(1.0, 4, 10, None, p, p, p, p, p, None) == 1
    assert "NULL" in lib.parsim_last_error(None).decode()
    assert 512 * 512 == MAX_CELLS
