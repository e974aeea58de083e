"""ctypes access to the oracle (oracle/libparsim_oracle.so, the plain-C restatement) and,
where it was built, to the unmodified reference (oracle/_ref/libparsim_ref_phases.so).
TEST INFRASTRUCTURE: imported only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libparsim_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libparsim_ref_phases.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "parsim")

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lp = C.POINTER(C.c_longlong)


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


class _Sim:
    """Common driver over the two libraries (prefix 'po_' = oracle, 'refph_' = reference)."""

    def __init__(self, lib, prefix, handle, n, ncells):
        self.lib, self.p, self.h, self.n, self.ncells = lib, prefix, handle, n, ncells

    def _f(self, name):
        return getattr(self.lib, self.p + name)

    def phase(self, k):
        self._f("phase")(self.h, k)

    def step(self):
        self._f("step")(self.h)

    def run(self, steps):
        for _ in range(steps):
            self.step()

    @property
    def collisions(self):
        return self._f("collisions")(self.h)

    def particles(self):
        n = self.n
        out = {k: np.zeros(n) for k in ("x", "y", "vx", "vy", "m", "fx", "fy")}
        alive = np.zeros(n, dtype=np.int32)
        cell = np.zeros(n, dtype=np.int32)
        self._f("particles")(self.h, *[_d(out[k]) for k in ("x", "y", "vx", "vy", "m", "fx", "fy")],
                             alive.ctypes.data_as(_ip), cell.ctypes.data_as(_ip))
        out["alive"], out["cell_index"] = alive, cell
        return out

    def cells(self):
        mx, my, m = (np.zeros(self.ncells) for _ in range(3))
        self._f("cells")(self.h, _d(mx), _d(my), _d(m))
        return mx, my, m

    def lists(self):
        start = np.zeros(self.ncells + 1, dtype=np.int64)
        tot = self._f("lists")(self.h, start.ctypes.data_as(_lp), None)
        ids = np.zeros(max(tot, 1), dtype=np.int64)
        self._f("lists")(self.h, start.ctypes.data_as(_lp), ids.ctypes.data_as(_lp))
        return start, ids[:tot]

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        self.close()


def _load_oracle():
    lib = C.CDLL(ORACLE_SO)
    lib.po_create.restype = C.c_void_p
    lib.po_create.argtypes = [C.c_long, C.c_double, C.c_long, C.c_longlong]
    lib.po_create_from.restype = C.c_void_p
    lib.po_create_from.argtypes = [C.c_double, C.c_long, C.c_longlong, _dp, _dp, _dp, _dp, _dp]
    lib.po_destroy.argtypes = [C.c_void_p]
    lib.po_set_semantics.argtypes = [C.c_void_p, C.c_int]
    lib.po_phase.argtypes = [C.c_void_p, C.c_int]
    lib.po_step.argtypes = [C.c_void_p]
    lib.po_run.argtypes = [C.c_void_p, C.c_long]
    lib.po_collisions.restype = C.c_longlong
    lib.po_collisions.argtypes = [C.c_void_p]
    lib.po_alive_pairs.restype = C.c_longlong
    lib.po_alive_pairs.argtypes = [C.c_void_p]
    lib.po_particles.argtypes = [C.c_void_p] + [_dp] * 7 + [_ip, _ip]
    lib.po_cells.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.po_lists.restype = C.c_longlong
    lib.po_lists.argtypes = [C.c_void_p, _lp, _lp]
    lib.po_init_particles.argtypes = [C.c_long, C.c_double, C.c_long, C.c_longlong] + [_dp] * 5
    lib.po_draws.argtypes = [C.c_long, C.c_int, _dp]
    return lib


def _load_ref():
    lib = C.CDLL(REF_SO)
    lib.refph_create.restype = C.c_void_p
    lib.refph_create.argtypes = [C.c_long, C.c_double, C.c_long, C.c_longlong]
    lib.refph_destroy.argtypes = [C.c_void_p]
    lib.refph_phase.argtypes = [C.c_void_p, C.c_int]
    lib.refph_step.argtypes = [C.c_void_p]
    lib.refph_collisions.restype = C.c_longlong
    lib.refph_collisions.argtypes = [C.c_void_p]
    lib.refph_particles.argtypes = [C.c_void_p] + [_dp] * 7 + [_ip, _ip]
    lib.refph_cells.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.refph_lists.restype = C.c_longlong
    lib.refph_lists.argtypes = [C.c_void_p, _lp, _lp]
    return lib


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        _oracle = _load_oracle()
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def oracle_sim(seed, side, ncside, n_part, semantics=0):
    """semantics: 0 = serial/parsim.cpp (the parity target), 1 = the OpenMP variant's list update"""
    lib = oracle_lib()
    h = lib.po_create(int(seed), float(side), int(ncside), int(n_part))
    if semantics:
        lib.po_set_semantics(h, int(semantics))
    return _Sim(lib, "po_", h, int(n_part), int(ncside) ** 2)


def oracle_sim_from(side, ncside, x, y, vx, vy, m):
    lib = oracle_lib()
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, vx, vy, m)]
    h = lib.po_create_from(float(side), int(ncside), len(arrs[0]), *[_d(a) for a in arrs])
    return _Sim(lib, "po_", h, len(arrs[0]), int(ncside) ** 2)


def ref_sim(seed, side, ncside, n_part):
    global _ref
    if _ref is None:
        _ref = _load_ref()
    h = _ref.refph_create(int(seed), float(side), int(ncside), int(n_part))
    return _Sim(_ref, "refph_", h, int(n_part), int(ncside) ** 2)


def oracle_init(seed, side, ncside, n_part):
    lib = oracle_lib()
    arrs = [np.zeros(int(n_part)) for _ in range(5)]
    lib.po_init_particles(int(seed), float(side), int(ncside), int(n_part), *[_d(a) for a in arrs])
    return arrs


def oracle_draws(seed, count):
    out = np.zeros(count)
    oracle_lib().po_draws(int(seed), count, _d(out))
    return out


def far_jump_case(n=4000, pairs=40, side=100.0, ncside=40, seed=11):
    """A hand-made initial condition with `pairs` close encounters (separations of a few 1e-6):
    their mutual force throws both partners tens of cells away in ONE step -- the reference just
    re-bins them wherever they land (serial/parsim.cpp:171-190).  Exercises the fmod / division
    slow paths and, on several GPUs, the far-record exchange."""
    rng = np.random.default_rng(seed)
    G, eps2 = 6.67408e-11, 0.005 * 0.005
    x = rng.random(n) * side
    y = rng.random(n) * side
    vx = (rng.random(n) - 0.5) * side / ncside / 5.0
    vy = (rng.random(n) - 0.5) * side / ncside / 5.0
    m = (0.25 + 0.5 * rng.random(n)) * 0.01 * (ncside * ncside) / n / G * eps2
    for k in range(pairs):
        a, b = 2 * k, 2 * k + 1
        d = 2.5e-6 + 4.5e-6 * rng.random()
        ang = 2 * np.pi * rng.random()
        x[b] = x[a] + d * np.cos(ang)
        y[b] = y[a] + d * np.sin(ang)
    x = np.clip(x, 0.0, np.nextafter(side, 0.0))
    y = np.clip(y, 0.0, np.nextafter(side, 0.0))
    return side, ncside, tuple(np.ascontiguousarray(v) for v in (x, y, vx, vy, m))
