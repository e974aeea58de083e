"""Host-side mirror of the reference's `ParticleSimulation` class (serial/parsim.cpp:197-455)
on top of the C ABI: same constructor arguments, `simulate(n)` and `print_result()`.

`init_particles` here is the numpy-free restatement of serial/parsim.cpp:18-48,:220-232 used
when the caller does not bring its own arrays; the C++ host (`parsim` CLI) has its own copy.
"""
import ctypes as C
import math

import numpy as np

from . import _lib

G = 6.67408e-11
EPSILON2 = 0.005 * 0.005


class ParsimError(RuntimeError):
    pass


def _uniform_stream(seed, count):
    """`count` draws of RandomGenerator::uniform01 (serial/parsim.cpp:26-32), vectorised:
    xorshift32 is sequential, so the states are produced by a tight integer loop in chunks."""
    state = (abs(int(seed)) + 987654321) & 0xFFFFFFFF
    states = np.empty(count + 1, dtype=np.uint32)
    states[0] = state
    s = state
    for i in range(1, count + 1):
        s ^= (s << 13) & 0xFFFFFFFF
        s ^= s >> 17
        s ^= (s << 5) & 0xFFFFFFFF
        states[i] = s
    tot = (states[:-1] + states[1:]).astype(np.uint32).view(np.int32)  # wrapping int add
    return 0.5 + 0.2328306e-09 * tot.astype(np.float64), int(s)


def init_particles(seed, side, ncside, n_part, out=None):
    """x, y, vx, vy, m exactly as ParticleSimulation::init_particles leaves them (host C++ code
    in libparsim_b200.so; needs no GPU).  `out` may hold five preallocated float64 arrays."""
    lib = _lib.load()
    n = int(n_part)
    arrs = out if out is not None else tuple(np.empty(n) for _ in range(5))
    rc = lib.parsim_host_init_particles(int(seed), float(side), int(ncside), n, 0, n,
                                        *[_dptr(a) for a in arrs])
    if rc:
        raise ParsimError(lib.parsim_last_error(None).decode())
    return arrs


def init_particles_py(seed, side, ncside, n_part):
    """Pure numpy/Python restatement of the same generator (cross-check for the tests)."""
    seed = int(np.int32(int(seed)))  # main stores the seed in an int (:465)
    n = int(n_part)
    if seed >= 0:
        r, _ = _uniform_stream(seed, 5 * n)
        r = r.reshape(n, 5)
    else:
        r = np.empty((n, 5))
        state = (abs(seed) + 987654321) & 0xFFFFFFFF

        def uni():
            nonlocal state
            b = state
            s = b
            s ^= (s << 13) & 0xFFFFFFFF
            s ^= s >> 17
            s ^= (s << 5) & 0xFFFFFFFF
            state = s
            t = (b + s) & 0xFFFFFFFF
            if t >= 0x80000000:
                t -= 0x100000000
            return 0.5 + 0.2328306e-09 * t

        flat = r.reshape(-1)
        for k in range(5 * n):
            while True:  # normal01, :34-43
                u1 = uni()
                u2 = uni()
                z = math.sqrt(-2 * math.log(u1)) * math.cos(2 * math.pi * u2)
                v = 0.5 + 0.15 * z
                if 0 <= v < 1:
                    break
            flat[k] = v
    x = r[:, 0] * side
    y = r[:, 1] * side
    vx = (r[:, 2] - 0.5) * side / ncside / 5.0
    vy = (r[:, 3] - 0.5) * side / ncside / 5.0
    m = r[:, 4] * 0.01 * float(ncside * ncside) / float(n) / G * EPSILON2
    return tuple(np.ascontiguousarray(a) for a in (x, y, vx, vy, m))


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class ParticleSimulation:
    """ParticleSimulation(seed, side, ncside, n_part) -> .simulate(n_steps) -> .print_result()."""

    def __init__(self, seed, side, ncside, n_part, n_gpus=1, arrays=None, rank=None, nranks=1,
                 device=0, nccl_id=None, device_init=False, options=None, upload=True, first_rows=None):
        """`first_rows` (n_strips + 1 row indices, or "balanced"): explicit / equal-particle strips
        instead of equal-row strips (multi-GPU only; "balanced" needs host arrays or a host init)."""
        self.lib = _lib.load()
        self.side, self.ncside, self.n_part = float(side), int(ncside), int(n_part)
        self._ctx = C.c_void_p()
        n_strips = n_gpus if rank is None else nranks
        if isinstance(first_rows, str):
            if first_rows != "balanced":
                raise ValueError("first_rows must be a sequence or 'balanced'")
            if arrays is None:
                arrays = init_particles(seed, side, ncside, n_part)
            from . import strips
            first_rows = strips.balanced_rows(strips.row_counts(arrays[1], side, ncside), n_strips)
        rows = None
        if first_rows is not None:
            if len(first_rows) != n_strips + 1:
                raise ValueError("first_rows needs n_strips + 1 entries")
            rows = (C.c_long * (n_strips + 1))(*[int(r) for r in first_rows])
        self.first_rows = None if first_rows is None else [int(r) for r in first_rows]
        if rank is None:
            rc = self.lib.parsim_create_rows(C.byref(self._ctx), self.side, self.ncside, self.n_part, n_gpus, rows)
        else:
            idbuf = None
            if nccl_id is not None:
                idbuf = (C.c_ubyte * 128).from_buffer_copy(bytes(nccl_id))
            rc = self.lib.parsim_create_rank_rows(C.byref(self._ctx), self.side, self.ncside, self.n_part,
                                                  rank, nranks, device, idbuf, rows)
        if rc:
            self._ctx = C.c_void_p()
            raise ParsimError(self.lib.parsim_last_error(None).decode())
        for k, v in (options or {}).items():
            self.set_option(k, v)
        if device_init:
            self._check(self.lib.parsim_init_device(self._ctx, int(seed)))
        elif upload:
            if arrays is None:
                arrays = init_particles(seed, side, ncside, n_part)
            self.upload(*arrays)

    def _check(self, rc):
        if rc:
            raise ParsimError(self.lib.parsim_last_error(self._ctx).decode())

    def upload(self, x, y, vx, vy, m):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, vx, vy, m)]
        for a in arrs:
            if a.shape != (self.n_part,):
                raise ValueError("arrays must have n_part elements")
        self._check(self.lib.parsim_upload(self._ctx, *[_dptr(a) for a in arrs]))

    def upload_ptrs(self, ptrs):
        """Upload from raw host addresses (e.g. pinned torch tensors' data_ptr())."""
        cast = [C.cast(C.c_void_p(p), C.POINTER(C.c_double)) for p in ptrs]
        self._check(self.lib.parsim_upload(self._ctx, *cast))

    def upload_strip_ptrs(self, count, id_ptr, ptrs):
        """Upload `count` particles with explicit ids from raw host addresses."""
        self._check(self.lib.parsim_upload_begin(self._ctx))
        cast = [C.cast(C.c_void_p(p), C.POINTER(C.c_double)) for p in ptrs]
        self._check(self.lib.parsim_upload_chunk_ids(
            self._ctx, int(count), C.cast(C.c_void_p(id_ptr), C.POINTER(C.c_uint)), *cast))
        self._check(self.lib.parsim_upload_end(self._ctx))

    def upload_whole_strip_ptrs(self, count, id_ptr, ptrs):
        """parsim_upload_strip: this rank's whole strip in one call (collective on multi-rank contexts)."""
        cast = [C.cast(C.c_void_p(p), C.POINTER(C.c_double)) for p in ptrs]
        self._check(self.lib.parsim_upload_strip(
            self._ctx, int(count), C.cast(C.c_void_p(id_ptr), C.POINTER(C.c_uint)), *cast))

    def download(self, capacity=None):
        """ids, x, y, vx, vy, m of the alive particles of this strip (device order)."""
        cnt = C.c_longlong()
        if capacity is None:  # ask for the count first (a capacity of 0 only reports it)
            one = np.zeros(1)
            oid = np.zeros(1, dtype=np.uint32)
            self.lib.parsim_download(self._ctx, 0, oid.ctypes.data_as(C.POINTER(C.c_uint)),
                                     *[_dptr(one)] * 5, C.byref(cnt))
            capacity = cnt.value
        cap = max(int(capacity), 1)
        ids = np.zeros(cap, dtype=np.uint32)
        arrs = [np.zeros(cap) for _ in range(5)]
        self._check(self.lib.parsim_download(self._ctx, cap, ids.ctypes.data_as(C.POINTER(C.c_uint)),
                                             *[_dptr(a) for a in arrs], C.byref(cnt)))
        n = cnt.value
        return (ids[:n],) + tuple(a[:n] for a in arrs)

    def reset(self):
        self._check(self.lib.parsim_reset(self._ctx))

    def fp64_peak_tflops(self):
        v = C.c_double()
        self._check(self.lib.parsim_fp64_peak(self._ctx, C.byref(v)))
        return v.value

    def pair_peak(self):
        v = C.c_double()
        self._check(self.lib.parsim_pair_peak(self._ctx, C.byref(v)))
        return v.value

    def simulate(self, n_time_steps):  # serial/parsim.cpp:422
        self._check(self.lib.parsim_run(self._ctx, int(n_time_steps)))

    def result(self):
        x0, y0, col = C.c_double(), C.c_double(), C.c_longlong()
        self._check(self.lib.parsim_result(self._ctx, C.byref(x0), C.byref(y0), C.byref(col)))
        return x0.value, y0.value, col.value

    def print_result(self):  # serial/parsim.cpp:450-453
        x0, y0, col = self.result()
        print(f"{x0:.3f} {y0:.3f}")
        print(col)

    # ---- test / measurement hooks ----
    def debug_stage(self, stage):
        self._check(self.lib.parsim_debug_stage(self._ctx, int(stage)))

    def dump(self):
        n = self.n_part
        out = {k: np.zeros(n) for k in ("x", "y", "vx", "vy", "m")}
        ints = {k: np.zeros(n, dtype=np.int32) for k in ("alive", "cell", "list_cell")}
        rank = np.zeros(n, dtype=np.uint64)
        owned = np.zeros(n, dtype=np.uint8)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
        self._check(self.lib.parsim_dump(
            self._ctx, _dptr(out["x"]), _dptr(out["y"]), _dptr(out["vx"]), _dptr(out["vy"]),
            _dptr(out["m"]), ip(ints["alive"]), ip(ints["cell"]), ip(ints["list_cell"]),
            rank.ctypes.data_as(C.POINTER(C.c_ulonglong)), owned.ctypes.data_as(C.POINTER(C.c_ubyte))))
        out.update(ints)
        out["list_rank"] = rank
        out["owned"] = owned
        return out

    def dump_cells(self):
        c = self.ncside * self.ncside
        mx, my, m = np.zeros(c), np.zeros(c), np.zeros(c)
        self._check(self.lib.parsim_dump_cells(self._ctx, _dptr(mx), _dptr(my), _dptr(m)))
        return mx, my, m

    def set_option(self, name, value):
        self._check(self.lib.parsim_set_option(self._ctx, name.encode(), int(value)))

    def stats(self):
        s = _lib.Stats()
        self._check(self.lib.parsim_stats(self._ctx, C.byref(s)))
        d = {f: getattr(s, f) for f, _ in _lib.Stats._fields_ if not f.startswith("kernel_")}
        d["kernels"] = {}
        for k in range(_lib.NKERNELS):
            name = self.lib.parsim_kernel_name(k)
            if name is None:
                break
            d["kernels"][name.decode()] = {"ms": s.kernel_ms[k], "launches": s.kernel_launches[k],
                                           "bytes": s.kernel_bytes[k]}
        return d

    def stats_reset(self):
        self._check(self.lib.parsim_stats_reset(self._ctx))

    def close(self):
        if self._ctx:
            self.lib.parsim_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
