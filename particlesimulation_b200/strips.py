"""Host side of the row-strip decomposition: which grid rows a rank owns, which
rank a particle starts on, and the small `torch.distributed` group the launcher (`bench.py`,
the tests) uses for the out-of-band plumbing -- the NCCL unique id, barriers, max-over-ranks
timing.  The data path (halo rows, emigrants, late and far records) is NCCL inside
libparsim_b200.so; nothing here touches particle data during a step.

The reference's MPI variant splits rows the same way (mpi/parsim-mpi.cpp:338-342: contiguous
row blocks, remainder spread one row each); here rank r owns rows [r*nc/R, (r+1)*nc/R), which
is what `parsim_create_rank` computes (csrc/parsim_rank.cu) and `parsim_strip_rows` exports.
"""
import os

import numpy as np


def strip_rows(ncside, nranks, rank):
    """(first_row, n_rows) of `rank`'s strip."""
    if not 0 <= rank < nranks:
        raise ValueError("rank out of range")
    if nranks > ncside:
        raise ValueError("more strips than grid rows")
    r0 = rank * ncside // nranks
    r1 = (rank + 1) * ncside // nranks
    return r0, r1 - r0


def strip_of_row(row, ncside, nranks, first_rows=None):
    """Owner of a grid row: the unique r with first_rows[r] <= row < first_rows[r+1]
    (equal-row strips, r*nc//R, when first_rows is None)."""
    if first_rows is None:
        first = np.array([r * int(ncside) // int(nranks) for r in range(int(nranks))], dtype=np.int64)
    else:
        first = np.asarray(first_rows[:-1], dtype=np.int64)
    return np.searchsorted(first, np.asarray(row, dtype=np.int64), side="right") - 1


def row_counts(y, side, ncside):
    """Particles per grid row (what parsim_host_row_counts accumulates)."""
    return np.bincount(cell_row(y, side, ncside), minlength=int(ncside)).astype(np.int64)


def balanced_rows(counts, nranks):
    """Equal-PARTICLE strips: boundary k is the row where the running count first
    reaches k/R of the total, moved as little as needed so that every strip keeps >= 1 row.
    Returns nranks + 1 row indices; same arithmetic as parsim_balance_rows (csrc/parsim_rank.cuh)."""
    counts = np.asarray(counts, dtype=np.int64)
    nc, R = len(counts), int(nranks)
    if not 1 <= R <= nc:
        raise ValueError("need 1 <= nranks <= rows")
    prefix = np.concatenate([[0], np.cumsum(counts)])  # prefix[r] = particles in rows < r
    total = int(prefix[-1])
    first = [0]
    for k in range(1, R):
        want = (total * k + R - 1) // R
        b = int(np.searchsorted(prefix, want, side="left"))  # smallest r with prefix[r] >= want
        b = min(max(b, first[-1] + 1), nc - (R - k))
        first.append(b)
    first.append(nc)
    return first


def cell_row(y, side, ncside):
    """Grid row of a y coordinate with the reference's binning arithmetic
    (serial/parsim.cpp:268-272: int(y / (side / ncside)), clamped into the grid)."""
    cw = float(side) / float(int(ncside))
    row = (np.asarray(y, dtype=np.float64) / cw).astype(np.int64)
    return np.clip(row, 0, int(ncside) - 1)


def strip_of_particles(y, side, ncside, nranks, first_rows=None):
    """Rank on which each particle is binned at step 0."""
    return strip_of_row(cell_row(y, side, ncside), int(ncside), int(nranks), first_rows)


def select_strip(rank, nranks, side, ncside, x, y, vx, vy, m, first_rows=None):
    """ids (ascending) and state of the particles that start on `rank`'s strip -- what a launcher
    hands to `parsim_upload_strip*` when every rank generates the full seeded stream on the host."""
    owner = strip_of_particles(y, side, ncside, nranks, first_rows)
    ids = np.nonzero(owner == rank)[0].astype(np.uint32)
    return (ids,) + tuple(np.ascontiguousarray(np.asarray(a)[ids]) for a in (x, y, vx, vy, m))


class RankGroup:
    """One process per GPU under `torch.distributed.run` (RANK / LOCAL_RANK / WORLD_SIZE /
    MASTER_* from the environment).  backend "nccl" on the GPU box, "gloo" in the CPU tests."""

    def __init__(self, backend="nccl", expect_world=None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.backend = backend
        self._dist = None
        if expect_world is not None and expect_world != self.world:
            raise SystemExit(f"--gpus {expect_world} but WORLD_SIZE={self.world}: launch with torch.distributed.run")
        if self.world > 1:
            import torch
            import torch.distributed as dist

            kw = {}
            if backend == "nccl":
                torch.cuda.set_device(self.local)
                kw["device_id"] = torch.device("cuda", self.local)
            dist.init_process_group(backend, **kw)
            self._dist = dist
        elif backend == "nccl":
            import torch

            torch.cuda.set_device(self.local)

    @property
    def _device(self):
        return "cuda" if self.backend == "nccl" else "cpu"

    def _sync(self):
        if self.backend == "nccl":
            import torch

            torch.cuda.synchronize()

    def broadcast_bytes(self, blob, src=0):
        """`blob` (bytes on `src`, ignored elsewhere) on every rank: how the 128-byte NCCL unique id
        of `parsim_nccl_unique_id` reaches `parsim_create_rank` on the other ranks."""
        if self.world == 1:
            return blob
        box = [blob if self.rank == src else None]
        self._dist.broadcast_object_list(box, src=src)
        return box[0]

    def barrier(self):
        self._sync()
        if self.world > 1:
            self._dist.barrier()
            self._sync()

    def _reduce(self, v, op):
        if self.world == 1:
            return v
        import torch

        t = torch.tensor([v], dtype=torch.float64, device=self._device)
        self._dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(self, v):
        """Multi-GPU times are the MAX over ranks (never a rank-0 wall clock)."""
        return self._reduce(v, self._dist.ReduceOp.MAX if self._dist else None)

    def sum_over_ranks(self, v):
        return self._reduce(v, self._dist.ReduceOp.SUM if self._dist else None)

    def gather_ints(self, v):
        """[v of rank 0, v of rank 1, ...] on every rank."""
        if self.world == 1:
            return [int(v)]
        import torch

        out = [torch.zeros(1, dtype=torch.int64, device=self._device) for _ in range(self.world)]
        self._dist.all_gather(out, torch.tensor([int(v)], dtype=torch.int64, device=self._device))
        return [int(t.item()) for t in out]

    def close(self):
        if self._dist is not None and self._dist.is_initialized():
            self._dist.destroy_process_group()
            self._dist = None
