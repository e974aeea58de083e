"""Worker of tests/test_multi_cpu.py: one process per (pretend) GPU under torch.distributed.run,
`gloo` backend, no CUDA.  Exercises the host side of the N>1 path exactly as bench.py drives
it: rendezvous, the 128-byte id broadcast, strip selection of the seeded particle stream,
count gathering, barrier + max-over-ranks timing.  Writes one JSON file per rank."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from particlesimulation_b200 import _lib, strips  # noqa: E402
from particlesimulation_b200.simulation import init_particles  # noqa: E402


def main():
    outdir, seed, side, nc, n = sys.argv[1], int(sys.argv[2]), float(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    grp = strips.RankGroup(backend="gloo")
    lib = _lib.load()

    # out-of-band id: rank 0 makes 128 bytes, everyone must hold the same blob afterwards
    blob = bytes((7 * i + 3) % 256 for i in range(128)) if grp.rank == 0 else None
    blob = grp.broadcast_bytes(blob)

    # every rank generates the full seeded stream on the host (as the CLI does) and keeps its strip
    x, y, vx, vy, m = init_particles(seed, side, nc, n)
    ids, sx, sy, svx, svy, sm = strips.select_strip(grp.rank, grp.world, side, nc, x, y, vx, vy, m)
    import ctypes as C

    r0, rows = C.c_long(), C.c_long()
    assert lib.parsim_strip_rows(nc, grp.world, grp.rank, C.byref(r0), C.byref(rows)) == 0
    assert (r0.value, rows.value) == strips.strip_rows(nc, grp.world, grp.rank)
    row = strips.cell_row(sy, side, nc)
    in_strip = bool(np.all((row >= r0.value) & (row < r0.value + rows.value)))

    counts = grp.gather_ints(len(ids))
    id_sum = grp.sum_over_ranks(float(ids.astype(np.float64).sum()))
    mass_sum = grp.sum_over_ranks(float(sm.sum()))

    grp.barrier()
    t0 = time.perf_counter()
    time.sleep(0.05 * (grp.rank + 1))  # rank 1 is the slow one: the reported time must be its
    mine = time.perf_counter() - t0
    slowest = grp.max_over_ranks(mine)
    grp.barrier()

    with open(os.path.join(outdir, f"rank{grp.rank}.json"), "w") as f:
        json.dump({"rank": grp.rank, "world": grp.world, "blob": blob.hex(), "count": len(ids),
                   "counts": counts, "ids_ascending": bool(np.all(np.diff(ids.astype(np.int64)) > 0)),
                   "in_strip": in_strip, "first_row": r0.value, "rows": rows.value, "id_sum": id_sum,
                   "mass_sum": mass_sum, "total_mass": float(np.sum(m)), "mine": mine, "slowest": slowest,
                   "first_ids": ids[:5].tolist()}, f)
    grp.close()


if __name__ == "__main__":
    main()
