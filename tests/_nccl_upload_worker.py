"""Worker of tests/test_gpu_multi.py::test_collective_strip_upload*: one process per GPU under
torch.distributed.run (NCCL).  Every rank generates the seeded particles on the host, keeps its strip
and hands it over with parsim_upload_strip (collective); mode "ok" runs and prints the result, mode
"fail" lets rank 1 pass a NULL id array -- every rank must come back with an error, nobody may hang."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from particlesimulation_b200 import ParticleSimulation, _lib, strips  # noqa: E402
from particlesimulation_b200.simulation import ParsimError, init_particles  # noqa: E402


def main():
    outdir, mode = sys.argv[1], sys.argv[2]
    seed, side, nc, n, steps = int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), int(sys.argv[7])
    grp = strips.RankGroup(backend="nccl")
    lib = _lib.load()
    blob = None
    if grp.rank == 0:
        import ctypes as C
        buf = (C.c_ubyte * 128)()
        assert lib.parsim_nccl_unique_id(buf) == 0
        blob = bytes(buf)
    blob = grp.broadcast_bytes(blob)
    sim = ParticleSimulation(seed, side, nc, n, rank=grp.rank, nranks=grp.world, device=grp.local, nccl_id=blob, upload=False)
    x, y, vx, vy, m = init_particles(seed, side, nc, n)
    ids, sx, sy, svx, svy, sm = strips.select_strip(grp.rank, grp.world, side, nc, x, y, vx, vy, m)
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (sx, sy, svx, svy, sm)]
    out = {"rank": grp.rank, "world": grp.world, "count": int(len(ids))}
    try:
        id_ptr = 0 if (mode == "fail" and grp.rank == 1) else ids.ctypes.data
        sim.upload_whole_strip_ptrs(len(ids), id_ptr, [a.ctypes.data for a in arrs])
        sim.simulate(steps)
        sim.upload_whole_strip_ptrs(len(ids), ids.ctypes.data, [a.ctypes.data for a in arrs])  # and once more from the start
        sim.simulate(steps)
        out["result"] = list(sim.result())
    except ParsimError as e:
        out["error"] = str(e)
    with open(os.path.join(outdir, f"rank{grp.rank}.json"), "w") as f:
        json.dump(out, f)
    if "error" not in out:
        sim.close()
        grp.close()
    else:
        os._exit(0)  # a failed group does not tear its communicator down collectively


if __name__ == "__main__":
    main()
