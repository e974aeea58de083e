#!/usr/bin/env python3
"""bench.py -- particle-steps/s of the parsim step loop on N B200s (BASELINE.json's metric).

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's own CPU code on the host cores

A bench "step" is ONE complete run of the workload: `n_step` simulation timesteps from the seeded
initial condition (the reference times exactly that: serial/parsim.cpp:475-477).  At N=1 the
workload is BASELINE.json configs[2] (seed=7 side=100 ncside=1000 n_part=1e8 n_step=10, the
HBM-roofline run; configs[1] is selectable with --workload config2).  At N>1 the same density
is weak-scaled: n_part = N*1e8, ncside = round(1000*sqrt(N)), side = ncside/10, row strips.

At N=1 the default line also carries `extra`: one short device-timed sub-line for each of the other
BASELINE configurations (configs[1], configs[3]'s density at 1e8 particles, configs[4]) with their
collision counts and pins; at N=2/4/8 `extra.config4` is the member of SURVEY 8(d)'s config-4
series for that GPU count (at 8 GPUs BASELINE configs[3] itself).  Expected collision counts come
from tests/golden/bench_pins.json.

value  : device-timed, initial state already resident in HBM (parsim_reset + parsim_run)
e2e    : host-timed through the C ABI with HOST buffers: parsim_upload from pinned memory,
         parsim_run, parsim_result -- host<->device copies inside the timed region
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (seed, side, ncside, n_part, n_step)
    "config3": (7, 100.0, 1000, 100_000_000, 10),   # BASELINE configs[2]
    "config2": (1, 10.0, 100, 1_000_000, 20),       # BASELINE configs[1]
    "config4w": (3, 354.0, 3536, 100_000_000, 10),  # BASELINE configs[3]'s density (8 per cell), 1e8 per GPU
    "config5": (5, 1.0, 20, 10_000_000, 50),        # BASELINE configs[4]
}
# SURVEY 8(d)'s weak-scaling series of config 4's density; at 8 GPUs it IS BASELINE configs[3]
CONFIG4_SERIES = {1: (354.0, 3536, 100_000_000), 2: (500.0, 5000, 200_000_000), 4: (707.0, 7071, 400_000_000),
                  8: (1000.0, 10000, 800_000_000)}


def pinned_collisions(seed, side, nc, n, steps):
    """Expected collision count of a full-size run and where it comes from
    (tests/golden/bench_pins.json: the unmodified serial reference where it fits the build
    container's RAM and patience, else a run of this library on fewer GPUs)."""
    try:
        pins = json.load(open(os.path.join(ROOT, "tests", "golden", "bench_pins.json")))
    except Exception:
        return None, None
    e = pins.get(f"{seed} {side:g} {nc} {n} {steps}")
    return (e["collisions"], e["source"]) if e else (None, None)
# bounded CPU sample of the same density and cell size (config 3: ppc = 100, cell = 0.1, 1/16 of the
# particles: 10-30 s on one host core, a few seconds with every core under OpenMP)
CPU_SAMPLE = {"config3": "7 25 250 6250000 10", "config2": "1 10 100 1000000 20",
              "config4w": "3 88.4 884 6250000 10", "config5": "5 1 20 100000 50"}


def workload_for(name, n_gpus, scale):
    seed, side, nc, n, steps = WORKLOADS[name]
    if scale != 1.0:  # smoke-test sizes: same density, fewer cells
        f = math.sqrt(scale)
        nc = max(3, round(nc * f))
        side = side / WORKLOADS[name][2] * nc
        n = max(10, int(n * scale))
    if n_gpus > 1 and name == "config4w" and scale == 1.0 and n_gpus in CONFIG4_SERIES:
        side, nc, n = CONFIG4_SERIES[n_gpus]
    elif n_gpus > 1:
        cellw = side / nc
        nc = round(nc * math.sqrt(n_gpus))
        side = nc * cellw
        n = n * n_gpus
    return seed, side, nc, n, steps, pinned_collisions(seed, side, nc, n, steps)[0]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        """The timed region starts here: earlier samples (warm-up) are dropped if later ones exist."""
        self.first = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        rows = self.rows[getattr(self, "first", 0):] or self.rows
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        load = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm
        load.sort()
        return {"sm_mhz": load[len(load) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


def run_reference_binary(binary, sample, threads, notes=None):
    """One run of a reference binary.  The OpenMP build races on its own lists and
    sometimes dies with SIGSEGV: a crashed run is repeated up to 3 times, then with half the
    threads, and what happened is written into `notes`."""
    tries = 0
    while True:
        env = dict(os.environ, OMP_NUM_THREADS=str(threads))
        t = time.time()
        p = subprocess.run([binary] + sample.split(), capture_output=True, text=True, env=env)
        wall = time.time() - t
        if p.returncode == 0 and len(p.stdout.split()) >= 3:
            break
        tries += 1
        if notes is not None:
            notes.append(f"{os.path.basename(binary)} with {threads} threads exited {p.returncode} (its own race); retried")
        if tries % 3 == 0:
            if threads == 1:
                raise SystemExit(f"{binary} {sample} fails even with one thread: {p.stderr[-300:]}")
            threads = max(1, threads // 2)
    try:
        secs = float(p.stderr.strip().splitlines()[-1].rstrip("s"))  # the reference's own timer
        if secs < 2.0:
            secs = wall  # "%.1lfs" is too coarse for short runs
    except Exception:
        secs = wall
    n, steps = int(sample.split()[3]), int(sample.split()[4])
    return n * steps / secs, secs, p.stdout.split()[-3:], threads  # x0 y0 collisions (the OpenMP build also prints PANIC lines)


def cpu_baseline(workload, threads=1):
    """The reference's own CPU code on this box's host cores, on a bounded sample of the workload
    (same particles-per-cell and cell size, 1/16 of the particles: 10-30 s on one core)."""
    sample = CPU_SAMPLE[workload]
    ref = os.path.join(ROOT, "oracle", "_ref", "parsim-omp" if threads > 1 else "parsim")
    kind = "reference"
    if not os.path.exists(ref):
        ref, kind, threads = os.path.join(ROOT, "oracle", "parsim_oracle"), "port", 1
        if not os.path.exists(ref):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    notes = []
    v, secs, out, threads = run_reference_binary(ref, sample, threads, notes)
    return {"value": v, "unit": "particle-steps/s", "cores": threads, "kind": kind, "same_config": False,
            **({"notes": notes} if notes else {}),
            "sample": f"{os.path.basename(ref)} {sample} ({secs:.1f} s; same particles per cell and cell size as the "
                      f"workload, {int(sample.split()[3]) / WORKLOADS[workload][3]:.4g} of its particles)",
            "stdout": out}


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    seed, side, nc, n, nstep, _ = workload_for(a.workload, a.gpus, a.scale)
    sample = CPU_SAMPLE[a.workload].split()
    work = int(sample[3]) * int(sample[4])
    secs = []
    base = None
    for i in range(a.warmup + a.steps):
        base = cpu_baseline(a.workload, threads)
        if i >= a.warmup:
            secs.append(work / base["value"])
    ms = 1e3 * sum(secs) / len(secs)
    value = work / (ms / 1e3)
    base["value"] = value
    print(json.dumps({
        "impl": "reference", "metric": "particle-steps/s", "value": value, "unit": "particle-steps/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded init_particles)",
        "config": {"workload": f"{a.workload}: seed={seed} side={side:g} ncside={nc} n_part={n} n_step={nstep}",
                   "measured_on": "bounded CPU sample " + CPU_SAMPLE[a.workload]},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "particle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))



def main_ours(a):
    import ctypes as C

    import numpy as np
    import torch

    from particlesimulation_b200 import ParticleSimulation, _lib
    from particlesimulation_b200.strips import RankGroup

    # stdout carries ONE JSON line: NCCL's banner / warnings (stdout by default) go to stderr, and
    # fd 1 itself points at stderr until the line is printed, whatever else a library writes
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"  # NCCL_DEBUG_FILE is honoured only above the VERSION level
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    grp = RankGroup(backend="nccl", expect_world=a.gpus)  # one process per GPU (tests: gloo, tests/test_multi_cpu.py)
    rank, world, local = grp.rank, grp.world, grp.local
    nccl_id = None
    if world > 1:
        blob = None
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            if _lib.load().parsim_nccl_unique_id(buf):
                raise SystemExit("parsim_nccl_unique_id failed")
            blob = bytes(buf)
        nccl_id = grp.broadcast_bytes(blob)
    barrier, max_over_ranks = grp.barrier, grp.max_over_ranks

    seed, side, nc, n, nstep, expect = workload_for(a.workload, a.gpus, a.scale)
    kw = dict(rank=rank, nranks=world, device=local, nccl_id=nccl_id) if world > 1 else {}
    sim = ParticleSimulation(seed, side, nc, n, device_init=True, options={"keep_initial": 1}, **kw)
    # host copies of this rank's inputs in pinned memory, for the end-to-end leg (untimed)
    ids, x, y, vx, vy, m = sim.download()
    cnt = len(ids)
    pinned = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for v in (x, y, vx, vy, m)]
    pinned_ids = torch.from_numpy(ids.astype(np.int64)).to(torch.int32).pin_memory() if world > 1 else None
    if world == 1:
        assert np.array_equal(ids, np.arange(n, dtype=np.uint32))  # upload order = id order
    del ids, x, y, vx, vy, m

    stream = torch.cuda.ExternalStream(sim.lib.parsim_stream(sim._ctx), device=torch.device("cuda", local))

    def one_run():
        sim.reset()
        sim.simulate(nstep)

    # ---- device-resident leg: `value` -----------------------------------------------------
    clocks = ClockSampler(local)
    clocks.start()  # before the warm-up: nvidia-smi takes a while to come up on an 8-GPU box
    for _ in range(a.warmup):
        one_run()
    barrier()
    clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(a.steps):
        one_run()
    e1.record(stream)
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clk = clocks.stop()
    x0, y0, collisions = sim.result()
    # the per-kernel table comes from a second, instrumented pass over the same K runs (CUDA events
    # around every launch and the pair counters cost a little, so they stay out of `value`)
    sim.set_option("stats", 1)
    sim.stats_reset()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record(stream)
    for _ in range(a.steps):
        one_run()
    s1.record(stream)
    barrier()
    ms_stats = max_over_ranks(s0.elapsed_time(s1))
    st = sim.stats()
    assert sim.result()[2] == collisions, "instrumented pass disagrees with the timed pass"
    sim.set_option("stats", 0)
    work = float(n) * nstep
    value = work * a.steps / (ms_total / 1e3)

    # ---- end-to-end leg: host buffers through the C ABI -------------------------------------
    upload_s = [0.0]

    def one_e2e():
        t_up = time.perf_counter()
        if world == 1:
            sim.upload_ptrs([t.data_ptr() for t in pinned])
        else:
            sim.upload_whole_strip_ptrs(cnt, pinned_ids.data_ptr(), [t.data_ptr() for t in pinned])  # collective
        upload_s[0] += time.perf_counter() - t_up  # parsim_upload returns when the records are on the device
        sim.simulate(nstep)
        return sim.result()

    one_e2e()
    barrier()
    upload_s[0] = 0.0
    t0 = time.perf_counter()
    for _ in range(a.steps):
        res = one_e2e()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = work * a.steps / e2e_s
    h2d_bytes = 40 * cnt + (4 * cnt if world > 1 else 0)
    h2d_gbs_slowest = h2d_bytes * a.steps / max_over_ranks(upload_s[0]) / 1e9
    # the platform's ceiling for that copy: the same pinned buffers through plain cudaMemcpyAsync into
    # scratch device memory, every rank at once, nothing else running (host DRAM / PCIe root complexes
    # are shared by the ranks of a node)
    scratch = [torch.empty(t.shape, dtype=t.dtype, device="cuda") for t in pinned]
    raw_s = None
    for rep in range(3):
        torch.cuda.synchronize()
        barrier()
        t_raw = time.perf_counter()
        for d_t, h_t in zip(scratch, pinned):
            d_t.copy_(h_t, non_blocking=True)
        torch.cuda.synchronize()
        dt_raw = time.perf_counter() - t_raw
        if rep:
            raw_s = dt_raw if raw_s is None else min(raw_s, dt_raw)
    h2d_raw_gbs_slowest = 40 * cnt / max_over_ranks(raw_s) / 1e9
    del scratch
    # same particles, same arithmetic: the uploaded run must end where the device-resident one did, bit for bit
    assert tuple(res) == (x0, y0, collisions), f"end-to-end run {tuple(res)} disagrees with the device-resident run {(x0, y0, collisions)}"

    # ---- the other BASELINE configurations, one short sub-line each (same code, same library) -----
    def side_run(label, tup, reps=3, pin_note=None):
        """Device-timed runs of another configuration on this rank set; returns the sub-line."""
        sd, sside, snc, sn, sstep = tup
        ids_kw = {}
        if world > 1:
            blob2 = None
            if rank == 0:
                buf2 = (C.c_ubyte * 128)()
                if _lib.load().parsim_nccl_unique_id(buf2):
                    raise SystemExit("parsim_nccl_unique_id failed")
                blob2 = bytes(buf2)
            ids_kw = dict(rank=rank, nranks=world, device=local, nccl_id=grp.broadcast_bytes(blob2))
        sim2 = ParticleSimulation(sd, sside, snc, sn, device_init=True, options={"keep_initial": 1}, **ids_kw)
        st2 = torch.cuda.ExternalStream(sim2.lib.parsim_stream(sim2._ctx), device=torch.device("cuda", local))
        sim2.simulate(sstep)  # warm-up run
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(st2)
        for _ in range(reps):
            sim2.reset()
            sim2.simulate(sstep)
        a1.record(st2)
        barrier()
        ms2 = max_over_ranks(a0.elapsed_time(a1)) / reps
        rx, ry, rc = sim2.result()
        sim2.close()
        exp, src = pinned_collisions(sd, sside, snc, sn, sstep)
        return {"workload": f"{label}: seed={sd} side={sside:g} ncside={snc} n_part={sn} n_step={sstep}",
                "ms_per_run": ms2, "value": float(sn) * sstep / (ms2 / 1e3), "unit": "particle-steps/s",
                "timing": f"CUDA events around {reps} reset+run repetitions, no instrumentation",
                "result": {"x0": round(rx, 3), "y0": round(ry, 3), "collisions": rc, "reference_collisions": exp,
                           "parity_ok": (rc == exp) if exp is not None else None,
                           "pin": src if exp is not None else (pin_note or "no pinned count for this tuple")}}

    # FP64 ceilings measured on this box (MEASURED_PEAKS.json has none): independent FMA chains, and the
    # force loop's own 13-instruction pair mix on register operands
    fp64_peak = sim.fp64_peak_tflops() if rank == 0 else None
    pair_peak = sim.pair_peak() if rank == 0 else None
    extra = {}
    if a.workload == "config3" and a.scale == 1.0 and not a.no_extras:
        sim.close()  # free the 24 GB of the main workload first
        if world == 1:
            extra["config2"] = side_run("BASELINE configs[1]", WORKLOADS["config2"], reps=5)
            extra["config4_density"] = side_run("BASELINE configs[3] density (8 per cell) at 1e8 particles on one GPU",
                                                WORKLOADS["config4w"])
            extra["config5"] = side_run(
                "BASELINE configs[4]", WORKLOADS["config5"],
                pin_note="unavailable: the unmodified serial reference dies with SIGSEGV at this size (a close "
                         "encounter throws a particle more than a box length, serial/parsim.cpp:171-177,:276-280); "
                         "pinned at 1/10 and 1/100 of the particles in tests/golden/reference_runs.jsonl")
        elif world in CONFIG4_SERIES:
            sside, snc, sn = CONFIG4_SERIES[world]
            extra["config4"] = side_run(
                "BASELINE configs[3]" + ("" if world == 8 else f" weak-scaling series member for {world} GPUs"),
                (3, sside, snc, sn, 10))
        sim = None

    if rank != 0:
        return
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    # ---- roofline of the dominant kernel + the per-kernel table -----------------------------
    hbm_peak, peak_src = peaks()
    kern = {k: v for k, v in st["kernels"].items() if v["launches"] > 0}
    dom = max(kern, key=lambda k: kern[k]["ms"])
    table = {}
    for k, v in kern.items():
        gbs = v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else 0.0
        table[k] = {"ms_per_step": v["ms"] / a.steps, "launches_per_step": v["launches"] / a.steps,
                    "algorithmic_GB_per_step": v["bytes"] / a.steps / 1e9, "achieved_GBps": gbs,
                    "hbm_frac": gbs / hbm_peak}
    d = kern[dom]
    achieved = d["bytes"] / (d["ms"] * 1e-3) / 1e9
    force_flops = 17.0 * st["force_pairs"] / 2 + 120.0 * st["alive_particle_steps"]
    coll_flops = 7.0 * st["collision_pairs"] / 2
    f_ms, c_ms = kern["force_integrate"]["ms"], kern["collide_find"]["ms"]
    # interactions the force kernel evaluates: ordered in-cell pairs + 8 centres of mass per alive particle
    interactions = st["force_pairs"] + 8.0 * st["alive_particle_steps"]
    pair_rate = interactions / (f_ms * 1e-3)
    fp64 = {"peak_tflops_measured_fma_chain": fp64_peak, "pair_mix_peak_per_s": pair_peak,
            "peaks_source": "measured live on this GPU by parsim_fp64_peak / parsim_pair_peak (the stated fallback "
                            "for peaks MEASURED_PEAKS.json does not hold)",
            "force_integrate": {"algorithmic_tflops": force_flops / (f_ms * 1e-3) / 1e12,
                                "frac": force_flops / (f_ms * 1e-3) / 1e12 / fp64_peak,
                                "interactions_per_s": pair_rate, "frac_of_pair_mix_peak": pair_rate / pair_peak,
                                "ordered_pairs_per_step": st["force_pairs"] / a.steps},
            "collide_find": {"algorithmic_tflops": coll_flops / (c_ms * 1e-3) / 1e12,
                             "frac": coll_flops / (c_ms * 1e-3) / 1e12 / fp64_peak,
                             "ordered_pairs_per_step": st["collision_pairs"] / a.steps}}
    # ncu-measured DRAM bytes per launch of the dominant kernel (one `--set full` capture of this
    # workload at N=1, profiles/traffic.json); null for any other workload / scale / GPU count
    traffic, traffic_info = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp) and world == 1 and a.scale == 1.0:
        traffic_info = json.load(open(tp)).get(a.workload, {}).get(dom)
        if traffic_info:
            traffic = traffic_info["dram_bytes_per_launch"] / 1e9  # GB per launch, like `achieved` x time
    hbm_part = {"achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "traffic_unit": "GB per launch (ncu dram read+write)",
                "algorithmic_GB_per_launch": d["bytes"] / max(d["launches"], 1) / 1e9, "ncu": traffic_info,
                "peak_source": peak_src}
    timing_note = ("CUDA events around every launch, on the launching stream, in a second pass over the same K runs "
                   "(ms_per_step_instrumented); the timed pass behind `value` runs without them")
    if dom == "force_integrate" and pair_rate / pair_peak > achieved / hbm_peak:
        # the in-cell pair loop decides: report the kernel against the FP64 roof, HBM beside it
        roofline = {"kernel": dom, "bound": "fp64", "achieved": pair_rate / 1e9, "peak": pair_peak / 1e9,
                    "unit": "G pair interactions/s (13 FP64 + 1 MUFU each)", "frac": pair_rate / pair_peak,
                    "traffic": traffic, "peak_source": "parsim_pair_peak measured live (not in MEASURED_PEAKS.json)",
                    "hbm": hbm_part, "timing": timing_note}
    else:
        roofline = {"kernel": dom, "bound": "hbm", **hbm_part, "timing": timing_note}
    wl_note = ""
    if world > 1:
        wl_note = (" (BASELINE configs[3] density weak-scaling series, SURVEY 8(d)" + ("; at 8 GPUs BASELINE configs[3] itself)" if world == 8 else ")")
                   if a.workload == "config4w" else " (BASELINE configs[2] density weak-scaled by row strips)")
    out = {
        "metric": "particle-steps/s", "value": value, "unit": "particle-steps/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_total / a.steps,
        "ms_per_step_instrumented": ms_stats / a.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": f"synthetic (seeded init_particles, seed {seed}, generated on the device)",
        "config": {"workload": f"{a.workload}: seed={seed} side={side:g} ncside={nc} n_part={n} n_step={nstep}" + wl_note,
                   "step": f"one full run of {nstep} timesteps from the initial condition",
                   "l2": f"inputs ({44e-9 * n / world:.2g} GB/GPU) " + ("are larger than L2; no flush needed" if 44 * n / world > 2.6e8
                                                                        else "fit L2: a small latency-bound case, not a bandwidth measurement"),
                   "parallelism": f"row strips x{world}" if world > 1 else "single GPU"},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "particle-steps/s", "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": 24, "ms_per_step": 1e3 * e2e_s / a.steps,
                "h2d_GBps_slowest_rank": h2d_gbs_slowest,
                "h2d_raw_GBps_slowest_rank": h2d_raw_gbs_slowest,
                "h2d_raw_note": "plain cudaMemcpyAsync of the same pinned buffers, all ranks at once, best of two: what the host "
                                "delivers to concurrent copies at that moment (a reference point for the upload rate, not a bound)",
                "inputs": "pinned host arrays of this rank's strip (x, y, vx, vy, m" + (", id" if world > 1 else "") + ")",
                "upload": ("parsim_upload_strip, collective" if world > 1 else "parsim_upload") +
                          ": x, y, m travel first, vx, vy last; binning, centres of mass and the first timestep's force "
                          "sums run under the velocity copy (PARSIM_VEL_LATE=0 switches that off)"},
        "gpu_launches": int(sum(v["launches"] for v in kern.values())),
        "roofline": roofline,
        "fp64": fp64, "kernels": table,
        "result": {"x0": round(x0, 3), "y0": round(y0, 3), "collisions": collisions,
                   "reference_collisions": expect, "parity_ok": (collisions == expect) if expect is not None else None,
                   "pin": pinned_collisions(seed, side, nc, n, nstep)[1]},
        "counters": {k: st[k] / a.steps for k in ("movers", "strays", "deaths", "migrants", "alive_particle_steps")},
    }
    if extra:
        out["extra"] = extra
    if world == 1 and not a.no_cpu_baseline:
        # north_star: the reference's serial, OpenMP and MPI builds on this box's host cores, same run
        cb = cpu_baseline(a.workload, 1)
        cb.pop("stdout", None)
        out["cpu_baseline"] = cb
        omp = cpu_baseline(a.workload, os.cpu_count() or 1)
        omp.pop("stdout", None)
        out["cpu_baseline_openmp"] = omp
        out["cpu_baseline_mpi"] = "unavailable on this box (no mpic++ / mpi.h / mpirun in the image)"
    print(json.dumps(out))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (smoke tests only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the sub-lines of the other BASELINE configurations")
    a = ap.parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
