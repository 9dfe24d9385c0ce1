#!/usr/bin/env python
"""bench.py -- headline benchmark of the force path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (configs[1]): direct all-pairs, N = 1,048,576, Leapfrog (Verlet_DKD), FP32, bodies
target-sharded over the GPUs with one NCCL all-gather of positions per force evaluation.  One "step" is
one Leapfrog step = half drift, all-gather, one force evaluation (N^2 interactions in total), kick + half
drift.  `value` = interactions per second of the whole job, inputs resident in HBM.  The same line carries
the Barnes-Hut metric (steps/s, N = 1,048,576, theta = 0.5, configs[2]) under "bh".

--impl reference: the reference has no CPU implementation (it is GPU-only), so this
arm times the OpenMP port of its loops (oracle/nbody_oracle.c) on the host cores, on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

N_BODIES = 1 << 20
EPS = 1.01
DT = 0.001
THETA = 0.5
FLOP_PER_INTERACTION = {3: 20.0, 2: 14.0}   # SURVEY.md section 8d / BASELINE.md section 4
SM_COUNT, FP32_LANES, MAX_MHZ = 148, 128, 1965.0


def shared_config(n, dim, world):
    """The `config` object both arms print (identical keys and values, so that the driver can match the lines)."""
    return {"workload": f"direct all-pairs N={n} uniform disk (R{dim}), Leapfrog (Verlet_DKD) step",
            "n": n, "softening": EPS, "dt": DT,
            "l2": "GPU arm: flushed between timed iterations (256 MB write); host arm: each sample streams all sources from DRAM",
            "parallelism": f"target-slice x{world}: targets sharded over {world} GPU(s), one NCCL all-gather of positions per "
                           f"step; the host arm runs on rank 0's cores only"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])), mx.append(float(r[2])), pw.append(float(r[3]))
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def run_reference_arm(args):
    """Host-core arm: OpenMP port of the reference loops (kind 'port': the reference itself is GPU-only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = graft.load_package()
    oracle = graft.load_oracle()
    # torchrun exports OMP_NUM_THREADS=1; the arm is meant to use every host core it can
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    pos = pkg.ics.uniform_disk(N_BODIES, seed=1, z_half_height=3000.0)
    sample_targets = 8192
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        oracle.direct_f32(pos, 3, EPS, t0=(it * sample_targets) % (N_BODIES - sample_targets), nt=sample_targets,
                          fast=True, threads=cores)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
    inter = float(sample_targets) * N_BODIES
    val = inter * len(times) / sum(times)
    sample = f"{sample_targets} targets x {N_BODIES} sources per step (1/128 of one force evaluation), OpenMP, FP32"
    line = {"impl": "reference", "metric": "interactions/s (direct, N=1M)", "value": val, "unit": "interactions/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": shared_config(N_BODIES, 3, args.gpus),
            "sampling": "each step evaluates a bounded slice of the targets against all sources (see cpu_baseline.sample)",
            "cpu_baseline": {"value": val, "unit": "interactions/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference is GPU-only; this arm is the OpenMP transcription of its direct loop (oracle/)"}
    emit(line)


def ref_bh_worker(out_path):
    """Child process of the bench: the reference's own Barnes-Hut kernels at the benchmark size, four evaluations."""
    import ctypes as C
    import torch
    pkg = graft.load_package()
    nb = N_BODIES
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libnbody_ref.so"))
    vp = C.c_void_p
    ref.ref_barnes_hut.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, C.c_float]
    dev = torch.device("cuda", 0)
    m = 3 * nb
    rp = torch.zeros(m, 4, device=dev)
    rp[:nb] = torch.from_numpy(pkg.ics.uniform_disk(nb, seed=1)).to(dev)
    ra = torch.zeros(nb, 4, device=dev)
    fmax = float(np.finfo(np.float32).max)
    rb = torch.tensor([fmax, fmax, -fmax, -fmax], device=dev)
    ri = torch.zeros(1, dtype=torch.int32, device=dev)
    rn = torch.zeros(4 * m, dtype=torch.int32, device=dev)
    rs = torch.zeros(nb, dtype=torch.int32, device=dev)
    rk = torch.zeros(nb, dtype=torch.int32, device=dev)
    tms = []
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ref.ref_barnes_hut(rp.data_ptr(), ra.data_ptr(), rb.data_ptr(), ri.data_ptr(), rn.data_ptr(), rs.data_ptr(),
                           rk.data_ptr(), nb, m, THETA, EPS)   # synchronises itself
        tms.append((time.perf_counter() - t0) * 1e3)
    np.save(out_path, ra[:, :2].cpu().numpy())
    print(json.dumps({"ms": tms}), flush=True)


_RESULT_FD = 1


def emit(line):
    """The ONE JSON line of the run, on the process's original stdout."""
    sys.stdout.flush()
    os.write(_RESULT_FD, (json.dumps(line) + "\n").encode())


def main():
    if len(sys.argv) == 3 and sys.argv[1] == "--ref-bh-worker":
        return ref_bh_worker(sys.argv[2])
    # libraries write to fd 1 too (NCCL's version banner): everything but the result line goes to stderr
    global _RESULT_FD
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=N_BODIES, help="bodies (parity / tuning runs only; default = headline)")
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--skip-bh", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-refcuda", action="store_true")
    ap.add_argument("--skip-calib", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = graft.load_package()
    n, dim = args.n, args.dim
    eng = pkg.Engine(local, n)
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- state: replicated positions, slice-local velocities / accelerations --------------------------
    pos_h = pkg.ics.uniform_disk(n, seed=1, z_half_height=3000.0 if dim == 3 else 0.0)
    begin, end = pkg.shard.slice_bounds(n, world, rank)
    count = end - begin
    pos = torch.from_numpy(pos_h).to(dev)
    acc = torch.zeros(count, 4, device=dev)
    vel = torch.zeros(count, 4, device=dev)
    def gather_positions():
        pkg.shard.gather_positions(pos, n, world, rank, dist)     # one in-place NCCL all-gather over NVLink

    def force():
        eng.direct_accel_slice(pos, begin, count, acc, dim, EPS)

    # initial velocities: tangential, from one force evaluation (DiskModel, src/InitialConditions.cu:160-192)
    force()
    torch.cuda.synchronize()
    vel.copy_(torch.from_numpy(pkg.ics.disk_velocities(pos_h[begin:end], acc.cpu().numpy(), space_r3=(dim == 3))))
    launches_force = eng.direct_last_launch()["launches"]

    def step():
        eng.verlet_dkd_drift(pos, vel, begin, count, DT)
        gather_positions()
        force()
        eng.verlet_dkd_kick_drift(pos, vel, acc, begin, count, DT)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in ev:
            flush.fill_(1)          # L2 flush between timed iterations (outside the event pair)
            a.record()
            fn()
            b.record()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    total_ms = timed(step, args.steps, args.warmup)
    # kernel-only time of the dominant kernel (the force evaluation of this rank's slice), same stream
    kern_ms = timed(force, args.steps, 1) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    interactions = float(n) * float(n)
    value = interactions * args.steps / (total_ms * 1e-3)

    # ---- end to end through the host-buffer C-ABI call ---------------------------------------------------
    pos_pin = torch.from_numpy(pos_h).pin_memory()
    acc_pin = torch.empty(n, 4).pin_memory()

    def e2e_step():
        # every rank uploads all positions, evaluates, downloads; the C ABI call synchronises
        eng.direct_accel_host(pos_pin, dim, EPS, out=acc_pin)

    if world == 1:
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        e2e = {"value": interactions * args.steps / (e2e_ms * 1e-3), "unit": "interactions/s",
               "h2d_bytes_per_step": n * 16, "d2h_bytes_per_step": n * 16,
               "call": "nb_direct_accel_host (pinned host float4[n] in, float4[n] out)"}
    else:
        # sharded end-to-end: upload the full positions, evaluate the rank's slice, download the slice
        acc_slice_pin = torch.empty(count, 4).pin_memory()

        def e2e_sharded():
            pos.copy_(pos_pin, non_blocking=True)
            force()
            acc_slice_pin.copy_(acc, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        for _ in range(2):
            e2e_sharded()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_sharded()
        barrier()
        t = torch.tensor([(time.perf_counter() - t0) * 1e3], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": interactions * args.steps / (float(t.item()) * 1e-3), "unit": "interactions/s",
               "h2d_bytes_per_step": n * 16, "d2h_bytes_per_step": count * 16,
               "call": "pinned H2D of all positions + nb_direct_accel_slice + pinned D2H of the slice, per rank"}

    # ---- roofline of the dominant kernel -----------------------------------------------------------------
    flop = FLOP_PER_INTERACTION[dim]
    slice_inter = float(count) * float(n)
    achieved = slice_inter * flop / (kern_ms * 1e-3) / 1e12
    peak = SM_COUNT * FP32_LANES * 2 * MAX_MHZ * 1e6 / 1e12
    roofline = {"bound": "fp32", "kernel": f"direct_kernel<{dim}D> (+pack, +partial reduce)", "achieved": achieved,
                "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                "peak_source": "148 SMs x 128 FP32 lanes x 2 x 1965 MHz (clocks.max.sm); MEASURED_PEAKS.json has no FP32 entry",
                "kernel_ms": kern_ms, "flop_per_interaction": flop,
                "fp32_pipe_ops_per_interaction": 12 if dim == 3 else 8,
                "launch": eng.direct_last_launch()}
    # the same kernel with the mass kept inside the sum for every tile (what any non-uniform mass distribution runs):
    # north_star's 70 % target is judged on this number; `frac` above profits from the disc's equal masses
    gen_ms = timed(lambda: eng.direct_accel_slice(pos, begin, count, acc, dim, EPS, pkg.NB_DIRECT_NO_HOIST), 2, 1) / 2
    roofline["kernel_ms_general_masses"] = gen_ms
    roofline["frac_general_masses"] = slice_inter * flop / (gen_ms * 1e-3) / 1e12 / peak
    tiles = pos_h[: (n // 1024) * 1024, 3].reshape(-1, 1024)
    roofline["hoisted_tile_fraction"] = float((tiles == pos_h[n // 2, 3]).all(axis=1).sum()) / max((n + 1023) // 1024, 1)
    try:   # DRAM traffic of the dominant kernel from the committed ncu --set full capture (per launch)
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02", "traffic.json")))
        if n == N_BODIES and dim == 3 and world == 1:
            roofline["traffic"] = tr["direct_kernel_n1048576"]["traffic_bytes"]
            roofline["traffic_source"] = tr["direct_kernel_n1048576"]["source"]
    except Exception:
        tr = {}
    if rank == 0 and clocks and clocks.get("sm_mhz"):
        roofline["frac_at_observed_clock"] = achieved / (SM_COUNT * FP32_LANES * 2 * clocks["sm_mhz"] * 1e6 / 1e12)

    extra = {}
    if rank == 0 and not args.skip_calib:
        try:
            cal = eng.calibrate()
            roofline["calibrated"] = cal
            roofline["frac_of_calibrated_ffma"] = achieved / max(cal["ffma_tflops"], cal["ffma2_tflops"])
        except Exception as e:  # calibration is informative only
            roofline["calibrated"] = str(e)

    # ---- the same workload in Space R2 (the reference's main.cu space; 2D force law m d/(eps+r^2), 14 flop) ----------
    if dim == 3 and not args.skip_bh:
        pos2d = torch.from_numpy(pkg.ics.uniform_disk(n, seed=1)).to(dev)
        ms2d = timed(lambda: eng.direct_accel_slice(pos2d, begin, count, acc, 2, EPS), 2, 1) / 2
        ach2d = float(count) * float(n) * FLOP_PER_INTERACTION[2] / (ms2d * 1e-3) / 1e12
        extra["direct_2d"] = {"metric": "interactions/s (direct 2D, N=%d)" % n, "value": interactions / (ms2d * 1e-3),
                              "kernel_ms": ms2d,
                              "roofline_frac_fp32": ach2d / peak,
                              "note": "2D is co-limited by the MUFU pipe (one rsqrt per 8 FP32 ops)"}
        del pos2d

    # ---- the reference's own CUDA kernel on the same box (informative; north_star parity target) ---------
    if rank == 0 and world == 1 and not args.skip_refcuda:
        so = os.path.join(ROOT, "oracle", "_ref", "libnbody_ref.so")
        if os.path.exists(so) and n % 128 == 0:
            import ctypes as C
            ref = C.CDLL(so)
            ref.ref_direct.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float]
            acc_ref = torch.zeros(n, 4, device=dev)
            ref.ref_direct(pos.data_ptr(), acc_ref.data_ptr(), n, dim, EPS)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ref.ref_direct(pos.data_ptr(), acc_ref.data_ptr(), n, dim, EPS)   # legacy default stream
            torch.cuda.synchronize()
            ref_ms = (time.perf_counter() - t0) * 1e3
            force()                                                # same positions, this engine (default mode)
            torch.cuda.synchronize()
            dd = (acc[:, :dim] - acc_ref[:, :dim]).norm(dim=1) / acc_ref[:, :dim].norm(dim=1)
            extra["reference_cuda_kernel"] = {"interactions_per_s": interactions / (ref_ms * 1e-3), "ms": ref_ms,
                                              "what": "unmodified src/DirectComputeKernels.cu built for sm_100, same box",
                                              "acc_rel_diff_median": float(dd.median()),
                                              "acc_rel_diff_p99": float(dd.quantile(0.99)) if n <= (1 << 24) else None,
                                              "acc_rel_diff_max": float(dd.max()),
                                              "acc_rel_diff_field": float((acc[:, :dim] - acc_ref[:, :dim]).norm() / acc_ref[:, :dim].norm()),
                                              "diff_note": "per body, all N bodies, default (hoisted, split) mode vs the reference "
                                                           "kernel's single FP32 accumulator per target; tests/test_direct_gpu.py::"
                                                           "test_direct_headline_size holds the FP64 comparison of both"}
            del acc_ref

    # ---- Barnes-Hut metric (configs[2]: N = 1,048,576, theta = 0.5, Leapfrog) -------------------------------------------
    # One step = one complete Verlet_DKD step of the whole system: half drift of all bodies fused into the bounds kernel,
    # tree build (redundant per GPU), walk of the rank's groups with kick + half drift in its epilogue, and -- sharded --
    # ONE all-gather of the (x', y', vx', vy') rows (16 B per body).  Nothing is left out of the timed region.
    if not args.skip_bh:
        try:
            nb = N_BODIES
            p2_h = pkg.ics.uniform_disk(nb, seed=1)
            p2 = torch.from_numpy(p2_h).to(dev)
            a2 = torch.zeros(nb, 4, device=dev)
            groups = (nb + 31) // 32
            g0, g1 = pkg.shard.group_bounds(nb, world, rank)
            eng.bh_accel(p2, a2, THETA, EPS)                      # initial state: accelerations -> tangential velocities
            torch.cuda.synchronize()
            a2_init = a2.clone()
            v2 = torch.from_numpy(pkg.ics.disk_velocities(p2_h, a2.cpu().numpy())).to(dev)
            p2p = world > 1 and os.environ.get("NB_BH_P2P", "1") != "0"
            stepper = pkg.integrators.FusedBarnesHut(eng, p2, v2, "Verlet_DKD", THETA, EPS, world, rank, dist, p2p=p2p)
            p2p = stepper.p2p      # False if the ranks could not map each other's memory (they then fall back together)
            words = p2p and getattr(stepper, "_sync", None) is not None
            coll_events = []
            if world > 1:
                plain_gather = stepper._exchange

                def timed_gather():
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    plain_gather()
                    e1.record()
                    coll_events.append((e0, e1))
                stepper._exchange = timed_gather
            ksteps = max(args.steps, 10)
            bh_ms = timed(lambda: stepper.step(DT, lazy=True), ksteps, 3) / ksteps
            coll_ms = sum(e0.elapsed_time(e1) for e0, e1 in coll_events[-ksteps:]) / ksteps if coll_events else 0.0
            stepper.materialize()
            torch.cuda.synchronize()

            def bh_eval():
                eng.bh_accel_groups(p2, a2, THETA, EPS, g0, g1 - g0)
            eval_ms = timed(bh_eval, ksteps, 3) / ksteps
            # this rank's own build + walk time (timed() above returns the slowest rank's): the spread is the load imbalance
            # of equal group counts per rank
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(ksteps):
                bh_eval()
            e1.record()
            torch.cuda.synchronize()
            own = torch.tensor([e0.elapsed_time(e1) / ksteps], device=dev, dtype=torch.float64)
            per_rank = [own.clone() for _ in range(world)]
            if world > 1:
                dist.all_gather(per_rank, own)
            per_rank = [float(t.item()) for t in per_rank]
            bh = {"metric": "BH Leapfrog steps/s (N=1M, theta=0.5)", "value": 1e3 / bh_ms, "unit": "steps/s",
                  "ms_per_step": bh_ms, "n": nb, "theta": THETA,
                  "step": "Verlet_DKD: drift (fused into the bounds kernel) + build + walk with kick/drift epilogue"
                          + ((" + rows (16 B/body) stored into every rank's array by the walk epilogue over NVLink peer memory, "
                              + ("evaluations ordered by arrival words the ranks store into each other's memory (no collective)"
                                 if words else "one 4-byte all-reduce as barrier")
                              if p2p else " + all-gather of 16 B/body rows (NCCL)") if world > 1 else ""),
                  "exchange": ("peer stores in the walk epilogue + " + ("arrival words" if words else "barrier")
                               if p2p else "NCCL all-gather") if world > 1 else None,
                  "collective_ms_per_step": coll_ms, "collective_bytes_per_step": nb * 16 if world > 1 else 0,
                  "force_evaluation_ms": eval_ms, "force_evaluation_ms_per_rank_warm_l2": per_rank,
                  "force_evaluation_note": "nb_bh_accel_groups alone on this rank's groups (build + walk, no kick/drift, no collective)"}
            if rank == 0:
                eng.bh_accel_groups(p2, a2, THETA, EPS, 0, groups, flags=pkg.NB_BH_COUNT_WALK)
                cnt = eng.bh_last_counters()
                b_lin = 144.0 * nb + 32.0 * cnt["nodes"]
                b_walk = 16.0 * (cnt["opened"] + cnt["examined"])
                bh["counters"] = cnt
                if world == 1:
                    gbs = (b_lin + b_walk) / (eval_ms * 1e-3) / 1e9
                    peak = 6551.0
                    try:
                        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
                        src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
                    except Exception:
                        src = "fallback 6551 GB/s (MEASURED_PEAKS.json absent on this box)"
                    bh["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                      "traffic": tr.get("bh_walk_n1048576_theta0.5", {}).get("traffic_bytes"),
                                      "traffic_note": "DRAM bytes of the walk kernel (the dominant kernel); the tree is L2-resident",
                                      "algorithmic_bytes": b_lin + b_walk, "b_lin": b_lin,
                                      "b_walk": b_walk, "peak_source": src, "over": "force_evaluation_ms",
                                      "lane_interactions_per_s": cnt["interactions"] / (eval_ms * 1e-3)}
                    # end to end through the host-buffer entry point
                    p2_pin = torch.from_numpy(p2_h).pin_memory()
                    a2_pin = torch.empty(nb, 4).pin_memory()
                    for _ in range(2):
                        eng.bh_accel_host(p2_pin, THETA, EPS, out=a2_pin)
                    t0 = time.perf_counter()
                    for _ in range(ksteps):
                        eng.bh_accel_host(p2_pin, THETA, EPS, out=a2_pin)
                    bh["e2e"] = {"value": ksteps / (time.perf_counter() - t0), "unit": "force evaluations/s",
                                 "h2d_bytes_per_step": nb * 16, "d2h_bytes_per_step": nb * 16,
                                 "call": "nb_bh_accel_host"}
                    # theta-error of the result against the FP64 direct sum on a sample of bodies, and the host baseline
                    # (initial positions: the state the reference's kernels are run on below)
                    if not args.skip_cpu:
                        oracle = graft.load_oracle()
                        sample = np.arange(0, nb, nb // 256)[:256]
                        ref64 = np.concatenate([oracle.direct_f64(p2_h, 2, EPS, t0=int(t), nt=1) for t in sample])[:, :2]
                        ours = a2_init[torch.from_numpy(sample).to(dev)].cpu().numpy()[:, :2].astype(np.float64)
                        err = np.linalg.norm(ours - ref64, axis=1) / np.linalg.norm(ref64, axis=1)
                        bh["theta_error_vs_fp64"] = {"median": float(np.median(err)), "p99": float(np.quantile(err, 0.99)),
                                                     "max": float(err.max()), "sample": 256}
                        t0 = time.perf_counter()
                        ob = oracle.bh_accel(p2_h, THETA, EPS)
                        dt_cpu = time.perf_counter() - t0
                        oerr = np.linalg.norm(ob["acc"][sample][:, :2] - ref64, axis=1) / np.linalg.norm(ref64, axis=1)
                        bh["cpu_baseline"] = {"value": 1.0 / dt_cpu, "unit": "force evaluations/s", "cores": oracle.max_threads(), "kind": "port",
                                              "sample": "one full force evaluation at N=1048576 (serial-equivalent tree insert, OpenMP walk)",
                                              "theta_error_median": float(np.median(oerr))}
                    so = os.path.join(ROOT, "oracle", "_ref", "libnbody_ref.so")
                    if os.path.exists(so) and not args.skip_refcuda:
                        # the reference's lock-based tree build can stall on sm_100 (spin on a slot with a divergent
                        # __syncthreads, src/BarnesHutKernels.cu:296-363): run it in a child process with a time limit
                        import tempfile
                        tmp = tempfile.mkdtemp()
                        ref_out = os.path.join(tmp, "ref_acc.npy")
                        try:
                            cp = subprocess.run([sys.executable, os.path.abspath(__file__), "--ref-bh-worker", ref_out],
                                                capture_output=True, text=True, timeout=120)
                            tms = json.loads(cp.stdout.strip().splitlines()[-1])["ms"] if cp.returncode == 0 else None
                        except (subprocess.TimeoutExpired, ValueError, IndexError, KeyError):
                            tms = None
                        if tms is None:
                            bh["reference_cuda_bh"] = {"unavailable": "the reference's Barnes-Hut kernels did not finish within 120 s "
                                                                      "(lock-based tree build stalled) or failed"}
                            extra["bh"] = bh
                            raise StopIteration
                        ra = torch.zeros(nb, 4, device=dev)
                        ra[:, :2] = torch.from_numpy(np.load(ref_out)).to(dev)
                        ref_ms = min(tms[1:])
                        d = (ra[:, :2] - a2_init[:, :2]).norm(dim=1) / ra[:, :2].norm(dim=1)
                        ref_err = None
                        if not args.skip_cpu:
                            rr = ra[torch.from_numpy(sample).to(dev)].cpu().numpy()[:, :2].astype(np.float64)
                            e2 = np.linalg.norm(rr - ref64, axis=1) / np.linalg.norm(ref64, axis=1)
                            ref_err = {"median": float(np.median(e2)), "p99": float(np.quantile(e2, 0.99)), "max": float(e2.max())}
                        bh["reference_cuda_bh"] = {"force_evaluations_per_s": 1e3 / ref_ms, "ms": ref_ms,
                                                   "speedup": ref_ms / eval_ms, "speedup_over": "force evaluation vs force evaluation",
                                                   "theta_error_vs_fp64": ref_err,
                                                   "acc_rel_diff_median": float(d.median()),
                                                   "acc_rel_diff_p99": float(d.quantile(0.99)),
                                                   "what": "unmodified src/BarnesHutKernels.cu (sm_100), knodes=3, same box"}
            extra["bh"] = bh
        except StopIteration:
            pass
        except pkg.NbodyError as e:
            extra["bh"] = {"unavailable": str(e)}

    # ---- Barnes-Hut in 3D (octree; beyond the reference): informative --------------------
    if not args.skip_bh and world == 1:
        try:
            p3 = torch.from_numpy(pkg.ics.uniform_disk(N_BODIES, seed=1, z_half_height=3000.0)).to(dev)
            a3 = torch.zeros(N_BODIES, 4, device=dev)
            ms3 = timed(lambda: eng.bh3d_accel(p3, a3, THETA, EPS), 5, 3) / 5
            eng.bh3d_accel(p3, a3, THETA, EPS, pkg.NB_BH_COUNT_WALK)
            c3 = eng.bh3d_last_counters()
            extra["bh3d"] = {"metric": "octree BH steps/s (N=1M thick disc, theta=0.5)", "value": 1e3 / ms3, "unit": "steps/s",
                             "ms_per_step": ms3, "counters": c3,
                             "lane_interactions_per_s": c3["interactions"] / (ms3 * 1e-3),
                             "note": "no reference counterpart (the reference forces Barnes-Hut to R2)"}
            del p3, a3
        except pkg.NbodyError as e:
            extra["bh3d"] = {"unavailable": str(e)}

    # ---- host baseline (rank 0, N=1 only) ----------------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        oracle = graft.load_oracle()
        nt = 4096
        oracle.direct_f32(pos_h, dim, EPS, t0=0, nt=256, fast=True)
        t0 = time.perf_counter()
        oracle.direct_f32(pos_h, dim, EPS, t0=0, nt=nt, fast=True)
        dt_cpu = time.perf_counter() - t0
        cpu_baseline = {"value": nt * float(n) / dt_cpu, "unit": "interactions/s", "cores": oracle.max_threads(),
                        "kind": "port",
                        "sample": f"{nt} targets x {n} sources (1/{n // nt} of one evaluation), OpenMP FP32 port of "
                                  "src/DirectComputeKernels.cu:60-68"}

    if rank == 0:
        line = {"metric": "interactions/s (direct, N=1M)", "value": value, "unit": "interactions/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": shared_config(n, dim, world),
                "clocks": clocks, "e2e": e2e, "gpu_launches": args.steps * (launches_force + 2),
                "roofline": roofline, "cpu_baseline": cpu_baseline}
        line.update(extra)
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
