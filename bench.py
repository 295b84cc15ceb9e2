#!/usr/bin/env python
"""bench.py — convex-MPC QP solves/s (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one pass of the hot path (SRBD build -> discretise -> condense -> QP -> GRF) over one
batch of 65 536 synthetic A1 trot problems, H=10 (the configuration the metric is quoted on).
`value` is device-resident throughput (inputs already in HBM), `e2e` goes through the public
host-buffer API (H2D + solve + D2H per step).  N>1: launched by torch.distributed.run, one
process per GPU; each rank owns its own 65 536-problem shard of the stream (weak scaling), no
data-path collective; time = max over ranks.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from legged_mpc_control_b200 import abi, presets, synth  # noqa: E402

METRIC = "convex_mpc_qp_solves_per_sec"
UNIT = "solves/s"
PRESET = "northstar_a1"
BATCH = 65536
SEED = 3
L2_FLUSH_BYTES = 256 << 20


# name -> (preset, seed, robot, gaits, per-problem params, label); BASELINE.json configs
WORKLOADS = {
    "headline": (PRESET, SEED, "a1", (abi.GAIT_TROT,), False, "a1_trot"),                                  # metric config
    "c3": ("northstar_go1_h20", 2, "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), False, "go1_mixed_gaits"),  # config 3
    "ref_h30": ("reference_a1", 6, "a1", (abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND), False,
                "a1_reference_preset_all_gaits"),                                                          # the reference's own H=30
    "sweep": (PRESET, 4, "a1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), True, "a1_param_sweep"),   # config 4
}
_WL = "headline"


def make_workload(B, start):
    preset_name, seed, robot, gaits, per, _ = WORKLOADS[_WL]
    p = presets.preset(preset_name)
    kw = {"gait_speed": (2.0, 4.5)} if per else {}
    return p, synth.make_batch(B, p.horizon, p.dt, seed=seed, robot=robot, gaits=gaits, start=start,
                               per_problem_params=per, base_params=p if per else None, **kw)


def workload_config(n_gpus, batch):
    preset_name, seed, _, _, _, label = WORKLOADS[_WL]
    p = presets.preset(preset_name)
    return {
        "workload": f"{label}_h{p.horizon}_dt{p.dt}_batch{batch}_seed{seed}",
        "preset": preset_name, "horizon": p.horizon, "dt": p.dt, "batch_per_gpu": batch,
        "global_batch": batch * n_gpus, "parallelism": f"batch-shard x{n_gpus} (no collective)",
        "l2": "flushed between timed steps (256 MiB write)",
    }


# ------------------------------------------------------------------ CPU arm (oracle port)
def cpu_sample(p, seconds_target=12.0, threads=None, n_max=None, keep=False):
    """Times the oracle (the CPU restatement of the reference path; kind 'port') on a bounded
    sample of the same workload with all host threads: the first n problems of the stream, n sized
    from a pilot so that the solve takes about `seconds_target` (capped at the batch).  The oracle
    runs its swing-eliminated form (the reduced QP the GPU path solves; its condensing skips the
    swing columns, so the baseline is not padded by work the solve discards).
    Returns (solves/s, cores, sample str, n, seconds, outputs or None)."""
    from oracle import binding as ob
    ob.build()
    threads = threads or (os.cpu_count() or 1)
    n_max = n_max or BATCH
    pilot_n = min(n_max, 32 * threads)
    _, b = make_workload(pilot_n, 0)
    t0 = time.perf_counter()
    ob.solve_batch(p, b, eliminate_swing=True, nthreads=threads, want_kkt=False)
    pilot = time.perf_counter() - t0
    n = int(min(n_max, max(pilot_n, pilot_n * seconds_target / max(pilot, 1e-3))))
    _, b = make_workload(n, 0)
    t0 = time.perf_counter()
    out = ob.solve_batch(p, b, eliminate_swing=True, nthreads=threads, want_kkt=False)
    dt = time.perf_counter() - t0
    assert out["failures"] == 0
    what = "the whole batch" if n == n_max else f"first {n} problems of the workload stream"
    return (n / dt, threads, f"{what}, {threads} threads, one problem per thread, swing legs eliminated, {dt:.1f} s",
            n, dt, (out, b) if keep else None)


def run_reference(args):
    """--impl reference: the reference's own solver (OSQP through osqp-eigen) cannot be built here
    (oracle/_ref compiles the reference's matrix construction against a capturing
    OSQP stub, which has no solve), so this arm times the oracle port on the host cores: rank 0 only.
    Every step solves the SAME bounded sample — the whole batch when the run fits in ~2.5 minutes —
    and `ms_per_step` is the measured time of that step, `problems_per_step` its size."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import binding as ob
    p = presets.preset(WORKLOADS[_WL][0])
    n_steps = args.warmup + args.steps
    budget = args.cpu_seconds * n_steps if args.cpu_seconds else 150.0
    # size the per-step sample once, from a pilot
    _, cores, _, n, dt, _ = cpu_sample(p, seconds_target=budget / n_steps, n_max=args.batch)
    _, b = make_workload(n, 0)
    times = []
    for i in range(n_steps):
        t0 = time.perf_counter()
        out = ob.solve_batch(p, b, eliminate_swing=True, nthreads=cores, want_kkt=False)
        t = time.perf_counter() - t0
        assert out["failures"] == 0
        if i >= args.warmup:
            times.append(t)
    step_s = float(np.mean(times))
    val = n / step_s
    what = "the whole batch" if n == args.batch else f"first {n} problems of the workload stream"
    sample = f"{what} per step, {cores} threads, one problem per thread, swing legs eliminated, {step_s:.2f} s per step"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * step_s, "problems_per_step": n,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args.gpus, args.batch),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "swing_eliminated": True,
                         "what": "oracle/lmpc_oracle.c: reference QP construction restated (pinned bit-for-bit to oracle/_ref, the reference's own "
                                 "ConvexQPSolver.cpp) + exact dual active-set solve; the reference's OSQP is not in this image"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks sampler
class ClockSampler(threading.Thread):
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                r = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                parts = [x.strip() for x in r.stdout.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        mhz = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": int(self.samples[0][1]) if self.samples[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.samples)}


# ------------------------------------------------------------------ GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from legged_mpc_control_b200.engine import ConvexMpcEngine, microbench_fp64

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        host_group = dist.new_group(backend="gloo")  # host-side waits and the final gather: no NCCL kernel spins on a GPU meanwhile

    # weak scaling: rank r solves problems [r*B, (r+1)*B) of the stream
    B = args.batch
    p, batch = make_workload(B, rank * B)
    H = p.horizon
    names = tuple(k for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact", "q", "r", "mu") if k in batch)
    tin = {k: torch.from_numpy(batch[k]).to(dev) for k in names}
    W = abi.active_words(H)
    tout = {"grf": torch.zeros(B, 12, dtype=torch.float64, device=dev),
            "status": torch.zeros(B, dtype=torch.int32, device=dev),
            "iters": torch.zeros(B, dtype=torch.int32, device=dev),
            "active_set": torch.zeros(B, W, dtype=torch.int32, device=dev)}
    pin = {k: v.data_ptr() for k, v in tin.items()}
    pout = {k: v.data_ptr() for k, v in tout.items()}
    eng = ConvexMpcEngine(p, B, device=local)
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.int32, device=dev)
    ts = torch.cuda.Stream(device=dev)  # the stream the kernels are launched on AND the events are recorded on
    stream = ts.cuda_stream

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        eng.solve_device(pin, pout, B, stream=stream, wait=True)

    for _ in range(args.warmup):
        device_step()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    n0 = eng.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kern_ms = []
    with torch.cuda.stream(ts):
        for i in range(args.steps):
            flush.fill_(i)  # L2 flush between timed iterations (outside the event pair)
            ev[i][0].record(ts)
            device_step()
            ev[i][1].record(ts)
            kern_ms.append(eng.last_kernel_ms)
    barrier()
    launches = eng.launch_count - n0
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    stats = eng.last_stats()
    bad = int((tout["status"] > 1).sum().item())
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * B * args.steps / (total_ms * 1e-3)

    # ---- e2e: public host-buffer API, pinned host memory, H2D + D2H inside the timed region
    hb = {k: torch.from_numpy(batch[k]).pin_memory().numpy() for k in names}
    for _ in range(max(1, args.warmup // 2)):
        eng.compute_grfs(hb, want_U=False, copy=False)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 10))
    e2e_lat = []
    for _ in range(e2e_steps):
        t1 = time.perf_counter()
        out = eng.compute_grfs(hb, want_U=False, copy=False)  # blocking: returns after the D2H copies into pinned buffers
        e2e_lat.append(time.perf_counter() - t1)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = world * B * e2e_steps / float(te.item())
    h2d = sum(hb[k].nbytes for k in names)
    d2h = out["grf"].nbytes + out["status"].nbytes + out["iters"].nbytes + out["active_set"].nbytes
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- warm-started re-solves (OSQP warm-start analogue): alternate between the batch and a
    #      slightly moved copy, so every solve starts from the faces of a near but different problem
    warm_info = None
    if not args.no_warm:
        x0b = tin["x0"] + 1e-3 * torch.randn(tin["x0"].shape, dtype=torch.float64, device=dev,
                                             generator=torch.Generator(device=dev).manual_seed(1234 + rank))
        pin_b = dict(pin)
        pin_b["x0"] = x0b.data_ptr()
        eng.set_warm_start(True, horizon_shift=0)
        eng.solve_device(pin, pout, B, stream=stream, wait=True)
        wev, hits, wfact = [], 0, 0.0
        with torch.cuda.stream(ts):
            for i in range(args.steps):
                flush.fill_(i)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(ts)
                eng.solve_device(pin_b if i % 2 == 0 else pin, pout, B, stream=stream, wait=True)
                e1.record(ts)
                wev.append((e0, e1))
                hits += eng.last_warm_hits
                wfact += eng.last_stats()["mean_factorizations"]
        barrier()
        wms = sum(a.elapsed_time(b) for a, b in wev)
        tw = torch.tensor([wms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        warm_info = {"value": world * B * args.steps / (float(tw.item()) * 1e-3), "unit": UNIT,
                     "warm_hit_fraction": hits / (B * args.steps), "mean_factorizations": wfact / args.steps,
                     "bad_status": int((tout["status"] > 1).sum().item()),
                     "what": "same workload, every solve warm-started from the stored faces of a copy of the batch moved by 1e-3 in x0"}
        eng.set_warm_start(False)

    # ---- strong scaling (N > 1): the SAME global batch of 65 536 cut into N contiguous shards
    strong_info = None
    if world > 1:
        Bs = B // world
        _, sb = make_workload(Bs, rank * Bs)
        sin_t = {k: torch.from_numpy(sb[k]).to(dev) for k in names}
        spin = {k: v.data_ptr() for k, v in sin_t.items()}
        for _ in range(3):
            eng.solve_device(spin, pout, Bs, stream=stream, wait=True)
        barrier()
        sev = []
        with torch.cuda.stream(ts):
            for i in range(args.steps):
                flush.fill_(i)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(ts)
                eng.solve_device(spin, pout, Bs, stream=stream, wait=True)
                e1.record(ts)
                sev.append((e0, e1))
        barrier()
        sms = torch.tensor([sum(a.elapsed_time(b) for a, b in sev)], dtype=torch.float64, device=dev)
        dist.all_reduce(sms, op=dist.ReduceOp.MAX)
        strong_info = {"value": world * Bs * args.steps / (float(sms.item()) * 1e-3), "unit": UNIT, "global_batch": world * Bs,
                       "batch_per_gpu": Bs, "ms_per_step": float(sms.item()) / args.steps,
                       "limit": "persistent-grid tail: %d problems over %d resident warps per GPU (%.1f waves), plus the fixed classify + launch latency"
                                % (Bs, 1184, Bs / 1184.0)}
        barrier()

    # ---- the product's single-process multi-GPU path (lmpc_solve_sharded: one host thread per
    #      device inside liblmpc.so, host pointers, no NCCL), timed by rank 0 over all N devices
    #      while the other ranks wait at the barrier
    sharded_info = None
    if rank == 0:
        from legged_mpc_control_b200.engine import ShardedConvexMpcEngine
        try:
            sh = ShardedConvexMpcEngine(p, world * B, list(range(world)))
            _, gb = make_workload(world * B, 0)
            ghb = {k: torch.from_numpy(gb[k]).pin_memory().numpy() for k in names}
            for _ in range(2):
                sh.compute_grfs(ghb, copy=False)
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                so = sh.compute_grfs(ghb, copy=False)
            ssec = time.perf_counter() - t0
            sharded_info = {"value": world * B * e2e_steps / ssec, "unit": UNIT, "devices": world, "steps": e2e_steps,
                            "global_batch": world * B, "not_converged": so["not_converged"],
                            "api": "ShardedConvexMpcEngine.compute_grfs -> lmpc_solve_sharded (host pointers, one thread per device, no NCCL)"}
            sh.close()
        except Exception as e:  # pragma: no cover
            sharded_info = {"error": str(e)[:200]}
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier(group=host_group)  # the other ranks wait here on the HOST: their GPUs are idle while rank 0 drives all N

    eng.solve_device(pin, pout, B, stream=stream, wait=True)  # tout = cold answers of this rank's batch (parity + gather below)
    bad = int((tout["status"] > 1).sum().item())

    # ---- final result gather (the only exchange the path has; outside the timed region; through host memory, no NCCL)
    if world > 1:
        from legged_mpc_control_b200 import shard
        shard._HOST_GROUP = host_group
        g = shard.gather_results(tout["grf"], world * B, rank, world)
        if rank == 0:
            assert g.shape[0] == world * B

    if rank == 0:
        # roofline: executed FP64 flops of the dominant kernel / its launch duration
        peaks = {}
        for name, kind in (("dfma", 0), ("dmma_m8n8k4", 1), ("dmma_m16n8k4", 2), ("dmma_m16n8k8", 3), ("dmma_m16n8k16", 4)):
            try:
                peaks[name] = microbench_fp64(kind, device=local, reps=3)
            except Exception as e:  # pragma: no cover
                peaks[name] = None
        peak = max(v for v in peaks.values() if v)
        kms = float(np.mean(kern_ms))
        achieved = stats["flops_executed"] / (kms * 1e-3) / 1e12
        n = 12 * H
        Kf = stats["mean_factorizations"]
        f_alg_dense = 1728 * H * (H - 1) + 3456 * H * (H + 1) * (H + 2) / 6 + (2 * n * n + 24 * n) + Kf * (n ** 3 / 3 + 4 * n * n)
        io_bytes = B * (h2d / B + 12 * 8 + 4 + 4 + 4 * W)
        cpu_val, cores, sample, n_cpu, _, kept = cpu_sample(p, keep=True)
        parity = _parity(p, kept, tout, n_cpu)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(world, B),
            "p50_batch_latency_ms": float(np.median([a.elapsed_time(b) for a, b in ev])),
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "p50_batch_latency_ms": 1e3 * float(np.median(e2e_lat)), "api": "ConvexMpcEngine.compute_grfs -> lmpc_solve (host pointers)"},
            "gpu_launches": int(launches),
            "roofline": {
                "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": _ncu_traffic() if _WL == "headline" else None,
                "traffic_source": "profiles/" + NCU_TRAFFIC_FILE + " (bytes per launch; input reads + write-backs of the L2-resident "
                                  "Hessian workspace, ~0.2% of HBM peak: the bound is FP64 latency / instruction fetch, not DRAM)",
                "kernel": "lmpc::solve_warp_kernel" if stats["mean_n_eff"] <= 63 else "lmpc::solve_tile_kernel / solve_kernel (by size class)",
                "kernel_ms": kms,
                "precision": "fp64", "peak_source": "measured on this GPU by lmpc_microbench_fp64 (MEASURED_PEAKS.json has no FP64 entry)",
                "fp64_peaks_tflops": peaks,
                "flops_per_launch_executed": stats["flops_executed"], "mean_n_eff": stats["mean_n_eff"],
                "mean_factorizations": Kf, "alg_flops_per_solve_dense_n12H": f_alg_dense,
                "note": "achieved counts the flops the kernel executes; the round-2 interior-point start needs 7.6 instead of 10.9 "
                        "factorizations per solve (round 1), i.e. ~30 % fewer executed flops per solve in ~16 % less time, so frac falls "
                        "while solves/s (the metric) rises; per-solve set-up, polish and instruction fetch are the remaining cost (DESIGN.md 6)",
                "hbm": {"algorithmic_bytes_per_launch": io_bytes, "achieved_GBps": io_bytes / (kms * 1e-3) / 1e9,
                        "peak_GBps": _hbm_peak()},
            },
            "cpu_baseline": {"value": cpu_val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "swing_eliminated": True},
            "parity": parity,
            "strong_scaling": strong_info,
            "e2e_sharded_single_process": sharded_info,
            "not_converged": bad,
            "warm_start": warm_info,
        }
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def run_rollout(args):
    """BASELINE config 5: closed-loop batched SRBD rollout, `--robots` robots per GPU x
    `--rollout-steps` control steps, warm-started QPs, state resident on the device (lmpc_rollout).
    One line per run: robot-steps/s (= MPC solves/s in closed loop) and per-step p50 for cold and warm."""
    import torch
    import torch.distributed as dist
    from legged_mpc_control_b200.engine import ConvexMpcEngine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    p = presets.preset(PRESET)
    B, S = args.robots, args.rollout_steps
    res = {}
    for mode, warm in (("cold", 0), ("warm", 1)):
        st = synth.make_rollout_state(B, seed=5, start=rank * B)
        fr = st.pop("foot_rel")
        t = {k: torch.from_numpy(v).to(dev) for k, v in st.items()}
        ptrs = {k: v.data_ptr() for k, v in t.items()}
        ptrs["batch"] = B
        eng = ConvexMpcEngine(p, B, device=local)
        eng.rollout(dict(ptrs), 20, fr, warm_start=warm, on_device=True)  # warm-up (state advances 20 steps)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        chunk, ms, out_tot = 50, [], {"solves": 0, "factorizations": 0, "warm_hits": 0, "not_converged": 0}
        n0 = eng.launch_count
        for c in range(0, S, chunk):
            o = eng.rollout(dict(ptrs), min(chunk, S - c), fr, warm_start=2 if warm else 0, on_device=True)
            ms.append(o["device_ms"] / min(chunk, S - c))
            for k in out_tot:
                out_tot[k] += o[k]
        total_ms = sum(m * min(chunk, S - c) for m, c in zip(ms, range(0, S, chunk)))
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        x = t["x"].cpu().numpy()
        res[mode] = {"robot_steps_per_s": world * B * S / (float(tt.item()) * 1e-3), "p50_step_ms": float(np.median(ms)),
                     "mean_factorizations": out_tot["factorizations"] / max(1, out_tot["solves"]),
                     "warm_hit_fraction": out_tot["warm_hits"] / max(1, out_tot["solves"]),
                     "not_converged": out_tot["not_converged"], "launches": eng.launch_count - n0,
                     "final_height_err_max": float(np.abs(x[:, 5] - st["cmd"][:, 2]).max()),
                     "final_tilt_max": float(np.abs(x[:, 0:2]).max())}
        eng.close()
    if rank == 0:
        print(json.dumps({"metric": "closed_loop_robot_steps_per_sec", "unit": "robot-steps/s", "value": res["warm"]["robot_steps_per_s"],
                          "n_gpus": world, "higher_is_better": True, "scaling": "weak", "dtype": "f64", "data": "synthetic",
                          "config": {"workload": f"config5_rollout_{B}robots_x_{S}steps_a1_trot_h{p.horizon}_dt{p.dt}", "preset": PRESET,
                                     "robots_per_gpu": B, "steps": S}, "cold": res["cold"], "warm": res["warm"]}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_c1(args):
    """BASELINE config 1: ONE robot, one tick (the reference's own use: ConvexMpc::grf_update at
    100 Hz).  p50 latency of the blocking host-pointer call for both presets, cold and warm, next
    to the oracle's single-thread time for the same problem."""
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    from oracle import binding as ob
    res = {}
    for name in ("northstar_a1", "reference_a1"):
        p = presets.preset(name)
        b = synth.make_batch(1, p.horizon, p.dt, seed=1)
        b = {k: b[k] for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
        eng = ConvexMpcEngine(p, 1, device=0)
        row = {}
        for mode in ("cold", "warm"):
            eng.set_warm_start(mode == "warm", 0)
            for _ in range(20):
                eng.compute_grfs(b, want_U=False, copy=False)
            lat, kms = [], []
            for _ in range(200):
                t0 = time.perf_counter()
                out = eng.compute_grfs(b, want_U=False, copy=False)
                lat.append(time.perf_counter() - t0)
                kms.append(eng.last_kernel_ms)
            row[mode] = {"p50_call_us": 1e6 * float(np.median(lat)), "p95_call_us": 1e6 * float(np.percentile(lat, 95)),
                         "p50_kernel_us": 1e3 * float(np.median(kms)), "status": int(out["status"][0]),
                         "factorizations": int(out["iters"][0] & 0xFFFF)}
        t0 = time.perf_counter()
        for _ in range(20):
            ob.solve_batch(p, b, eliminate_swing=True, nthreads=1, want_kkt=False)
        row["cpu_oracle_us"] = 1e6 * (time.perf_counter() - t0) / 20
        res[name] = row
        eng.close()
    print(json.dumps({"metric": "single_robot_tick_latency", "unit": "us", "value": res["reference_a1"]["warm"]["p50_call_us"],
                      "higher_is_better": False, "n_gpus": 1, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": "c1_single_solve_batch1", "api": "ConvexMpcEngine.compute_grfs -> lmpc_solve (host pointers)"},
                      "presets": res}), flush=True)


def _parity(p, kept, tout, n):
    """CUDA results of the timed batch against the CPU oracle's answers for its first n problems
    (the very problems cpu_baseline solved): tests/parity_util.py rules."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import parity_util as pu
    ref, b = kept
    got = {"grf": tout["grf"][:n].cpu().numpy(), "status": tout["status"][:n].cpu().numpy(),
           "active_set": tout["active_set"][:n].cpu().numpy().view(np.uint32)}
    bad, forgiven, rows = pu.active_mismatch(p, got["active_set"], ref["active_set"], ref["U"], b.get("mu"))
    return {"problems": int(n), "against": "oracle (cpu_baseline's own outputs)",
            "grf_max_rel_elementwise": float(pu.elem_err(got["grf"], ref["grf"]).max()),
            "grf_max_rel_normwise": float(pu.norm_err(got["grf"], ref["grf"]).max()),
            "active_mismatch_rows": bad, "active_tie_rows_forgiven": forgiven, "active_rows": rows,
            "status_counts": np.bincount(got["status"], minlength=4).tolist(), "tolerance": pu.REL_TOL}


NCU_TRAFFIC_FILE = "r02_v13_ncu_full_solve_warp_kernel.csv"  # one 65 536-problem launch of the headline workload, `ncu --set full`


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one solve_warp_kernel launch of the headline
    workload, from the committed `ncu --set full` capture (bench.py never runs under a profiler)."""
    path = os.path.join(ROOT, "profiles", NCU_TRAFFIC_FILE)
    try:
        tot = 0.0
        for ln in open(path):
            f = ln.strip().split(",")
            if f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(f[2]) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[f[1]]
        return tot or None
    except Exception:
        return None


def _hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"]
    except Exception:
        return 6650.0  # fallback stated in B200_PROFILING.md


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--cpu-seconds", type=float, default=None, help="CPU work per step of --impl reference (default: 60 s over the run)")
    ap.add_argument("--no-warm", action="store_true", help="skip the warm-started re-solve measurement")
    ap.add_argument("--workload", default="headline", choices=["headline", "c1", "c3", "ref_h30", "sweep", "rollout"],
                    help="headline: the metric's config (default); c3 / sweep: BASELINE configs 3 / 4; ref_h30: the reference's own "
                         "H=30 preset; rollout: config 5 (16384 robots x --rollout-steps control steps, closed loop on device)")
    ap.add_argument("--rollout-steps", type=int, default=1000)
    ap.add_argument("--robots", type=int, default=16384)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: W >= 3
    global _WL
    if args.workload not in ("rollout", "c1"):
        _WL = args.workload
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "rollout":
        run_rollout(args)
    elif args.workload == "c1":
        run_c1(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
