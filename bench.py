#!/usr/bin/env python
"""bench.py - IK-converged retargeted frames/s of the batched retargeting hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): unitree_g1 <- 4096 synthetic LAFAN1-shaped BVH clips x 300
frames per GPU, warm-started per clip, cold start (qpos0) per clip like the reference's dataset
driver (scripts/bvh_to_robot_dataset.py:84-102).  One "step" = the whole batch through the hot
path once: gmr_batch_reset + gmr_batch_run (one fused clip-kernel launch).  Clips shard across
ranks with no data-path collective (weak scaling: every rank owns its own 4096 clips).

`value`   : frames/s, inputs resident in HBM, CUDA events on the launching stream, max over ranks.
`e2e`     : same metric through BatchedRetargeter.retarget_clips_host() = one gmr_batch_run_host call:
            pinned host stream -> H2D -> reset + run -> D2H of qpos, all inside the timed region
            (frame ranges pipelined over three streams).
`roofline`: HBM view (algorithmic bytes / clip-kernel time vs measured copy bandwidth) as the
            contract asks, plus `compute`: algorithmic flops vs the measured FP64 FMA peak - the
            pipe this latency-bound kernel actually runs on.
`cpu_baseline` / --impl reference: the C restatement of the reference loop (oracle/gmr_oracle.c;
            the reference itself is Python over mink/mujoco/daqp, none installable here) on the
            box's host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SRC, ROBOT = "bvh", "unitree_g1"
METRIC, UNIT = "ik_converged_retargeted_frames_per_sec", "frames/s"
SEED = 0xB200 + 2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--clips", type=int, default=4096, help="clips per GPU")
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra blocks (other configs, strong scaling, B=1 shim)")
    ap.add_argument("--strong-clips", type=int, default=33334, help="clips of the strong-scaling block (configs[4])")
    ap.add_argument("--e2e-chunks", type=int, default=0,
                    help="gmr_batch_run_host mode: 0 = zero-copy on the pinned host buffers, k >= 1 = k staged frame ranges")
    return ap.parse_args()


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def workload_name(clips, frames):
    return f"unitree_g1 <- {clips} synthetic LAFAN1-shaped BVH clips x {frames} frames per GPU (BASELINE.json configs[1])"


# ---------------------------------------------------------------- flop / byte model
def flops_per_iteration(nv, m, nb, dbar=14.0):
    """SURVEY.md 8d dense count of one IK iteration (1 FMA = 2 flops)."""
    return (2 * 6 * m * nv * (nv + 1) / 2 + 2 * 6 * m * nv + 2 * 36 * nv * m + nv ** 3 / 3 + 2 * nv ** 2
            + 150 * nb + 450 * m + 40 * m * dbar)


def bytes_per_frame(h, nq):
    """human stream read once (float32) + qpos written once (float64)."""
    return h * 7 * 4 + nq * 8


# ---------------------------------------------------------------- clocks sampler
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.thread.join(timeout=5)
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); smax.append(float(r[1])); power.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------- CPU legs (oracle = checker, timed as the baseline)
def build_cpu_case(clips, frames, workers):
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    m, t = model_pack.build_tables(SRC, ROBOT)
    human, heights, names = synth.synth_human_clips(m, t, SRC, B=clips, T=frames, seed=SEED, workers=workers)
    return ModelDesc(m, t, names), human, synth.clip_scales(t, heights)


def time_oracle(desc, human, scale, nthreads):
    from oracle import oracle as oc
    t0 = time.perf_counter()
    res = oc.run_clips(desc, human, scale=scale, nthreads=nthreads, diag=True)
    dt = time.perf_counter() - t0
    return human.shape[0] * human.shape[1] / dt, dt, res


def cpu_baseline(desc, human, scale, budget_s, cores):
    """Oracle C port on all host cores, on the first n clips of the same workload (n sized for ~budget_s)."""
    from oracle import oracle as oc
    oc.build()
    T = human.shape[1]
    probe = min(human.shape[0], 2 * cores)
    rate, _, _ = time_oracle(desc, human[:probe], scale[:probe], cores)
    n = int(min(human.shape[0], max(probe, (rate * budget_s / T) // cores * cores)))
    rate, dt, res = time_oracle(desc, human[:n], scale[:n], cores)
    # configs[0] on the CPU side: ONE clip x 1000 frames on ONE thread (the reference's loop is single-threaded)
    from gmr_b200 import model_pack, synth
    m1, t1 = model_pack.build_tables(SRC, ROBOT)
    h1, hg1, _ = synth.synth_human_clips(m1, t1, SRC, B=1, T=1000, seed=SEED - 1)
    one, _, _ = time_oracle(desc, h1, synth.clip_scales(t1, hg1), 1)
    return {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"first {n} clips x {T} frames of the same workload, oracle/gmr_oracle.c with {cores} pthreads, {dt:.1f} s",
            "iters_per_frame": float(res["iters"].sum(-1).mean()),
            "single_thread_frames_per_s": one,
            "single_thread_sample": "BASELINE.json configs[0]: 1 synthetic LAFAN1-shaped clip x 1000 frames, 1 thread",
            "published_reference_fps": "60-70 FPS on a Threadripper 7960X, 35-45 FPS on an i9-13900K (README.md:615-621; the "
                                       "real mink/mujoco/daqp loop incl. viewer, one thread; not runnable here)"}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (C restatement; mink/mujoco/daqp are not
    installable here) with every host thread, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    T = args.frames
    n_steps = args.steps + args.warmup
    per_step_s = max(2.0, min(25.0, 200.0 / max(1, n_steps)))
    # the full config when one pass over it fits the per-step budget (same_config), else a prefix sized from a probe
    desc, human, scale = build_cpu_case(args.clips, T, cores)
    from oracle import oracle as oc
    oc.build()
    probe = min(human.shape[0], 2 * cores)
    rate, _, _ = time_oracle(desc, human[:probe], scale[:probe], cores)
    n = int(min(human.shape[0], max(cores, (rate * per_step_s / T) // cores * cores)))
    if n >= 0.9 * human.shape[0]:
        n = human.shape[0]
    for _ in range(args.warmup):
        time_oracle(desc, human[:n], scale[:n], cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, _, res = time_oracle(desc, human[:n], scale[:n], cores)
    dt = time.perf_counter() - t0
    value = n * T * args.steps / dt
    full = n == human.shape[0]
    sample = (f"all {n} clips x {T} frames of the workload per step (same config as the GPU arm)" if full else
              f"first {n} clips x {T} frames of the workload per step (a rate: the prefix is the same seeded clips)") + \
        f", oracle/gmr_oracle.c with {cores} pthreads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.clips, T), "robot": ROBOT, "source": SRC, "frames_per_clip": T,
                       "clips_per_step_sample": n, "same_config": full},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "iters_per_frame": float(res["iters"].sum(-1).mean()),
            "published_reference_fps": "60-70 FPS (README.md:615-621: the real mink loop, one thread, Threadripper 7960X)",
            "note": "reference = Python over mink/mujoco/daqp (absent, no network); this arm times its C restatement"}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from gmr_b200 import model_pack, synth
    from gmr_b200.engine import BatchedRetargeter, DeviceModel, probe_fma_peak

    B, T, K, W = args.clips, args.frames, args.steps, max(args.warmup, 0)
    cores = host_cores()
    m, t = model_pack.build_tables(SRC, ROBOT)
    names = synth.human_body_names(SRC, t)
    h = len(names)
    # this rank's shard of the clip set: clips [rank*B, (rank+1)*B) of one seeded stream, generated
    # straight into pinned host memory (the e2e leg copies from there)
    human_host = torch.empty((B, T, h, 7), dtype=torch.float32, pin_memory=True)
    t0 = time.perf_counter()
    _, heights, _ = synth.synth_human_clips(m, t, SRC, B=B, T=T, seed=SEED + 7919 * rank,
                                            workers=max(1, cores // max(1, world)), out=human_host.numpy())
    t_synth = time.perf_counter() - t0

    dm = DeviceModel.load(SRC, ROBOT, names, device=local)
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights)
    human_dev = human_host.to(dev)
    qpos_dev = torch.empty((B, T, br.nq_max), dtype=torch.float64, device=dev)
    qpos_host = torch.empty((B, T, br.nq_max), dtype=torch.float64, pin_memory=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x: float) -> float:
        if world == 1:
            return x
        v = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        return float(v.item())

    # ---- warm-up (untimed), with diagnostics once: iteration counts and the convergence fraction
    br.reset()
    _, diag = br.run(human_dev, diag=True, out=qpos_dev)
    iters = diag["iters"].double()
    iters_per_frame = float(iters.sum(-1).mean().item())
    hit_max = float((diag["iters"] >= dm.desc.c.max_iter + 1).any(-1).double().mean().item())
    sweeps_per_solve = float((diag["nfact"].double().sum() / iters.sum()).item())
    del diag
    for _ in range(max(W - 1, 0)):
        br.reset()
        br.run(human_dev, out=qpos_dev)
    torch.cuda.synchronize()
    if not torch.isfinite(qpos_dev).all().item():
        raise SystemExit("bench.py: non-finite qpos")

    # ---- timed: K steps, inputs resident in HBM (826 MB stream per step >> 126 MB L2: no flush needed)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    ev[0].record()
    for k in range(K):
        br.reset()
        kev[k][0].record()
        br.run(human_dev, out=qpos_dev)
        kev[k][1].record()
    ev[1].record()
    barrier()
    ms_total = maxr(ev[0].elapsed_time(ev[1]))
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    value = world * B * T * K / (ms_total * 1e-3)

    # ---- e2e: host buffers in, host result out, every step
    for _ in range(min(W, 2)):
        br.retarget_clips_host(human_host, qpos_host, n_chunks=args.e2e_chunks)
    barrier()
    ev[0].record()
    for k in range(K):
        br.retarget_clips_host(human_host, qpos_host, n_chunks=args.e2e_chunks)
    ev[1].record()
    barrier()
    clk = clocks.stop()            # sampled over both timed regions (device-resident and end-to-end)
    ms_e2e = maxr(ev[0].elapsed_time(ev[1]))
    e2e_value = world * B * T * K / (ms_e2e * 1e-3)
    checksum = float(qpos_host[:, -1, :3].sum().item())

    # ---- extra blocks: strong scaling on every rank count, the other configs at N = 1
    extras = {}
    if not args.no_extras:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_extras
        del qpos_dev
        extras["strong"] = bench_extras.strong_block(lambda n: BatchedRetargeter(dm, n), human_dev, rank, world, maxr, barrier,
                                                      total_clips=args.strong_clips)
        if world == 1:
            del human_dev
            torch.cuda.empty_cache()
            extras["configs0_shim"] = bench_extras.shim_block()
            extras["configs2_hands"] = bench_extras.hands_block(local, cores)
            extras["configs3_heterogeneous"] = bench_extras.hetero_block(local, cores)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant (only) kernel: gmr::clip_kernel<float>
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as f:
            hbm_peak, peak_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    frames_per_launch = B * T
    bpf = bytes_per_frame(h, dm.desc.c.nq)
    achieved_gbs = bpf * frames_per_launch / (kernel_ms * 1e-3) / 1e9
    fpi = flops_per_iteration(dm.desc.c.nv, max(dm.desc.c.n_task[0], dm.desc.c.n_task[1]), dm.desc.c.nbody - 1)
    achieved_tf = fpi * iters_per_frame * frames_per_launch / (kernel_ms * 1e-3) / 1e12
    fp64_peak = probe_fma_peak(local, True)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "clip_kernel_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    # executed work: ncu's per-opcode thread counts of one launch of this workload (profiles/clip_kernel_ops.json),
    # turned into flops per frame and divided by the launch time measured here
    executed = None
    opath = os.path.join(ROOT, "profiles", "clip_kernel_ops.json")
    if os.path.exists(opath):
        with open(opath) as f:
            ops = json.load(f)
        fl_frame = (2.0 * ops["dfma"] + ops["dmul"] + ops["dadd"]) / ops["frames"]
        ex_tf = fl_frame * frames_per_launch / (kernel_ms * 1e-3) / 1e12
        executed = {"tflops": ex_tf, "frac_of_fp64_peak": ex_tf / fp64_peak if fp64_peak > 0 else None,
                    "flops_per_frame": fl_frame, "warp_instructions_per_frame": ops["inst_executed"] / ops["frames"],
                    "issue_active_pct": ops.get("issue_active_pct"), "pipe_fp64_pct": ops.get("pipe_fp64_pct"),
                    "source": ops.get("source"),
                    "note": "smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd}_pred_on of one ncu-profiled launch of this "
                            "workload / its frames, x frames per launch / the launch time measured in this run"}
    roofline = {"bound": "fp64_fma", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                "frac": achieved_tf / fp64_peak if fp64_peak > 0 else None, "traffic": traffic,
                "peak_source": "measured here (gmr_probe_fma_peak, DFMA chains on all SMs; MEASURED_PEAKS.json has no FP64 entry)",
                "kernel": f"gmr::clip_kernel<float, true, {br.launch_info().get('layout_id', -1)}>", "kernel_ms_per_launch": kernel_ms,
                "algorithmic_flops_per_iteration": fpi, "iterations_per_frame": iters_per_frame,
                "note": "per-instance 6x6 contractions on a 35-dof tree: FP64 CUDA-core work on dependent chains, neither "
                        "HBM nor tensor cores bind it.  `achieved` counts SURVEY.md 8d's DENSE algorithm (84x35 Jacobian, "
                        "J'W2J, Cholesky) at the measured iteration count - the contract figure between rounds; the kernel "
                        "runs an O(n) articulated-body recursion instead, so `executed` (ncu opcode counts) is what the "
                        "FP64 pipe really does and issue_active / pipe_fp64 say how busy the SM is",
                "executed": executed,
                "hbm": {"achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": achieved_gbs / hbm_peak,
                        "algorithmic_bytes_per_frame": bpf, "peak_source": peak_src}}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(B, T), "robot": ROBOT, "source": SRC, "clips_per_gpu": B,
                       "frames_per_clip": T, "human_bodies": h, "parallelism": f"clips sharded over {world} GPU(s), no collective",
                       "l2": "inputs larger than L2 (826 MB stream per step)", "launch": br.launch_info(),
                       "synth_seconds": round(t_synth, 1)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": human_host.numel() * 4 * world,
                    "d2h_bytes_per_step": qpos_host.numel() * 8 * world, "ms_per_step": ms_e2e / K,
                    "api": "BatchedRetargeter.retarget_clips_host -> gmr_batch_run_host, pinned host in / pinned host out; " + (
                        "zero-copy: the kernel's warps pull each frame over PCIe from the pinned input and push each qpos row to the pinned output inside the one launch (no device staging)"
                        if args.e2e_chunks == 0 else
                        f"{args.e2e_chunks} frame ranges, H2D / retarget / D2H overlapped on 3 streams through device staging")},
            "gpu_launches": 2 * K, "gpu_launches_note": "per step: fill_rows_kernel (reset to qpos0) + clip_kernel",
            "clocks": clk, "roofline": roofline,
            "convergence": {"iters_per_frame": iters_per_frame, "frames_hitting_max_iter": hit_max,
                            "qp_sweeps_per_solve": sweeps_per_solve}, "checksum": checksum}
    line.update(extras)
    if not args.no_cpu_baseline and world == 1:
        desc = dm.desc
        hn = human_host.numpy()
        line["cpu_baseline"] = cpu_baseline(desc, hn, synth.clip_scales(t, heights), args.cpu_seconds, cores)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
