#!/usr/bin/env python
"""bench.py — labelled env-steps/sec of the batched air-hockey engine (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs-per-gpu E] [--impl ours|reference]

A "step" is one pass of the hot path over the whole batch: physics + programmed AI + reward/terminal/labels for every
env and the new 3x128x128 uint8 frame of every env rasterised into the HBM ring — one fused kernel (csrc/ahk_step.cu).
Workload: BASELINE.json configs[1] — 4096 envs per GPU, AI on both sides, auto-reset on goals and every 200 steps
(ai/supervised/generate_data.py:44-58).  N > 1 (torchrun): envs shard by global id, no collective on the step path;
statistics are all-reduced once after the timed region (NCCL).  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

BYTES_PER_STEP = 49152 + 64 + 64 + 7  # frame written + state read + state written + labels/reward/done
RASTER_BYTES_PER_FRAME = 49152 + 16   # the raster kernel alone: one frame written + one 16-byte draw record read
# dram__bytes_read.sum + dram__bytes_write.sum of one step_kernel<1> launch at 4096 envs, ncu --set full
# (profiles/step_kernel_r1j_ncu.txt).  Below the algorithmic 201.9 MB because a launch ends with ~50 MB of its frames
# still dirty in the 126 MB L2 (they reach DRAM during the next launch): nothing is re-read, nothing written twice.
STEP_TRAFFIC_NCU_4096 = 142.62e6  # mean of three captured launches: 142.23 MB written + 0.39 MB read
METRIC = "labeled env-steps/sec (physics+render+label)"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def cpu_baseline(n_envs, budget_s, threads):
    """The oracle port (oracle/ahk_oracle.c) doing the reference's per-step work — native 450x700 draw, NEAREST
    resize, float32 (9,128,128) observation, reward, labels — on `threads` host threads, for about budget_s seconds."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    from gym_air_hockey_b200.assets import make_table_image
    orc.build()
    table = make_table_image()
    envs = [orc.Env(seed=0, env_id=i, table=table) for i in range(n_envs)]
    for e in envs:
        e.reset()
    orc.run_batch(envs, 2, 200, threads)  # warm-up
    t0 = time.perf_counter()
    orc.run_batch(envs, 4, 200, threads)
    per_step = (time.perf_counter() - t0) / 4
    steps = max(4, int(budget_s / max(per_step, 1e-9)))
    t0 = time.perf_counter()
    orc.run_batch(envs, steps, 200, threads)
    dt = time.perf_counter() - t0
    return n_envs * steps / dt, "%d envs x %d steps, %d threads, %.1f s" % (n_envs, steps, threads, dt)


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path on this box's host cores.  The reference's own
    simulator cannot run anywhere (missing submodule), so this times the oracle port with all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_envs = max(threads * 4, 64)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    from gym_air_hockey_b200.assets import make_table_image
    orc.build()
    table = make_table_image()
    envs = [orc.Env(seed=0, env_id=i, table=table) for i in range(n_envs)]
    for e in envs:
        e.reset()
    sub = 8  # sim steps per env per bench "step" (a bounded sample of the workload)
    for _ in range(max(args.warmup, 1)):
        orc.run_batch(envs, sub, 200, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.run_batch(envs, sub, 200, threads)
    dt = time.perf_counter() - t0
    value = n_envs * sub * args.steps / dt
    sample = "%d envs x %d sim-steps per bench step, %d threads" % (n_envs, sub, threads)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "oracle port of configs[1] (AI vs AI, auto-reset 200), host cores only", "sample": sample},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--envs-per-gpu", type=int, default=4096)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from gym_air_hockey_b200 import BatchedAirHockeyEnv

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.envs_per_gpu
    env = BatchedAirHockeyEnv(n, device=local, seed=0, env_id_base=rank * n, auto_reset=True, max_episode_steps=200)
    env.reset()
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value"): K steps, CUDA events on the launching stream
    for _ in range(max(args.warmup, 3)):
        env.step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = env.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        env.step()
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = env.launches - launches0
    clocks = sampler.summary() if sampler else None

    # ---- the dominant kernel alone: CUDA events recorded on the launching stream around the step's kernel(s)
    # (ahk_step_profiled), averaged over a separate pass of the same workload.  Default path: ONE fused kernel per
    # step (step_kernel<1>: physics + AI + labels + raster), reported as raster_ms with env_ms == 0; with AHK_PATH=split
    # the env kernel and the raster kernel are timed separately.
    prof_steps = min(max(args.steps // 4, 20), 400)
    env_ms = raster_ms = 0.0
    for _ in range(prof_steps):
        a_ms, b_ms = env.step_profiled()
        env_ms += a_ms
        raster_ms += b_ms
    env_ms /= prof_steps
    raster_ms /= prof_steps

    # ---- end to end through the host-buffer call (actions from pinned host memory, results back to host)
    out = env.host_buffers(want_frame=False)
    act = torch.full((n,), 255, dtype=torch.uint8).pin_memory().numpy()
    e2e_steps = max(args.steps // 4, 10)
    for _ in range(3):
        env.step_host(act, act, out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        env.step_host(act, act, out)
    barrier()
    e2e_s = time.perf_counter() - t0

    # ---- the same call pipelined (ahk_step_host_begin/_wait, 4 steps in flight, each with its own host buffers): valid
    # for this workload because the AI plays both sides — the next actions do not depend on this step's results
    depth = 4
    outs = [env.host_buffers(want_frame=False) for _ in range(depth)]
    acts = [torch.full((n,), 255, dtype=torch.uint8).pin_memory().numpy() for _ in range(depth)]
    tickets = [env.step_host_begin(acts[k], acts[k], outs[k]) for k in range(depth)]
    for k in range(depth):
        env.step_host_wait(tickets[k])
    barrier()
    t0 = time.perf_counter()
    tickets = [env.step_host_begin(acts[k], acts[k], outs[k]) for k in range(depth)]
    for i in range(e2e_steps):
        k = i % depth
        env.step_host_wait(tickets[k])
        checksum = float(outs[k]['reward'][0])  # the host reads the step's result
        if i + depth < e2e_steps:
            tickets[k] = env.step_host_begin(acts[k], acts[k], outs[k])
    barrier()
    pipe_s = time.perf_counter() - t0

    # ---- and with the new frame copied to the host every step (49 152 B per env over PCIe): the dataset-writer case
    outf = env.host_buffers(want_frame=True)
    frame_steps = max(min(e2e_steps // 20, 50), 5)
    env.step_host(act, act, outf)
    barrier()
    t0 = time.perf_counter()
    for _ in range(frame_steps):
        env.step_host(act, act, outf)
    barrier()
    frames_s = (time.perf_counter() - t0) / frame_steps

    t = torch.tensor([ms, e2e_s, raster_ms, env_ms, pipe_s, frames_s], dtype=torch.float64, device="cuda")
    stats = env.stats.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)  # episode statistics: the only collective, off the step path
    ms, e2e_s, raster_ms, env_ms, pipe_s, frames_s = (v.item() for v in t)

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        total_envs = n * world
        value = total_envs * args.steps / (ms * 1e-3)
        # roofline of the dominant kernel: algorithmic bytes of ONE raster launch = n frames x (49152 B written + 16 B
        # draw record read) over its event-timed duration; the whole step (49287 B/env-step) beside it
        fused = env_ms == 0.0
        kernel_bytes = n * (BYTES_PER_STEP if fused else RASTER_BYTES_PER_FRAME)
        kernel_name = "step_kernel<1> (fused physics+AI+label+raster)" if fused else \
            ("raster_staged_kernel" if os.environ.get("AHK_RASTER") == "staged" else "step_kernel<2> (raster only)")
        # Fused path: the timed region IS K launches of the one kernel, back to back on one stream, so its average
        # launch duration is region / K.  (Consecutive launches overlap by programmatic dependent launch: the next
        # launch's prologue and game tick run under this one's tail.  raster_ms, taken with events BETWEEN launches,
        # cannot overlap and is reported as isolated_kernel_ms.)
        kernel_ms = ms / args.steps if fused else raster_ms
        achieved = kernel_bytes / (kernel_ms * 1e-3) / 1e9
        step_gbs = n * BYTES_PER_STEP / (ms / args.steps * 1e-3) / 1e9
        write_peak = None
        wp = os.path.join(ROOT, "profiles", "wprobe_r1b.json")  # tools/wprobe.cu: pure write streams over the same footprint
        if os.path.exists(wp):
            with open(wp) as f:
                write_peak = max(r["steady_GBs"] for r in json.load(f)["results"])
        line = {
            "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: %d envs/GPU, physics+AI label+render 3x128x128 u8, AI vs AI, auto-reset 200" % n,
                       "envs_per_gpu": n, "total_envs": total_envs, "mode": "rgb",
                       "l2": "each step writes %.0f MB of frames per GPU (> 126 MB L2), ring of 3" % (n * 49152 / 1e6)},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": STEP_TRAFFIC_NCU_4096 if (fused and n == 4096) else None,
                         "peak_kind": peak_kind + " copy (read+write)",
                         "kernel": kernel_name, "kernel_ms": kernel_ms, "isolated_kernel_ms": raster_ms, "env_kernel_ms": env_ms,
                         "bytes_per_launch": kernel_bytes,
                         "whole_step": {"achieved": step_gbs, "frac": step_gbs / peaks["hbm_gbs"], "bytes_per_env_step": BYTES_PER_STEP},
                         "write_only_peak": write_peak,
                         "frac_of_write_only_peak": (achieved / write_peak) if write_peak else None,
                         "note": "the path only WRITES; write_only_peak = best steady-state pure write stream (st.global.cs.v4 / cp.async.bulk) over the same 3 x 201 MB ring, tools/wprobe.cu -> profiles/wprobe_r1b.json; peak = MEASURED_PEAKS.json copy (read+write)"},
            "e2e": {"value": total_envs * e2e_steps / e2e_s, "unit": "env-steps/s", "h2d_bytes_per_step": 2 * n,
                    "d2h_bytes_per_step": 7 * n,
                    "note": "ahk_step_host, one synchronous call per step: actions (2 B/env) read from pinned host memory and reward/done/labels (7 B/env) written to pinned host memory by the step kernel itself (mapped, zero-copy), then the host waits; frames stay in the HBM ring",
                    "pipelined": {"value": total_envs * e2e_steps / pipe_s, "depth": depth, "note": "ahk_step_host_begin/_wait, same bytes per step; legal when actions do not depend on results (AI vs AI data generation)"},
                    "with_frames_to_host": {"value": total_envs / frames_s, "d2h_bytes_per_step": (7 + 49152) * n, "note": "new uint8 frame of every env copied to pinned host memory each step (PCIe-bound)"}},
            "stats": {"steps": int(stats[0]), "episodes": int(stats[1]), "goals_top": int(stats[2]), "goals_bottom": int(stats[3])},
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, sample = cpu_baseline(max(threads * 2, 32), args.cpu_seconds, threads)
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
