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


_REAL_STDOUT = None


def quiet_stdout():
    """Everything libraries print to stdout (NCCL's version banner, for one) goes to stderr; emit() writes the one JSON line
    the driver parses to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    data = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


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
    emit(({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "oracle port of configs[1] (AI vs AI, auto-reset 200), host cores only", "sample": sample},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_rollout(args):
    """--workload rollout: BASELINE.json configs[3] — policy-in-the-loop rollout (ai/reinforcement/train.py:87-96 at batch
    n): every step = fused env step with the robot's action from the CNN + the CNN forward on the new frames (six tcgen05
    kernels reading the uint8 ring).  65 536 envs over 8 GPUs = 8192 per GPU; the human side is the programmed AI;
    episode_length 256 (train.py:45); random-init weights (Glorot, seed 0).  N > 1: a flat 2 403 961-float32 gradient
    buffer is all-reduced over NCCL every 25 steps (training_interval, train.py:47) — the buffer is synthetic: the
    backward pass is SURVEY.md §8f N2, not built."""
    import torch
    import torch.distributed as dist
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import FLOPS_PER_SAMPLE, LAYER_MACS, PolicyNetwork

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.envs_per_gpu or 8192
    steps, warmup = min(args.steps, 512), max(min(args.warmup, 20), 3)
    env = BatchedAirHockeyEnv(n, device=local, seed=0, env_id_base=rank * n, auto_reset=True, max_episode_steps=256)
    pol = PolicyNetwork(max_batch=n, device=local)
    env.reset()
    stream = torch.cuda.current_stream()
    logits = torch.empty((n, 10), dtype=torch.float32, device="cuda")
    action = torch.full((n,), 4, dtype=torch.uint8, device="cuda")
    grad = torch.zeros(2403961, dtype=torch.float32, device="cuda")
    pending = [None]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def cycle(t):
        env.step(robot_action=action)
        pol.forward(env, logits, action)
        if world > 1 and t % 25 == 24:
            if pending[0] is not None:
                pending[0].wait()
            pending[0] = dist.all_reduce(grad, async_op=True)

    for t in range(warmup):
        cycle(t)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = env.launches + pol.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for t in range(steps):
        cycle(t)
    if pending[0] is not None:
        pending[0].wait()
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = env.launches + pol.launches - l0
    clocks = sampler.summary() if sampler else None

    layers = [0.0] * 6
    prof = 20
    for _ in range(prof):
        env.step(robot_action=action)
        for i, v in enumerate(pol.forward_profiled(env, logits, action)):
            layers[i] += v / prof

    # end to end: every step the host supplies an exploration override (epsilon-greedy, rl_agents.py:50-57: 255 = keep the
    # policy's action) from pinned memory and reads reward + done back
    override = torch.full((n,), 255, dtype=torch.uint8).pin_memory()
    override[::10] = 4
    rew_h, done_h = torch.empty(n, dtype=torch.float32).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    e2e_steps = max(steps // 4, 10)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ov = override.to("cuda", non_blocking=True)
        act = torch.where(ov == 255, action, ov)
        _, reward, done, _ = env.step(robot_action=act)
        pol.forward(env, logits, action)
        rew_h.copy_(reward, non_blocking=True)
        done_h.copy_(env.done, non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    e2e_s = time.perf_counter() - t0

    t = torch.tensor([ms, e2e_s] + layers, dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_s = t[0].item(), t[1].item()
    layers = [v.item() for v in t[2:]]
    if rank == 0:
        peaks, peak_kind = measured_peaks()
        total = n * world
        names = ("conv1", "conv2", "conv3", "dense1", "dense2", "out")
        fwd_ms = sum(layers)
        top = max(range(6), key=lambda i: layers[i])
        achieved = 2 * LAYER_MACS[top] * n / (layers[top] * 1e-3) / 1e12
        peak = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
        emit(({
            "metric": "labeled env-steps/sec (physics+render+label+CNN policy)", "value": total * steps / (ms * 1e-3), "unit": "env-steps/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "configs[3]: %d envs/GPU policy-in-the-loop rollout: fused env step + CNN forward (models.py:20-65, random init), robot = argmax, human = programmed AI, episode_length 256%s"
                                   % (n, ", 9.6 MB gradient all-reduce (synthetic buffer) every 25 steps" if world > 1 else ""),
                       "envs_per_gpu": n, "total_envs": total, "mode": "rgb",
                       "l2": "each step writes %.0f MB of frames and reads the 3-frame stack (%.0f MB) per GPU (> 126 MB L2)" % (n * 49152 / 1e6, n * 147456 / 1e6)},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                         "peak_kind": peak_kind + " cuBLAS bf16 sustained", "kernel": names[top] + " (tcgen05, Toeplitz form)" if top == 0 else names[top],
                         "kernel_ms": layers[top], "flops_per_launch": 2 * LAYER_MACS[top] * n,
                         "layer_ms": dict(zip(names, layers)),
                         "forward": {"ms": fwd_ms, "tflops": n * FLOPS_PER_SAMPLE / (fwd_ms * 1e-3) / 1e12, "frac": n * FLOPS_PER_SAMPLE / (fwd_ms * 1e-3) / 1e12 / peak,
                                     "input_GBs": n * 147456 / (fwd_ms * 1e-3) / 1e9, "frac_of_hbm": n * 147456 / (fwd_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]},
                         "note": "algorithmic flops (SURVEY.md 8d); the network sits below the bf16 ridge at uint8 input (139 FLOP/B), and conv1 (N = 5 outputs) executes 3.6x its algorithmic MACs in Toeplitz form"},
            "e2e": {"value": total * e2e_steps / e2e_s, "unit": "env-steps/s", "h2d_bytes_per_step": n, "d2h_bytes_per_step": 5 * n,
                    "note": "per step: exploration override (1 B/env) from pinned host memory, env step + CNN forward on the device, reward + done (5 B/env) back to pinned host memory, host waits"},
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--workload", default="datagen", choices=["datagen", "rollout"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "rollout":
        return run_rollout(args)
    args.envs_per_gpu = args.envs_per_gpu or 4096

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

    # ---- the same loop through ahk_step_host_results: the call returns when reward / done / labels of the step are in the
    # host buffers (the kernel signals the end of its game ticks through mapped memory) and the frames are still being
    # rasterised; the host reads the result, then makes the next call — so the next actions MAY depend on this step's
    # results (policy in the loop), yet the kernels run back to back.  host_sync (all frames written) is inside the timing.
    for _ in range(3):
        env.step_host_results(act, act, out)
    env.host_sync()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        env.step_host_results(act, act, out)
        checksum = float(out['reward'][0])  # the host reads the step's result before the next call
    env.host_sync()
    barrier()
    res_s = time.perf_counter() - t0

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
    frames_s = float("nan")
    if n * 49152 <= (1 << 30):  # a pinned frame buffer beyond 1 GB is not worth measuring PCIe with
        outf = env.host_buffers(want_frame=True)
        frame_steps = max(min(e2e_steps // 20, 50), 5)
        env.step_host(act, act, outf)
        barrier()
        t0 = time.perf_counter()
        for _ in range(frame_steps):
            env.step_host(act, act, outf)
        barrier()
        frames_s = (time.perf_counter() - t0) / frame_steps

    t = torch.tensor([ms, e2e_s, raster_ms, env_ms, pipe_s, frames_s, res_s], dtype=torch.float64, device="cuda")
    stats = env.stats.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)  # episode statistics: the only collective, off the step path
    ms, e2e_s, raster_ms, env_ms, pipe_s, frames_s, res_s = (v.item() for v in t)

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
            "e2e": {"value": total_envs * e2e_steps / res_s, "unit": "env-steps/s", "h2d_bytes_per_step": 2 * n,
                    "d2h_bytes_per_step": 7 * n,
                    "note": "ahk_step_host_results, one call per step, the host reads the step's result before the next call: actions (2 B/env) read from pinned host memory and reward/done/labels (7 B/env) written to pinned host memory by the step kernel itself (mapped, zero-copy); the call returns when the results have landed (game ticks over), the frames stay in the HBM ring and are complete at the host_sync that closes the timed region",
                    "full_wait": {"value": total_envs * e2e_steps / e2e_s, "note": "ahk_step_host: the same bytes, but every call waits for the whole kernel (frames included) before returning"},
                    "pipelined": {"value": total_envs * e2e_steps / pipe_s, "depth": depth, "note": "ahk_step_host_begin/_wait, same bytes per step; legal when actions do not depend on results (AI vs AI data generation)"},
                    "with_frames_to_host": {"value": (total_envs / frames_s) if frames_s == frames_s else None, "d2h_bytes_per_step": (7 + 49152) * n, "note": "new uint8 frame of every env copied to pinned host memory each step (PCIe-bound)"}},
            "stats": {"steps": int(stats[0]), "episodes": int(stats[1]), "goals_top": int(stats[2]), "goals_bottom": int(stats[3])},
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, sample = cpu_baseline(max(threads * 2, 32), args.cpu_seconds, threads)
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
