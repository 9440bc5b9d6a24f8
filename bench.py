#!/usr/bin/env python
"""bench.py — labelled env-steps/sec of the batched air-hockey engine (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs-per-gpu E] [--impl ours|reference]

Workload: BASELINE.json configs[1] — 4096 envs per GPU, AI on both sides, auto-reset on goals and every 200 steps
(ai/supervised/generate_data.py:44-58), 'rgb' mode.  One ENV step = physics + programmed AI + reward/terminal/labels for
every env and the new 3x128x128 uint8 frame of every env written to the HBM ring (csrc/ahk_step.cu).  One BENCH step (what
--steps / --warmup count) = a chunk of `config.env_steps_per_bench_step` consecutive env steps issued by ONE
ahk_step_many call, so that the driver's `--steps 20 --warmup 5` times 0.3 s of steady state, not 1 ms.
N > 1 (torchrun): envs shard by global id, no collective on the step path; statistics are all-reduced once after the
timed region (NCCL).  The line also carries a `rollout` sub-record: BASELINE.json configs[3] (8192 envs per GPU, CNN policy
in the loop, DDQN fit every 25 steps with the real gradient all-reduce of gym_air_hockey_b200/rl.py).  See DESIGN.md §5.
"""
import argparse
import glob
import json
import os
import re
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

BYTES_PER_STEP = 49152 + 64 + 64 + 7  # frame written + state read + state written + labels/reward/done
RASTER_BYTES_PER_FRAME = 49152 + 16   # the raster kernel alone (A/B split path): one frame written + one draw record read
METRIC = "labeled env-steps/sec (physics+render+label)"
MIN_REGION_S = 0.25                   # every timed region lasts at least this long


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


def ncu_traffic(pattern):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the LAST (by name) committed ncu summary matching `pattern`
    (tools/ncu_summary.py output under profiles/), or None.  Returns (bytes, file name).  A `_T<k>_` in the file name says
    the captured launches ran k env steps each (the persistent ahk_step_many kernel)."""
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", pattern)))  # by name: the snapshot on the GPU box does not keep mtimes
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for path in reversed(files):
        total, seen = 0.0, 0
        with open(path) as f:
            for line in f:
                m = re.match(r"dram__bytes_(read|write)\.sum\s+(\w+)\s+(.*)", line)
                if m and m.group(2) in unit:
                    vals = [float(v.replace(",", "")) for v in m.group(3).split()]
                    total += unit[m.group(2)] * sorted(vals)[len(vals) // 2]
                    seen += 1
        if seen == 2:
            return total, os.path.basename(path)
    return None, None


def forward_traffic(n):
    """DRAM bytes one 'rgb' CNN forward of n samples moves, from the committed ncu --set full captures of its conv kernels at 8192
    samples (dram__bytes_read.sum + dram__bytes_write.sum, scaled by n / 8192); the three dense layers are added with their
    algorithmic bytes (activations in and out + weights once).  -> (bytes, source) or (None, None)."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def columns(name):
        cols = None
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                for line in f:
                    m = re.match(r"dram__bytes_(read|write)\.sum\s+(\w+)\s+(.*)", line)
                    if m and m.group(2) in unit:
                        vals = [unit[m.group(2)] * float(v.replace(",", "")) for v in m.group(3).split()]
                        cols = vals if cols is None else [a + b for a, b in zip(cols, vals)]
        except OSError:
            return None
        return cols
    conv1, conv23 = columns("conv1t_roles_r2.txt"), columns("conv23_r2s5n_ncu.txt")  # conv23_kernel: conv2 -> conv3 fused, act2 stays on the SM
    if not conv1 or not conv23:
        return None, None
    dense = 8192 * (4096 + 512 + 512 + 512 + 512) * 2 + 8192 * 10 * 4 + (4096 * 512 + 512 * 512 + 512 * 16) * 2
    return (conv1[0] + conv23[0] + dense) * n / 8192.0, "conv1t_roles_r2.txt + conv23_r2s5n_ncu.txt (conv2 -> conv3 fused) + algorithmic bytes of the dense layers"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region, read in-process through NVML (no fork next to the launch
    loop).  start() it at least 0.3 s before the region."""

    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, torch_index):
        super().__init__(daemon=True)
        self.samples, self.stop_flag, self.handle, self.nv, self.max_mhz = [], threading.Event(), None, None, None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(torch_index).uuid)
            try:
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(torch_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nv = pynvml
        except Exception as e:  # recorded in the line; the bench still runs
            self.error = repr(e)

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag.is_set():
            try:
                self.samples.append((time.perf_counter(), nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM), int(reasons_fn(self.handle))))
            except Exception:
                pass
            self.stop_flag.wait(0.05)

    def summary(self, t0=None, t1=None):
        self.stop_flag.set()
        if self.is_alive():
            self.join(timeout=2)
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "error": getattr(self, "error", "NVML unavailable")}
        inside = [s for s in self.samples if (t0 is None or s[0] >= t0) and (t1 is None or s[0] <= t1)] or self.samples
        sm = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": [name for bit, name in self.REASONS if bits & bit], "samples": len(inside), "how": "NVML in-process, 50 ms"}


# ------------------------------------------------------------------------------------------------------- CPU arm
class CpuArm(object):
    """The reference's per-step work for this path on the host cores: the oracle port (oracle/ahk_oracle.c — native 450x700
    draw, NEAREST resize, float32 (9,128,128) observation, reward, labels; the reference's own Pygame simulator cannot run
    anywhere) through the ahk_cpu_* twin of the C ABI (include/ahk_cpu.h: one ahk_cpu_step_many call per bench
    step with the float32 observation bound, host buffers in the engine's layouts) on a persistent pool of threads with
    outputs allocated once.  The reference arm
    (--impl reference) and the cpu_baseline leg both go through this class, so their numbers agree."""

    def __init__(self, n_envs=4096, threads=None):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        from gym_air_hockey_b200.assets import make_table_image
        orc.build()
        self.threads = threads or os.cpu_count() or 1
        per_env = 700 * 450 * 3 + 3 * 49152 + 9 * 128 * 128 * 4 + 4096  # native scratch + ring + float32 observation
        avail = 8 << 30
        try:
            with open("/proc/meminfo") as f:
                for line in f:
                    if line.startswith("MemAvailable"):
                        avail = int(line.split()[1]) * 1024
        except OSError:
            pass
        while n_envs > 64 and n_envs * per_env > avail // 3:
            n_envs //= 2
        self.n = n_envs
        # the CPU twin of the engine's C ABI (include/ahk_cpu.h): same structs, buffers and call protocol as the CUDA arm
        self.eng = orc.CpuEngine(self.n, seed=0, table=make_table_image(), auto_reset=True, max_episode_steps=200)
        self.threads = self.eng.threads
        self.eng.reset()
        self.obs = np.empty((self.n, 9, 128, 128), np.float32)
        # what the reference's env.step() hands back besides reward / terminal / labels: the float32 observation, written by the
        # worker that just stepped the env (ahk_cpu_bind_observation)
        self.eng.bind_observation(1, self.obs)
        self._steps(1)  # touches every output page
        t0 = time.perf_counter()
        self._steps(1)
        per_step = max(time.perf_counter() - t0, 1e-6)
        self.sub = max(1, int(round(1.2 * MIN_REGION_S / per_step)))  # sim steps per env per timed call: >= 0.25 s of work

    def _steps(self, k):
        # one ahk_cpu_step_many call: every worker runs k consecutive steps of each of its envs (env-major, as a CTA of the CUDA
        # kernel does and as the reference's own single-env loop does: state, ring and native frame stay in cache)
        self.eng.step_many(k)

    def run(self, calls):
        t0 = time.perf_counter()
        for _ in range(calls):
            self._steps(self.sub)
        return time.perf_counter() - t0

    def sample(self, calls):
        return "%d envs x %d sim-steps per call x %d calls, %d threads (one ahk_cpu_step_many call, float32 observation of every step bound; persistent pool, outputs preallocated)" % (self.n, self.sub, calls, self.threads)


def cpu_reference_wrapper():
    """profiles/cpu_reference_wrapper.json: the reference's OWN CPU code timed in the build container (tools/cpu_reference.py:
    processor.py:31-66 on oracle frames, and the unmodified AirHockeyEnv + DataProcessor over oracle/air_hockey for 10 000
    frames — BASELINE.md §4 B2/B3).  /root/reference does not exist on the GPU box, so the committed record is quoted."""
    path = os.path.join(ROOT, "profiles", "cpu_reference_wrapper.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)
    return None


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path on this box's host cores, all threads, same workload
    (AI vs AI, auto-reset 200, 4096 envs when the host memory allows).  One bench step = one pool call of >= 0.25 s."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(args.envs_per_gpu or 4096)
    arm.run(max(args.warmup, 1))
    dt = arm.run(args.steps)
    value = arm.n * arm.sub * args.steps / dt
    sample = arm.sample(args.steps)
    emit({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1] on the host cores: %d envs, physics+AI label+render (native 450x700 draw, NEAREST resize, float32 9x128x128 observation), AI vs AI, auto-reset 200" % arm.n,
                   "envs": arm.n, "env_steps_per_bench_step": arm.n * arm.sub, "mode": "rgb",
                   "note": "oracle port of the path (kind=port): the reference's Pygame simulator is a missing submodule and cannot run (SURVEY.md F1)"},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": arm.threads, "kind": "port", "sample": sample},
        "reference_wrapper_in_build_container": cpu_reference_wrapper(),
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


# ------------------------------------------------------------------------------------------------------- rollout (configs[3])
def measure_rollout(args, world, rank, local, barrier, min_s=MIN_REGION_S):
    """BASELINE.json configs[3]: 8192 envs per GPU (65 536 over 8), policy-in-the-loop rollout as DDQN training
    (ai/reinforcement/train.py:76-118 at batch n): act = CNN forward on the uint8 ring + epsilon-greedy, env step with the
    robot's action (human = programmed AI), store, a DDQN fit every 25 steps whose gradients (and BatchNorm statistics) are
    all-reduced over NCCL by gym_air_hockey_b200/rl.py — the real collective, not a synthetic buffer.  Returns the sub-record."""
    import torch
    import torch.distributed as dist
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import FLOPS_PER_SAMPLE, LAYER_MACS
    from gym_air_hockey_b200.rl import DDQNAgent

    n = args.rollout_envs
    # ring_depth = replay capacity: the rasteriser writes every frame once, into the slot replay sampling reads it from
    env = BatchedAirHockeyEnv(n, device=local, seed=0, env_id_base=rank * n, auto_reset=True, max_episode_steps=256, ring_depth=32)
    env.reset()
    agent = DDQNAgent(env, buffer_steps=32, batch_size=32, discount_rate=0.98, eps_min=0.05, eps_max=0.99, eps_decay_steps=100000,
                      seed=0, group=None)
    interval = 25

    def loop(steps, train=True):
        fits = 0
        for _ in range(steps):
            action = agent.act()
            _, reward, done, _ = env.step(action)
            agent.store_experience(action, reward, done)
            if train and agent.iteration % interval == 0:
                agent.train()
                fits += 1
        return fits

    agent.replay.record_observation(env)
    loop(2 * interval)  # warm-up: fills the store, first fits (cuDNN/NCCL set-up)
    barrier()
    stream = torch.cuda.current_stream()
    # calibrate the step count for >= min_s
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    loop(interval)
    ev1.record(stream)
    torch.cuda.synchronize()
    per = max(ev0.elapsed_time(ev1) * 1e-3 / interval, 1e-6)
    steps = max(2 * interval, int(-(-1.2 * min_s / per // interval)) * interval)
    if world > 1:
        t = torch.tensor([steps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        steps = int(t.item())
    barrier()
    l0 = env.launches + agent.policy.launches
    ev0.record(stream)
    fits = loop(steps)
    ev1.record(stream)
    barrier()
    train_ms = ev0.elapsed_time(ev1)
    launches = env.launches + agent.policy.launches - l0
    barrier()
    ev0.record(stream)
    loop(steps, train=False)
    ev1.record(stream)
    barrier()
    roll_ms = ev0.elapsed_time(ev1)

    layers = [0.0] * 6
    prof = 10
    for _ in range(prof):
        env.step(robot_action=agent.greedy)
        for i, v in enumerate(agent.policy.forward_profiled(env, agent.logits, agent.greedy)):
            layers[i] += v / prof
    t = torch.tensor([train_ms, roll_ms] + layers, dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    train_ms, roll_ms = t[0].item(), t[1].item()
    layers = [v.item() for v in t[2:]]
    peaks, peak_kind = measured_peaks()
    names = ("conv1", "conv2", "conv3", "dense1", "dense2", "out")
    fwd_ms = sum(layers)
    tensor_peak = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
    tflops = n * FLOPS_PER_SAMPLE / (fwd_ms * 1e-3) / 1e12
    in_gbs = n * 147456 / (fwd_ms * 1e-3) / 1e9
    fwd_traffic, fwd_traffic_src = forward_traffic(n)
    top = max(range(6), key=lambda i: layers[i])
    total = n * world
    rec = {
        "metric": "labeled env-steps/sec (physics+render+label+CNN policy, DDQN fit every 25 steps)",
        "value": total * steps / (train_ms * 1e-3), "unit": "env-steps/s", "steps": steps, "ms_per_step": train_ms / steps,
        "rollout_only": {"value": total * steps / (roll_ms * 1e-3), "ms_per_step": roll_ms / steps},
        "fits": fits, "fit_batch": 32, "fit_interval": interval,
        "config": {"workload": "configs[3]: %d envs/GPU, act = CNN forward (models.py:20-65, random init) on the uint8 ring + epsilon-greedy, env step (human = programmed AI), replay memory = the engine's 32-deep frame ring (no copies), DDQN fit every 25 steps (rl_agents.py:177-191)%s"
                               % (n, ", gradients + BatchNorm statistics all-reduced over NCCL (2 403 961 float32)" if world > 1 else ""),
                   "envs_per_gpu": n, "total_envs": total, "mode": "rgb", "episode_length": 256},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": in_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": in_gbs / peaks["hbm_gbs"], "traffic": fwd_traffic,
                     "traffic_source": fwd_traffic_src,
                     "traffic_GBs": (fwd_traffic / (fwd_ms * 1e-3) / 1e9) if fwd_traffic else None,
                     "frac_of_peak_by_traffic": (fwd_traffic / (fwd_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if fwd_traffic else None,
                     "traffic_note": "act1, act3, h1 and h2 travel through HBM between the kernels (act2 stays on the SM: conv2 -> conv3 are one kernel): the forward MOVES this many bytes (ncu) for 147 456 algorithmic bytes per sample; conv1 alone is 1.41 GB of it per 8192 samples and runs at 0.76 of the copy peak (profiles/conv1t_role_removal_r2s5e.txt); layer_ms come from the profiled forward, which runs conv2 and conv3 as two kernels",
                     "what": "CNN forward: its algorithmic input, the uint8 3-frame stack (147 456 B/sample), over the summed layer-kernel times; the network sits below the bf16 ridge at uint8 input (139 FLOP/B)",
                     "forward_ms": fwd_ms, "layer_ms": dict(zip(names, layers)),
                     "tensor": {"achieved": tflops, "peak": tensor_peak, "unit": "TFLOP/s", "frac": tflops / tensor_peak,
                                "peak_kind": peak_kind + " cuBLAS bf16 sustained", "flops_per_sample": FLOPS_PER_SAMPLE,
                                "slowest_layer": names[top], "slowest_layer_tflops": 2 * LAYER_MACS[top] * n / (layers[top] * 1e-3) / 1e12}},
    }
    agent.policy.close()
    agent.target_policy.close()
    env.close()
    del agent, env
    torch.cuda.empty_cache()
    return rec


# ------------------------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40, help="timed bench steps (each = --chunk env steps)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--chunk", type=int, default=512, help="env steps per bench step (config.env_steps_per_bench_step)")
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--rollout-envs", type=int, default=8192)
    ap.add_argument("--no-rollout", action="store_true")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-calls", type=int, default=24, help="timed calls of the cpu_baseline leg (each >= 0.25 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)
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
    n, chunk, steps, warmup = args.envs_per_gpu, max(args.chunk, 1), max(args.steps, 1), max(args.warmup, 3)
    env = BatchedAirHockeyEnv(n, device=local, seed=0, env_id_base=rank * n, auto_reset=True, max_episode_steps=200)
    env.reset()
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()  # well ahead of the timed region: the warm-up below lasts >= 0.3 s

    # ---- device-resident throughput ("value"): K bench steps of `chunk` env steps each, CUDA events on the launching
    # stream, one event per bench step (so the spread over bench steps is reported too)
    t_w = time.perf_counter()
    for _ in range(warmup):
        env.step_many(chunk)
    while time.perf_counter() - t_w < 0.35:
        env.step_many(chunk)
    barrier()
    launches0 = env.launches
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    t_region0 = time.perf_counter()
    evs[0].record(stream)
    t_host = 0.0
    for k in range(steps):
        h0 = time.perf_counter()
        env.step_many(chunk)
        t_host += time.perf_counter() - h0
        evs[k + 1].record(stream)
    t_host_calls = steps
    barrier()
    t_region1 = time.perf_counter()
    ms = evs[0].elapsed_time(evs[steps])
    per_step = sorted(evs[k].elapsed_time(evs[k + 1]) for k in range(steps))
    launches = env.launches - launches0
    clocks = sampler.summary(t_region0, t_region1) if sampler else None

    # ---- host cost of the public per-step call (launch-rate-boundness): env.step() in a Python loop on an idle stream
    barrier()
    h0 = time.perf_counter()
    for _ in range(256):
        env.step()
    host_us_step = (time.perf_counter() - h0) / 256 * 1e6
    barrier()

    # ---- the dominant kernel alone: CUDA events around single launches (no overlap with neighbours): isolated_kernel_ms
    prof_steps = 100
    env_ms = raster_ms = 0.0
    for _ in range(prof_steps):
        a_ms, b_ms = env.step_profiled()
        env_ms += a_ms
        raster_ms += b_ms
    env_ms /= prof_steps
    raster_ms /= prof_steps

    # ---- end to end through the host-buffer call, one call per ENV step: actions from pinned host memory, reward / done /
    # labels back in pinned host memory, read by the host before the next call (ahk_step_host_results)
    us_env_step = ms * 1e3 / (steps * chunk)
    e2e_steps = max(int(1.2 * MIN_REGION_S / (us_env_step * 1e-6)), 1000)
    out = env.host_buffers(want_frame=False)
    act = torch.full((n,), 255, dtype=torch.uint8).pin_memory().numpy()
    for _ in range(50):
        env.step_host_results(act, act, out)
    env.host_sync()
    barrier()
    t0 = time.perf_counter()
    checksum = 0.0
    for _ in range(e2e_steps):
        env.step_host_results(act, act, out)
        checksum += float(out['reward'][0])  # the host reads the step's result before the next call
    env.host_sync()
    barrier()
    res_s = time.perf_counter() - t0

    full_steps = max(e2e_steps // 4, 500)
    for _ in range(10):
        env.step_host(act, act, out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(full_steps):
        env.step_host(act, act, out)
    barrier()
    e2e_s = time.perf_counter() - t0

    # the same call pipelined (ahk_step_host_begin/_wait, 4 steps in flight, each with its own host buffers): valid for
    # this workload because the AI plays both sides — the next actions do not depend on this step's results
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
        checksum += float(outs[k]['reward'][0])
        if i + depth < e2e_steps:
            tickets[k] = env.step_host_begin(acts[k], acts[k], outs[k])
    barrier()
    pipe_s = time.perf_counter() - t0

    # and with the new frame copied to the host every step (49 152 B per env over PCIe): the dataset-writer case
    frames_s = float("nan")
    if n * 49152 <= (1 << 30):
        outf = env.host_buffers(want_frame=True)
        frame_steps = 12
        env.step_host(act, act, outf)
        barrier()
        t0 = time.perf_counter()
        for _ in range(frame_steps):
            env.step_host(act, act, outf)
        barrier()
        frames_s = (time.perf_counter() - t0) / frame_steps

    t = torch.tensor([ms, e2e_s / full_steps, raster_ms, env_ms, pipe_s, frames_s, res_s, per_step[0], per_step[-1], per_step[len(per_step) // 2]],
                     dtype=torch.float64, device="cuda")
    stats = env.stats.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)  # episode statistics: the only collective, off the step path
    ms, full_s_per, raster_ms, env_ms, pipe_s, frames_s, res_s, step_min, step_max, step_med = (v.item() for v in t)
    env.close()
    del env
    torch.cuda.empty_cache()

    rollout = None
    if not args.no_rollout:
        try:
            rollout = measure_rollout(args, world, rank, local, barrier)
        except Exception as e:  # the headline line must not be lost to the sub-record
            rollout = {"error": repr(e)}
            if world > 1:
                raise

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        total_envs = n * world
        env_steps = steps * chunk
        value = total_envs * env_steps / (ms * 1e-3)
        fused = env_ms == 0.0
        persistent = fused and launches == steps  # ahk_step_many took the persistent kernel: one launch per bench step
        steps_per_launch = env_steps / max(launches, 1) if fused else 1
        kernel_bytes = n * (BYTES_PER_STEP if fused else RASTER_BYTES_PER_FRAME) * steps_per_launch
        kernel_name = ("step_many_kernel (persistent: %d env steps per launch, fused physics+AI+label+raster)" % steps_per_launch) if persistent else \
            "step_kernel<1> (fused physics+AI+label+raster)" if fused else \
            ("raster_staged_kernel" if os.environ.get("AHK_RASTER") == "staged" else "step_kernel<2> (raster only)")
        # Fused path: the timed region consists of exactly `launches` launches of the one kernel, back to back on one
        # stream, so its average launch duration is region / launches.  (isolated_kernel_ms: ONE single-step launch of
        # step_kernel<1> with events on both sides — what ahk_step costs when nothing overlaps it.)
        kernel_ms = ms / max(launches, 1) if fused else raster_ms
        achieved = kernel_bytes / (kernel_ms * 1e-3) / 1e9
        step_gbs = n * BYTES_PER_STEP / (ms / env_steps * 1e-3) / 1e9
        write_peak = None
        wps = sorted(glob.glob(os.path.join(ROOT, "profiles", "wprobe_*.json")), key=os.path.getmtime)  # tools/wprobe.cu
        if wps:
            with open(wps[-1]) as f:
                write_peak = max(r["steady_GBs"] for r in json.load(f)["results"])
        traffic, traffic_src = ncu_traffic("step_many_T*_ncu.txt") if (fused and n == 4096) else (None, None)
        if traffic is not None:  # per captured launch of T env steps -> per launch of this run
            t_cap = int(re.search(r"_T(\d+)_", traffic_src).group(1))
            traffic = traffic / t_cap * (env_steps / max(launches, 1))
        e2e_value = total_envs * e2e_steps / res_s
        line = {
            "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: %d envs/GPU, physics+AI label+render 3x128x128 u8, AI vs AI, auto-reset 200" % n,
                       "envs_per_gpu": n, "total_envs": total_envs, "mode": "rgb", "env_steps_per_bench_step": chunk,
                       "bench_step": "one ahk_step_many call of %d consecutive env steps; --steps/--warmup count these" % chunk,
                       "l2": "each env step writes %.0f MB of frames per GPU (> 126 MB L2), ring of 3" % (n * 49152 / 1e6)},
            "us_per_env_step": ms * 1e3 / env_steps,
            "bench_step_ms": {"median": step_med, "min": step_min, "max": step_max},
            "timed_region_s": ms * 1e-3,
            "gpu_launches": launches,
            "host": {"us_per_step_many_call": t_host / t_host_calls * 1e6, "us_per_env_step_call_python": host_us_step,
                     "note": "host time to ENQUEUE work: one ahk_step_many call (%d env steps in one kernel launch) and one env.step() call from Python (one launch per env step, on an idle stream); the latter must stay below us_per_env_step or a step-by-step loop starves the stream" % chunk},
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_src,
                         "peak_kind": peak_kind + " copy (read+write)",
                         "kernel": kernel_name, "kernel_ms": kernel_ms, "isolated_kernel_ms": raster_ms, "env_kernel_ms": env_ms,
                         "bytes_per_launch": kernel_bytes,
                         "whole_step": {"achieved": step_gbs, "frac": step_gbs / peaks["hbm_gbs"], "bytes_per_env_step": BYTES_PER_STEP},
                         "write_only_peak": write_peak,
                         "frac_of_write_only_peak": (achieved / write_peak) if write_peak else None,
                         "env_steps_per_launch": steps_per_launch,
                         "note": "the path only WRITES; write_only_peak = best steady-state pure write stream over the same 3 x 201 MB ring (tools/wprobe.cu -> profiles/wprobe_*.json); peak = MEASURED_PEAKS.json copy (read+write); traffic = ncu dram bytes per launch scaled from the newest profiles/step_many_T*_ncu.txt capture"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": 2 * n, "d2h_bytes_per_step": 7 * n,
                    "calls": e2e_steps, "timed_region_s": res_s, "le_value": bool(e2e_value <= value * 1.02),
                    "note": "ahk_step_host_results, ONE call per ENV step, the host reads the step's result before the next call: actions (2 B/env) read from pinned host memory and reward/done/labels (7 B/env) written to pinned host memory by the step kernel itself (mapped, zero-copy); the call returns when the results have landed (game ticks over), the frames stay in the HBM ring and are complete at the host_sync that closes the timed region",
                    "full_wait": {"value": total_envs / full_s_per, "note": "ahk_step_host: the same bytes, but every call waits for the whole kernel (frames included) before returning"},
                    "pipelined": {"value": total_envs * e2e_steps / pipe_s, "depth": depth, "note": "ahk_step_host_begin/_wait, same bytes per step; legal when actions do not depend on results (AI vs AI data generation)"},
                    "with_frames_to_host": {"value": (total_envs / frames_s) if frames_s == frames_s else None, "d2h_bytes_per_step": (7 + 49152) * n, "note": "new uint8 frame of every env copied to pinned host memory each step (PCIe-bound)"}},
            "stats": {"steps": int(stats[0]), "episodes": int(stats[1]), "goals_top": int(stats[2]), "goals_bottom": int(stats[3])},
            "rollout": rollout,
        }
        if world == 1 and not args.no_cpu_baseline:
            # the CPU leg runs as `bench.py --impl reference` in a fresh process: inside this one (CUDA context, pinned
            # mappings, torch / NVML threads) the same pool ran up to 1.5x slower than the reference arm the driver starts
            import subprocess
            cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", str(args.cpu_calls), "--warmup", "2"]
            proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""), timeout=900)
            ref_line = None
            for row in proc.stdout.decode().splitlines():
                if row.startswith("{"):
                    ref_line = json.loads(row)
            if ref_line is None:
                raise RuntimeError("cpu_baseline leg failed: %s" % proc.stderr.decode()[-400:])
            line["cpu_baseline"] = dict(ref_line["cpu_baseline"], seconds=ref_line["ms_per_step"] * 1e-3 * ref_line["steps"],
                                        how="bench.py --impl reference in a fresh process",
                                        reference_wrapper_in_build_container=cpu_reference_wrapper())
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
