"""CNN policy forward (csrc/ahk_policy.cu) and the policy-in-the-loop rollout (BASELINE.json configs[3]) timed on one
GPU with CUDA events on the launching stream.  Prints one JSON object.

    python tools/policy_bench.py [--envs 8192] [--iters 50] [--rollout-steps 256]

forward:  policy.forward(env) alone, frames resident in the ring (reads 3 x 49152 B/sample of uint8).
rollout:  env.step(robot_action = previous argmax) -> policy.forward -> ... (ai/reinforcement/train.py:87-96 at batch n;
          ai/supervised/keras/test.py:36-51), the human side played by the programmed AI.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))

import torch  # noqa: E402

from gym_air_hockey_b200 import BatchedAirHockeyEnv  # noqa: E402
from gym_air_hockey_b200.policy import FLOPS_PER_SAMPLE, LAYER_MACS, PolicyNetwork  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--rollout-steps", type=int, default=256)
    ap.add_argument("--lib", default=None, help="A/B: another build of libahk.so (tools/build_variant.sh)")
    args = ap.parse_args()
    if args.lib:
        from gym_air_hockey_b200 import _lib
        _lib.LIB_PATH = os.path.abspath(args.lib)
    n = args.envs
    env = BatchedAirHockeyEnv(n, seed=0, auto_reset=True, max_episode_steps=256)
    pol = PolicyNetwork(max_batch=n)
    env.reset()
    for _ in range(3):
        env.step()
    logits = torch.empty((n, 10), dtype=torch.float32, device="cuda")
    action = torch.empty((n,), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()

    def timed(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(iters):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    fwd_ms = timed(lambda: pol.forward(env, logits, action), args.iters)
    step_ms = timed(lambda: env.step(), args.iters)

    def cycle():
        env.step(robot_action=action)
        pol.forward(env, logits, action)

    roll_ms = timed(cycle, args.rollout_steps)
    layers = [0.0] * 6
    for _ in range(20):
        for i, v in enumerate(pol.forward_profiled(env, logits, action)):
            layers[i] += v / 20
    peaks = {}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            peaks = json.load(f)
    tf = n * FLOPS_PER_SAMPLE / (fwd_ms * 1e-3) / 1e12
    print(json.dumps({
        "envs": n, "forward_ms": fwd_ms, "forward_samples_per_s": n / (fwd_ms * 1e-3), "forward_tflops": tf,
        "forward_frac_of_bf16_sustained": tf / peaks.get("bf16_tflops_sustained", 1392.5),
        "forward_input_GBs": n * 3 * 49152 / (fwd_ms * 1e-3) / 1e9,
        "step_only_ms": step_ms, "rollout_ms_per_step": roll_ms, "rollout_env_steps_per_s": n / (roll_ms * 1e-3),
        "layer_ms": dict(zip(("conv1", "conv2", "conv3", "dense1", "dense2", "out"), layers)),
        "layer_tflops": dict(zip(("conv1", "conv2", "conv3", "dense1", "dense2", "out"), [2 * m * n / (t * 1e-3) / 1e12 for m, t in zip(LAYER_MACS, layers)])),
        "action_hist": torch.bincount(action.long(), minlength=10).tolist(),
    }))


if __name__ == "__main__":
    main()
