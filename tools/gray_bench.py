"""'gray-diff' network forward (models.py:67-116) timed per layer on one GPU.  python tools/gray_bench.py [--envs 8192]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import torch  # noqa: E402
from gym_air_hockey_b200 import BatchedAirHockeyEnv  # noqa: E402
from gym_air_hockey_b200.policy import GRAY_FLOPS_PER_SAMPLE, PolicyNetwork  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=8192)
args = ap.parse_args()
n = args.envs
env = BatchedAirHockeyEnv(n, seed=0, max_episode_steps=256)
pol = PolicyNetwork(max_batch=n, mode='gray-diff')
env.reset()
for _ in range(3):
    env.step()
logits = torch.empty((n, 10), dtype=torch.float32, device="cuda")
action = torch.empty((n,), dtype=torch.uint8, device="cuda")
for _ in range(3):
    pol.forward(env, logits, action)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    pol.forward(env, logits, action)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
layers = [0.0] * 6
for _ in range(20):
    for i, v in enumerate(pol.forward_profiled(env, logits, action)):
        layers[i] += v / 20
print(json.dumps({"envs": n, "mode": "gray-diff", "forward_ms": ms, "samples_per_s": n / (ms * 1e-3),
                  "tflops": n * GRAY_FLOPS_PER_SAMPLE / (ms * 1e-3) / 1e12,
                  "layer_ms": dict(zip(("input+conv1", "conv2", "conv3", "conv4", "dense1", "out"), layers)),
                  "action_hist": torch.bincount(action.long(), minlength=10).tolist()}))
