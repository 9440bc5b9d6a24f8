#!/usr/bin/env python
"""How the step kernel is launched matters at 4096 envs (a ~40 us kernel): CPU enqueue cost per step, GPU time per step
for a plain stream loop, and for a CUDA graph of 60 steps replayed.  Run on the GPU box; prints one JSON object."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import torch
from gym_air_hockey_b200 import BatchedAirHockeyEnv

n = int(os.environ.get("ENVS", "4096"))
env = BatchedAirHockeyEnv(n, seed=0, max_episode_steps=200)
env.reset()
for _ in range(50):
    env.step()
torch.cuda.synchronize()
out = {"envs": n, "pdl": os.environ.get("AHK_PDL", "1")}
K = 1200
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(K):
    env.step()
e1.record(); t1 = time.perf_counter()
torch.cuda.synchronize()
out["stream_cpu_enqueue_us_per_step"] = (t1 - t0) / K * 1e6
out["stream_gpu_us_per_step"] = e0.elapsed_time(e1) / K * 1e3
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for _ in range(60):
            env.step()
    for _ in range(3):
        g.replay()
    s.synchronize()
    e0.record(s)
    for _ in range(20):
        g.replay()
    e1.record(s)
    s.synchronize()
out["graph_gpu_us_per_step"] = e0.elapsed_time(e1) / (20 * 60) * 1e3
print(json.dumps(out))
