#!/usr/bin/env python
"""Phase timeline of the fused step kernel from its own %globaltimer stamps (ahk_set_trace): where inside a launch the
time goes — game tick, staging, rasterising — per CTA and warp.  Run on the GPU box:
    python tools/phase_trace.py [--envs 4096] [--steps 30] > gpurun_out/phase_trace.json"""
import argparse, ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import numpy as np
import torch
from gym_air_hockey_b200 import BatchedAirHockeyEnv, _lib

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=4096)
ap.add_argument("--steps", type=int, default=30)
args = ap.parse_args()
env = BatchedAirHockeyEnv(args.envs, seed=0, max_episode_steps=200)
env.reset()
for _ in range(20):
    env.step()
ctas, warps = C.c_int(), C.c_int()
_lib.check(env.lib.ahk_trace_shape(env._h, C.byref(ctas), C.byref(warps)))
trace2 = torch.zeros((2, ctas.value, warps.value, 4), dtype=torch.int64, device=env.device)
trace = trace2
_lib.check(env.lib.ahk_set_trace(env._h, C.c_void_p(trace.data_ptr())))
rows = []
for _ in range(args.steps):
    trace.zero_()
    torch.cuda.synchronize()
    env.step()
    torch.cuda.synchronize()
    t = trace2[(env.launches - 1) & 1].cpu().numpy().astype(np.float64)
    t0 = t[:, :, 0].min()
    rows.append({
        "cta_start_spread_us": (t[:, 0, 0].max() - t0) / 1e3,
        "tick_done_us_mean": (t[:, :, 1].max(axis=1) - t[:, 0, 0]).mean() / 1e3,
        "tick_done_us_max": (t[:, :, 1].max(axis=1) - t[:, 0, 0]).max() / 1e3,
        "staged_us_mean": (t[:, :, 2].max(axis=1) - t[:, 0, 0]).mean() / 1e3,
        "raster_us_mean": (t[:, :, 3].max(axis=1) - t[:, :, 2].min(axis=1)).mean() / 1e3,
        "warp_finish_spread_us_mean": (t[:, :, 3].max(axis=1) - t[:, :, 3].min(axis=1)).mean() / 1e3,
        "cta_issue_end_us_mean": (t[:, :, 3].max(axis=1) - t0).mean() / 1e3,
        "cta_issue_end_us_max": (t[:, :, 3].max() - t0) / 1e3,
        "cta_issue_end_us_min": (t[:, :, 3].max(axis=1) - t0).min() / 1e3,
    })
_lib.check(env.lib.ahk_set_trace(env._h, None))
med = {k: float(np.median([r[k] for r in rows])) for k in rows[0]}
print(json.dumps({"envs": args.envs, "ctas": ctas.value, "warps": warps.value, "steps": args.steps, "median": med}, indent=1))
