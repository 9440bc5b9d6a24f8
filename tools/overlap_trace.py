#!/usr/bin/env python
"""Timeline of two CONSECUTIVE overlapped step launches (programmatic dependent launch) from their own %globaltimer
stamps: when launch t+1's CTAs start, finish their game tick and begin to rasterise, relative to launch t's CTAs."""
import argparse, ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import numpy as np
import torch
from gym_air_hockey_b200 import BatchedAirHockeyEnv, _lib

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=4096)
args = ap.parse_args()
env = BatchedAirHockeyEnv(args.envs, seed=0, max_episode_steps=200)
env.reset()
for _ in range(40):
    env.step()
ctas, warps = C.c_int(), C.c_int()
_lib.check(env.lib.ahk_trace_shape(env._h, C.byref(ctas), C.byref(warps)))
tr = torch.zeros((2, ctas.value, warps.value, 4), dtype=torch.int64, device=env.device)
_lib.check(env.lib.ahk_set_trace(env._h, C.c_void_p(tr.data_ptr())))
rows = []
for rep in range(10):
    torch.cuda.synchronize()
    for _ in range(6):
        env.step()
    torch.cuda.synchronize()
    t = tr.cpu().numpy().astype(np.float64)
    last = (env.launches - 1) & 1
    a, b = t[last ^ 1], t[last]          # launch t, launch t+1
    t0 = a[:, :, 0].min()
    us = lambda x: float(x - t0) / 1e3
    rows.append({
        "t_first_cta_end": us(a[:, :, 3].max(axis=1).min()), "t_mean_cta_end": us(a[:, :, 3].max(axis=1).mean()), "t_last_cta_end": us(a[:, :, 3].max()),
        "t1_first_cta_start": us(b[:, :, 0].min()), "t1_mean_cta_start": us(b[:, 0, 0].mean()), "t1_last_cta_start": us(b[:, 0, 0].max()),
        "t1_mean_tick_done": us(b[:, :, 1].max(axis=1).mean()),
        "t1_first_raster_start": us(b[:, :, 2].min()), "t1_mean_raster_start": us(b[:, :, 2].max(axis=1).mean()), "t1_last_raster_start": us(b[:, :, 2].max()),
        "t1_mean_cta_end": us(b[:, :, 3].max(axis=1).mean()), "t1_last_cta_end": us(b[:, :, 3].max()),
        "period": us(b[:, :, 3].max()) - us(a[:, :, 3].max()),
    })
_lib.check(env.lib.ahk_set_trace(env._h, None))
print(json.dumps({"envs": args.envs, "pdl": os.environ.get("AHK_PDL", "1"), "median_us": {k: round(float(np.median([r[k] for r in rows])), 2) for k in rows[0]}}))
