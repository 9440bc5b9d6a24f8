#!/usr/bin/env python
"""Why do some CTAs of the fused step kernel finish later than others?  Per-CTA end times (ahk_set_trace) against what
the CTA had to draw (noise frames, reset frames) and across two different env populations (same SMs, other games)."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import numpy as np
import torch
from gym_air_hockey_b200 import BatchedAirHockeyEnv, _lib

def run(seed, steps=40, n=4096):
    env = BatchedAirHockeyEnv(n, seed=seed, max_episode_steps=200)
    env.reset()
    for _ in range(30):
        env.step()
    ctas, warps = C.c_int(), C.c_int()
    env.lib.ahk_trace_shape(env._h, C.byref(ctas), C.byref(warps))
    tr = torch.zeros((2, ctas.value, warps.value, 4), dtype=torch.int64, device='cuda')
    env.lib.ahk_set_trace(env._h, C.c_void_p(tr.data_ptr()))
    ends, noise = [], []
    bounds = [b * n // ctas.value for b in range(ctas.value + 1)]
    for _ in range(steps):
        torch.cuda.synchronize(); env.step(); torch.cuda.synchronize()
        t = tr[(env.launches - 1) & 1].cpu().numpy().astype(np.float64)
        ends.append((t[:, :, 3].max(axis=1) - t[:, :, 0].min()) / 1e3)
        ep = env.get_state()[1]['episode']
        noise.append([int((ep[bounds[b]:bounds[b + 1]] == 2).sum()) for b in range(ctas.value)])
    env.lib.ahk_set_trace(env._h, None)
    return np.array(ends), np.array(noise)

e0, n0 = run(0)
e1, n1 = run(12345)
m0, m1 = e0.mean(axis=0), e1.mean(axis=0)
out = {
    "cta_end_us": {"mean": float(m0.mean()), "min": float(m0.min()), "max": float(m0.max()), "per_step_max_mean": float(e0.max(axis=1).mean())},
    "corr_end_vs_noise_frames_seed0": float(np.corrcoef(m0, n0.mean(axis=0))[0, 1]),
    "corr_end_vs_noise_frames_seed1": float(np.corrcoef(m1, n1.mean(axis=0))[0, 1]),
    "corr_end_seed0_vs_seed1 (same SMs, other games)": float(np.corrcoef(m0, m1)[0, 1]),
    "end_by_noise_count_seed0": {str(k): float(m0[np.round(n0.mean(axis=0)) == k].mean()) for k in range(0, 6) if (np.round(n0.mean(axis=0)) == k).any()},
}
print(json.dumps(out, indent=1))
