"""Adversarial initial states for BASELINE.json configs[4] (collision-stress sweep) — TEST INFRASTRUCTURE.

Shared by tests/test_env_gpu.py and tools/stress_sweep.py.  States are a function of the GLOBAL env id only, so any
sharding of the ids over GPUs sweeps the same set of trajectories.
"""
import numpy as np

KINDS = ("corner arc", "grazing side wall", "goal post", "mallet rim", "coincident with mallet", "anywhere, top speed")


def stress_state(env_id, seed=5):
    """float32[12] puck(x,y,vx,vy) | bottom mallet | top mallet for global env `env_id`: puck at (or beyond) top speed
    aimed at a corner arc, grazing a side wall, at a goal post, on a mallet's rim, exactly on a mallet, or anywhere."""
    rs = np.random.RandomState((seed * 1000003 + int(env_id)) & 0x7FFFFFFF)
    s = np.zeros(12, dtype=np.float32)
    s[4:6] = (225, 600)
    s[8:10] = (225, 100)
    kind = int(env_id) % 6
    ang = rs.uniform(0, 2 * np.pi)
    v = 18.0 * np.array([np.cos(ang), np.sin(ang)])
    if kind == 0:    # next to a corner arc
        cx, cy = rs.choice([85, 365]), rs.choice([85, 615])
        p = np.array([cx, cy]) + rs.uniform(-1, 1, 2) * 30
    elif kind == 1:  # grazing a side wall
        p = np.array([rs.choice([44, 406]), rs.uniform(100, 600)])
        v = np.array([rs.choice([-1, 1]) * 0.3, rs.choice([-18, 18])])
    elif kind == 2:  # at a goal post
        p = np.array([rs.choice([163, 287]) + rs.uniform(-3, 3), rs.choice([50, 650])])
        v = np.array([rs.uniform(-2, 2), 17 if p[1] > 350 else -17])
    elif kind == 3:  # on a mallet's rim
        m = s[4:6] if (int(env_id) // 6) % 2 else s[8:10]
        p = m + 42.0 * np.array([np.cos(ang), np.sin(ang)])
        v = -v
    elif kind == 4:  # exactly on top of a mallet (coincident centres)
        p = s[4:6].copy()
        v = np.zeros(2)
    else:
        p = np.array([rs.uniform(43, 407), rs.uniform(43, 657)])
    s[0:2] = p
    s[2:4] = v
    return s


def stress_states(lo, hi, seed=5):
    """float32[hi - lo, 12] for global env ids lo .. hi-1."""
    return np.stack([stress_state(i, seed) for i in range(lo, hi)])
