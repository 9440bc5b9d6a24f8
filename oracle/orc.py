"""ctypes binding of oracle/libahk_oracle.so — TEST INFRASTRUCTURE (see the header of ahk_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libahk_oracle.so")

W, H, DIM, FRAME = 450, 700, 128, 3 * 128 * 128
USE_ARM, USE_PUCK, USE_TOP, USE_ALL = 1, 2, 4, 7
EP_NORMAL, EP_MISSING, EP_RANDOM = 0, 1, 2
ACT_NONE = 255
SCORED = {0: None, 1: "top", 2: "bottom"}


def build(force=False):
    src = os.path.join(HERE, "ahk_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "-s", "CC=gcc"] + (["-B"] if force else []))
    return LIB_PATH


class Info(C.Structure):  # orc_info
    _fields_ = [("scored", C.c_int32), ("robot_action", C.c_int8 * 2), ("human_action", C.c_int8 * 2),
                ("puck_was_hit", C.c_uint8), ("hit_the_border", C.c_uint8), ("puck_is_at_the_bottom", C.c_uint8),
                ("distance_decreased", C.c_uint8), ("in_the_target", C.c_uint8)]


class StepOut(C.Structure):  # orc_step_out
    _fields_ = [("reward", C.c_float), ("terminal", C.c_uint8), ("robot_label", C.c_uint8),
                ("human_label", C.c_uint8), ("scored", C.c_uint8), ("flags", C.c_uint8)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    L = C.CDLL(LIB_PATH)
    vp, u8p, f32p, u32p = C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    L.orc_philox.argtypes = [u32p, u32p, u32p]
    L.orc_sizeof_sim.restype = C.c_int
    L.orc_sim_init.argtypes = [vp, C.c_uint64, C.c_uint32]
    L.orc_sim_reset.argtypes = [vp]
    L.orc_sim_step.argtypes = [vp, C.POINTER(C.c_int8), C.POINTER(C.c_int8), C.c_int, C.POINTER(Info)]
    L.orc_draw.argtypes = [vp, C.c_int, u8p, u8p]
    L.orc_resize_nearest_chw.argtypes = [u8p, C.c_int, C.c_int, u8p]
    L.orc_push_frame.argtypes = [f32p, u8p]
    L.orc_observe_rgb.argtypes = [f32p, f32p]
    L.orc_observe_gray_diff.argtypes = [f32p, C.POINTER(C.c_double)]
    L.orc_noise_frame.argtypes = [u32p, C.c_uint32, C.c_uint32, u8p]
    L.orc_action_to_label.argtypes = [C.POINTER(C.c_int8)]
    L.orc_action_to_label.restype = C.c_int
    L.orc_reward.argtypes = [C.c_int, C.POINTER(Info), C.c_int]
    L.orc_reward.restype = C.c_float
    L.orc_env_create.argtypes = [C.c_uint64, C.c_uint32, u8p]
    L.orc_env_create.restype = vp
    L.orc_env_destroy.argtypes = [vp]
    L.orc_env_reset.argtypes = [vp, C.c_int, C.c_int]
    L.orc_env_step.argtypes = [vp, C.c_int, C.c_int, C.POINTER(StepOut)]
    L.orc_env_step_auto.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(StepOut)]
    L.orc_env_set_pending.argtypes = [vp, C.c_int]
    L.orc_env_ring.argtypes = [vp]
    L.orc_env_ring.restype = u8p
    L.orc_env_native.argtypes = [vp]
    L.orc_env_native.restype = u8p
    L.orc_env_sim.argtypes = [vp]
    L.orc_env_sim.restype = vp
    L.orc_env_get_state.argtypes = [vp, f32p, u32p]
    L.orc_env_set_state.argtypes = [vp, f32p, u32p]
    L.orc_env_observe_rgb.argtypes = [vp, f32p]
    L.orc_run_batch.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_int, f32p, f32p, u8p, u8p]
    L.orc_run_batch.restype = C.c_uint64
    L.orc_pool_create.argtypes = [C.c_int]
    L.orc_pool_create.restype = vp
    L.orc_pool_destroy.argtypes = [vp]
    L.orc_pool_run_batch.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int, C.c_int, f32p, f32p, u8p, u8p]
    L.orc_pool_run_batch.restype = C.c_uint64
    L.orc_run_trace.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p, u8p, u32p, u32p]
    _lib = L
    return L


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib().orc_philox(_p(c, C.c_uint32), _p(k, C.c_uint32), _p(out, C.c_uint32))
    return out


def resize_nearest_chw(frame):
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    out = np.empty((3, DIM, DIM), dtype=np.uint8)
    lib().orc_resize_nearest_chw(_p(frame, C.c_uint8), frame.shape[0], frame.shape[1], _p(out, C.c_uint8))
    return out


def noise_frame(seed, env_id, noise_count):
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)
    out = np.empty((3, DIM, DIM), dtype=np.uint8)
    lib().orc_noise_frame(_p(key, C.c_uint32), env_id, noise_count, _p(out, C.c_uint8))
    return out


class Stack:
    """DataProcessor's float32 3-frame shift register (processor.py:23,47-64)."""

    def __init__(self):
        self.buf = np.zeros((9, DIM, DIM), dtype=np.float32)

    def push(self, chw):
        chw = np.ascontiguousarray(chw, dtype=np.uint8)
        lib().orc_push_frame(_p(self.buf, C.c_float), _p(chw, C.c_uint8))

    def rgb(self):
        out = np.empty((9, DIM, DIM), dtype=np.float32)
        lib().orc_observe_rgb(_p(self.buf, C.c_float), _p(out, C.c_float))
        return out

    def gray_diff(self):
        out = np.empty((1, DIM, DIM), dtype=np.float64)
        lib().orc_observe_gray_diff(_p(self.buf, C.c_float), _p(out, C.c_double))
        return out


class Sim:
    """The bare game (≙ the missing `air_hockey.AirHockey`): SPEC.md §5-§7."""

    def __init__(self, seed=0, env_id=0, table=None):
        L = lib()
        self._buf = C.create_string_buffer(L.orc_sizeof_sim())
        self._p = C.cast(self._buf, C.c_void_p)
        L.orc_sim_init(self._p, seed, env_id)
        self.table = np.ascontiguousarray(table, dtype=np.uint8) if table is not None else None

    @property
    def state(self):
        """float32[12] view: puck(x,y,vx,vy), bottom, top."""
        return np.frombuffer(self._buf, dtype=np.float32, count=12)

    @property
    def serve_count(self):
        return int(np.frombuffer(self._buf, dtype=np.uint32, count=14)[12])

    @serve_count.setter
    def serve_count(self, v):
        np.frombuffer(self._buf, dtype=np.uint32, count=14)[12] = v

    def reset(self):
        lib().orc_sim_reset(self._p)

    def step(self, robot=None, human=None, use=USE_ALL):
        info = Info()
        r = (C.c_int8 * 2)(*[int(v) for v in robot]) if robot is not None else None
        h = (C.c_int8 * 2)(*[int(v) for v in human]) if human is not None else None
        lib().orc_sim_step(self._p, r, h, use, C.byref(info))
        return info

    def draw(self, use=USE_ALL):
        frame = np.empty((H, W, 3), dtype=np.uint8)
        lib().orc_draw(self._p, use, _p(self.table, C.c_uint8), _p(frame, C.c_uint8))
        return frame


class Env:
    """One env with the reference wrapper's semantics (SPEC.md §9), uint8 ring instead of the float stack."""

    def __init__(self, seed=0, env_id=0, table=None):
        self.table = np.ascontiguousarray(table, dtype=np.uint8)
        assert self.table.shape == (H, W, 3)
        self._h = lib().orc_env_create(seed, env_id, _p(self.table, C.c_uint8))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_env_destroy(self._h)
            self._h = None

    def reset(self, episode=-1, use=USE_ALL):
        lib().orc_env_reset(self._h, episode, use)

    def step(self, robot=ACT_NONE, human=ACT_NONE):
        o = StepOut()
        lib().orc_env_step(self._h, robot, human, C.byref(o))
        return o

    def step_auto(self, robot=ACT_NONE, human=ACT_NONE, max_episode_steps=0):
        o = StepOut()
        lib().orc_env_step_auto(self._h, robot, human, max_episode_steps, C.byref(o))
        return o

    def ring(self):
        """uint8[3][3][128][128], newest first (copy)."""
        p = lib().orc_env_ring(self._h)
        return np.ctypeslib.as_array(p, shape=(3, 3, DIM, DIM)).copy()

    def get_state(self):
        s = np.empty(12, dtype=np.float32)
        m = np.empty(8, dtype=np.uint32)
        lib().orc_env_get_state(self._h, _p(s, C.c_float), _p(m, C.c_uint32))
        return s, m

    def set_state(self, s, m):
        s = np.ascontiguousarray(s, dtype=np.float32)
        m = np.ascontiguousarray(m, dtype=np.uint32)
        lib().orc_env_set_state(self._h, _p(s, C.c_float), _p(m, C.c_uint32))

    def observe_rgb(self):
        out = np.empty((9, DIM, DIM), dtype=np.float32)
        lib().orc_env_observe_rgb(self._h, _p(out, C.c_float))
        return out


def run_batch(envs, steps, max_episode_steps=200, threads=1, want_obs=True):
    """Timed CPU baseline: every env does `steps` auto-reset AI-vs-AI steps; returns (checksum, obs, reward, done, labels)."""
    n = len(envs)
    handles = (C.c_void_p * n)(*[e._h for e in envs])
    obs = np.empty((n, 9, DIM, DIM), dtype=np.float32) if want_obs else None
    reward = np.empty(n, dtype=np.float32)
    done = np.empty(n, dtype=np.uint8)
    labels = np.empty((n, 2), dtype=np.uint8)
    chk = lib().orc_run_batch(handles, n, steps, max_episode_steps, threads,
                              _p(obs, C.c_float) if want_obs else None, _p(reward, C.c_float),
                              _p(done, C.c_uint8), _p(labels, C.c_uint8))
    return chk, obs, reward, done, labels


class Pool:
    """A persistent pool of worker threads plus preallocated outputs: the timed CPU arm (bench.py) calls `run` in a loop
    without creating threads or arrays inside the timed region; `trace` logs long horizons for the stress sweep."""

    def __init__(self, envs, threads=1, want_obs=True):
        self.envs, self.n = envs, len(envs)
        self.threads = max(1, min(int(threads), self.n))
        self._p = lib().orc_pool_create(self.threads)
        self._handles = (C.c_void_p * self.n)(*[e._h for e in envs])
        self.obs = np.empty((self.n, 9, DIM, DIM), dtype=np.float32) if want_obs else None
        self.reward = np.empty(self.n, dtype=np.float32)
        self.done = np.empty(self.n, dtype=np.uint8)
        self.labels = np.empty((self.n, 2), dtype=np.uint8)

    def run(self, steps, max_episode_steps=200):
        """Every env does `steps` auto-reset AI-vs-AI steps, each producing what the reference's env.step() hands back
        (float32 (9,128,128) observation, reward, terminal, two labels).  -> checksum."""
        return lib().orc_pool_run_batch(self._p, self._handles, self.n, steps, max_episode_steps,
                                        _p(self.obs, C.c_float) if self.obs is not None else None,
                                        _p(self.reward, C.c_float), _p(self.done, C.c_uint8), _p(self.labels, C.c_uint8))

    def trace(self, steps, max_episode_steps=0, auto_reset=True, draw_last=3, want_state=True, want_out=True,
              robot=None, human=None):
        """robot / human: optional uint8[steps, n] action labels (255 = programmed AI).
        -> (state uint32[steps,16,n] in the engine's SoA word layout, out uint32[steps,n] packed step results)."""
        state = np.empty((steps, 16, self.n), dtype=np.uint32) if want_state else None
        out = np.empty((steps, self.n), dtype=np.uint32) if want_out else None
        r = None if robot is None else np.ascontiguousarray(robot, dtype=np.uint8).reshape(steps, self.n)
        h = None if human is None else np.ascontiguousarray(human, dtype=np.uint8).reshape(steps, self.n)
        lib().orc_run_trace(self._p, self._handles, self.n, steps, max_episode_steps, int(auto_reset), draw_last,
                            None if r is None else _p(r, C.c_uint8), None if h is None else _p(h, C.c_uint8),
                            _p(state, C.c_uint32) if want_state else None, _p(out, C.c_uint32) if want_out else None)
        return state, out

    def close(self):
        if getattr(self, "_p", None):
            lib().orc_pool_destroy(self._p)
            self._p = None

    def __del__(self):
        self.close()


def unpack_out(out):
    """Fields of Pool.trace's packed step results."""
    return {"reward": ((out & 7).astype(np.float32) - 2.0) * 0.5, "terminal": (out >> 3) & 1, "robot_label": (out >> 4) & 15,
            "human_label": (out >> 8) & 15, "flags": (out >> 12) & 31, "scored": (out >> 17) & 3}


# ------------------------------------------------------------------------------------------------ ahk_cpu_* (include/ahk_cpu.h)
class CpuConfig(C.Structure):  # ahk_config (include/ahk.h)
    _fields_ = [("n_envs", C.c_int32), ("device", C.c_int32), ("seed", C.c_uint64), ("env_id_base", C.c_uint32),
                ("auto_reset", C.c_int32), ("max_episode_steps", C.c_int32), ("ring_depth", C.c_int32), ("reserved", C.c_int32 * 4)]


class CpuBuffers(C.Structure):  # ahk_buffers
    _fields_ = [(name, C.c_void_p) for name in
                ("state", "frames", "reward", "done", "robot_label", "human_label", "info", "stats", "drawlist", "warm")]


class CpuStepLog(C.Structure):  # ahk_step_log
    _fields_ = [("robot_label", C.c_void_p), ("human_label", C.c_void_p), ("done", C.c_void_p), ("reward", C.c_void_p)]


# every symbol include/ahk_cpu.h declares: name -> (restype, argtypes); the twin of gym_air_hockey_b200._lib.SYMBOLS
_vp, _i = C.c_void_p, C.c_int
CPU_SYMBOLS = {
    "ahk_cpu_create": (_i, [C.POINTER(CpuConfig), C.POINTER(_vp)]),
    "ahk_cpu_destroy": (None, [_vp]),
    "ahk_cpu_threads": (_i, [_vp]),
    "ahk_cpu_set_table_image": (_i, [_vp, _vp, _i, _i]),
    "ahk_cpu_bind_buffers": (_i, [_vp, C.POINTER(CpuBuffers)]),
    "ahk_cpu_reset": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "ahk_cpu_step": (_i, [_vp, _vp, _vp, _vp]),
    "ahk_cpu_step_many": (_i, [_vp, _i, C.POINTER(CpuStepLog), _vp]),
    "ahk_cpu_observe": (_i, [_vp, _i, _vp, _vp]),
    "ahk_cpu_bind_observation": (_i, [_vp, _i, _vp]),
    "ahk_cpu_head": (_i, [_vp]),
    "ahk_cpu_last_error": (C.c_char_p, []),
}


class CpuEngine(object):
    """The batched engine's call protocol on the host cores, through the ahk_cpu_* twins of the C ABI: numpy arrays in
    the layouts of ahk_buffers (state [16, n] uint32, frames [3, n, 3, 128, 128] uint8 ring with the newest frame in
    slot `head`, reward / done / labels [n])."""

    def __init__(self, n_envs, seed=0, table=None, auto_reset=True, max_episode_steps=0, env_id_base=0):
        L = lib()
        for name, (res, args) in CPU_SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        self.L, self.n = L, int(n_envs)
        cfg = CpuConfig(n_envs=self.n, device=-1, seed=seed, env_id_base=env_id_base, auto_reset=int(auto_reset),
                        max_episode_steps=int(max_episode_steps), ring_depth=3)
        h = C.c_void_p()
        self._check(L.ahk_cpu_create(C.byref(cfg), C.byref(h)))
        self._h = h
        table = np.ascontiguousarray(table, dtype=np.uint8)
        self._check(L.ahk_cpu_set_table_image(h, table.ctypes.data_as(_vp), table.shape[0], table.shape[1]))
        n = self.n
        self.state = np.zeros((16, n), np.uint32)
        self.frames = np.zeros((3, n, 3, 128, 128), np.uint8)
        self.reward = np.zeros(n, np.float32)
        self.done, self.robot_label, self.human_label, self.info = (np.zeros(n, np.uint8) for _ in range(4))
        self.stats = np.zeros(32, np.uint64)
        b = CpuBuffers(**{k: getattr(self, k).ctypes.data_as(_vp) for k in
                          ("state", "frames", "reward", "done", "robot_label", "human_label", "info", "stats")})
        self._check(L.ahk_cpu_bind_buffers(h, C.byref(b)))

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("ahk_cpu: %s (%d)" % (self.L.ahk_cpu_last_error().decode(), rc))

    @property
    def head(self):
        return self.L.ahk_cpu_head(self._h)

    @property
    def threads(self):
        return self.L.ahk_cpu_threads(self._h)

    @staticmethod
    def _u8(a):
        return None if a is None else np.ascontiguousarray(a, dtype=np.uint8)

    def reset(self, mask=None, episode=None, use=None):
        m, e, u = self._u8(mask), self._u8(episode), self._u8(use)
        self._check(self.L.ahk_cpu_reset(self._h, *(None if a is None else a.ctypes.data_as(_vp) for a in (m, e, u)), None))

    def step(self, robot=None, human=None):
        r, h = self._u8(robot), self._u8(human)
        self._check(self.L.ahk_cpu_step(self._h, *(None if a is None else a.ctypes.data_as(_vp) for a in (r, h)), None))

    def step_many(self, steps, log=None):
        lg = None
        if log is not None:
            lg = C.byref(CpuStepLog(**{k: v.ctypes.data_as(_vp) for k, v in log.items()}))
        self._check(self.L.ahk_cpu_step_many(self._h, int(steps), lg, None))

    def observe(self, fmt, out=None):
        """fmt 0: uint8 [n, 9, 128, 128]; 1: float32 (x - 128) / 128; 2: float64 gray-diff [n, 1, 128, 128]."""
        shape, dt = [((self.n, 9, 128, 128), np.uint8), ((self.n, 9, 128, 128), np.float32), ((self.n, 1, 128, 128), np.float64)][fmt]
        out = np.empty(shape, dt) if out is None else out
        self._check(self.L.ahk_cpu_observe(self._h, int(fmt), out.ctypes.data_as(_vp), None))
        return out

    def bind_observation(self, fmt, out):
        """Every step also writes each env's observation (format as observe()) into `out`; out None unbinds."""
        self._obs = out  # keeps the buffer alive
        self._check(self.L.ahk_cpu_bind_observation(self._h, int(fmt), None if out is None else out.ctypes.data_as(_vp)))

    def close(self):
        if getattr(self, "_h", None):
            self.L.ahk_cpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
