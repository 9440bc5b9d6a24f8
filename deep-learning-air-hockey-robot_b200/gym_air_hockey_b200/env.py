"""Batched air-hockey environments on one B200, behind the reference's gym-air-hockey Env API.

`BatchedAirHockeyEnv` ≙ N copies of `AirHockeyEnv` (gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py:8-155)
stepped by two CUDA kernels (csrc/ahk_kernels.cu) through the C ABI in include/ahk.h.  PyTorch only owns the
device memory and the stream.  `AirHockeyEnv` is the n_envs == 1 adapter with the reference's exact return types,
so `ai/supervised/generate_data.py:44-58` and `ai/reinforcement/train.py:80-118` style loops run unchanged.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .assets import make_table_image
from .processor import DataProcessor

EPISODES = ('normal', 'missing', 'random')  # air_hockey_env.py:28


class Discrete(object):
    """≙ gym.spaces.Discrete (gym itself is not a dependency of the engine)."""

    def __init__(self, n):
        self.n = n


class Box(object):
    """≙ gym.spaces.Box as the reference uses it (air_hockey_env.py:17)."""

    def __init__(self, low, high, shape):
        self.low, self.high, self.shape = low, high, shape


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class BatchedAirHockeyEnv(object):
    """N independent environments on one GPU.

    reset()/step() keep the reference's names and meaning; actions are uint8 labels 0..9 per env (a CUDA tensor,
    a numpy array, or None = the programmed AI plays that side, air_hockey_env.py:86-89); results are CUDA tensors:
    reward float32[N], terminal bool[N], info {'robot_action': uint8[N], 'human_action': uint8[N]}.
    The observation lives in `self.frames` — uint8 [3, N, 3, 128, 128], a ring whose slot `self.head` is the newest
    frame — and is never copied on the step path; `observation(fmt)` materialises the stacked forms.

    ring_depth > 3 makes the ring DEEP: frames [ring_depth, N, 3, 128, 128], slot = step mod ring_depth, every frame
    written once and kept for ring_depth steps — the ring then IS the replay memory of the RL loop (rl_agents.py:59-90;
    `rl.ReplayStore` becomes a view of it) and the staging area of `step_many`.  A reset then files its three warm-up
    frames in `self.warm` [3, N, 3, 128, 128] instead of rewriting the two older slots (they hold the previous episode's
    last frames), so the 3-frame stack of an env is found through `stack_view()` / `observation()` / the policy kernels,
    not by indexing `frames` alone.
    """

    def __init__(self, n_envs, mode='rgb', device=0, seed=0, env_id_base=0, auto_reset=True, max_episode_steps=200,
                 table=None, obs='ring', ring_depth=3):
        self.lib = _lib.load()
        self._ptr_cache = {}
        if not torch.cuda.is_available():
            raise RuntimeError('BatchedAirHockeyEnv needs a CUDA device; there is no CPU fallback')
        self.processor = DataProcessor(mode=mode)
        self.processor.check_mode()
        self.n_envs = int(n_envs)
        self.env_id_base = int(env_id_base)
        self.device = torch.device('cuda', device if isinstance(device, int) else torch.device(device).index or 0)
        self.obs_format = obs
        self.ring_depth = int(ring_depth)
        if self.ring_depth < 3:
            raise ValueError('ring_depth must be at least 3 (the 3-frame stack, processor.py:23)')
        self.height = self.width = self.processor.dim
        self.action_space = Discrete(len(self.processor.actions))
        self.observation_space = Box(0, 255, (9, self.height, self.width))
        self.reward_range = (-1.0, 1.0)
        self.n_actions = self.action_space.n

        n, dev = self.n_envs, self.device
        with torch.cuda.device(dev):
            self.state = torch.zeros((_lib.STATE_WORDS, n), dtype=torch.int32, device=dev)
            self.frames = torch.zeros((self.ring_depth, n, 3, 128, 128), dtype=torch.uint8, device=dev)
            self.warm = torch.zeros((3, n, 3, 128, 128), dtype=torch.uint8, device=dev) if self.ring_depth > 3 else None
            self.reward = torch.zeros(n, dtype=torch.float32, device=dev)
            self.done = torch.zeros(n, dtype=torch.uint8, device=dev)
            self.robot_label = torch.zeros(n, dtype=torch.uint8, device=dev)
            self.human_label = torch.zeros(n, dtype=torch.uint8, device=dev)
            self.info_bits = torch.zeros(n, dtype=torch.uint8, device=dev)
            self.stats = torch.zeros(_lib.STATS_WORDS, dtype=torch.int64, device=dev)
            self.drawlist = torch.zeros((3, n, 4), dtype=torch.int32, device=dev)

        cfg = _lib.Config(n_envs=n, device=self.device.index, seed=seed, env_id_base=env_id_base,
                          auto_reset=int(bool(auto_reset)), max_episode_steps=int(max_episode_steps),
                          ring_depth=self.ring_depth)
        handle = C.c_void_p()
        _lib.check(self.lib.ahk_create(C.byref(cfg), C.byref(handle)))
        self._h = handle
        table = make_table_image() if table is None else np.ascontiguousarray(table, dtype=np.uint8)
        self.table = table
        _lib.check(self.lib.ahk_set_table_image(self._h, table.ctypes.data_as(C.c_void_p), table.shape[0], table.shape[1]))
        bufs = _lib.Buffers(state=self.state.data_ptr(), frames=self.frames.data_ptr(), reward=self.reward.data_ptr(),
                            done=self.done.data_ptr(), robot_label=self.robot_label.data_ptr(),
                            human_label=self.human_label.data_ptr(), info=self.info_bits.data_ptr(),
                            stats=self.stats.data_ptr(), drawlist=self.drawlist.data_ptr(),
                            warm=self.warm.data_ptr() if self.warm is not None else None)
        _lib.check(self.lib.ahk_bind_buffers(self._h, C.byref(bufs)))

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, '_h', None):
            self.lib.ahk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _labels(self, a):
        """None | tensor | array of labels -> uint8 CUDA tensor (kept alive by the caller's frame) or None."""
        if a is None:
            return None
        if not torch.is_tensor(a):
            a = torch.as_tensor(np.asarray(a, dtype=np.uint8))
        a = a.to(device=self.device, dtype=torch.uint8, non_blocking=True).contiguous()
        if a.numel() != self.n_envs:
            raise ValueError('expected %d action labels, got %d' % (self.n_envs, a.numel()))
        return a

    @property
    def head(self):
        return self.lib.ahk_head(self._h)

    def stack_view(self):
        """ahk_stack: where the current 3-frame stack of every env lives (ring + warm-up side buffer + episode steps)."""
        st = _lib.Stack()
        _lib.check(self.lib.ahk_stack_view(self._h, C.byref(st)))
        return st

    @property
    def launches(self):
        k = C.c_uint64()
        _lib.check(self.lib.ahk_launch_count(self._h, C.byref(k)))
        return k.value

    # ------------------------------------------------------------------ the Env API
    def reset(self, mask=None, episode=None, use_object=None):
        """≙ AirHockeyEnv.reset (air_hockey_env.py:26-76) for every env (or those where mask != 0).

        episode / use_object: optional uint8[N] to force the episode type (0 normal, 1 missing, 2 random, 255 sample)
        and the object mask (bit0 arm, bit1 puck, bit2 top_mallet) instead of sampling them (SPEC.md §4)."""
        m, e, u = self._labels(mask), self._labels(episode), self._labels(use_object)
        _lib.check(self.lib.ahk_reset(self._h, _ptr(m), _ptr(e), _ptr(u), self._stream()))
        return self._obs()

    def step(self, robot_action=None, human_action=None, debug=False):
        """≙ AirHockeyEnv.step (air_hockey_env.py:84-155)."""
        r, h = self._labels(robot_action), self._labels(human_action)
        _lib.check(self.lib.ahk_step(self._h, _ptr(r), _ptr(h), self._stream()))
        info = {'robot_action': self.robot_label, 'human_action': self.human_label}
        return self._obs(), self.reward, self.done.view(torch.bool), info

    def step_many(self, steps, log=None):
        """`steps` consecutive action-free steps in ONE call (ahk_step_many) — the data-generation loop
        (generate_data.py:44-58), where the programmed AI plays both sides.  log: optional dict of preallocated CUDA
        tensors [steps, N] under 'robot_action', 'human_action' (uint8), 'done' (uint8), 'reward' (float32) that receive
        every step's outputs.  Returns like step(): the outputs of the LAST step."""
        st = None
        if log:
            st = _lib.StepLog()
            for key, field, dt in (('robot_action', 'robot_label', torch.uint8), ('human_action', 'human_label', torch.uint8),
                                   ('done', 'done', torch.uint8), ('reward', 'reward', torch.float32)):
                t = log.get(key)
                if t is not None:
                    if t.dtype != dt or tuple(t.shape) != (int(steps), self.n_envs) or not t.is_contiguous() or t.device != self.device:
                        raise ValueError("log[%r] must be a contiguous %s CUDA tensor [steps, n_envs]" % (key, dt))
                    setattr(st, field, t.data_ptr())
        _lib.check(self.lib.ahk_step_many(self._h, int(steps), C.byref(st) if st is not None else None, self._stream()))
        info = {'robot_action': self.robot_label, 'human_action': self.human_label}
        return self._obs(), self.reward, self.done.view(torch.bool), info

    def step_profiled(self, robot_action=None, human_action=None):
        """step() with CUDA events around each kernel: -> (env_kernel ms, raster kernel ms).  For bench.py's roofline."""
        r, h = self._labels(robot_action), self._labels(human_action)
        env_ms, raster_ms = C.c_float(), C.c_float()
        _lib.check(self.lib.ahk_step_profiled(self._h, _ptr(r), _ptr(h), self._stream(), C.byref(env_ms), C.byref(raster_ms)))
        return env_ms.value, raster_ms.value

    def step_host(self, robot_action=None, human_action=None, out=None, want_frame=False):
        """The same step for a caller on the CPU: numpy uint8 labels in (or None), numpy results out, one C call
        (ahk_step_host) that copies in, steps, copies out and waits.  `out` is a dict of preallocated (ideally
        pinned) arrays 'reward','done','robot_action','human_action'[,'frame'] reused across calls."""
        if out is None:
            out = self.host_buffers(want_frame)
        p = self._host_ptr
        rc = self.lib.ahk_step_host(self._h, p(robot_action), p(human_action), p(out['reward']), p(out['done']),
                                    p(out['robot_action']), p(out['human_action']), p(out.get('frame')))
        if rc:
            _lib.check(rc)
        return out

    def step_host_results(self, robot_action=None, human_action=None, out=None):
        """step_host that returns as soon as reward / done / labels are in `out` (ahk_step_host_results): the kernel signals
        the end of its game ticks through mapped host memory and goes on rasterising.  For loops whose next actions depend
        on this step's results but which leave the frames on the device.  host_sync() waits for the frames too."""
        if out is None:
            out = self.host_buffers()
        p = self._host_ptr
        rc = self.lib.ahk_step_host_results(self._h, p(robot_action), p(human_action), p(out['reward']), p(out['done']),
                                            p(out['robot_action']), p(out['human_action']))
        if rc:
            _lib.check(rc)
        return out

    def host_sync(self):
        _lib.check(self.lib.ahk_host_sync(self._h))

    def _host_ptr(self, a):
        """Address of a host array, cached per array object: building a ctypes pointer costs more than the C call."""
        if a is None:
            return None
        hit = self._ptr_cache.get(id(a))
        if hit is None or hit[0] is not a:
            if len(self._ptr_cache) > 64:
                self._ptr_cache.clear()
            hit = (a, a.ctypes.data)
            self._ptr_cache[id(a)] = hit
        return hit[1]

    def step_host_begin(self, robot_action=None, human_action=None, out=None):
        """Enqueue a host-buffer step without waiting (ahk_step_host_begin) -> ticket for step_host_wait.  For loops whose
        next actions do not depend on this step's results (generate_data.py:51: the AI plays both sides): up to 8 steps
        may be in flight, each with its own `out` buffers."""
        p = self._host_ptr
        t = self.lib.ahk_step_host_begin(self._h, p(robot_action), p(human_action), p(out['reward']), p(out['done']),
                                         p(out['robot_action']), p(out['human_action']), p(out.get('frame')))
        if t < 0:
            _lib.check(t)
        return t

    def step_host_wait(self, ticket):
        rc = self.lib.ahk_step_host_wait(self._h, ticket)
        if rc:
            _lib.check(rc)

    def host_buffers(self, want_frame=False):
        """Pinned host arrays for step_host."""
        n = self.n_envs
        pin = lambda shape, dt: torch.empty(shape, dtype=dt).pin_memory().numpy()
        out = {'reward': pin((n,), torch.float32), 'done': pin((n,), torch.uint8),
               'robot_action': pin((n,), torch.uint8), 'human_action': pin((n,), torch.uint8)}
        if want_frame:
            out['frame'] = pin((n, 3, 128, 128), torch.uint8)
        return out

    def set_pending(self, pending):
        """Arm (1) or disarm (0) the auto-reset of each env: an armed env restarts at the beginning of its next step
        (ahk_set_pending) — what `if reset: env.reset()` does per env in the reference's loops (generate_data.py:47-49)."""
        p = self._labels(pending)
        _lib.check(self.lib.ahk_set_pending(self._h, _ptr(p), self._stream()))

    def update(self, video_file=None, mode=None):
        """≙ AirHockeyEnv.update (air_hockey_env.py:78-82).  A new mode means a new DataProcessor, whose frame stack
        starts as zeros (processor.py:23) — so the ring is cleared, as in the reference."""
        if video_file is not None:
            raise NotImplementedError('video recording lived in the missing simulator; the engine is headless')
        if mode is not None:
            self.processor = DataProcessor(mode=mode)
            self.processor.check_mode()
            self.frames.zero_()
            if self.warm is not None:
                self.warm.zero_()

    def observation(self, fmt=None):
        """Materialise the stacked observation: 'u8' uint8[N,9,128,128] (newest frame first, processor.py:47-49),
        'rgb' float32[N,9,128,128] = (x-128)/128 (processor.py:52), 'gray-diff' float64[N,1,128,128] (processor.py:53-64)."""
        fmt = fmt or self.processor.mode
        n, dev = self.n_envs, self.device
        if fmt == 'u8':
            out, code = torch.empty((n, 9, 128, 128), dtype=torch.uint8, device=dev), _lib.OBS_U8
        elif fmt == 'rgb':
            out, code = torch.empty((n, 9, 128, 128), dtype=torch.float32, device=dev), _lib.OBS_RGB_F32
        elif fmt == 'gray-diff':
            out, code = torch.empty((n, 1, 128, 128), dtype=torch.float64, device=dev), _lib.OBS_GRAY_DIFF_F64
        else:
            raise ValueError('Processing mode is not supported')
        _lib.check(self.lib.ahk_observe(self._h, code, _ptr(out), self._stream()))
        return out

    def _obs(self):
        if self.obs_format == 'ring':
            return self.frames
        return self.observation(self.obs_format if self.obs_format != 'mode' else None)

    # ------------------------------------------------------------------ state injection (tests, checkpoints)
    def get_state(self):
        """-> (float32[N,12] puck|bottom|top, dict of uint32[N] counters and meta fields) on the host."""
        st = self.state.cpu().numpy().view(np.uint32)
        meta = st[15]
        return (st[:12].view(np.float32).T.copy(),
                {'serve_count': st[12].copy(), 'reset_count': st[13].copy(), 'noise_count': st[14].copy(),
                 'episode_steps': meta & 0xFFFF, 'episode': (meta >> 16) & 3, 'use': (meta >> 18) & 7,
                 'pending': (meta >> 21) & 1})

    def set_state(self, bodies, serve_count=None, reset_count=None, noise_count=None, episode_steps=None,
                  episode=None, use=None, pending=None):
        cur_b, cur = self.get_state()
        n = self.n_envs
        pick = lambda v, k: np.broadcast_to(np.asarray(cur[k] if v is None else v, dtype=np.uint32), (n,))
        st = np.empty((_lib.STATE_WORDS, n), dtype=np.uint32)
        st[:12] = np.ascontiguousarray(np.asarray(cur_b if bodies is None else bodies, dtype=np.float32).reshape(n, 12).T).view(np.uint32)
        st[12], st[13], st[14] = pick(serve_count, 'serve_count'), pick(reset_count, 'reset_count'), pick(noise_count, 'noise_count')
        st[15] = (np.minimum(pick(episode_steps, 'episode_steps'), 0xFFFF) | pick(episode, 'episode') << 16 |
                  pick(use, 'use') << 18 | pick(pending, 'pending') << 21)
        self.state.copy_(torch.from_numpy(st.view(np.int32)))
        torch.cuda.current_stream(self.device).synchronize()  # written outside the engine: complete before the next launch

    def partition(self):
        """Env boundaries of the step kernel's CTAs right now (int32[ctas + 1]; the adaptive partition)."""
        buf = (C.c_int32 * 1024)()
        g = self.lib.ahk_get_partition(self._h, buf, 1024)
        if g < 0:
            _lib.check(g)
        return np.array(buf[:g + 1], dtype=np.int32)

    def statistics(self):
        """Running sums since construction (this GPU only; gather across ranks with parallel.gather_statistics)."""
        s = self.stats.cpu().numpy()
        return {'steps': int(s[0]), 'episodes': int(s[1]), 'goals_top': int(s[2]), 'goals_bottom': int(s[3]),
                'reward_sum': 0.5 * float(s[4]), 'robot_label_hist': s[5:15].tolist(), 'human_label_hist': s[15:25].tolist()}


class AirHockeyEnv(object):
    """Drop-in for the reference's single env (air_hockey_env.py:8-155): numpy in, numpy out, same types.

    state: float32 (9,128,128) in 'rgb' mode, float64 (1,128,128) in 'gray-diff' (processor.py:52,64);
    reward: float; terminal: bool; info: {'robot_action': int, 'human_action': int}.
    Differences, all forced by running on a GPU: episode type / object mask / serve are drawn from the engine's
    Philox stream instead of the global np.random one (SPEC.md §4); the `random` episode's noise likewise.
    """

    def __init__(self, device=0, seed=0, env_id=0, table=None, verbose=True):
        self._b = BatchedAirHockeyEnv(1, device=device, seed=seed, env_id_base=env_id, auto_reset=False,
                                      max_episode_steps=0, table=table)
        self.game = self._b  # the reference exposes the simulator as .game
        self.processor = self._b.processor
        self.height = self.width = self.processor.dim
        self.action_space = self._b.action_space
        self.observation_space = self._b.observation_space
        self.reward_range = (-1.0, 1.0)
        self.n_actions = self.action_space.n
        self.viewer = None
        self.verbose = verbose
        self.episode = None
        self.use_object = None
        self.reset()

    def _sync_episode(self):
        _, meta = self._b.get_state()
        self.episode = EPISODES[int(meta['episode'][0])]
        use = int(meta['use'][0])
        self.use_object = {'arm': bool(use & 1), 'puck': bool(use & 2), 'top_mallet': bool(use & 4)}

    def _state(self):
        return self._b.observation(self.processor.mode)[0].cpu().numpy()

    def reset(self, episode=None, use_object=None):
        e = None if episode is None else np.array([EPISODES.index(episode)], dtype=np.uint8)
        u = None
        if use_object is not None:
            u = np.array([(1 if use_object['arm'] else 0) | (2 if use_object['puck'] else 0) |
                          (4 if use_object['top_mallet'] else 0)], dtype=np.uint8)
        self._b.reset(episode=e, use_object=u)
        self._sync_episode()
        if self.verbose:
            print('Episode type: {}'.format(self.episode))  # air_hockey_env.py:30
        if self.episode == 'random':
            return self._b.observation('u8')[0].cpu().numpy()  # raw uint8 noise, air_hockey_env.py:74
        return self._state()

    def update(self, video_file=None, mode=None):
        self._b.update(video_file=video_file, mode=mode)
        self.processor = self._b.processor

    def step(self, robot_action=None, human_action=None, debug=False):
        r = None if robot_action is None else np.array([robot_action], dtype=np.uint8)
        h = None if human_action is None else np.array([human_action], dtype=np.uint8)
        _, reward, done, info = self._b.step(r, h)
        reward = float(np.clip(reward[0].item(), -1.0, 1.0))
        terminal = bool(done[0].item())
        labels = {'robot_action': int(info['robot_action'][0].item()), 'human_action': int(info['human_action'][0].item())}
        if self.episode == 'random':  # air_hockey_env.py:147-151: standardised noise, never terminal
            state = self._b.observation('u8')[0].cpu().numpy().astype(np.float64)
            state -= np.mean(state)
            state /= np.std(state)
            terminal = False
        else:
            state = self._state()
        return state, reward, terminal, labels

    def close(self):
        self._b.close()


_REGISTRY = {'AirHockey-v0': AirHockeyEnv}  # gym_air_hockey/__init__.py:3-6


def make(name, **kwargs):
    """≙ gym.make('AirHockey-v0')."""
    return _REGISTRY[name](**kwargs)
