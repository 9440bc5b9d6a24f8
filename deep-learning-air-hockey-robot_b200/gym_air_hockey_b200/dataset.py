"""Supervised dataset generation: the reference's `ai/supervised/generate_data.py:35-73` loop run
over N environments at once, written in the reference's dataset layout (`ai/supervised/utils/data_utils.py:3-13`):
three arrays `states`, `robot_actions`, `human_actions`.

Reference: one env, `env.step()` with no actions (the programmed AI plays both sides, `generate_data.py:51`), reset on
a goal or every 200 steps (`:57-58`), states as float32 (n, 9, 128, 128) normalised `(x-128)/128`, labels int8, one
HDF5 file `<time>_<mode>_<n>.h5` (`:68-73`).  Here: the batched engine with `auto_reset` and `max_episode_steps=200`
plays the same protocol per env; every step the stacked observation of all envs is materialised on the GPU
(`ahk_observe`), copied to one of two pinned host buffers on a side stream while the next step runs, copied once more into
the output by a writer thread.  The container is the reference's: one HDF5 file holding the three datasets, written by
the package's own HDF5 writer (h5lite.py; `h5py` is not in this image) — the datasets are declared with their final
shape and every pinned block is written straight to its place in the file by positional writes (no shard buffer, no
concatenation).  `shard_states > 0` writes several such files, `container='npz'` NumPy archives instead (`load_data`
reads all of them).

Sample order is step-major: sample `t * n_envs + e` is env `e` after step `t`.  `states='u8'` keeps the uint8 stack
(4x smaller; `load_data(..., normalise=True)` applies `(x-128)/128` on load) — the float32 form costs 589 824 B per
sample of PCIe and disk, which is what bounds this path, not the GPU.
"""
import datetime
import glob
import os
import queue
import threading
import time

import numpy as np
import torch

from . import h5lite


def data_file_name(mode, n_states, when=None, ext=None):
    """`<YYYY-mm-dd-HH-MM-SS>_<mode>_<n>.h5` (generate_data.py:62-63,72)."""
    when = int(time.time()) if when is None else when
    stamp = datetime.datetime.fromtimestamp(when).strftime('%Y-%m-%d-%H-%M-%S')
    return '%s_%s_%d.%s' % (stamp, mode, n_states, ext or 'h5')


def save_data(data_file, states, robot_actions, human_actions):
    """≙ data_utils.save_data (data_utils.py:3-7)."""
    if not data_file.endswith('.npz'):
        with h5lite.File(data_file, 'w') as f:
            f.create_dataset('states', data=states)
            f.create_dataset('robot_actions', data=robot_actions)
            f.create_dataset('human_actions', data=human_actions)
    else:
        with open(data_file, 'wb') as f:
            np.savez(f, states=states, robot_actions=robot_actions, human_actions=human_actions)


def load_data(data_file, normalise=False):
    """≙ data_utils.load_data (data_utils.py:9-13).  `data_file` may be one file or a glob of shards (concatenated in
    name order).  normalise=True turns uint8 states into the reference's float32 `(x-128)/128` (processor.py:39-42)."""
    files = sorted(glob.glob(data_file)) or [data_file]
    parts = []
    for name in files:
        if h5lite.is_hdf5(name):
            with h5lite.File(name, 'r') as f:
                parts.append((f['states'][:], f['robot_actions'][:], f['human_actions'][:]))
        else:
            with np.load(name) as f:
                parts.append((f['states'], f['robot_actions'], f['human_actions']))
    states, robot, human = (np.concatenate([p[k] for p in parts]) if len(parts) > 1 else parts[0][k] for k in range(3))
    if normalise and states.dtype == np.uint8:
        states = (states.astype(np.float32) - 128.0) / 128.0
    return states, robot, human


class _Writer(threading.Thread):
    """Takes (states, robot, human) blocks — views of the pinned host buffers — and puts them into the output.

    HDF5 (default): the current file's three datasets were declared with their final shape when the file was opened, so
    a block goes straight from the pinned buffer to its place in the file — its state rows split over `savers` threads
    of positional writes (`os.pwrite` releases the GIL) — and the buffer is released as soon as those return.
    `.npz`: blocks are copied into a shard buffer and full shards handed to a pool of saver threads (`np.savez`: CRC32 +
    file I/O), at most `savers` shards in flight."""

    def __init__(self, out_dir, mode, shard_states, when, ext, total_states, savers=4):
        super().__init__(daemon=True)
        self.out_dir, self.mode, self.shard_states, self.when, self.ext = out_dir, mode, shard_states, when, ext or 'h5'
        self.total = int(total_states)
        self.q = queue.Queue(maxsize=2)
        self.files, self.error = [], None
        self._named = {}
        self._buf, self._fill, self._shard, self._done = None, 0, 0, 0
        self._n_savers = max(1, int(savers))
        self._slots = threading.Semaphore(self._n_savers)
        self._savers = []
        self._lock = threading.Lock()
        self._h5 = None

    def _capacity(self):
        left = self.total - self._done
        return min(self.shard_states, left) if self.shard_states else left

    def _path(self, shard, n):
        name = data_file_name(self.mode, n, self.when, self.ext)
        if shard or self.shard_states:
            base, ext = os.path.splitext(name)
            name = '%s_part%04d%s' % (base, shard, ext)
        return os.path.join(self.out_dir, name)

    # -- .npz shards ------------------------------------------------------------------------------------------------
    def _save(self, shard, states, robot, human):
        try:
            path = self._path(shard, len(states))
            save_data(path, states, robot, human)
            with self._lock:
                self._named[shard] = path
        except Exception as e:  # surfaced by generate()
            self.error = e
        finally:
            self._slots.release()

    def _flush(self):
        if self._buf is None or self._fill == 0:
            return
        states, robot, human = (a[:self._fill] for a in self._buf)
        self._slots.acquire()
        t = threading.Thread(target=self._save, args=(self._shard, states, robot, human), daemon=True)
        t.start()
        self._savers.append(t)
        self._done += self._fill
        self._buf, self._fill, self._shard = None, 0, self._shard + 1

    def _put_npz(self, states, robot, human):
        at = 0
        while at < len(robot):
            if self._buf is None:
                cap = self._capacity()
                self._buf = (np.empty((cap,) + states.shape[1:], dtype=states.dtype),
                             np.empty((cap,), dtype=robot.dtype), np.empty((cap,), dtype=human.dtype))
            k = min(len(robot) - at, len(self._buf[1]) - self._fill)
            for dst, src in zip(self._buf, (states, robot, human)):
                dst[self._fill:self._fill + k] = src[at:at + k]
            self._fill += k
            at += k
            if self._fill == len(self._buf[1]):
                self._flush()

    # -- HDF5, written in place -------------------------------------------------------------------------------------
    def _close_h5(self):
        if self._h5 is not None:
            f, _ds, cap = self._h5
            f.close()
            self._named[self._shard] = f.filename
            self._done += cap
            self._h5, self._fill, self._shard = None, 0, self._shard + 1

    def _put_h5(self, states, robot, human, pool):
        at = 0
        while at < len(robot):
            if self._h5 is None:
                cap = self._capacity()
                f = h5lite.File(self._path(self._shard, cap), 'w')
                ds = (f.create_dataset('states', (cap,) + states.shape[1:], dtype=states.dtype),
                      f.create_dataset('robot_actions', (cap,), dtype=robot.dtype),
                      f.create_dataset('human_actions', (cap,), dtype=human.dtype))
                self._h5, self._fill = (f, ds, cap), 0
            _f, ds, cap = self._h5
            k = min(len(robot) - at, cap - self._fill)
            per = (k + self._n_savers - 1) // self._n_savers
            jobs = [pool.submit(ds[0].write_rows, self._fill + lo, states[at + lo:at + min(lo + per, k)])
                    for lo in range(0, k, per)]
            ds[1].write_rows(self._fill, robot[at:at + k])
            ds[2].write_rows(self._fill, human[at:at + k])
            for j in jobs:
                j.result()
            self._fill += k
            at += k
            if self._fill == cap:
                self._close_h5()

    def run(self):
        from concurrent.futures import ThreadPoolExecutor
        pool = None
        try:
            if self.ext != 'npz':
                pool = ThreadPoolExecutor(self._n_savers)
            while True:
                item = self.q.get()
                if item is None:
                    break
                states, robot, human, release = item
                if pool is not None:
                    self._put_h5(states, robot, human, pool)
                else:
                    self._put_npz(states, robot, human)
                release.set()  # the pinned buffer may be refilled
            self._flush()
            self._close_h5()
            for t in self._savers:
                t.join()
            self.files = [self._named[k] for k in sorted(self._named)]
        except Exception as e:  # surfaced by generate()
            self.error = e
            while self.q.get() is not None:
                pass
        finally:
            if pool is not None:
                pool.shutdown()


def generate(env, n_states, out_dir, states='f32', shard_states=0, when=None, keep_in_memory=False,
             label_random_as_undefined=False, writer_threads=4, container='h5'):
    """Run the data-generation loop on `env` (a BatchedAirHockeyEnv built with auto_reset=True, max_episode_steps=200)
    until at least n_states samples exist (whole steps: ceil(n_states / n_envs) * n_envs samples are produced and the
    surplus dropped), writing `states` / `robot_actions` / `human_actions`.

    states: 'f32' = the reference's normalised float32 stack (mode 'rgb') — or float64 gray-diff when env.processor.mode
    is 'gray-diff' (processor.py:53-64); 'u8' = the raw uint8 stack.  shard_states: samples per output file (0 = one file).
    label_random_as_undefined: samples of `random` episodes (pure-noise frames, air_hockey_env.py:73-74,147-151; one reset
    in twenty, :28) get label 9 = "Undefined" on both sides — the class the reference only obtains by mixing foreign
    images into the data afterwards (mix_random_images_with_data.py:44-71).
    writer_threads: concurrent positional writes per block (HDF5) / shards written concurrently (.npz).
    container: 'h5' (the reference's, data_utils.py:3-7) or 'npz'.
    -> dict(files, samples, seconds, samples_per_s, bytes_to_host).  keep_in_memory also returns the arrays (tests)."""
    if states not in ('f32', 'u8'):
        raise ValueError("states must be 'f32' or 'u8'")
    if container not in ('h5', 'npz'):
        raise ValueError("container must be 'h5' or 'npz'")
    os.makedirs(out_dir, exist_ok=True)
    n = env.n_envs
    mode = env.processor.mode
    fmt = 'u8' if states == 'u8' else mode
    steps = (int(n_states) + n - 1) // n
    dev = env.device
    copy_stream = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream(dev)

    def pinned(shape, dtype):
        return torch.empty(shape, dtype=dtype).pin_memory()

    probe = env.observation(fmt)
    slots = []
    for _ in range(2):
        slots.append({'obs': pinned(tuple(probe.shape), probe.dtype),
                      'robot': pinned((n,), torch.uint8), 'human': pinned((n,), torch.uint8),
                      'ready': torch.cuda.Event(), 'free': threading.Event()})
        slots[-1]['free'].set()
    del probe
    writer = _Writer(out_dir, mode, int(shard_states), when, container, int(n_states), savers=writer_threads)
    writer.start()
    kept = []
    t0 = time.perf_counter()
    pending = None
    for t in range(steps + 1):
        if t < steps:
            s = slots[t & 1]
            s['free'].wait()       # the writer has consumed this slot's previous contents
            s['free'].clear()
            env.step()
            # the stacked observation must be taken before the next step overwrites the oldest ring slot
            obs = env.observation(fmt)
            rl, hl = env.robot_label.clone(), env.human_label.clone()
            if label_random_as_undefined:
                noise = ((env.state[15] >> 16) & 3) == 2  # episode type in the meta word
                rl[noise] = 9
                hl[noise] = 9
            done_step = torch.cuda.Event()
            done_step.record(main)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(done_step)
                s['obs'].copy_(obs, non_blocking=True)
                s['robot'].copy_(rl, non_blocking=True)
                s['human'].copy_(hl, non_blocking=True)
                s['ready'].record(copy_stream)
            obs.record_stream(copy_stream)
            rl.record_stream(copy_stream)
            hl.record_stream(copy_stream)
        if pending is not None:    # hand the previous step's buffers to the writer while this step runs
            p = pending
            p['ready'].synchronize()
            take = min(n, int(n_states) - (t - 1) * n)
            item = (p['obs'].numpy()[:take], p['robot'].numpy()[:take].view(np.int8), p['human'].numpy()[:take].view(np.int8), p['free'])
            if keep_in_memory:
                kept.append(tuple(a.copy() for a in item[:3]))
            writer.q.put(item)
        pending = slots[t & 1] if t < steps else None
    writer.q.put(None)
    writer.join()
    if writer.error is not None:
        raise writer.error
    dt = time.perf_counter() - t0
    samples = int(n_states)
    per_sample = int(np.prod(slots[0]['obs'].shape[1:])) * slots[0]['obs'].element_size() + 2
    out = {'files': writer.files, 'samples': samples, 'seconds': dt, 'samples_per_s': samples / dt,
           'bytes_to_host': samples * per_sample, 'steps': steps}
    if keep_in_memory:
        out['arrays'] = tuple(np.concatenate([k[i] for k in kept]) for i in range(3))
    return out
