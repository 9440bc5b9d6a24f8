"""Generate tests/golden/*.npz by running the REFERENCE's own code (unmodified, imported from /root/reference):

    gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py   AirHockeyEnv  (reset / step / reward table / labels)
    gym-air-hockey/gym_air_hockey/processor.py             DataProcessor (resize, CHW, stack, normalise, gray-diff)

on top of oracle/air_hockey (the stand-in for the MISSING simulator submodule, backed by oracle/ahk_oracle.c) and
oracle/gym_shim (a 20-line `gym`).  Run it in the build container only — /root/reference does not exist on the GPU
box; the fixtures it writes are committed.

    python tests/golden/make_golden.py

Pillow: the reference calls Image.resize((128,128)) with no filter (processor.py:35).  When it was written (Pillow
< 7.0) that meant NEAREST; Pillow 12 would pick BICUBIC.  The default is pinned back to NEAREST here, by wrapping
Image.Image.resize — the reference source is not touched.
"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "gym_shim"),
                "/root/reference/gym-air-hockey", os.path.join(ROOT, "deep-learning-air-hockey-robot_b200")]
OUT = os.path.dirname(os.path.abspath(__file__))

from PIL import Image  # noqa: E402

_resize = Image.Image.resize


def _resize_2018(self, size, resample=None, *args, **kwargs):
    return _resize(self, size, Image.NEAREST if resample is None else resample, *args, **kwargs)


Image.Image.resize = _resize_2018

import air_hockey  # noqa: E402  (oracle stand-in)
import orc  # noqa: E402
from gym_air_hockey_b200.assets import make_table_image  # noqa: E402  (data: the table picture)

SEED, ENV_ID = 20181104, 7
air_hockey.SEED, air_hockey.ENV_ID = SEED, ENV_ID
air_hockey.TABLE = make_table_image()

import gym  # noqa: E402  (shim)
import gym_air_hockey  # noqa: E402  (THE REFERENCE)
from gym_air_hockey import DataProcessor  # noqa: E402


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def to_u8(state):
    """(x-128)/128 float32 -> uint8, asserting it is lossless."""
    u = np.rint(state.astype(np.float64) * 128.0 + 128.0).astype(np.uint8)
    assert np.array_equal((u.astype(np.float32) - 128.0) / 128.0, state)
    return u


def use_mask(d):
    return (1 if d['arm'] else 0) | (2 if d['puck'] else 0) | (4 if d['top_mallet'] else 0)


EP = {'normal': 0, 'missing': 1, 'random': 2}


# puck | bottom mallet | top mallet (x, y, vx, vy): two shots nobody can stop, so that the fixtures contain goals
INJECT = {
    101: np.array([225, 640, 0, 15, 100, 500, 0, 0, 225, 100, 0, 0], dtype=np.float32),   # into the bottom goal: 'top' scores
    131: np.array([230, 60, 0.5, -15, 225, 600, 0, 0, 80, 300, 0, 0], dtype=np.float32),  # into the top goal: 'bottom' scores
    301: np.array([200, 645, -1, 16, 380, 420, 0, 0, 225, 100, 0, 0], dtype=np.float32),
}


def wrapper_trajectory(mode, steps, reset_every, keep, inject=()):
    """The generate_data.py:44-58 loop, plus caller-chosen actions in later phases (reinforcement/train.py:94-96)."""
    np.random.seed(SEED & 0xFFFFFFFF)
    env = gym.make('AirHockey-v0')     # AirHockeyEnv.__init__ -> AirHockey() + reset()
    resets = [(-1, EP[env.episode], use_mask(env.use_object))]  # (before step index, episode, mask)
    env.update(mode=mode)              # generate_data.py:36 — zeroes the frame stack (quirk B.2)
    rng = np.random.RandomState(99)
    rec = {k: [] for k in ('robot_in', 'human_in', 'reward', 'terminal', 'robot', 'human', 'crc', 'is_random', 'state12')}
    kept, kept_at = [], []
    reset = False
    for i in range(steps):
        if reset:
            reset = False
            env.reset()
            resets.append((i, EP[env.episode], use_mask(env.use_object)))
        if i in inject:
            env.game.sim.state[:] = INJECT[i]
        phase = i * 3 // steps
        robot_in = None if phase == 0 else int(rng.randint(10))
        human_in = None if phase < 2 else int(rng.randint(10))
        state, reward, terminal, info = env.step(robot_action=robot_in, human_action=human_in)
        rec['robot_in'].append(255 if robot_in is None else robot_in)
        rec['human_in'].append(255 if human_in is None else human_in)
        rec['reward'].append(reward)
        rec['terminal'].append(terminal)
        rec['robot'].append(info['robot_action'])
        rec['human'].append(info['human_action'])
        rec['is_random'].append(env.episode == 'random')
        rec['crc'].append(0 if env.episode == 'random' else crc(state))
        rec['state12'].append(env.game.sim.state.copy())
        if i in keep and env.episode != 'random':
            kept.append(to_u8(state) if mode == 'rgb' else state.copy())
            kept_at.append(i)
        if terminal or i % reset_every == 0:
            reset = True
    assert state.dtype == (np.float32 if mode == 'rgb' else np.float64)
    return dict(seed=SEED, env_id=ENV_ID, mode=mode, resets=np.array(resets, dtype=np.int32),
                robot_in=np.array(rec['robot_in'], dtype=np.uint8), human_in=np.array(rec['human_in'], dtype=np.uint8),
                reward=np.array(rec['reward'], dtype=np.float64), terminal=np.array(rec['terminal'], dtype=np.uint8),
                robot=np.array(rec['robot'], dtype=np.uint8), human=np.array(rec['human'], dtype=np.uint8),
                is_random=np.array(rec['is_random'], dtype=np.uint8), crc=np.array(rec['crc'], dtype=np.uint32),
                state12=np.array(rec['state12'], dtype=np.float32),
                kept=np.array(kept), kept_at=np.array(kept_at, dtype=np.int32),
                inject_at=np.array(sorted(inject), dtype=np.int32),
                inject_state=np.array([INJECT[i] for i in sorted(inject)], dtype=np.float32).reshape(-1, 12))


class _FakeInfo(object):
    pass


class _FakeGame(object):
    """Feeds chosen GameInfo values to the reference's step() so its reward table can be enumerated."""

    def __init__(self):
        self.next = None

    def step(self, robot_action=None, human_action=None, use_object=None):
        return self.next


def reward_table():
    """air_hockey_env.py:95-148 enumerated: episode x scored x 5 flags x 9 reachable robot labels."""
    env = gym.make('AirHockey-v0')
    game = _FakeGame()
    env.game = game
    frame = np.zeros((8, 8, 3), dtype=np.uint8)
    rows = []
    for episode in ('normal', 'missing', 'random'):
        for scored in (None, 'top', 'bottom'):
            for flags in range(32):
                for label in range(9):
                    g = _FakeInfo()
                    g.frame = frame
                    g.scored = scored
                    g.robot_action = env.processor.actions[label]
                    g.human_action = env.processor.actions[4]
                    g.puck_was_hit = bool(flags & 1)
                    g.hit_the_border = bool(flags & 2)
                    g.puck_is_at_the_bottom = bool(flags & 4)
                    g.distance_decreased = bool(flags & 8)
                    g.in_the_target = bool(flags & 16)
                    game.next = g
                    env.episode = episode
                    _, reward, terminal, info = env.step()
                    assert info['robot_action'] == label
                    rows.append((EP[episode], {None: 0, 'top': 1, 'bottom': 2}[scored], flags, label, reward, terminal))
    return np.array(rows, dtype=np.float64)


def processor_vectors():
    """DataProcessor on random native-size frames: both modes, 5 pushes each (processor.py:31-66)."""
    rs = np.random.RandomState(4242)
    frames = rs.randint(0, 256, size=(5, 700, 450, 3)).astype(np.uint8)
    out = {'frames_seed': 4242}
    for mode in ('rgb', 'gray-diff'):
        p = DataProcessor(mode=mode)
        crcs, last = [], None
        for f in frames:
            last = p.process_observation(f)
            crcs.append(crc(last))
        out[mode + '_crc'] = np.array(crcs, dtype=np.uint32)
        out[mode + '_last'] = to_u8(last) if mode == 'rgb' else last
        out[mode + '_dtype'] = str(last.dtype)
    p = DataProcessor()
    out['labels'] = np.array([p.action_to_label(p.process_action(l)) for l in range(10)], dtype=np.int32)
    out['actions'] = p.actions.copy()
    # a small odd-sized frame, to pin the resize index formula away from 450x700
    small = rs.randint(0, 256, size=(37, 53, 3)).astype(np.uint8)
    out['small'] = small
    out['small_resized'] = np.array(Image.fromarray(small).resize((128, 128))).transpose(2, 0, 1)
    return out


if __name__ == '__main__':
    import io
    import contextlib
    with contextlib.redirect_stdout(io.StringIO()):  # the reference prints the episode type on every reset
        rgb = wrapper_trajectory('rgb', steps=360, reset_every=40, keep={0, 1, 2, 45, 104, 200, 359}, inject=(101, 131, 301))
        gray = wrapper_trajectory('gray-diff', steps=60, reset_every=25, keep={3, 59})
        table = reward_table()
        proc = processor_vectors()
    np.savez_compressed(os.path.join(OUT, 'wrapper_rgb.npz'), **rgb)
    np.savez_compressed(os.path.join(OUT, 'wrapper_gray_diff.npz'), **gray)
    np.savez_compressed(os.path.join(OUT, 'reward_table.npz'), table=table)
    np.savez_compressed(os.path.join(OUT, 'processor.npz'), **proc)
    for f in sorted(os.listdir(OUT)):
        if f.endswith('.npz'):
            print(f, os.path.getsize(os.path.join(OUT, f)))
    print('episodes seen (rgb):', rgb['resets'][:, 1].tolist())
    print('goals (rgb):', int(rgb['terminal'].sum()), 'labels robot:', np.bincount(rgb['robot'], minlength=10).tolist())
