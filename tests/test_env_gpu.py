"""GPU parity tests: the CUDA engine, called through the C ABI (include/ahk.h), against the CPU oracle on the same
seeded inputs — bit-exact for labels, rewards, terminals, flags, pixels AND the float32 state (SPEC.md §1)."""
import zlib

import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def make_env(n, table, **kw):
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    return BatchedAirHockeyEnv(n, table=table, **kw)


def oracle_envs(orc, n, table, seed, base=0):
    return [orc.Env(seed=seed, env_id=base + i, table=table) for i in range(n)]


def ring_newest_first(env):
    """uint8 [N, 3(age), 3, 128, 128] from the device ring."""
    fr = env.frames.cpu().numpy()
    h = env.head
    return np.stack([fr[(h + 3 - age) % 3] for age in range(3)], axis=1)


def assert_same_state(env, oracles, where=""):
    bodies, meta = env.get_state()
    for i, o in enumerate(oracles):
        s, m = o.get_state()
        assert np.array_equal(bodies[i].view(np.uint32), s.view(np.uint32)), (where, i, bodies[i], s)
        got = (meta['serve_count'][i], meta['reset_count'][i], meta['noise_count'][i], meta['episode_steps'][i],
               meta['episode'][i], meta['use'][i], meta['pending'][i])
        assert got == tuple(int(v) for v in m[:7]), (where, i, got, m)


def run_lockstep(env, oracles, steps, actions=None, max_episode_steps=0, auto=True, check_frames_every=1):
    n = len(oracles)
    goals = 0
    for t in range(steps):
        robot = human = None
        if actions is not None:
            robot, human = actions[t]
        _, reward, done, info = env.step(robot, human)
        reward, done = reward.cpu().numpy(), done.cpu().numpy()
        rl, hl, bits = info['robot_action'].cpu().numpy(), info['human_action'].cpu().numpy(), env.info_bits.cpu().numpy()
        newest = env.frames[env.head].cpu().numpy() if t % check_frames_every == 0 else None
        for i, o in enumerate(oracles):
            r = 255 if robot is None else int(robot[i])
            h = 255 if human is None else int(human[i])
            out = o.step_auto(r, h, max_episode_steps) if auto else o.step(r, h)
            assert (reward[i], bool(done[i]), rl[i], hl[i]) == (out.reward, bool(out.terminal), out.robot_label, out.human_label), (t, i)
            assert bits[i] == (out.flags | out.scored << 5), (t, i)
            goals += out.terminal
            if newest is not None:
                assert np.array_equal(newest[i], o.ring()[0]), (t, i)
        assert_same_state(env, oracles, "step %d" % t)
    return goals


def test_lockstep_ai_vs_ai(cuda, orc, table):
    """configs[1] in small: AI on both sides, auto-reset on goals and every 40 steps, sampled episode types."""
    n, seed = 96, 11
    env = make_env(n, table, seed=seed, auto_reset=True, max_episode_steps=40)
    oracles = oracle_envs(orc, n, table, seed)
    env.reset()
    for o in oracles:
        o.reset()
    assert_same_state(env, oracles, "reset")
    ring = ring_newest_first(env)
    for i, o in enumerate(oracles):
        assert np.array_equal(ring[i], o.ring()), i
    run_lockstep(env, oracles, 130, max_episode_steps=40)
    ring = ring_newest_first(env)
    for i, o in enumerate(oracles):
        assert np.array_equal(ring[i], o.ring()), i
    _, meta = env.get_state()
    assert set(meta['episode'].tolist()) == {0, 1, 2}      # all three episode types occurred
    st = env.statistics()
    assert st['steps'] == n * 130 and sum(st['robot_label_hist']) == n * 130


def test_lockstep_caller_actions(cuda, orc, noise_table):
    """Caller-supplied labels 0..9 for both sides (reinforcement/train.py:94-96), noise table image, no auto-reset."""
    n, seed, steps = 64, 3, 120
    rs = np.random.RandomState(1)
    env = make_env(n, noise_table, seed=seed, auto_reset=False, max_episode_steps=0)
    oracles = oracle_envs(orc, n, noise_table, seed)
    ep = np.zeros(n, dtype=np.uint8)
    use = np.full(n, 7, dtype=np.uint8)
    env.reset(episode=ep, use_object=use)
    for o in oracles:
        o.reset(0, 7)
    actions = []
    for t in range(steps):
        robot = rs.randint(0, 10, size=n).astype(np.uint8)
        human = rs.randint(0, 10, size=n).astype(np.uint8) if t % 3 else None
        actions.append((robot, human))
    run_lockstep(env, oracles, steps, actions=actions, auto=False)


def test_forced_episodes_and_masks(cuda, orc, noise_table):
    """Every episode type and every object mask (air_hockey_env.py:32-63), including the warm-up quirk B.1."""
    combos = [(0, 7)] + [(1, u) for u in (0, 4, 2, 6, 1, 5)] + [(2, 7)]
    n = len(combos)
    env = make_env(n, noise_table, seed=21, auto_reset=False, max_episode_steps=0)
    oracles = oracle_envs(orc, n, noise_table, 21)
    env.reset(episode=np.array([c[0] for c in combos], dtype=np.uint8), use_object=np.array([c[1] for c in combos], dtype=np.uint8))
    for o, (e, u) in zip(oracles, combos):
        o.reset(e, u)
    ring = ring_newest_first(env)
    for i, o in enumerate(oracles):
        assert np.array_equal(ring[i], o.ring()), combos[i]
    run_lockstep(env, oracles, 40, auto=False)
    # masked reset: only env 0 and 3 restart
    mask = np.zeros(n, dtype=np.uint8)
    mask[[0, 3]] = 1
    env.reset(mask=mask)
    for i in (0, 3):
        oracles[i].reset()
    assert_same_state(env, oracles, "masked reset")
    ring = ring_newest_first(env)
    for i, o in enumerate(oracles):
        assert np.array_equal(ring[i], o.ring()), i


@pytest.mark.parametrize("n,limit,steps", [(333, 17, 45), (1184, 5, 42), (1184, 1, 30)])
def test_back_to_back_steps_overlap(cuda, orc, noise_table, n, limit, steps):
    """Step kernels launched back to back overlap (the next launch's game tick runs under this launch's rasterising,
    csrc/ahk_step.cu): no host synchronisation for `steps` steps, then state, the whole 3-frame ring, the last step's
    outputs and the statistics must equal the oracle's.  333 envs: a grid smaller than the device, wait before the first
    store; 1184 envs: one CTA per SM, the wait deferred to the end of the kernel, and a reset every 5 steps so that the
    frames rewriting older ring slots (the ones that must still wait) are exercised in every launch; limit 1: every env
    resets in every step, so every frame of every launch is one that has to wait."""
    seed = 21
    env = make_env(n, noise_table, seed=seed, max_episode_steps=limit)
    oracles = oracle_envs(orc, n, noise_table, seed)
    env.reset()
    for _ in range(steps):
        _, reward, done, info = env.step()
    cuda.cuda.synchronize()
    outs = []
    for o in oracles:
        o.reset()
        for _ in range(steps):
            out = o.step_auto(255, 255, limit)
        outs.append(out)
    assert_same_state(env, oracles, "after %d unsynchronised steps" % steps)
    ring = ring_newest_first(env)
    for i, o in enumerate(oracles):
        assert np.array_equal(ring[i], o.ring()), i
    assert np.array_equal(reward.cpu().numpy(), np.array([o.reward for o in outs], dtype=np.float32))
    assert np.array_equal(done.cpu().numpy(), np.array([bool(o.terminal) for o in outs]))
    assert np.array_equal(info['robot_action'].cpu().numpy(), np.array([o.robot_label for o in outs], dtype=np.uint8))
    assert env.statistics()['steps'] == n * steps


def stress_states(n, rs):
    """configs[4]: pucks at top speed aimed at corners, walls at grazing angles, posts and mallet edges."""
    s = np.zeros((n, 12), dtype=np.float32)
    s[:, 4:6] = (225, 600)
    s[:, 8:10] = (225, 100)
    for i in range(n):
        kind = i % 6
        ang = rs.uniform(0, 2 * np.pi)
        v = 18.0 * np.array([np.cos(ang), np.sin(ang)])
        if kind == 0:   # next to a corner arc
            cx, cy = rs.choice([85, 365]), rs.choice([85, 615])
            p = np.array([cx, cy]) + rs.uniform(-1, 1, 2) * 30
        elif kind == 1:  # grazing a side wall
            p = np.array([rs.choice([44, 406]), rs.uniform(100, 600)]); v = np.array([rs.choice([-1, 1]) * 0.3, rs.choice([-18, 18])])
        elif kind == 2:  # at a goal post
            p = np.array([rs.choice([163, 287]) + rs.uniform(-3, 3), rs.choice([50, 650])]); v = np.array([rs.uniform(-2, 2), 17 if p[1] > 350 else -17])
        elif kind == 3:  # on a mallet's rim
            m = s[i, 4:6] if i % 2 else s[i, 8:10]
            p = m + 42.0 * np.array([np.cos(ang), np.sin(ang)]); v = -v
        elif kind == 4:  # exactly on top of a mallet (coincident centres)
            p = s[i, 4:6].copy(); v = np.zeros(2)
        else:
            p = np.array([rs.uniform(43, 407), rs.uniform(43, 657)])
        s[i, 0:2] = p
        s[i, 2:4] = v
    return s


def test_collision_stress(cuda, orc, table):
    n, seed = 60, 8
    rs = np.random.RandomState(5)
    env = make_env(n, table, seed=seed, auto_reset=False, max_episode_steps=0)
    oracles = oracle_envs(orc, n, table, seed)
    env.reset(episode=np.zeros(n, dtype=np.uint8), use_object=np.full(n, 7, dtype=np.uint8))
    s = stress_states(n, rs)
    env.set_state(s)
    for i, o in enumerate(oracles):
        o.reset(0, 7)
        _, m = o.get_state()
        o.set_state(s[i], m)
    goals = run_lockstep(env, oracles, 250, auto=False, check_frames_every=25)
    assert goals >= 1


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


@pytest.mark.parametrize("name", ["wrapper_rgb.npz", "wrapper_gray_diff.npz"])
def test_drop_in_env_golden(cuda, table, name):
    """The n_envs == 1 adapter replays the REFERENCE wrapper's recorded run (tests/golden/make_golden.py): same
    float observations (CRC every step), rewards, terminals and labels, same dtypes."""
    from gym_air_hockey_b200 import make
    g = golden(name)
    mode = str(g["mode"])
    ep_names = ('normal', 'missing', 'random')
    resets = {int(i): (ep_names[int(e)], {'arm': bool(u & 1), 'puck': bool(u & 2), 'top_mallet': bool(u & 4)}) for i, e, u in g["resets"]}
    inject = {int(i): s for i, s in zip(g["inject_at"], g["inject_state"])}
    kept = {int(i): k for i, k in zip(g["kept_at"], g["kept"])}
    env = make('AirHockey-v0', seed=int(g["seed"]), env_id=int(g["env_id"]), table=table, verbose=False)
    # the constructor's own reset sampled from the engine's stream; redo it with the reference's recorded draw.
    # (serve_count must restart from the constructor's single serve for the puck placements to line up)
    env._b.set_state(None, serve_count=1, reset_count=0, noise_count=0)
    env.reset(*resets[-1])
    env.update(mode=mode)
    for i in range(len(g["reward"])):
        if i in resets:
            env.reset(*resets[i])
        if i in inject:
            env._b.set_state(inject[i][None, :])
        r, h = int(g["robot_in"][i]), int(g["human_in"][i])
        state, reward, terminal, info = env.step(None if r == 255 else r, None if h == 255 else h)
        assert isinstance(reward, float) and isinstance(terminal, bool) and isinstance(info['robot_action'], int)
        assert (reward, terminal, info['robot_action'], info['human_action']) == (g["reward"][i], bool(g["terminal"][i]), g["robot"][i], g["human"][i]), i
        bodies, _ = env._b.get_state()
        assert np.array_equal(bodies[0].view(np.uint32), g["state12"][i].view(np.uint32)), i
        if g["is_random"][i]:
            assert state.dtype == np.float64 and state.shape == (9, 128, 128) and abs(state.mean()) < 1e-9
            continue
        assert state.dtype == (np.float32 if mode == 'rgb' else np.float64)
        assert state.shape == ((9, 128, 128) if mode == 'rgb' else (1, 128, 128))
        assert crc(state) == g["crc"][i], i
        if i in kept:
            want = (kept[i].astype(np.float32) - 128.0) / 128.0 if mode == 'rgb' else kept[i]
            assert np.array_equal(state, want)
    env.close()


def test_observation_forms(cuda, orc, noise_table):
    """ahk_observe: u8 stack, (x-128)/128 float32, gray-diff float64 == the oracle's restatement of processor.py:52-64."""
    n = 5
    env = make_env(n, noise_table, seed=2, auto_reset=False, max_episode_steps=0)
    env.reset(episode=np.zeros(n, dtype=np.uint8))
    for _ in range(4):
        env.step()
    ring = ring_newest_first(env)
    u8 = env.observation('u8').cpu().numpy()
    assert np.array_equal(u8, ring.reshape(n, 9, 128, 128))
    f32 = env.observation('rgb').cpu().numpy()
    assert f32.dtype == np.float32 and np.array_equal(f32, (u8.astype(np.float32) - 128.0) / 128.0)
    g64 = env.observation('gray-diff').cpu().numpy()
    assert g64.dtype == np.float64 and g64.shape == (n, 1, 128, 128)
    for i in range(n):
        st = orc.Stack()
        for age in (2, 1, 0):
            st.push(ring[i, age])
        assert np.array_equal(g64[i], st.gray_diff())
    with pytest.raises(ValueError, match='Processing mode is not supported'):
        env.observation('hsv')


def test_step_host_matches_device_path(cuda, table):
    """ahk_step_host (host buffers in and out) gives the same results as the stream-ordered device call."""
    n = 300
    a = make_env(n, table, seed=4, max_episode_steps=30)
    b = make_env(n, table, seed=4, max_episode_steps=30)
    a.reset(); b.reset()
    rs = np.random.RandomState(0)
    out = b.host_buffers(want_frame=True)
    for t in range(45):
        robot = rs.randint(0, 10, size=n).astype(np.uint8) if t % 2 else None
        _, reward, done, info = a.step(robot, None)
        b.step_host(robot, None, out)
        assert np.array_equal(reward.cpu().numpy(), out['reward'])
        assert np.array_equal(done.cpu().numpy().astype(np.uint8), out['done'])
        assert np.array_equal(info['robot_action'].cpu().numpy(), out['robot_action'])
        assert np.array_equal(info['human_action'].cpu().numpy(), out['human_action'])
        assert np.array_equal(a.frames[a.head].cpu().numpy(), out['frame'])


@pytest.mark.parametrize("n", [300, 1500])
def test_step_host_results_early_return(cuda, table, n):
    """ahk_step_host_results returns when reward / done / labels are in the host buffers (the kernel's game ticks are
    over, its frames may not be): every step's results, fed back as the next actions, equal the device path's; after
    host_sync the ring and the state do too.  300 envs: a grid smaller than the device; 1500: one CTA per SM."""
    a = make_env(n, table, seed=13, max_episode_steps=20)
    b = make_env(n, table, seed=13, max_episode_steps=20)
    a.reset(); b.reset()
    cuda.cuda.synchronize()
    out = b.host_buffers()
    act = cuda.zeros(n, dtype=cuda.uint8).pin_memory().numpy()
    act[:] = 4
    for t in range(60):
        _, reward, done, info = a.step(act.copy(), None)
        b.step_host_results(act, None, out)
        assert np.array_equal(reward.cpu().numpy(), out['reward']), t
        assert np.array_equal(done.cpu().numpy().astype(np.uint8), out['done']), t
        assert np.array_equal(info['robot_action'].cpu().numpy(), out['robot_action']), t
        assert np.array_equal(info['human_action'].cpu().numpy(), out['human_action']), t
        act[:] = (out['human_action'] + out['robot_action'] + t) % 10   # the next actions depend on this step's results
    # a device-path call on another stream is ordered behind the launch still rasterising on the engine's stream
    _, reward, _, _ = a.step(None, None)
    _, reward_b, _, _ = b.step(None, None)
    obs_b = b.observation('u8')
    assert np.array_equal(reward.cpu().numpy(), reward_b.cpu().numpy())
    assert np.array_equal(a.observation('u8').cpu().numpy(), obs_b.cpu().numpy())
    b.host_sync()
    cuda.cuda.synchronize()
    assert b.head == a.head
    assert np.array_equal(a.frames.cpu().numpy(), b.frames.cpu().numpy())
    assert np.array_equal(a.get_state()[0].view(np.uint32), b.get_state()[0].view(np.uint32))
    pageable = {k: np.empty_like(v) for k, v in out.items()}              # pageable buffers: falls back to the full wait
    _, reward, _, _ = a.step(None, None)
    b.step_host_results(None, None, pageable)
    assert np.array_equal(reward.cpu().numpy(), pageable['reward'])


def test_step_host_pipelined_and_pageable(cuda, table):
    """ahk_step_host_begin/_wait with four steps in flight (pinned buffers: the kernel reads the actions and writes the
    results through mapped host memory) and plain pageable numpy buffers (copy engine) both replay the device path."""
    n = 700
    a = make_env(n, table, seed=11, max_episode_steps=25)
    b = make_env(n, table, seed=11, max_episode_steps=25)
    c = make_env(n, table, seed=11, max_episode_steps=25)
    for e in (a, b, c):
        e.reset()
    rs = np.random.RandomState(3)
    depth, steps = 4, 40
    outs = [b.host_buffers() for _ in range(depth)]
    acts = [cuda.empty(n, dtype=cuda.uint8).pin_memory().numpy() for _ in range(depth)]
    pageable = {k: np.empty_like(v) for k, v in outs[0].items()}
    expect, tickets = [], []
    for t in range(steps + depth):
        if t >= depth:  # retire the oldest step in flight
            k = (t - depth) % depth
            b.step_host_wait(tickets[t - depth])
            for key, want in zip(('reward', 'done', 'robot_action', 'human_action'), expect[t - depth]):
                assert np.array_equal(outs[k][key], want), (t, key)
        if t < steps:
            k = t % depth
            acts[k][:] = rs.randint(0, 10, size=n)
            _, reward, done, info = a.step(acts[k].copy(), acts[k].copy())
            expect.append((reward.cpu().numpy(), done.cpu().numpy().astype(np.uint8), info['robot_action'].cpu().numpy(), info['human_action'].cpu().numpy()))
            tickets.append(b.step_host_begin(acts[k], acts[k], outs[k]))
            c.step_host(acts[k].copy(), acts[k].copy(), pageable)
            for key, want in zip(('reward', 'done', 'robot_action', 'human_action'), expect[-1]):
                assert np.array_equal(pageable[key], want), (t, key)
    assert np.array_equal(a.frames.cpu().numpy(), b.frames.cpu().numpy())
    assert np.array_equal(a.get_state()[0].view(np.uint32), b.get_state()[0].view(np.uint32))
    assert np.array_equal(b.reward.cpu().numpy(), expect[-1][0])  # the device copies are written as well


def test_full_size_properties(cuda, table):
    """configs[1] at full size (4096 envs): properties that need no oracle — determinism, independence of sharding
    (global env ids, SPEC.md §4), bodies inside the rink, statistics that add up."""
    n, steps = 4096, 260
    whole = make_env(n, table, seed=77, max_episode_steps=200)
    whole.reset()
    for _ in range(steps):
        whole.step()
    bodies, meta = whole.get_state()
    # two shards with global ids continue the same sequences
    parts = [make_env(n // 2, table, seed=77, env_id_base=base, max_episode_steps=200) for base in (0, n // 2)]
    for p in parts:
        p.reset()
        for _ in range(steps):
            p.step()
    b2 = np.concatenate([p.get_state()[0] for p in parts])
    assert np.array_equal(bodies.view(np.uint32), b2.view(np.uint32))
    f_whole = whole.frames[whole.head].cpu().numpy()
    f_parts = np.concatenate([p.frames[p.head].cpu().numpy() for p in parts])
    assert np.array_equal(f_whole, f_parts)
    # geometry invariants
    assert (bodies[:, 4] >= 49).all() and (bodies[:, 4] <= 401).all() and (bodies[:, 5] >= 374).all() and (bodies[:, 5] <= 651).all()
    assert (bodies[:, 8] >= 49).all() and (bodies[:, 8] <= 401).all() and (bodies[:, 9] >= 49).all() and (bodies[:, 9] <= 326).all()
    assert (bodies[:, 2] ** 2 + bodies[:, 3] ** 2 <= 18.0 ** 2 * 1.0001).all()
    st = whole.statistics()
    assert st['steps'] == n * steps
    assert sum(st['robot_label_hist']) == n * steps and sum(st['human_label_hist']) == n * steps
    assert st['episodes'] >= 2 * n           # initial reset + the 200-step limit
    assert st['goals_top'] + st['goals_bottom'] > 0
    assert st['robot_label_hist'][9] == 0    # label 9 never comes from the simulator
    import os
    assert whole.launches == (2 if os.environ.get('AHK_PATH') == 'split' else 1) * (steps + 1)
