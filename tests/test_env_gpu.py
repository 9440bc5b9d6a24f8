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


def test_collision_stress(cuda, orc, table):
    """configs[4] in small, with every frame checked: pucks at top speed aimed at corners, walls at grazing angles, posts
    and mallet edges (oracle/stress.py)."""
    import stress
    n, seed = 60, 8
    env = make_env(n, table, seed=seed, auto_reset=False, max_episode_steps=0)
    oracles = oracle_envs(orc, n, table, seed)
    env.reset(episode=np.zeros(n, dtype=np.uint8), use_object=np.full(n, 7, dtype=np.uint8))
    s = stress.stress_states(0, n)
    env.set_state(s)
    for i, o in enumerate(oracles):
        o.reset(0, 7)
        _, m = o.get_state()
        o.set_state(s[i], m)
    goals = run_lockstep(env, oracles, 250, auto=False, check_frames_every=25)
    assert goals >= 1


def test_stress_sweep_long_horizon(cuda, orc):
    """configs[4] as tools/stress_sweep.py runs it (there: 384 envs x 10 000 steps at 1/2/4/8 GPUs, profiles/stress_sweep_N*.json),
    here 96 envs x 2 000 steps: state compared with the oracle at every step, ring at every chunk end, and the per-env
    checksums of a 2-shard run (global ids) equal those of the whole."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("stress_sweep", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "stress_sweep.py"))
    sw = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sw)
    whole = sw.sweep(96, 2000, 500, threads=8)
    assert whole["first_divergence_step"] is None and whole["max_abs_delta"] == 0.0
    assert whole["output_mismatches"] == 0 and whole["frame_mismatches"] == 0 and whole["frames_checked"] == 4 * 96 * 3
    assert whole["goals"] > 0 and whole["env_resets"] > 96
    parts = [sw.sweep(96, 2000, 1000, rank=r, world=2, threads=8) for r in range(2)]
    assert all(p["first_divergence_step"] is None for p in parts)
    assert np.array_equal(np.concatenate([p["checksums"] for p in parts]), whole["checksums"])


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


def state_words(env, ids):
    """uint32 [16, len(ids)] of the engine's SoA state for the sampled local env ids."""
    return env.state.cpu().numpy().view(np.uint32)[:, ids]


def test_configs1_4096_envs_vs_oracle(cuda, orc, table):
    """BASELINE.json configs[1] at its own size: 4096 envs (28 per CTA — the benchmarked shape), AI on both sides,
    auto-reset on goals and every 200 steps (generate_data.py:44-58), 130 + 140 steps launched back to back with no host
    synchronisation in between.  Trajectories are independent and keyed by the global env id, so the oracle only runs a
    sample of 320 ids (first and last env, both sides of CTA boundaries 27|28, 55|56, ..., and 300 random ones): state
    bits, the whole 3-slot ring and the last step's outputs must be equal at both checkpoints."""
    n, seed, limit = 4096, 20181104, 200
    rs = np.random.RandomState(4096)
    ids = np.unique(np.concatenate([[0, 1, 27, 28, 55, 56, 2047, 2048, 4067, 4068, n - 2, n - 1], rs.choice(n, 300, replace=False)]))
    assert len(ids) >= 256
    env = make_env(n, table, seed=seed, auto_reset=True, max_episode_steps=limit)
    oracles = [orc.Env(seed=seed, env_id=int(i), table=table) for i in ids]
    pool = orc.Pool(oracles, threads=8, want_obs=False)
    env.reset()
    for o in oracles:
        o.reset()
    done_steps = 0
    for steps in (130, 140):
        for _ in range(steps):
            _, reward, done, info = env.step()
        cuda.cuda.synchronize()
        done_steps += steps
        st, out = pool.trace(steps, limit, auto_reset=True, draw_last=3)
        assert np.array_equal(state_words(env, ids), st[-1]), "state after %d steps" % done_steps
        u = orc.unpack_out(out[-1])
        assert np.array_equal(reward.cpu().numpy()[ids], u['reward'])
        assert np.array_equal(done.cpu().numpy()[ids].astype(np.uint32), u['terminal'])
        assert np.array_equal(info['robot_action'].cpu().numpy()[ids].astype(np.uint32), u['robot_label'])
        assert np.array_equal(info['human_action'].cpu().numpy()[ids].astype(np.uint32), u['human_label'])
        assert np.array_equal(env.info_bits.cpu().numpy()[ids].astype(np.uint32), u['flags'] | u['scored'] << 5)
        fr, h = env.frames[:, ids].cpu().numpy(), env.head
        for k, o in enumerate(oracles):
            ring = o.ring()
            for age in range(3):
                assert np.array_equal(fr[(h + 3 - age) % 3, k], ring[age]), (done_steps, int(ids[k]), age)
    _, meta = env.get_state()
    assert set(meta['episode'][ids].tolist()) == {0, 1, 2}  # the sample saw all three episode types
    assert env.statistics()['steps'] == n * done_steps
    pool.close()


def test_producer_kernel_before_step_two_engines_one_stream(cuda, orc, noise_table):
    """Actions written by a kernel immediately before every unsynchronised step, and two engines alternating on ONE stream
    (ADVICE r1): the step kernel must wait for its predecessor before it loads caller-produced inputs
    (griddepcontrol.wait), and the deferred wait of the step-to-step chain must not be used when another engine launched
    in between.  1184 envs = one CTA per SM (the shape that takes the deferred wait when it is legal)."""
    torch = cuda
    n, steps, limit = 1184, 36, 7
    seeds = (5, 6)
    envs = [make_env(n, noise_table, seed=s, auto_reset=True, max_episode_steps=limit) for s in seeds]
    iota = torch.arange(n, device='cuda', dtype=torch.int32)
    acts = [torch.empty(n, dtype=torch.uint8, device='cuda') for _ in envs]
    for e in envs:
        e.reset()
    for t in range(steps):
        for k, e in enumerate(envs):
            # produced on the device right before the launch that consumes it, rewritten every step in place
            acts[k].copy_(((iota * (3 + k) + t * 7) % 10).to(torch.uint8))
            e.step(robot_action=acts[k], human_action=None if t % 2 else acts[k])
    torch.cuda.synchronize()
    i_np = np.arange(n)
    for k, (e, s) in enumerate(zip(envs, seeds)):
        oracles = oracle_envs(orc, n, noise_table, s)
        for o in oracles:
            o.reset()
        pool = orc.Pool(oracles, threads=8, want_obs=False)
        robot = np.stack([((i_np * (3 + k) + t * 7) % 10).astype(np.uint8) for t in range(steps)])
        human = np.stack([robot[t] if t % 2 == 0 else np.full(n, 255, np.uint8) for t in range(steps)])
        st, out = pool.trace(steps, limit, auto_reset=True, draw_last=3, robot=robot, human=human)
        assert np.array_equal(e.state.cpu().numpy().view(np.uint32), st[-1]), k
        ring = ring_newest_first(e)
        for i, o in enumerate(oracles):
            assert np.array_equal(ring[i], o.ring()), (k, i)
        assert np.array_equal(e.reward.cpu().numpy(), orc.unpack_out(out[-1])['reward'])
        pool.close()


@pytest.mark.parametrize("n,T,limit", [(333, 1, 9), (1184, 7, 5), (4096, 64, 23), (1184, 40, 1), (4096, 64, 2), (2500, 48, 3), (300, 700, 37)])
def test_step_many_matches_oracle(cuda, orc, table, n, T, limit):
    """ahk_step_many(T): T action-free steps in one call, resets inside the window (limit < T), every step's outputs
    logged.  The kernel runs its steps without block barriers (warps up to two steps apart on the 3-slot ring): limit 1 / 2 / 3
    make every env rewrite the older ring slots every / every other / every third step, T = 700 spans two launches.  A sample of envs against the oracle: state, ring, and the whole [T, n] log; twice in a row, so that the second
    window starts from envs armed for a reset."""
    torch = cuda
    seed = 1000 + T
    env = make_env(n, table, seed=seed, auto_reset=True, max_episode_steps=limit)
    rs = np.random.RandomState(T)
    ids = np.unique(np.concatenate([[0, n - 1, n // 2], rs.choice(n, min(n, 96), replace=False)]))
    oracles = [orc.Env(seed=seed, env_id=int(i), table=table) for i in ids]
    pool = orc.Pool(oracles, threads=8, want_obs=False)
    env.reset()
    for o in oracles:
        o.reset()
    log = {'robot_action': torch.zeros((T, n), dtype=torch.uint8, device='cuda'), 'human_action': torch.zeros((T, n), dtype=torch.uint8, device='cuda'),
           'done': torch.zeros((T, n), dtype=torch.uint8, device='cuda'), 'reward': torch.zeros((T, n), dtype=torch.float32, device='cuda')}
    for window in range(2):
        _, reward, done, info = env.step_many(T, log)
        torch.cuda.synchronize()
        st, out = pool.trace(T, limit, auto_reset=True, draw_last=3)
        u = orc.unpack_out(out)
        assert np.array_equal(state_words(env, ids), st[-1]), window
        assert np.array_equal(log['reward'].cpu().numpy()[:, ids], u['reward'])
        assert np.array_equal(log['done'].cpu().numpy()[:, ids].astype(np.uint32), u['terminal'])
        assert np.array_equal(log['robot_action'].cpu().numpy()[:, ids].astype(np.uint32), u['robot_label'])
        assert np.array_equal(log['human_action'].cpu().numpy()[:, ids].astype(np.uint32), u['human_label'])
        assert np.array_equal(reward.cpu().numpy()[ids], u['reward'][-1])
        fr, h = env.frames[:, ids].cpu().numpy(), env.head
        for k, o in enumerate(oracles):
            ring = o.ring()
            for age in range(3):
                assert np.array_equal(fr[(h + 3 - age) % 3, k], ring[age]), (window, int(ids[k]), age)
    assert env.statistics()['steps'] == 2 * T * n
    pool.close()


@pytest.mark.parametrize("depth", [3, 8, 67])
def test_ring_depth_matches_oracle(cuda, orc, noise_table, depth):
    """ahk_config.ring_depth: the ring as replay memory / staging area.  Every step's newest frame goes to slot
    step mod depth and stays there for `depth` steps (written once); a reset files its warm-up frames in the side buffer
    instead of rewriting older slots, and the 3-frame stack found through ahk_stack_view (observation('u8')) is the
    oracle's ring at every step — single steps, step_many windows, explicit masked resets, resets one step apart
    (limit 1 on the last stretch is covered by max_episode_steps=2 here: every other step resets)."""
    torch = cuda
    n, seed, limit = 333, 77, 2 if depth == 8 else 11
    env = make_env(n, noise_table, seed=seed, auto_reset=True, max_episode_steps=limit, ring_depth=depth)
    oracles = oracle_envs(orc, n, noise_table, seed)
    assert env.frames.shape[0] == depth and (env.warm is None) == (depth == 3)
    env.reset()
    for o in oracles:
        o.reset()
    newest_log = {}                                     # absolute step -> the oracle's newest frame of every env

    def check(step):
        u8 = env.observation('u8').cpu().numpy().reshape(n, 3, 3, 128, 128)
        for i, o in enumerate(oracles):
            assert np.array_equal(u8[i], o.ring()), (step, i)
        assert_same_state(env, oracles, "step %d" % step)

    check(0)
    step = 0
    for chunk in (1, 1, 1, 5, 1, depth + 2, 1, 2):
        if chunk == 1:
            env.step()
        else:
            env.step_many(chunk)
        for _ in range(chunk):
            step += 1
            for o in oracles:
                o.step_auto(255, 255, limit)
            newest_log[step] = np.stack([o.ring()[0] for o in oracles[:40]])
        check(step)
        if step == 8:                                   # an explicit masked reset in the middle
            mask = (np.arange(n) % 3 == 0).astype(np.uint8)
            env.reset(mask=mask)
            for i in np.nonzero(mask)[0]:
                oracles[i].reset()
            check(step)
    # the deep ring still holds the newest frames of the last `depth` steps, each where it was written
    fr = env.frames[:, :40].cpu().numpy()
    h = env.head
    _, meta = env.get_state()
    for back in range(min(depth, step)):
        if depth == 3 and back > 0:
            break                                       # the 3-slot ring rewrites older slots on a reset: only the newest is certain
        assert np.array_equal(fr[(h - back) % depth], newest_log[step - back]), back
    assert env.statistics()['steps'] == n * step


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
