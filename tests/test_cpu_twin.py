"""The ahk_cpu_* twin of the C ABI (include/ahk_cpu.h): same structs, layouts and call protocol as the
CUDA library, on the host cores — the timed CPU baseline and a whole-buffer checker.  CPU tests pin it to the per-env
oracle; the GPU test compares every buffer of the two arms byte for byte."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_binding_and_library_agree(orc):
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "ahk_cpu.h")).read(), flags=re.S)
    declared = set(re.findall(r"\b(ahk_cpu_[a-z_]+)\s*\(", text))
    assert declared == set(orc.CPU_SYMBOLS) and len(declared) == 12
    L = orc.lib()
    for name in declared:
        assert hasattr(L, name), name
    # every twin mirrors an entry point of the product header, argument for argument
    from gym_air_hockey_b200 import _lib
    for name, (res, args) in orc.CPU_SYMBOLS.items():
        twin = name.replace("ahk_cpu_", "ahk_")
        if twin in _lib.SYMBOLS:
            pres, pargs = _lib.SYMBOLS[twin]
            assert res == pres and len(args) == len(pargs), name
    assert C.sizeof(orc.CpuConfig) == C.sizeof(_lib.Config) and C.sizeof(orc.CpuBuffers) == C.sizeof(_lib.Buffers)
    assert C.sizeof(orc.CpuStepLog) == C.sizeof(_lib.StepLog)
    # the product library is the CUDA path only
    product = _lib.load()
    assert not hasattr(product, "ahk_cpu_step")


def test_twin_matches_per_env_oracle(orc, noise_table):
    n, seed, limit = 13, 31, 7
    eng = orc.CpuEngine(n, seed=seed, table=noise_table, auto_reset=True, max_episode_steps=limit, env_id_base=5)
    envs = [orc.Env(seed=seed, env_id=5 + i, table=noise_table) for i in range(n)]
    episode = np.array([0, 1, 2, 255] * 4, np.uint8)[:n]
    use = np.array([7, 5, 7, 7] * 4, np.uint8)[:n]
    eng.reset(episode=episode, use=use)
    for i, o in enumerate(envs):
        o.reset(-1 if episode[i] == 255 else int(episode[i]), int(use[i]))
    assert eng.head == 0 and eng.threads >= 1
    rs = np.random.RandomState(3)
    log = {k: np.zeros((4, n), np.float32 if k == 'reward' else np.uint8) for k in ('robot_label', 'human_label', 'done', 'reward')}
    for t in range(30):
        if t == 20:
            eng.step_many(4, log)
            outs = [[o.step_auto(255, 255, limit) for o in envs] for _ in range(4)]
            assert np.array_equal(log['robot_label'], np.array([[x.robot_label for x in row] for row in outs]))
            assert np.array_equal(log['reward'], np.array([[x.reward for x in row] for row in outs], np.float32))
            out = outs[-1]
        else:
            robot = rs.randint(0, 10, n).astype(np.uint8) if t % 3 == 0 else None
            human = np.where(rs.rand(n) < 0.5, 255, rs.randint(0, 10, n)).astype(np.uint8) if t % 4 == 1 else None
            eng.step(robot, human)
            out = [o.step_auto(255 if robot is None else int(robot[i]), 255 if human is None else int(human[i]), limit) for i, o in enumerate(envs)]
        h = eng.head
        for i, o in enumerate(envs):
            assert (eng.reward[i], eng.done[i], eng.robot_label[i], eng.human_label[i]) == (out[i].reward, out[i].terminal, out[i].robot_label, out[i].human_label)
            assert eng.info[i] == (out[i].flags | out[i].scored << 5)
            s, m = o.get_state()
            assert np.array_equal(eng.state[:12, i], s.view(np.uint32))
            assert tuple(eng.state[12:15, i]) == tuple(int(v) for v in m[:3])
            assert eng.state[15, i] == (int(m[3]) | int(m[4]) << 16 | int(m[5]) << 18 | int(m[6]) << 21)
            ring = o.ring()
            for age in range(3):
                assert np.array_equal(eng.frames[(h + 3 - age) % 3, i], ring[age]), (t, i, age)
    assert eng.stats[0] == n * 33 and eng.stats[5:15].sum() == n * 33
    steps_so_far = int(eng.stats[0])
    u8, f32, gray = eng.observe(0), eng.observe(1), eng.observe(2)
    for i, o in enumerate(envs):
        assert np.array_equal(u8[i].reshape(3, 3, 128, 128), o.ring())
        assert np.array_equal(f32[i], o.observe_rgb())
        st = orc.Stack()
        for age in (2, 1, 0):
            st.push(o.ring()[age])
        assert np.array_equal(gray[i], st.gray_diff())
    bound = np.zeros((n, 9, 128, 128), np.float32)   # the observation written by the step itself
    eng.bind_observation(1, bound)
    eng.step()
    assert np.array_equal(bound, eng.observe(1))
    eng.bind_observation(1, None)
    for o in envs:
        o.step_auto(255, 255, limit)
    with pytest.raises(RuntimeError, match="3-slot ring"):
        L = orc.lib()
        cfg = orc.CpuConfig(n_envs=2, ring_depth=8)
        h = C.c_void_p()
        rc = L.ahk_cpu_create(C.byref(cfg), C.byref(h))
        raise RuntimeError(L.ahk_cpu_last_error().decode()) if rc else AssertionError("accepted")
    eng.close()


@pytest.mark.gpu
def test_cuda_engine_equals_cpu_twin_buffer_for_buffer(cuda, orc, table):
    """Both arms of the C ABI on the same seeds and calls: every bound buffer — SoA state, the whole frame ring with the
    same head, reward, done, labels, info, statistics, and the three observation formats — is byte-identical."""
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    n, seed, limit = 301, 11, 9
    gpu = BatchedAirHockeyEnv(n, table=table, seed=seed, auto_reset=True, max_episode_steps=limit, env_id_base=1000)
    cpu = orc.CpuEngine(n, seed=seed, table=table, auto_reset=True, max_episode_steps=limit, env_id_base=1000)
    gpu.reset()
    cpu.reset()
    rs = np.random.RandomState(0)
    for t in range(40):
        robot = rs.randint(0, 10, n).astype(np.uint8) if t % 2 else None
        if t == 25:
            gpu.step_many(6)
            cpu.step_many(6)
        else:
            gpu.step(None if robot is None else cuda.from_numpy(robot).cuda(), None)
            cpu.step(robot, None)
        assert gpu.head == cpu.head
        assert np.array_equal(gpu.state.cpu().numpy().view(np.uint32), cpu.state), t
        assert np.array_equal(gpu.frames.cpu().numpy(), cpu.frames), t
        for name in ('reward', 'done', 'robot_label', 'human_label'):
            assert np.array_equal(getattr(gpu, name).cpu().numpy().view(getattr(cpu, name).dtype), getattr(cpu, name)), (t, name)
        assert np.array_equal(gpu.info_bits.cpu().numpy(), cpu.info), t
    st = gpu.statistics()
    assert st['steps'] == int(cpu.stats[0]) and st['episodes'] == int(cpu.stats[1])
    assert st['goals_top'] == int(cpu.stats[2]) and st['goals_bottom'] == int(cpu.stats[3])
    assert np.array_equal(gpu.observation('u8').cpu().numpy().reshape(n, 9, 128, 128), cpu.observe(0))
    assert np.array_equal(gpu.observation('rgb').cpu().numpy(), cpu.observe(1))
    assert np.array_equal(gpu.observation('gray-diff').cpu().numpy(), cpu.observe(2))
