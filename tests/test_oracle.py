"""CPU tests: pin the oracle (oracle/ahk_oracle.c) against known answers, Pillow, and the golden fixtures produced by
the REFERENCE's own AirHockeyEnv + DataProcessor (tests/golden/make_golden.py)."""
import zlib

import numpy as np
import pytest

from conftest import golden


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


# ---- Philox4x32-10 known-answer vectors (Random123 kat_vectors: philox4x32 10)
@pytest.mark.parametrize("ctr,key,want", [
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0], [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
])
def test_philox_known_answers(orc, ctr, key, want):
    assert orc.philox(ctr, key).tolist() == want


def test_resize_matches_pillow_nearest(orc):
    """SPEC.md §8: the oracle's resize and the closed form ((2j+1)*W)>>8 both equal Pillow's NEAREST."""
    from PIL import Image
    rs = np.random.RandomState(0)
    for h, w in ((700, 450), (37, 53), (128, 128), (1000, 3)):
        img = rs.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
        want = np.array(Image.fromarray(img).resize((128, 128), Image.NEAREST)).transpose(2, 0, 1)
        assert np.array_equal(orc.resize_nearest_chw(img), want), (h, w)
    j = np.arange(128)
    xs, ys = ((2 * j + 1) * 450) >> 8, ((2 * j + 1) * 700) >> 8
    img = rs.randint(0, 256, size=(700, 450, 3)).astype(np.uint8)
    want = np.array(Image.fromarray(img).resize((128, 128), Image.NEAREST)).transpose(2, 0, 1)
    assert np.array_equal(img[ys][:, xs].transpose(2, 0, 1), want)


def test_processor_golden(orc):
    """oracle resize + stack + normalise + gray-diff == reference DataProcessor (processor.py:31-66)."""
    g = golden("processor.npz")
    frames = np.random.RandomState(int(g["frames_seed"])).randint(0, 256, size=(5, 700, 450, 3)).astype(np.uint8)
    st = orc.Stack()
    rgb_crc, gray_crc = [], []
    for f in frames:
        st.push(orc.resize_nearest_chw(f))
        rgb_crc.append(crc(st.rgb()))
        gray_crc.append(crc(st.gray_diff()))
    assert rgb_crc == g["rgb_crc"].tolist()
    assert gray_crc == g["gray-diff_crc"].tolist()
    assert st.rgb().dtype == np.float32 and str(g["rgb_dtype"]) == "float32"
    assert st.gray_diff().dtype == np.float64 and str(g["gray-diff_dtype"]) == "float64"
    assert np.array_equal(st.gray_diff(), g["gray-diff_last"])
    assert np.array_equal(st.rgb(), (g["rgb_last"].astype(np.float32) - 128.0) / 128.0)
    assert np.array_equal(orc.resize_nearest_chw(g["small"]), g["small_resized"])


def test_action_table_golden(orc):
    import ctypes as C
    g = golden("processor.npz")
    assert g["labels"].tolist() == [0, 1, 2, 3, 4, 5, 6, 7, 8, 4]  # label 9 maps back to 4
    for row, want in zip(g["actions"], g["labels"]):
        a = (C.c_int8 * 2)(int(row[0]), int(row[1]))
        assert orc.lib().orc_action_to_label(a) == want


def test_reward_table_golden(orc):
    """orc_reward == the reference's reward chain (air_hockey_env.py:95-145) on all 2592 enumerated cases."""
    import ctypes as C
    rows = golden("reward_table.npz")["table"]
    assert len(rows) == 3 * 3 * 32 * 9
    for episode, scored, flags, label, reward, terminal in rows:
        flags = int(flags)
        info = orc.Info(scored=int(scored), puck_was_hit=flags & 1, hit_the_border=(flags >> 1) & 1,
                        puck_is_at_the_bottom=(flags >> 2) & 1, distance_decreased=(flags >> 3) & 1,
                        in_the_target=(flags >> 4) & 1)
        assert orc.lib().orc_reward(int(episode), C.byref(info), int(label)) == reward
        assert bool(terminal) == (scored != 0 and episode != 2)


def replay_oracle(orc, g, table):
    """Replay a golden wrapper trajectory on the oracle's own env (orc.Env): yields per-step outputs."""
    env = orc.Env(seed=int(g["seed"]), env_id=int(g["env_id"]), table=table)
    resets = {int(i): (int(e), int(u)) for i, e, u in g["resets"]}
    inject = {int(i): s for i, s in zip(g["inject_at"], g["inject_state"])}
    env.reset(*resets[-1])
    # env.update(mode=...) zeroes the stack (quirk B.2): the oracle env exposes the ring only, so track a Stack beside it
    st = orc.Stack()
    for i in range(len(g["reward"])):
        if i in resets:
            env.reset(*resets[i])
            ring = env.ring()
            for age in (2, 1, 0):
                st.push(ring[age])
        if i in inject:
            s, m = env.get_state()
            env.set_state(inject[i], m)
        o = env.step(int(g["robot_in"][i]), int(g["human_in"][i]))
        st.push(env.ring()[0])
        yield i, o, env, st


@pytest.mark.parametrize("name", ["wrapper_rgb.npz", "wrapper_gray_diff.npz"])
def test_wrapper_golden(orc, table, name):
    """The oracle's env wrapper == the reference's AirHockeyEnv over the same game: rewards, terminals, labels, the
    float observation (CRC of every step + full frames) and the continuous state."""
    g = golden(name)
    mode = str(g["mode"])
    kept = {int(i): k for i, k in zip(g["kept_at"], g["kept"])}
    n_checked = 0
    for i, o, env, st in replay_oracle(orc, g, table):
        assert o.reward == g["reward"][i], i
        assert o.terminal == g["terminal"][i], i
        assert (o.robot_label, o.human_label) == (g["robot"][i], g["human"][i]), i
        s, _ = env.get_state()
        assert np.array_equal(s.view(np.uint32), g["state12"][i].view(np.uint32)), i
        if not g["is_random"][i]:
            obs = st.rgb() if mode == "rgb" else st.gray_diff()
            assert crc(obs) == g["crc"][i], i
            if i in kept:
                want = (kept[i].astype(np.float32) - 128.0) / 128.0 if mode == "rgb" else kept[i]
                assert np.array_equal(obs, want)
                n_checked += 1
    assert n_checked == len(kept)
    assert g["terminal"].sum() >= (2 if mode == "rgb" else 0)


def test_oracle_invariants(orc, table):
    """Long AI-vs-AI run: bodies stay inside the rink, speeds clamped, labels in range, goals happen."""
    goals = 0
    for env_id in range(4):
        env = orc.Env(seed=5, env_id=env_id, table=table)
        env.reset(0, 7)
        for _ in range(1500):
            o = env.step()
            s, _ = env.get_state()
            assert 0 <= o.robot_label <= 8 and 0 <= o.human_label <= 8
            assert 25 + 24 <= s[4] <= 425 - 24 and 350 + 24 <= s[5] <= 675 - 24
            assert 25 + 24 <= s[8] <= 425 - 24 and 25 + 24 <= s[9] <= 350 - 24
            assert 25 + 18 <= s[0] <= 425 - 18 and 25 <= s[1] <= 675
            assert s[2] ** 2 + s[3] ** 2 <= 18.0 ** 2 * 1.0001
            goals += o.terminal
    assert goals >= 0


def test_noise_frame_range(orc):
    f = orc.noise_frame(3, 9, 0)
    assert f.max() <= 254 and f.min() == 0 and f.std() > 60   # randint(0, 255): 255 never appears
    assert not np.array_equal(f, orc.noise_frame(3, 9, 1))
    assert not np.array_equal(f, orc.noise_frame(3, 10, 0))


def test_pool_and_trace_replay_stepping(orc, table):
    """The persistent pool (bench.py's CPU arm) and the long-horizon trace (tools/stress_sweep.py, the 4096-env parity test)
    give what stepping the same envs one call at a time gives: checksums, logged state words, packed outputs, ring."""
    n, steps, limit = 12, 45, 16
    a = [orc.Env(seed=3, env_id=i, table=table) for i in range(n)]
    b = [orc.Env(seed=3, env_id=i, table=table) for i in range(n)]
    for e in a + b:
        e.reset()
    pool = orc.Pool(a, threads=3)
    assert pool.run(7, limit) == orc.run_batch(b, 7, limit, 2)[0]
    rs = np.random.RandomState(0)
    robot = rs.randint(0, 10, size=(steps, n)).astype(np.uint8)
    human = np.where(rs.rand(steps, n) < 0.5, 255, rs.randint(0, 10, size=(steps, n))).astype(np.uint8)
    st, out = pool.trace(steps, limit, auto_reset=True, draw_last=3, robot=robot, human=human)
    u = orc.unpack_out(out)
    for i, e in enumerate(b):
        for t in range(steps):
            o = e.step_auto(int(robot[t, i]), int(human[t, i]), limit)
            s, m = e.get_state()
            assert np.array_equal(st[t, :12, i], s.view(np.uint32)), (t, i)
            assert (st[t, 12, i], st[t, 13, i], st[t, 14, i]) == (m[0], m[1], m[2])
            assert st[t, 15, i] == (m[3] | m[4] << 16 | m[5] << 18 | m[6] << 21)
            assert (u['reward'][t, i], u['terminal'][t, i], u['robot_label'][t, i], u['human_label'][t, i], u['flags'][t, i], u['scored'][t, i]) == \
                   (o.reward, o.terminal, o.robot_label, o.human_label, o.flags, o.scored)
        assert np.array_equal(a[i].ring(), e.ring()), i  # frames skipped on the way, the last three drawn
    pool.close()


def test_stress_states_are_a_function_of_the_global_id():
    import stress
    whole = stress.stress_states(0, 24)
    assert whole.shape == (24, 12) and whole.dtype == np.float32
    assert np.array_equal(np.concatenate([stress.stress_states(0, 10), stress.stress_states(10, 24)]), whole)
    assert np.array_equal(whole[4, 0:2], whole[4, 4:6])          # kind 4: the puck exactly on the bottom mallet
    assert np.hypot(whole[0, 2], whole[0, 3]) > 17.9             # kind 0: top speed
