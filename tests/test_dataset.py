"""Supervised dataset generation (ai/supervised/generate_data.py:35-73, utils/data_utils.py:3-13)."""
import glob
import os

import numpy as np
import pytest


def test_file_name_and_round_trip(tmp_path):
    """CPU: the container keeps the reference's three array names, dtypes and file-name pattern."""
    from gym_air_hockey_b200 import dataset
    name = dataset.data_file_name('rgb', 3000, when=1500000000, ext='h5')
    assert name.endswith('_rgb_3000.h5') and len(name.split('_')[0]) == len('2017-07-14-02-40-00')
    rs = np.random.RandomState(0)
    states = rs.randint(0, 256, size=(5, 9, 128, 128)).astype(np.uint8)
    robot, human = rs.randint(0, 9, size=5).astype(np.int8), rs.randint(0, 9, size=5).astype(np.int8)
    for k in range(2):
        dataset.save_data(str(tmp_path / ('x_part%04d.npz' % k)), states[k * 2:k * 2 + 3], robot[k * 2:k * 2 + 3], human[k * 2:k * 2 + 3])
    s, r, h = dataset.load_data(str(tmp_path / 'x_part*.npz'), normalise=True)
    assert s.dtype == np.float32 and s.shape == (6, 9, 128, 128) and r.dtype == np.int8
    assert np.array_equal(s[:3], (states[:3].astype(np.float32) - 128.0) / 128.0)   # processor.py:39-42
    assert np.array_equal(r, np.concatenate([robot[0:3], robot[2:5]])) and np.array_equal(h[3:], human[2:5])


@pytest.mark.gpu
@pytest.mark.parametrize("states,container", [("f32", "h5"), ("u8", "h5"), ("u8", "npz")])
def test_generate_matches_oracle(cuda, orc, table, tmp_path, states, container):
    """The written samples are what the ORACLE produces for the same seeds — the reference's data-generation loop
    (generate_data.py:44-58: env.step() with no actions, reset on a goal or at the step limit) over the oracle env, whose
    observation is the reference processor's float32 (x-128)/128 stack (processor.py:39-52) — in step-major order and in
    the reference's dtypes: float32 normalised (n, 9, 128, 128) states, int8 labels (data_utils.py:3-7)."""
    from gym_air_hockey_b200 import BatchedAirHockeyEnv, dataset
    n, total, seed, limit = 6, 50, 21, 20
    env = BatchedAirHockeyEnv(n, table=table, seed=seed, max_episode_steps=limit)
    env.reset()
    out = dataset.generate(env, total, str(tmp_path), states=states, shard_states=24, when=1500000000, container=container)
    assert out['samples'] == total and out['steps'] == 9 and len(out['files']) == 3 and out['files'][0].endswith(container)
    s, r, h = dataset.load_data(os.path.join(str(tmp_path), '*' + os.path.splitext(out['files'][0])[1]))
    assert s.shape == (total, 9, 128, 128) and r.shape == (total,) and r.dtype == np.int8 and h.dtype == np.int8
    assert s.dtype == (np.float32 if states == 'f32' else np.uint8)

    oracles = [orc.Env(seed=seed, env_id=i, table=table) for i in range(n)]
    for o in oracles:
        o.reset()
    want_s, want_r, want_h = [], [], []
    for _ in range(9):
        for o in oracles:
            res = o.step_auto(255, 255, limit)
            want_s.append(o.observe_rgb() if states == 'f32' else o.ring().reshape(9, 128, 128))
            want_r.append(res.robot_label)
            want_h.append(res.human_label)
    assert np.array_equal(s, np.stack(want_s)[:total])
    assert np.array_equal(r, np.array(want_r, dtype=np.int8)[:total]) and np.array_equal(h, np.array(want_h, dtype=np.int8)[:total])
    assert set(np.unique(r)) <= set(range(9))   # label 9 never comes from the simulator


@pytest.mark.gpu
def test_random_episodes_labelled_undefined(cuda, table, tmp_path):
    """N4: noise frames of `random` episodes carry label 9 when asked to; every other sample keeps its 0..8 label."""
    from gym_air_hockey_b200 import BatchedAirHockeyEnv, dataset
    n = 64
    env = BatchedAirHockeyEnv(n, table=table, seed=5, max_episode_steps=20)
    env.reset(episode=np.where(np.arange(n) % 4 == 0, 2, 0).astype(np.uint8))
    out = dataset.generate(env, 3 * n, str(tmp_path), states='u8', keep_in_memory=True, label_random_as_undefined=True)
    s, r, h = out['arrays']
    first = r[:n]
    assert (first[::4] == 9).all() and (np.delete(first, np.arange(0, n, 4)) < 9).all()
    assert ((r == 9) == (h == 9)).all()


def test_weights_from_keras_list():
    """N3: model.get_weights() order == ahk_policy_weights order; shape errors are reported by name."""
    from gym_air_hockey_b200.policy import SHAPES, keras_default_init, weights_from_keras, save_weights_npz, load_weights
    w = keras_default_init(seed=1, randomize_bn=True)
    got = weights_from_keras([w[name] for name, _ in SHAPES])
    assert all(np.array_equal(got[k], w[k]) for k in w)
    with pytest.raises(ValueError, match='conv2_w'):
        weights_from_keras([w[name] if name != 'conv2_w' else w[name][:3] for name, _ in SHAPES])
    import tempfile, os
    with tempfile.TemporaryDirectory() as d:
        save_weights_npz(os.path.join(d, 'm.npz'), w)
        back = load_weights(os.path.join(d, 'm.npz'))
        np.savez(os.path.join(d, 'l.npz'), *[w[name] for name, _ in SHAPES])
        back2 = load_weights(os.path.join(d, 'l.npz'))
    assert all(np.array_equal(back[k], w[k]) and np.array_equal(back2[k], w[k]) for k in w)
