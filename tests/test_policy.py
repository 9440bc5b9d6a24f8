"""CNN policy forward (ai/supervised/keras/models.py:20-65): the tcgen05 kernels against the PyTorch float32
restatement (oracle/cnn_ref.py).

Tolerance (stated here because north_star gives none for the network): operands are bfloat16 (8-bit mantissa), the
accumulation float32.  Against the float32 model the logits must agree to 3e-2 of the logit range of the batch;
against the same network with bfloat16 rounding at the kernels' rounding points (emulate_bf16) to 6e-3 — what is left
there is accumulation order, which can flip a bfloat16 rounding (1 ulp = 0.4 %) of an activation.  Actions (argmax) must agree wherever the float32 top-2 margin exceeds the tolerance.
"""
import numpy as np
import pytest


def test_reference_restatement_shapes_and_consistency():
    """CPU: the float32 restatement has the layer shapes models.py implies and the folded bf16 variant tracks it."""
    import cnn_ref
    from gym_air_hockey_b200.policy import FLOPS_PER_SAMPLE, SHAPES, keras_default_init
    w = keras_default_init(seed=3, randomize_bn=True)
    assert sum(int(np.prod(s)) for _, s in SHAPES) == 2404163          # parameter count
    assert FLOPS_PER_SAMPLE == 20454016
    obs = np.random.RandomState(0).randint(0, 256, size=(3, 9, 128, 128)).astype(np.uint8)
    a = cnn_ref.forward(w, obs)
    b = cnn_ref.forward(w, obs, emulate_bf16=True)
    assert a.shape == (3, 10) and a.dtype == np.float32
    scale = np.abs(a).max()
    assert np.abs(a - b).max() <= 3e-2 * scale


def test_keras_default_init_statistics():
    from gym_air_hockey_b200.policy import keras_default_init
    w = keras_default_init(seed=0)
    assert np.all(w['bn1_gamma'] == 1) and np.all(w['bn2_var'] == 1) and np.all(w['conv1_b'] == 0)
    lim = np.sqrt(6.0 / (8 * 8 * 9 + 8 * 8 * 5))
    assert np.abs(w['conv1_w']).max() <= lim and np.abs(w['conv1_w']).max() > 0.9 * lim


@pytest.mark.gpu
@pytest.mark.parametrize("n,source,max_batch", [(1, "noise", 256), (37, "noise", 256), (200, "game", 256), (37, "noise", 37), (3, "noise", 3)])
def test_policy_forward_matches_reference(cuda, table, n, source, max_batch):
    """max_batch 37 / 3: an ODD number of samples fills the activation buffers to their last byte — conv3's item copies are
    tensor-map rows of two act2 pixels, and the last item of an odd batch reaches past the last sample."""
    import cnn_ref
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init
    torch = cuda
    w = keras_default_init(seed=5, randomize_bn=True)
    net = PolicyNetwork(w, max_batch=max_batch)
    if source == "game":
        env = BatchedAirHockeyEnv(n, table=table, seed=9, max_episode_steps=50)
        env.reset()
        for _ in range(7):
            env.step()
        frames, head = env.frames, env.head
    else:
        frames = torch.from_numpy(np.random.RandomState(n).randint(0, 256, size=(3, n, 3, 128, 128)).astype(np.uint8)).cuda()
        head = 1
    logits, action = net.forward_ring(frames, head)
    torch.cuda.synchronize()
    logits, action = logits.cpu().numpy(), action.cpu().numpy()
    fr = frames.cpu().numpy()
    obs = np.concatenate([fr[(head + 3 - age) % 3] for age in range(3)], axis=1)  # newest first (processor.py:47-49)
    ref32 = cnn_ref.forward(w, obs)
    ref16 = cnn_ref.forward(w, obs, emulate_bf16=True)
    scale = max(np.abs(ref32).max(), 1e-6)
    assert np.isfinite(logits).all()
    assert np.abs(logits - ref16).max() <= 6e-3 * scale, np.abs(logits - ref16).max() / scale
    assert np.abs(logits - ref32).max() <= 3e-2 * scale, np.abs(logits - ref32).max() / scale
    assert np.array_equal(action, logits.argmax(axis=1))
    top2 = np.sort(ref32, axis=1)
    clear = (top2[:, -1] - top2[:, -2]) > 6e-2 * scale
    assert np.array_equal(action[clear], ref32.argmax(axis=1)[clear])


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["rgb", "gray-diff"])
def test_policy_forward_8192_samples_matches_reference(cuda, table, mode):
    """BASELINE.json configs[3]'s per-GPU batch: PolicyNetwork(max_batch=8192) on 8192 GAME frames (mixed episode types,
    resets in flight), both modes.  The persistent multi-tile paths of every kernel run at full grid here (conv1: 57 344
    work items over 148 CTAs; conv2: 23 104 tiles; dense: 64 row tiles x 2 N halves).  640 sampled rows — the first and
    last 128-row tile of the dense layers, both sides of tile boundaries, the first and last sample, random others — are
    compared with oracle/cnn_ref.py at the same tolerances as the small tests."""
    import cnn_ref
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import GRAY_SHAPES, SHAPES, PolicyNetwork, keras_default_init
    torch = cuda
    n = 8192
    w = keras_default_init(seed=7, randomize_bn=True, shapes=SHAPES if mode == 'rgb' else GRAY_SHAPES)
    net = PolicyNetwork(w, max_batch=n, mode=mode)
    env = BatchedAirHockeyEnv(n, table=table, seed=31, max_episode_steps=6)
    env.reset()
    for _ in range(8):
        env.step()
    logits, action = net.forward(env)
    torch.cuda.synchronize()
    rs = np.random.RandomState(8192)
    ids = np.unique(np.concatenate([np.arange(0, 128), np.arange(n - 128, n), [127, 128, 129, 255, 256, 4095, 4096, 4097],
                                    rs.choice(n, 400, replace=False)]))
    assert len(ids) >= 512
    fr = env.frames[:, ids].cpu().numpy()
    head = env.head
    obs = np.concatenate([fr[(head + 3 - age) % 3] for age in range(3)], axis=1)
    fwd = cnn_ref.forward if mode == 'rgb' else cnn_ref.forward_gray
    ref32, ref16 = fwd(w, obs), fwd(w, obs, emulate_bf16=True)
    got, act = logits.cpu().numpy()[ids], action.cpu().numpy()[ids]
    scale = max(np.abs(ref32).max(), 1e-6)
    assert np.isfinite(logits.cpu().numpy()).all()
    assert np.abs(got - ref16).max() <= 6e-3 * scale, np.abs(got - ref16).max() / scale
    assert np.abs(got - ref32).max() <= 3e-2 * scale, np.abs(got - ref32).max() / scale
    assert np.array_equal(action.cpu().numpy(), logits.cpu().numpy().argmax(axis=1))
    top2 = np.sort(ref32, axis=1)
    clear = (top2[:, -1] - top2[:, -2]) > 6e-2 * scale
    assert np.array_equal(act[clear], ref32.argmax(axis=1)[clear])
    # a second pass over the same ring gives the same bits (no cross-tile state left behind by the persistent kernels)
    logits2, _ = net.forward(env)
    assert torch.equal(logits, logits2)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["rgb", "gray-diff"])
def test_policy_forward_on_a_deep_ring(cuda, table, mode):
    """ahk_policy_forward_stack: the network on an engine whose ring is deep (ring_depth 9) — envs one and two steps past a
    reset take their older frames from the warm-up side buffer.  The logits must be the bits the same network gives on
    the materialised stack laid out as a plain 3-slot ring, and match the float32 restatement."""
    import cnn_ref
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import GRAY_SHAPES, SHAPES, PolicyNetwork, keras_default_init
    torch = cuda
    n = 150
    w = keras_default_init(seed=8, randomize_bn=True, shapes=SHAPES if mode == 'rgb' else GRAY_SHAPES)
    net = PolicyNetwork(w, max_batch=256, mode=mode)
    env = BatchedAirHockeyEnv(n, table=table, seed=17, max_episode_steps=50, ring_depth=9)
    env.reset()
    for t in range(11):
        if t in (8, 9, 10):  # a third of the envs restarts before each of the last three steps: episode_steps 3, 2, 1 at the end
            env.set_pending(cuda.from_numpy((np.arange(n) % 3 == t - 8).astype(np.uint8)).cuda() | ((env.state[15] >> 21) & 1).to(cuda.uint8))
        env.step()
    _, meta = env.get_state()
    assert {1, 2, 3} <= set(meta['episode_steps'].tolist())      # stacks with two, one and no warm-up frames
    logits, action = net.forward(env)
    u8 = env.observation('u8')                                    # [n, 9, 128, 128] newest first
    ring = u8.reshape(n, 3, 3, 128, 128).permute(1, 0, 2, 3, 4).flip(0).contiguous()  # slots oldest..newest, head = 2
    logits3, action3 = net.forward_ring(ring, 2)
    torch.cuda.synchronize()
    assert torch.equal(logits, logits3) and torch.equal(action, action3)
    fwd = cnn_ref.forward if mode == 'rgb' else cnn_ref.forward_gray
    ref16 = fwd(w, u8.cpu().numpy(), emulate_bf16=True)
    scale = max(np.abs(ref16).max(), 1e-6)
    assert np.abs(logits.cpu().numpy() - ref16).max() <= 6e-3 * scale


@pytest.mark.gpu
def test_policy_in_the_loop(cuda, table):
    """keras/test.py:36-51 as a batched loop: frames -> CNN -> argmax -> env.step, all on the device."""
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import PolicyNetwork
    n = 512
    env = BatchedAirHockeyEnv(n, table=table, seed=1, max_episode_steps=64)
    net = PolicyNetwork(max_batch=n)
    env.reset()
    for _ in range(20):
        _, action = net.forward(env)
        _, reward, done, info = env.step(action, None)
        assert cuda.equal(info['robot_action'], cuda.where(action == 9, cuda.full_like(action, 4), action))  # label 9 -> stand -> 4
    assert env.statistics()['steps'] == 20 * n


def test_gray_diff_restatement_matches_reference_processor():
    """CPU: the gray-diff input of the restatement is what the reference's own DataProcessor returns (golden fixture
    recorded from processor.py:53-64), and the network has the shapes models.py:67-116 implies."""
    import cnn_ref
    from conftest import golden
    from gym_air_hockey_b200.policy import GRAY_SHAPES, GRAY_FLOPS_PER_SAMPLE, keras_default_init
    g = golden("wrapper_gray_diff.npz")
    w = keras_default_init(seed=2, randomize_bn=True, shapes=GRAY_SHAPES)
    assert GRAY_FLOPS_PER_SAMPLE == 13651968
    obs = np.random.RandomState(1).randint(0, 256, size=(2, 9, 128, 128)).astype(np.uint8)
    x = cnn_ref.gray_diff_input(obs)
    assert x.dtype == np.float64 and x.shape == (2, 1, 128, 128)
    d = (obs[0, 6:9].astype(np.int64) - obs[0, 3:6] - obs[0, 0:3]) % 256
    assert np.allclose(x[0, 0], ((0.2989 * d[0] + 0.5870 * d[1] + 0.1140 * d[2]) - 128.0) / 128.0, atol=1e-12)
    a = cnn_ref.forward_gray(w, obs)
    b = cnn_ref.forward_gray(w, obs, emulate_bf16=True)
    assert a.shape == (2, 10) and np.abs(a - b).max() <= 3e-2 * np.abs(a).max()
    assert len(g.files) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,source", [(1, "noise"), (130, "game")])
def test_gray_policy_forward_matches_reference(cuda, table, n, source):
    """The 'gray-diff' network (models.py:67-116) on the ring, against the float32 restatement; same tolerances."""
    import cnn_ref
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import GRAY_SHAPES, PolicyNetwork, keras_default_init
    torch = cuda
    w = keras_default_init(seed=6, randomize_bn=True, shapes=GRAY_SHAPES)
    net = PolicyNetwork(w, max_batch=256, mode='gray-diff')
    if source == "game":
        env = BatchedAirHockeyEnv(n, table=table, seed=3, max_episode_steps=50)
        env.reset()
        for _ in range(5):
            env.step()
        frames, head = env.frames, env.head
    else:
        frames = torch.from_numpy(np.random.RandomState(n).randint(0, 256, size=(3, n, 3, 128, 128)).astype(np.uint8)).cuda()
        head = 2
    logits, action = net.forward_ring(frames, head)
    torch.cuda.synchronize()
    logits, action = logits.cpu().numpy(), action.cpu().numpy()
    fr = frames.cpu().numpy()
    obs = np.concatenate([fr[(head + 3 - age) % 3] for age in range(3)], axis=1)
    ref32 = cnn_ref.forward_gray(w, obs)
    ref16 = cnn_ref.forward_gray(w, obs, emulate_bf16=True)
    scale = max(np.abs(ref32).max(), 1e-6)
    assert np.isfinite(logits).all()
    assert np.abs(logits - ref16).max() <= 6e-3 * scale, np.abs(logits - ref16).max() / scale
    assert np.abs(logits - ref32).max() <= 3e-2 * scale, np.abs(logits - ref32).max() / scale
    assert np.array_equal(action, logits.argmax(axis=1))


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 37, 300, 2051])
def test_fused_conv23_is_bit_identical(cuda, table, n, monkeypatch):
    """conv23_kernel (the default of inference forwards: conv2 -> conv3 in one kernel, act2 in shared memory) does the arithmetic
    of the two conv_tm_kernel launches it replaces (AHK_FUSE23=0): same logits bit for bit, also on a second call (stale
    accumulators / act2 of the previous pair), for odd batches and for more pairs than CTAs."""
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init
    torch = cuda
    w = keras_default_init(seed=6, randomize_bn=True)
    env = BatchedAirHockeyEnv(n, table=table, seed=4, max_episode_steps=50)
    env.reset()
    for _ in range(5):
        env.step()
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("AHK_FUSE23", mode)
        net = PolicyNetwork(w, max_batch=n)
        first = net.forward_ring(env.frames, env.head)[0].clone()
        second = net.forward_ring(env.frames, env.head)[0].clone()
        torch.cuda.synchronize()
        assert torch.equal(first, second)
        out[mode] = first
        net.close()
    assert torch.isfinite(out["1"]).all() and torch.equal(out["0"], out["1"])
