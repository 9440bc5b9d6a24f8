"""Batched RL step: csrc/ahk_rl.cu and gym_air_hockey_b200/rl.py against oracle/rl_ref.py, the numpy
restatement of ai/reinforcement/rl_agents.py:64-77,177-191,246-276.

Tolerances: DDQN targets are bit-exact (one float64 -> float32 rounding, as numpy's assignment); losses and output
gradients agree to 1e-5 relative (float32 sums in a different order); standardised returns to 2e-6 absolute (float64
arithmetic, summation order differs from numpy's pairwise mean/std, result rounded to float32).
"""
import numpy as np
import pytest


# ----------------------------------------------------------------------------------------------------------- CPU
def test_trainable_network_matches_restatement_and_round_trips():
    import cnn_ref
    import torch
    from gym_air_hockey_b200.policy import SHAPES, keras_default_init
    from train_ref import TrainableNetwork  # the float32 autograd checker (oracle/)
    w = keras_default_init(seed=11, randomize_bn=True)
    net = TrainableNetwork(w).eval()
    obs = np.random.RandomState(1).randint(0, 256, size=(4, 9, 128, 128)).astype(np.uint8)
    with torch.no_grad():
        got = net(torch.from_numpy(obs)).numpy()
    want = cnn_ref.forward(w, obs)
    assert np.abs(got - want).max() <= 1e-4 * max(np.abs(want).max(), 1.0)
    back = net.keras_weights()
    for name, shape in SHAPES:
        assert back[name].shape == shape and np.array_equal(back[name], w[name]), name


def test_product_has_no_autograd_fit():
    """The fit is hand-written (csrc/ahk_train.cu): the product package must not define a torch module, call backward() or
    import the optimizers — the float32 autograd restatement lives in oracle/train_ref.py as the checker."""
    import os
    import re
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'deep-learning-air-hockey-robot_b200', 'gym_air_hockey_b200')
    for name in sorted(os.listdir(pkg)):
        if name.endswith('.py'):
            src = open(os.path.join(pkg, name)).read()
            code = '\n'.join(line.split('#')[0] for line in src.split('\n') if not line.lstrip().startswith(('"', '*', '`')))
            assert not re.search(r'torch\.nn|nn\.Module|logits\.backward|loss\.backward|torch\.optim|torch\.autograd|requires_grad', code), name
    assert 'ahk_train_backward' in open(os.path.join(pkg, 'train.py')).read()


def test_valid_transitions_window_logic():
    import torch
    from gym_air_hockey_b200.rl import valid_transitions
    cap, n = 10, 2
    pending = torch.zeros((cap, n), dtype=torch.uint8)
    pending[4, 1] = 1  # env 1 armed after step 4 -> step 5 is a reset
    # 8 steps recorded into slots 0..7, cursor at 8
    ok = valid_transitions(pending, cursor=8, filled=8)
    assert ok[:, 0].tolist() == [False, False, True, True, True, True, True, False, False, False]  # t-2 >= 0, t+1 <= 7
    assert ok[:, 1].tolist() == [False, False, True, True, False, False, False, False, False, False]  # 4,5,6 see the reset
    # full and wrapped: cursor 3 -> newest slot 2, oldest slot 3
    ok = valid_transitions(torch.zeros((cap, n), dtype=torch.uint8), cursor=3, filled=cap)
    assert ok[:, 0].tolist() == [True, True, False, False, False, True, True, True, True, True]
    # states only (PG fit): the newest slot is usable too
    ok = valid_transitions(torch.zeros((cap, n), dtype=torch.uint8), cursor=8, filled=8, need_next=False)
    assert ok[:, 0].tolist() == [False, False, True, True, True, True, True, True, False, False]


def test_discount_rewards_restatement_single_episode():
    import rl_ref
    r = [0.5, -0.5, 0.0, 1.0]
    g = rl_ref.discount_rewards(r, 0.98)
    raw = np.array([0.5 + 0.98 * (-0.5 + 0.98 * (0.0 + 0.98 * 1.0)), -0.5 + 0.98 * (0.98 * 1.0), 0.98 * 1.0, 1.0])
    assert np.allclose(g, (raw - raw.mean()) / raw.std())
    assert np.array_equal(rl_ref.discount_rewards([0.5], 0.98), [0.0])  # std == 0 -> zeros (rl_agents.py:265-266)


# ----------------------------------------------------------------------------------------------------------- GPU
def _lib_and_ptr():
    import ctypes as C
    from gym_air_hockey_b200 import _lib
    return _lib, _lib.load(), (lambda t: C.c_void_p(t.data_ptr()) if t is not None else None)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 32, 1000])
def test_ddqn_targets_kernel(cuda, n):
    import rl_ref
    torch = cuda
    _lib, lib, ptr = _lib_and_ptr()
    rs = np.random.RandomState(n)
    q, qo, qt, qfit = [(rs.randn(n, 10) * 0.8).astype(np.float32) for _ in range(4)]
    qo[0, 3] = qo[0, 7] = 5.0  # a tie: np.argmax takes the first
    action = rs.randint(0, 10, size=n).astype(np.uint8)
    reward = rs.choice([-1.0, -0.5, 0.0, 0.5, 1.0], size=n)
    done = rs.rand(n) < 0.3
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dq, dqo, dqt, dfit, da = d(q), d(qo), d(qt), d(qfit), d(action)
    dr, dd = d(reward.astype(np.float32)), d(done.astype(np.uint8))
    targets, grad, loss = torch.empty_like(dq), torch.empty_like(dq), torch.zeros(1, device='cuda')
    _lib.check(lib.ahk_rl_ddqn_targets(ptr(dq), ptr(dqo), ptr(dqt), ptr(da), ptr(dr), ptr(dd), n, 0.98, ptr(dfit), ptr(targets),
                                       ptr(grad), ptr(loss), None))
    torch.cuda.synchronize()
    want = rl_ref.ddqn_targets(q, qo, qt, action.astype(np.int64), reward, done, 0.98)
    assert np.array_equal(targets.cpu().numpy(), want)                      # bit-exact
    wl, wg = rl_ref.mse_loss_and_grad(qfit, want)
    assert abs(loss.item() - wl) <= 1e-5 * max(wl, 1e-6)
    assert np.allclose(grad.cpu().numpy(), wg, rtol=1e-5, atol=1e-9)
    # q_fit = NULL: the fitted output is q itself
    _lib.check(lib.ahk_rl_ddqn_targets(ptr(dq), ptr(dqo), ptr(dqt), ptr(da), ptr(dr), ptr(dd), n, 0.98, None, None, ptr(grad),
                                       ptr(loss), None))
    wl, wg = rl_ref.mse_loss_and_grad(q, want)
    assert abs(loss.item() - wl) <= 1e-5 * max(wl, 1e-6) and np.allclose(grad.cpu().numpy(), wg, rtol=1e-5, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("T,n", [(1, 3), (17, 5), (256, 300)])
def test_pg_returns_and_targets_kernels(cuda, T, n):
    import rl_ref
    torch = cuda
    _lib, lib, ptr = _lib_and_ptr()
    rs = np.random.RandomState(T + n)
    reward = rs.choice([-1.0, -0.5, 0.5, 1.0], size=(T, n)).astype(np.float32)
    reward[:, 0] = 0.5                                   # a constant-reward env: std != 0 through discounting only
    ends = (rs.rand(T, n) < 0.05).astype(np.uint8)
    if T > 2:
        ends[1, 1] = ends[2, 1] = 1                      # a one-step episode: std == 0 -> zero
    returns = torch.empty((T, n), device='cuda')
    dr, de = torch.from_numpy(reward).cuda(), torch.from_numpy(ends).cuda()
    _lib.check(lib.ahk_rl_pg_returns(ptr(dr), ptr(de), T, n, 0.98, ptr(returns), None))
    torch.cuda.synchronize()
    want = rl_ref.discount_rewards_batched(reward, ends, 0.98)
    got = returns.cpu().numpy()
    assert np.abs(got - want).max() <= 2e-6, np.abs(got - want).max()
    if T > 2:
        assert got[2, 1] == 0.0

    m = T * n
    logits = rs.randn(m, 10).astype(np.float32)
    prob = np.exp(logits) / np.exp(logits).sum(axis=1, keepdims=True)
    fit_logits = logits + 0.1 * rs.randn(m, 10).astype(np.float32)
    prob_fit = (np.exp(fit_logits) / np.exp(fit_logits).sum(axis=1, keepdims=True)).astype(np.float32)
    action = rs.randint(0, 10, size=m).astype(np.uint8)
    dp, dpf, da = torch.from_numpy(prob).cuda(), torch.from_numpy(prob_fit).cuda(), torch.from_numpy(action).cuda()
    targets, grad, loss = torch.empty_like(dp), torch.empty_like(dp), torch.zeros(1, device='cuda')
    flat = returns.reshape(-1).contiguous()
    _lib.check(lib.ahk_rl_pg_targets(ptr(dp), ptr(da), ptr(flat), m, 0.01, ptr(dpf), ptr(targets), ptr(grad), ptr(loss), None))
    torch.cuda.synchronize()
    wt = rl_ref.pg_targets(prob, action.astype(np.int64), got.reshape(-1), np.float32(0.01))
    assert np.allclose(targets.cpu().numpy(), wt, rtol=1e-6, atol=1e-7)
    wl, wg = rl_ref.crossentropy_loss_and_grad(prob_fit, targets.cpu().numpy())
    assert abs(loss.item() - wl) <= 1e-4 * abs(wl)
    assert np.allclose(grad.cpu().numpy(), wg, rtol=1e-4, atol=1e-8)


@pytest.mark.gpu
def test_epsilon_greedy_kernel(cuda):
    """rl_agents.py:48-56 per env: epsilon 0 -> greedy, 1 -> uniform over the 10 actions, in between the explored fraction
    is epsilon; draws are keyed by the GLOBAL env id (two shards reproduce the whole) and by the iteration."""
    torch = cuda
    _lib, lib, ptr = _lib_and_ptr()
    n = 200000
    greedy = torch.full((n,), 7, dtype=torch.uint8, device='cuda')

    def draw(eps, base=0, count=n, it=5, hist=None, g=greedy):
        out = torch.empty(count, dtype=torch.uint8, device='cuda')
        _lib.check(lib.ahk_rl_epsilon_greedy(ptr(g), count, eps, 1234, base, it, ptr(out), ptr(hist), None))
        return out

    assert bool((draw(0.0) == 7).all())
    hist = torch.zeros(10, dtype=torch.int64, device='cuda')
    a = draw(1.0, hist=hist)
    counts = torch.bincount(a.long(), minlength=10)
    assert torch.equal(counts, hist) and int(a.max()) == 9
    assert (counts.float() / n - 0.1).abs().max().item() < 0.004          # uniform: 5 sigma at n = 200 000
    b = draw(0.3)
    explored = (b != 7).float().mean().item()                              # 0.3 * 0.9 of the draws leave the greedy action
    assert abs(explored - 0.27) < 0.005
    whole = draw(0.5)
    halves = torch.cat([draw(0.5, 0, n // 2, g=greedy[:n // 2]), draw(0.5, n // 2, n - n // 2, g=greedy[n // 2:])])
    assert torch.equal(whole, halves)                                       # independent of the sharding
    assert not torch.equal(whole, draw(0.5, it=6))                          # a new iteration, new draws
    assert torch.equal(whole, draw(0.5))                                    # deterministic


@pytest.mark.gpu
@pytest.mark.parametrize("view", [False, True])
def test_replay_store_windows_are_the_ring(cuda, table, view):
    """Every sampled transition's gathered frames equal the observation the env actually showed at that step (and the
    next one); windows that would straddle a reset are never sampled.  view: the store is the engine's deep ring itself
    (ring_depth = capacity): no frame is copied, and sampling must give exactly what the copying store gives."""
    torch = cuda
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.rl import ReplayStore, stacked
    n, cap, steps = 24, 16, 41
    env = BatchedAirHockeyEnv(n, table=table, seed=21, max_episode_steps=9, ring_depth=cap if view else 3)
    env.reset()
    store = ReplayStore(env, cap)
    assert store.view == view and (store.frames.data_ptr() == env.frames.data_ptr()) == view
    obs_log, act_log, rew_log, done_log = {}, {}, {}, {}
    store.record_observation(env)
    obs_log[0] = env.observation('u8').cpu().numpy()
    gen = torch.Generator(device='cuda'); gen.manual_seed(3)
    for s in range(1, steps + 1):
        action = torch.randint(0, 10, (n,), device='cuda', generator=gen).to(torch.uint8)
        _, reward, done, _ = env.step(action)
        store.record(env, action, reward, done)
        obs_log[s] = env.observation('u8').cpu().numpy()
        act_log[s], rew_log[s], done_log[s] = action.cpu().numpy(), reward.cpu().numpy().copy(), done.cpu().numpy().copy()
    assert store.filled == cap
    if not view:
        assert store.cursor == (steps + 1) % cap
    batch = store.sample(10 ** 6, gen)          # everything that is valid
    b = batch['t'].numel()
    assert 0 < b < cap * n
    g = batch['frames']
    s_t, s_next = stacked(g[0:3]).cpu().numpy(), stacked(g[1:4]).cpu().numpy()
    t_idx, e_idx = batch['t'].cpu().numpy(), batch['env'].cpu().numpy()
    newest = steps                                # absolute step held by slot (cursor - 1)
    for k in range(b):
        age = (store.cursor - 1 - t_idx[k]) % cap
        step = newest - age
        assert np.array_equal(s_t[k], obs_log[step][e_idx[k]]), (k, step)
        assert np.array_equal(s_next[k], obs_log[step + 1][e_idx[k]]), (k, step)
        assert batch['action'][k].item() == act_log[step + 1][e_idx[k]]
        assert batch['reward'][k].item() == rew_log[step + 1][e_idx[k]]
        assert bool(batch['done'][k].item()) == bool(done_log[step + 1][e_idx[k]])
    # resets happened (max_episode_steps 9), so some windows had to be excluded
    assert b < (cap - 3) * n


@pytest.mark.gpu
def test_set_weights_reaches_the_forward_kernels(cuda, table):
    import cnn_ref
    torch = cuda
    from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init
    import ctypes as C
    from gym_air_hockey_b200 import _lib
    from gym_air_hockey_b200.policy import SHAPES
    w0, w1 = keras_default_init(seed=1, randomize_bn=True), keras_default_init(seed=2, randomize_bn=True)
    net = PolicyNetwork(w0, max_batch=64)
    frames = torch.from_numpy(np.random.RandomState(0).randint(0, 256, size=(3, 40, 3, 128, 128)).astype(np.uint8)).cuda()
    before = net.forward_ring(frames, 2)[0].cpu().numpy()
    st = _lib.PolicyWeights()
    for name, _ in SHAPES:
        setattr(st, name, w1[name].ctypes.data_as(C.c_void_p))
    _lib.check(net.lib.ahk_policy_set_weights(net._h, C.byref(st)))
    after = net.forward_ring(frames, 2)[0].cpu().numpy()
    fr = frames.cpu().numpy()
    obs = np.concatenate([fr[2], fr[1], fr[0]], axis=1)
    ref = cnn_ref.forward(w1, obs, emulate_bf16=True)
    scale = np.abs(ref).max()
    assert np.abs(after - ref).max() <= 6e-3 * scale
    assert np.abs(before - ref).max() > 0.1 * scale


@pytest.mark.gpu
def test_set_weights_dev_equals_host_path(cuda):
    """ahk_policy_set_weights_dev (fold + pack by kernels, from device tensors) leaves the forward kernels with the same
    network as the host path: weights are bit-identical bf16, biases differ at most by the order of a float64 sum."""
    import ctypes as C
    torch = cuda
    from gym_air_hockey_b200 import _lib
    from gym_air_hockey_b200.policy import SHAPES, PolicyNetwork, keras_default_init
    w0, w1 = keras_default_init(seed=1, randomize_bn=True), keras_default_init(seed=2, randomize_bn=True)
    host, dev = PolicyNetwork(w1, max_batch=64), PolicyNetwork(w0, max_batch=64)
    wd = {name: torch.from_numpy(w1[name]).cuda() for name, _ in SHAPES}
    st = _lib.PolicyWeights()
    for name, _ in SHAPES:
        setattr(st, name, C.c_void_p(wd[name].data_ptr()))
    _lib.check(dev.lib.ahk_policy_set_weights_dev(dev._h, C.byref(st), None))
    frames = torch.from_numpy(np.random.RandomState(0).randint(0, 256, size=(3, 50, 3, 128, 128)).astype(np.uint8)).cuda()
    a = host.forward_ring(frames, 1)[0].cpu().numpy()
    b = dev.forward_ring(frames, 1)[0].cpu().numpy()
    assert np.abs(a - b).max() <= 1e-5 * np.abs(a).max(), np.abs(a - b).max()


@pytest.mark.gpu
@pytest.mark.parametrize("ring_depth", [3, 16])
def test_ddqn_training_loop(cuda, table, ring_depth):
    """train.py:76-118 batched: act / step / store / train every 5 iterations / target copy.  The fitted weights must
    reach the forward kernels (their Q values track the float32 module) and the losses must be finite.  ring_depth 16: the
    replay memory is the engine's deep ring (no frame copies), acting reads the stack through ahk_stack_view."""
    torch = cuda
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    import train_ref
    from gym_air_hockey_b200.rl import DDQNAgent, run_training, stacked
    env = BatchedAirHockeyEnv(48, table=table, seed=5, max_episode_steps=40, ring_depth=ring_depth)
    env.reset()
    agent = DDQNAgent(env, buffer_steps=16, batch_size=32, discount_rate=0.98, eps_min=0.05, eps_max=0.99, eps_decay_steps=100, seed=7)
    w_before = agent.trainer.keras_weights()
    losses = run_training(agent, 40, training_start=8, training_interval=5, copy_interval=20)
    assert len(losses) == 7 and all(np.isfinite(l) and l >= 0 for l in losses)
    assert agent.use_target                                      # iteration 40 > 20 and 40 % 20 == 0
    assert agent.trainer.iterations == 7
    w_after = agent.trainer.keras_weights()
    assert not np.array_equal(w_before['dense2_w'], w_after['dense2_w'])
    assert not np.array_equal(w_before['bn1_mean'], w_after['bn1_mean'])   # fit() moved the BatchNorm statistics
    assert int(agent.action_distribution.sum().item()) == 40 * 48
    # the kernels now hold the fitted network
    batch = agent.replay.sample(32, agent.gen)
    q_kernel = agent.policy.forward_ring(batch['frames'][0:3], 2)[0]
    with torch.no_grad():  # the float32 checker on the fitted weights (inference mode: moving statistics)
        q_module = train_ref.TrainableNetwork(w_after).eval()(stacked(batch['frames'][0:3]).cpu())
    scale = q_module.abs().max().item()
    assert (q_kernel.cpu() - q_module).abs().max().item() <= 3e-2 * scale


@pytest.mark.gpu
def test_model_files_round_trip(cuda, table, tmp_path):
    """rl_agents.py:103 `self.model.save(self.model_file)` at every target copy and :153-154 `load_model(pretrained_model_file)`:
    the agent writes a Keras-layout HDF5 file (h5lite.py) and a second agent started from it holds the same network."""
    torch = cuda
    from gym_air_hockey_b200 import BatchedAirHockeyEnv, h5lite
    from gym_air_hockey_b200.rl import DDQNAgent, PGAgent, run_training
    env = BatchedAirHockeyEnv(32, table=table, seed=9, max_episode_steps=40, ring_depth=16)
    env.reset()
    path = str(tmp_path / 'ddqn_model.h5')
    agent = DDQNAgent(env, batch_size=32, seed=3, model_file=path)
    run_training(agent, 20, training_start=8, training_interval=5, copy_interval=10)
    assert agent.use_target and h5lite.is_hdf5(path)
    with h5lite.File(path) as f:
        assert f.attrs['layer_names'][-1] == b'output' and f['output/output/kernel:0'].shape == (512, 10)
    saved = agent.trainer.keras_weights()
    again = PGAgent(env, pretrained_model_file=path, episode_length=8, fit_batch=32)
    loaded = again.trainer.keras_weights()
    assert all(np.array_equal(saved[k], loaded[k]) for k in saved)
    logits_a = agent.policy.forward(env)[0].clone()
    logits_b = again.policy.forward(env)[0]
    assert torch.equal(logits_a, logits_b)


@pytest.mark.gpu
@pytest.mark.parametrize("ring_depth", [3, 13])
def test_pg_training_loop(cuda, table, ring_depth):
    torch = cuda
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.rl import PGAgent, run_training
    env = BatchedAirHockeyEnv(16, table=table, seed=6, max_episode_steps=12, ring_depth=ring_depth)
    env.reset()
    agent = PGAgent(env, episode_length=10, fit_batch=32, discount_rate=0.98, eps_min=0.05, eps_max=0.99, eps_decay_steps=100, seed=8)
    w_before = agent.trainer.keras_weights()
    losses = run_training(agent, 25)
    assert len(losses) == 2 and all(np.isfinite(l) for l in losses)
    assert agent.replay.filled <= 6                                # buffer reset after each train (rl_agents.py:280)
    assert not np.array_equal(w_before['out_w'], agent.trainer.keras_weights()['out_w'])
