"""model.fit on the device (rl_agents.py:189,276): the hand-written training-mode forward, backward and
optimizer kernels (csrc/ahk_train.cu, ahk_policy_forward_train) against oracle/train_ref.py — float32 PyTorch autograd on
the CPU — and against the Keras optimizer restatements (oracle/rl_ref.py, oracle/train_ref.py).

Tolerances.  The forward kernels keep activations in bf16 (8 significant bits) and accumulate in float32; the backward
kernels accumulate in float32 over those activations with the float32 master weights.  Two comparisons:
  * given the SAME activations (train_ref.fit_gradients_given_activations: autograd whose ReLU outputs take the kernels'
    values and masks), every gradient tensor agrees to 1e-4 of its largest entry (measured 1e-6: float32 summation order)
    with the dense data gradients on the float32 functor GEMM (AHK_TRAIN_DENSE=simt) — this pins the backward arithmetic —
    and to 3e-2 (measured 1.5e-2) in the default configuration, where those two GEMMs run on tcgen05 with bf16 operands;
  * against a pure float32 forward + backward the logits agree to 2e-2 of the logit range and the batch statistics to
    1e-2, but a bf16 forward flips the sign of ~0.3 % of the near-zero pre-activations, and every flipped ReLU is a whole
    gradient element: per-tensor L2 differences of 8-25 % at random initialisation (tools/train_parity.py,
    profiles/train_parity_r2_*.json).  The test bounds the whole gradient vector at 0.2 (measured 0.09) and each tensor at 0.6,
    so that a structural error still fails.
Optimizer steps are float32 restatements of float64 formulas: 2e-5 relative.
"""
import numpy as np
import pytest


def _frames(torch, table, n, seed=9, steps=7):
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    env = BatchedAirHockeyEnv(n, table=table, seed=seed, max_episode_steps=50)
    env.reset()
    for _ in range(steps):
        env.step()
    fr = env.frames
    order = [(env.head + 1 + i) % 3 for i in range(3)]  # oldest .. newest -> ring layout with head 2
    return torch.stack([fr[order[0]], fr[order[1]], fr[order[2]]]).contiguous()


GRAD_NAMES = ("conv1_w", "conv1_b", "bn1_gamma", "bn1_beta", "conv2_w", "conv2_b", "bn2_gamma", "bn2_beta", "conv3_w", "conv3_b",
              "bn3_gamma", "bn3_beta", "dense1_w", "dense1_b", "dense2_w", "dense2_b", "out_w", "out_b")


@pytest.mark.gpu
@pytest.mark.parametrize("n,dense", [(32, "tcgen05"), (200, "tcgen05"), (32, "simt"), (200, "simt")])
def test_forward_backward_match_autograd(cuda, table, n, dense, monkeypatch):
    """dense: the two dense data-gradient GEMMs on the tcgen05 GEMM (bf16 operands: the default) or on the float32 functor
    GEMM (AHK_TRAIN_DENSE=simt, the A/B switch) — the latter pins the arithmetic to 1e-4, the former is bounded at 3e-2."""
    import train_ref
    monkeypatch.setenv("AHK_TRAIN_DENSE", dense)
    from gym_air_hockey_b200.policy import keras_default_init
    from gym_air_hockey_b200.train import NativeTrainer
    torch = cuda
    w = keras_default_init(seed=5, randomize_bn=True)
    tr = NativeTrainer(w, max_batch=256)
    frames3 = _frames(torch, table, n)
    dlogits = (np.random.RandomState(3).randn(n, 10) * 0.05).astype(np.float32)
    logits = tr.forward(frames3).clone()
    tr.backward(torch.from_numpy(dlogits).cuda())
    torch.cuda.synchronize()
    want_logits, want, stats = train_ref.fit_gradients(w, frames3.cpu().numpy(), dlogits)
    scale = np.abs(want_logits).max()
    assert np.abs(logits.cpu().numpy() - want_logits).max() <= 2e-2 * scale
    bs = tr.batch_stats.cpu().numpy()
    for name, at, c in (("bn1", 0, 5), ("bn2", 8, 32), ("bn3", 40, 64)):
        assert np.allclose(bs[at:at + c], stats[name + "_mean"], rtol=1e-2, atol=1e-3), name
        assert np.allclose(bs[104 + at:104 + at + c], stats[name + "_var"], rtol=2e-2, atol=1e-4), name
    acts = [a.cpu().numpy() for a in tr.activations()]
    given = train_ref.fit_gradients_given_activations(w, frames3.cpu().numpy(), dlogits, acts)
    tight, loose = {}, {}
    for name in GRAD_NAMES:
        got = tr.tensor(name, tr.grads).cpu().numpy()
        tight[name] = float(np.abs(got - given[name]).max() / max(np.abs(given[name]).max(), 1e-30))
        loose[name] = float(np.linalg.norm(got - want[name]) / max(np.linalg.norm(want[name]), 1e-30))
    bound = 1e-4 if dense == "simt" else 3e-2  # measured 6e-6 / 1.5e-2
    assert not {k: v for k, v in tight.items() if not v <= bound}, tight
    assert not {k: v for k, v in loose.items() if not v <= 0.6}, loose  # 5-element sums of cancelling terms (d beta1) reach 0.36
    flat_got = np.concatenate([tr.tensor(name, tr.grads).cpu().numpy().ravel() for name in GRAD_NAMES])
    flat_ref = np.concatenate([want[name].ravel() for name in GRAD_NAMES])
    assert np.linalg.norm(flat_got - flat_ref) <= 0.2 * np.linalg.norm(flat_ref)  # the whole gradient: measured 0.09


@pytest.mark.gpu
def test_sgd_and_adam_steps_match_keras_restatement(cuda):
    import rl_ref
    import train_ref
    from gym_air_hockey_b200.policy import SHAPES, keras_default_init
    from gym_air_hockey_b200.train import NativeTrainer
    torch = cuda
    w = keras_default_init(seed=2, randomize_bn=True)
    names = [name for name, _ in SHAPES if not name.endswith(("_mean", "_var"))]
    rs = np.random.RandomState(4)
    for kind in ("sgd", "adam"):
        tr = NativeTrainer(w, max_batch=8, optimizer=kind, l2=1e-4, l2_exclude=("out_w",))
        steps = []
        for scale in (1e-3, 0.5, 20.0):  # SGD: the last two clip
            g = {name: (rs.randn(*dict(SHAPES)[name]) * scale / np.sqrt(np.prod(dict(SHAPES)[name]))).astype(np.float32) for name in names}
            steps.append(g)
        params = [w[name].astype(np.float64) for name in names]
        raw_steps = []
        for g in steps:
            tr.grads.zero_()
            for name in names:
                tr.tensor(name, tr.grads).copy_(torch.from_numpy(g[name]))
            before = tr.keras_weights()
            loss = torch.zeros(1, dtype=torch.float32, device="cuda")
            tr.batch_stats.fill_(0.25)
            tr.step(loss)
            torch.cuda.synchronize()
            # the regulariser's share: g + 2 l2 W on kernels except the excluded one; penalty l2 sum W^2
            reg = [g[name] + (2e-4 * before[name] if name.endswith("_w") and name != "out_w" else 0) for name in names]
            raw_steps.append(reg)
            penalty = sum(1e-4 * np.sum(before[name].astype(np.float64) ** 2) for name in names if name.endswith("_w") and name != "out_w")
            assert abs(loss.item() - penalty) <= 1e-5 * penalty
            after = tr.keras_weights()
            for bn in ("bn1", "bn2", "bn3"):  # moving statistics: 0.99 * old + 0.01 * batch
                assert np.allclose(after[bn + "_mean"], 0.99 * before[bn + "_mean"] + 0.01 * 0.25, rtol=1e-6, atol=1e-7)
                assert np.allclose(after[bn + "_var"], 0.99 * before[bn + "_var"] + 0.01 * 0.25, rtol=1e-6, atol=1e-7)
        want = rl_ref.keras_sgd_steps(params, raw_steps) if kind == "sgd" else train_ref.keras_adam_steps(params, raw_steps)
        got = tr.keras_weights()
        for name, wv in zip(names, want):
            assert np.allclose(got[name], wv, rtol=2e-5, atol=2e-7), (kind, name, np.abs(got[name] - wv).max())
        assert tr.iterations == 3
