"""ahk_policy_forward_static (include/ahk.h): the allocation-free variant of the CNN forward — activations between the layer
kernels in a static array of the library's device image — against the existing entry point ahk_policy_forward, through
the C ABI.  No counterpart in the reference.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# tests/test_policy.py's tolerance for the 'rgb' network: 3e-2 of the logit range (bf16 operands, fp32 accumulation).
# Both entry points run the same kernels, so the difference is expected to be zero; the bound is the suite's.
TOL = 3e-2


def _forward(lib, fn, handle, frames, n, head, torch):
    from gym_air_hockey_b200 import _lib
    logits = torch.full((n, 10), float('nan'), dtype=torch.float32, device='cuda')
    action = torch.full((n,), 255, dtype=torch.uint8, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(fn(handle, C.c_void_p(frames.data_ptr()), n, head, C.c_void_p(logits.data_ptr()), C.c_void_p(action.data_ptr()), stream))
    torch.cuda.synchronize()
    return logits.cpu().numpy(), action.cpu().numpy()


def test_static_forward_matches_forward(cuda, table):
    torch = cuda
    from gym_air_hockey_b200 import BatchedAirHockeyEnv, _lib
    from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init
    lib = _lib.load()
    limit = 64                                       # AHK_POLICY_STATIC_MAX_BATCH
    env = BatchedAirHockeyEnv(limit, table=table, seed=31, max_episode_steps=50)
    env.reset()
    for _ in range(4):
        env.step()
    pol = PolicyNetwork(keras_default_init(seed=2, randomize_bn=True), max_batch=256)

    for call, n in enumerate((limit, 37)):           # two calls in a row: the second finds the first one's activations in the scratch
        frames = env.frames[:, :n].contiguous()      # the ring [3][n][3][128][128] of the first n envs
        want_logits, want_action = _forward(lib, lib.ahk_policy_forward, pol._h, frames, n, env.head, torch)
        got_logits, got_action = _forward(lib, lib.ahk_policy_forward_static, pol._h, frames, n, env.head, torch)
        assert np.isfinite(got_logits).all() and np.isfinite(want_logits).all()
        scale = float(np.abs(want_logits).max())
        assert scale > 0
        assert np.abs(got_logits - want_logits).max() <= TOL * scale, (call, n)
        assert np.array_equal(got_action, got_logits.argmax(axis=1)), (call, n)   # each call's action is the argmax of its own logits
        clear = np.sort(want_logits, axis=1)
        decided = (clear[:, -1] - clear[:, -2]) > 2 * TOL * scale                   # rows whose best action is not a near-tie
        assert np.array_equal(got_action[decided], want_action[decided]), (call, n)
        env.step()
        env.step()
    # a call allocates nothing: set-up (symbol address, tensor maps, lazily loaded kernels) is behind us, so the device's free
    # memory must be the same before and after one more call
    frames = env.frames[:, :limit].contiguous()
    logits = torch.empty((limit, 10), dtype=torch.float32, device='cuda')
    action = torch.empty((limit,), dtype=torch.uint8, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(lib.ahk_policy_forward_static(pol._h, C.c_void_p(frames.data_ptr()), limit, env.head, C.c_void_p(logits.data_ptr()),
                                             C.c_void_p(action.data_ptr()), stream))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before

    # sizes beyond the maximum are rejected with a clear error, nothing is launched
    big = torch.zeros((3, limit + 1, 3, 128, 128), dtype=torch.uint8, device='cuda')
    rc = lib.ahk_policy_forward_static(pol._h, C.c_void_p(big.data_ptr()), limit + 1, 0, C.c_void_p(logits.data_ptr()),
                                       C.c_void_p(action.data_ptr()), stream)
    assert rc == -1 and b"AHK_POLICY_STATIC_MAX_BATCH = 64" in lib.ahk_last_error()
    rc = lib.ahk_policy_forward_static(pol._h, C.c_void_p(big.data_ptr()), 0, 0, C.c_void_p(logits.data_ptr()), C.c_void_p(action.data_ptr()), stream)
    assert rc == -1
    gray = PolicyNetwork(max_batch=64, mode='gray-diff')
    rc = lib.ahk_policy_forward_static(gray._h, C.c_void_p(frames.data_ptr()), 8, 0, C.c_void_p(logits.data_ptr()), C.c_void_p(action.data_ptr()), stream)
    assert rc == -1 and b"gray-diff" in lib.ahk_last_error()
    pol.close()
    gray.close()
