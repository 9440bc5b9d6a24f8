"""conv23_kernel (AHK_FUSE23=1: conv2 -> conv3 in one kernel, act2 in shared memory) against the two-kernel path: logits must be
bit-identical (same arithmetic), then both are timed.   python tools/fuse_ab.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import torch  # noqa: E402
from gym_air_hockey_b200 import BatchedAirHockeyEnv  # noqa: E402
from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init  # noqa: E402

w = keras_default_init(seed=5, randomize_bn=True)
res = {"identical": {}, "forward_ms": {}}
stream = torch.cuda.current_stream()
for n in (8192, 37, 1, 300):
    env = BatchedAirHockeyEnv(n, seed=3, max_episode_steps=50)
    env.reset()
    for _ in range(5):
        env.step()
    outs = {}
    for mode in ("0", "1"):
        os.environ["AHK_FUSE23"] = mode
        net = PolicyNetwork(w, max_batch=n)
        logits, action = net.forward_ring(env.frames, env.head)
        torch.cuda.synchronize()
        outs[mode] = (logits.clone(), action.clone())
        if mode == "1":  # a second call: stale act2 / accumulators from the first must not matter
            l2, _ = net.forward_ring(env.frames, env.head)
            torch.cuda.synchronize()
            res["identical"]["%d repeat" % n] = bool(torch.equal(l2, outs["1"][0]))
        if n == 8192:
            for _ in range(3):
                net.forward_ring(env.frames, env.head, logits, action)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for rep in range(2):
                e0.record(stream)
                for _ in range(50):
                    net.forward_ring(env.frames, env.head, logits, action)
                e1.record(stream)
                torch.cuda.synchronize()
                res["forward_ms"].setdefault("fuse23=" + mode, []).append(e0.elapsed_time(e1) / 50)
        net.close()
    res["identical"][str(n)] = bool(torch.equal(outs["0"][0], outs["1"][0]) and torch.equal(outs["0"][1], outs["1"][1]))
    res.setdefault("max_abs_diff", {})[str(n)] = float((outs["0"][0] - outs["1"][0]).abs().max())
    env.close()
print(json.dumps(res))
