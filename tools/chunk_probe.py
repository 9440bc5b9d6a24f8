"""Does the CNN forward gain from running in sub-batches whose activations stay in L2 (126 MB)?  8192 samples as 1 x 8192,
2 x 4096, 4 x 2048, 8 x 1024 calls, each chunk on its own frames (same input traffic), the policy handle's activation
buffers (max_batch = chunk) reused by every call.   python tools/chunk_probe.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
import torch  # noqa: E402
from gym_air_hockey_b200 import BatchedAirHockeyEnv  # noqa: E402
from gym_air_hockey_b200.policy import PolicyNetwork  # noqa: E402

total = 8192
env = BatchedAirHockeyEnv(total, seed=0, max_episode_steps=256)
env.reset()
for _ in range(3):
    env.step()
stream = torch.cuda.current_stream()
out = {}
for chunk in (8192, 4096, 2048, 1024, 8192):
    k = total // chunk
    rings = [env.frames[:, c * chunk:(c + 1) * chunk].contiguous() for c in range(k)]
    pol = PolicyNetwork(max_batch=chunk)
    logits = torch.empty((chunk, 10), dtype=torch.float32, device="cuda")
    action = torch.empty((chunk,), dtype=torch.uint8, device="cuda")

    def run():
        for r in rings:
            pol.forward_ring(r, env.head, logits, action)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(30):
        run()
    e1.record(stream)
    torch.cuda.synchronize()
    out.setdefault("%d x %d" % (k, chunk), []).append(e0.elapsed_time(e1) / 30)
    pol.close()
    del rings
print(json.dumps({"ms_per_8192_samples": out}))
