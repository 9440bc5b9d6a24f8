"""Per-tensor gradient error of the hand-written backward pass against float32 autograd (oracle/train_ref.py, CPU).
    python tools/train_parity.py [--n 32] -> one JSON object (max |g - g_ref| / max |g_ref| per tensor)"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
import train_ref
from gym_air_hockey_b200 import BatchedAirHockeyEnv
from gym_air_hockey_b200.policy import keras_default_init
from gym_air_hockey_b200.train import NativeTrainer

ap = argparse.ArgumentParser(); ap.add_argument("--n", type=int, default=32); args = ap.parse_args()
n = args.n
w = keras_default_init(seed=5, randomize_bn=True)
tr = NativeTrainer(w, max_batch=max(n, 32))
env = BatchedAirHockeyEnv(n, seed=9, max_episode_steps=50); env.reset()
for _ in range(7): env.step()
order = [(env.head + 1 + i) % 3 for i in range(3)]
frames3 = torch.stack([env.frames[o] for o in order]).contiguous()
dlogits = (np.random.RandomState(3).randn(n, 10) * 0.05).astype(np.float32)
logits = tr.forward(frames3).clone(); tr.backward(torch.from_numpy(dlogits).cuda()); torch.cuda.synchronize()
want_logits, want, stats = train_ref.fit_gradients(w, frames3.cpu().numpy(), dlogits)
acts = [a.cpu().numpy() for a in tr.activations()]
given = train_ref.fit_gradients_given_activations(w, frames3.cpu().numpy(), dlogits, acts)
rep = {"n": n, "logits": float(np.abs(logits.cpu().numpy() - want_logits).max() / np.abs(want_logits).max())}
for name in want:
    got = tr.tensor(name, tr.grads).cpu().numpy(); ref = want[name]
    rep[name] = {"fp32_autograd_l2": round(float(np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30)), 5),
                 "same_activations_l2": round(float(np.linalg.norm(got - given[name]) / max(np.linalg.norm(given[name]), 1e-30)), 6),
                 "same_activations_max": round(float(np.abs(got - given[name]).max() / max(np.abs(given[name]).max(), 1e-30)), 6)}
print(json.dumps(rep))
