import os, sys
sys.path.insert(0, "deep-learning-air-hockey-robot_b200")
import torch
from gym_air_hockey_b200.train import NativeTrainer
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
frames3 = torch.randint(0, 256, (3, b, 3, 128, 128), dtype=torch.uint8, device="cuda")
dlogits = torch.randn(b, 10, device="cuda") * 0.01
tr = NativeTrainer(max_batch=b)
loss = torch.zeros(1, device="cuda")
for _ in range(2):
    tr.forward(frames3); tr.backward(dlogits); tr.step(loss)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(5):
    tr.backward(dlogits)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("backward: host enqueue %.3f ms per call, with final sync %.3f ms per call" % ((t1 - t0) / 5 * 1e3, (t2 - t0) / 5 * 1e3))
