"""One `model.fit` batch on the device, timed by phase with CUDA events (csrc/ahk_train.cu; rl_agents.py:189,276):
    python tools/fit_probe.py [--batch 4096] [--library]
-> JSON: training-mode forward, backward, optimizer step, weight push, in ms.  --library also times the float32 PyTorch
autograd checker (oracle/train_ref.py: cuDNN / cuBLAS) on the same batch — what round 1 shipped as the fit."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import torch  # noqa: E402

from gym_air_hockey_b200.policy import PolicyNetwork  # noqa: E402
from gym_air_hockey_b200.train import NativeTrainer  # noqa: E402


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--library", action="store_true")
    args = ap.parse_args()
    b = args.batch
    frames3 = torch.randint(0, 256, (3, b, 3, 128, 128), dtype=torch.uint8, device="cuda")
    dlogits = torch.randn(b, 10, device="cuda") * 0.01
    out = {"batch": b}
    for kind in ("sgd", "adam"):
        tr = NativeTrainer(max_batch=b, optimizer=kind)
        pol = PolicyNetwork(max_batch=64)
        loss = torch.zeros(1, device="cuda")
        rec = {"forward_train_ms": timed(lambda: tr.forward(frames3)), "backward_ms": timed(lambda: tr.backward(dlogits)),
               "step_ms": timed(lambda: tr.step(loss)), "push_weights_ms": timed(lambda: tr.push(pol))}

        def fit():
            tr.forward(frames3)
            tr.backward(dlogits)
            tr.step(loss)
            tr.push(pol)
        rec["fit_ms"] = timed(fit)
        rec["samples_per_s"] = b / rec["fit_ms"] * 1e3
        out[kind] = rec
        del tr
    if args.library:
        import train_ref
        net = train_ref.TrainableNetwork().cuda().train().to(memory_format=torch.channels_last)
        obs = train_ref.stacked(frames3)

        def fb():
            for p in net.parameters():
                p.grad = None
            x = ((obs.float() - 128.0) / 128.0).contiguous(memory_format=torch.channels_last)
            y = net.out(torch.relu(net.dense2(torch.relu(net.dense1(net.bn3(torch.relu(net.conv3(net.bn2(torch.relu(net.conv2(net.bn1(torch.relu(net.conv1(x))))))))).flatten(1))))))
            y.backward(dlogits)
        out["library_float32_autograd_fwd_bwd_ms"] = timed(fb, 5)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
