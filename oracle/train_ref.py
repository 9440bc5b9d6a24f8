"""CHECKER (test infrastructure, never on the product path): `model.fit`'s arithmetic restated in float32 PyTorch autograd.

`TrainableNetwork` is models.py:20-65 as a torch module — Conv2D(+ReLU) -> BatchNormalization(axis=1) x3, Flatten (C,H,W),
Dense 512 ReLU x2, Dense 10 (linear: logits) — with the Keras BatchNorm defaults (epsilon 1e-3, momentum 0.99; in
training mode it normalises with the batch's mean and biased variance, as `fit` does).  `fit_gradients` runs one
training-mode forward + backward and returns logits, gradients (Keras layouts) and the batch statistics the hand-written
kernels (csrc/ahk_train.cu, gym_air_hockey_b200/train.py) are compared with.  `keras_adam_steps` restates
keras.optimizers.Adam.get_updates (Keras 2.x, models.py:119); SGD is in rl_ref.keras_sgd_steps.

Only tests/ and tools/ import this file.  It was the product's backward pass in round 1 (PyTorch autograd on cuDNN/cuBLAS).
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'deep-learning-air-hockey-robot_b200'))
from gym_air_hockey_b200.policy import SHAPES, keras_default_init  # noqa: E402  (shapes and the Keras initialiser only: no kernels)


class TrainableNetwork(nn.Module):
    def __init__(self, weights=None):
        super().__init__()
        self.conv1, self.bn1 = nn.Conv2d(9, 5, 8, stride=3), nn.BatchNorm2d(5, eps=1e-3, momentum=0.01)
        self.conv2, self.bn2 = nn.Conv2d(5, 32, 4, stride=2), nn.BatchNorm2d(32, eps=1e-3, momentum=0.01)
        self.conv3, self.bn3 = nn.Conv2d(32, 64, 4, stride=2), nn.BatchNorm2d(64, eps=1e-3, momentum=0.01)
        self.dense1, self.dense2, self.out = nn.Linear(4096, 512), nn.Linear(512, 512), nn.Linear(512, 10)
        self.load_keras(keras_default_init() if weights is None else weights)

    def forward(self, obs_u8):
        """obs_u8: uint8 [B, 9, 128, 128], newest frame first (processor.py:47-49) -> logits float32 [B, 10]."""
        x = (obs_u8.float() - 128.0) / 128.0  # processor.py:39-42
        x = self.bn1(F.relu(self.conv1(x)))
        x = self.bn2(F.relu(self.conv2(x)))
        x = self.bn3(F.relu(self.conv3(x)))
        x = x.flatten(1)
        x = F.relu(self.dense1(x))
        x = F.relu(self.dense2(x))
        return self.out(x)

    def load_keras(self, w):
        with torch.no_grad():
            for i in (1, 2, 3):
                conv, bn = getattr(self, 'conv%d' % i), getattr(self, 'bn%d' % i)
                conv.weight.copy_(torch.from_numpy(np.ascontiguousarray(np.asarray(w['conv%d_w' % i], dtype=np.float32).transpose(3, 2, 0, 1))))
                conv.bias.copy_(torch.from_numpy(np.asarray(w['conv%d_b' % i], dtype=np.float32)))
                bn.weight.copy_(torch.from_numpy(np.asarray(w['bn%d_gamma' % i], dtype=np.float32)))
                bn.bias.copy_(torch.from_numpy(np.asarray(w['bn%d_beta' % i], dtype=np.float32)))
                bn.running_mean.copy_(torch.from_numpy(np.asarray(w['bn%d_mean' % i], dtype=np.float32)))
                bn.running_var.copy_(torch.from_numpy(np.asarray(w['bn%d_var' % i], dtype=np.float32)))
            for name in ('dense1', 'dense2', 'out'):
                lin = getattr(self, name)
                lin.weight.copy_(torch.from_numpy(np.ascontiguousarray(np.asarray(w[name + '_w'], dtype=np.float32).T)))
                lin.bias.copy_(torch.from_numpy(np.asarray(w[name + '_b'], dtype=np.float32)))

    def keras_tensors(self, grad=False):
        """-> dict of float32 numpy arrays in Keras layouts (policy.SHAPES): the parameters, or their gradients."""
        pick = (lambda p: p.grad) if grad else (lambda p: p.detach())
        w = {}
        for i in (1, 2, 3):
            conv, bn = getattr(self, 'conv%d' % i), getattr(self, 'bn%d' % i)
            w['conv%d_w' % i] = pick(conv.weight).permute(2, 3, 1, 0).contiguous().cpu().numpy()
            w['conv%d_b' % i] = pick(conv.bias).cpu().numpy().copy()
            w['bn%d_gamma' % i], w['bn%d_beta' % i] = pick(bn.weight).cpu().numpy().copy(), pick(bn.bias).cpu().numpy().copy()
            if not grad:
                w['bn%d_mean' % i], w['bn%d_var' % i] = bn.running_mean.cpu().numpy().copy(), bn.running_var.cpu().numpy().copy()
        for name in ('dense1', 'dense2', 'out'):
            lin = getattr(self, name)
            w[name + '_w'] = pick(lin.weight).t().contiguous().cpu().numpy()
            w[name + '_b'] = pick(lin.bias).cpu().numpy().copy()
        return {k: np.ascontiguousarray(v, dtype=np.float32) for k, v in w.items()}

    def keras_weights(self):
        w = self.keras_tensors()
        return {name: w[name] for name, _ in SHAPES}


def stacked(frames3):
    """uint8 [3, B, 3, 128, 128] oldest..newest (ring layout, head 2) -> the (9,128,128) stack per sample, newest first."""
    return torch.cat((frames3[2], frames3[1], frames3[0]), dim=1)


def fit_gradients(weights, frames3, dlogits):
    """One training-mode forward + backward in float32 on the CPU.  frames3: uint8 numpy [3, B, 3, 128, 128] (head 2);
    dlogits float32 [B, 10].  -> logits, dict of gradients (Keras layouts), dict of batch statistics (mean, biased var)."""
    net = TrainableNetwork(weights).train()
    stats = {}

    def hook(name):
        def fn(mod, inp, out):
            x = inp[0].detach()
            stats[name + '_mean'] = x.mean(dim=(0, 2, 3)).numpy()
            stats[name + '_var'] = x.var(dim=(0, 2, 3), unbiased=False).numpy()
        return fn
    for name in ('bn1', 'bn2', 'bn3'):
        getattr(net, name).register_forward_hook(hook(name))
    logits = net(stacked(torch.from_numpy(frames3)))
    logits.backward(torch.from_numpy(dlogits))
    return logits.detach().numpy(), net.keras_tensors(grad=True), stats


def keras_adam_steps(params, grads_per_step, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7):
    """keras.optimizers.Adam.get_updates (Keras 2.x): lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t); p -= lr_t * m / (sqrt(v) + eps)."""
    params = [np.array(p, dtype=np.float64) for p in params]
    m = [np.zeros_like(p) for p in params]
    v = [np.zeros_like(p) for p in params]
    for it, grads in enumerate(grads_per_step):
        t = it + 1
        lr_t = lr * np.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
        for p, mi, vi, g in zip(params, m, v, grads):
            g = np.asarray(g, dtype=np.float64)
            mi *= beta1; mi += (1 - beta1) * g
            vi *= beta2; vi += (1 - beta2) * g * g
            p -= lr_t * mi / (np.sqrt(vi) + eps)
    return params


class _ReluAs(torch.autograd.Function):
    """relu whose forward VALUE is a given tensor (the kernels' bf16 activation) and whose backward uses that tensor's mask:
    isolates the backward arithmetic from the forward's rounding (a bf16 forward flips the sign of a fraction of a percent
    of near-zero pre-activations against a float32 forward, and each flip is a whole gradient element)."""

    @staticmethod
    def forward(ctx, z, value):
        ctx.save_for_backward(value)
        return value.clone()

    @staticmethod
    def backward(ctx, g):
        (value,) = ctx.saved_tensors
        return g * (value > 0).to(g.dtype), None


def fit_gradients_given_activations(weights, frames3, dlogits, acts):
    """The same step with the forward VALUES of every ReLU output replaced by `acts` = (r1 [B,1681,8], r2 [B,361,32],
    r3 [B,64(px),64(c)], h1, h2) as float32 numpy arrays — what the forward kernels produced.  BatchNorm then normalises
    those values with their own batch statistics, exactly as the kernels do.  -> gradients (Keras layouts)."""
    net = TrainableNetwork(weights).train()
    b = frames3.shape[1]
    r1 = torch.from_numpy(acts[0][:, :, :5]).permute(0, 2, 1).reshape(b, 5, 41, 41).contiguous()
    r2 = torch.from_numpy(acts[1]).permute(0, 2, 1).reshape(b, 32, 19, 19).contiguous()
    r3 = torch.from_numpy(acts[2]).permute(0, 2, 1).reshape(b, 64, 8, 8).contiguous()
    h1, h2 = torch.from_numpy(acts[3]), torch.from_numpy(acts[4])
    x = (stacked(torch.from_numpy(frames3)).float() - 128.0) / 128.0
    a1 = net.bn1(_ReluAs.apply(net.conv1(x), r1))
    a2 = net.bn2(_ReluAs.apply(net.conv2(a1), r2))
    a3 = net.bn3(_ReluAs.apply(net.conv3(a2), r3))
    y1 = _ReluAs.apply(net.dense1(a3.flatten(1)), h1)
    y2 = _ReluAs.apply(net.dense2(y1), h2)
    net.out(y2).backward(torch.from_numpy(dlogits))
    return net.keras_tensors(grad=True)
