"""Plain PyTorch float32 restatement of the reference's 'rgb' CNN — TEST INFRASTRUCTURE, never on the product path.

Follows ai/supervised/keras/models.py:20-65 with the Keras defaults that file relies on: data_format
channels_first, padding 'valid', bias on, activation applied inside the layer (so BatchNormalization, axis=1,
epsilon 1e-3, inference mode, comes AFTER the ReLU), Flatten in (C, H, W) order, softmax only on 'out'.
Keras itself is not installed on either box, so parity for the network is pinned to this restatement
plus the layer shapes that follow from models.py: (5,41,41) -> (32,19,19) -> (64,8,8) -> 4096.

`forward(..., emulate_bf16=True)` rounds the operands at exactly the points where the tcgen05 kernels do (weights after
folding, activations after each ReLU: bfloat16 everywhere except conv1, whose default kernel multiplies the exact uint8
pixels with UNSCALED float16 weights and applies 1/128 and the folded bias to the float32 sum; `conv1='bf16'` is the
older kernel kept behind AHK_CONV1=n144: bfloat16 weights / 128), so the CUDA result can be checked tightly, while
`emulate_bf16=False` is the float32 model the stated tolerance refers to.
"""
import numpy as np
import torch
import torch.nn.functional as F

EPS = 1e-3  # keras.layers.BatchNormalization default epsilon


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).double()


def forward(w, obs_u8, emulate_bf16=False, conv1=None):
    """w: dict of Keras-layout float32 arrays (gym_air_hockey_b200.policy.SHAPES); obs_u8: uint8 [n, 9, 128, 128]
    stacked newest-first (processor.py:47-49).  Returns float32 logits [n, 10] (pre-softmax)."""
    x = torch.from_numpy(np.ascontiguousarray(obs_u8)).double()
    if not emulate_bf16:
        x = ((x.float() - 128.0) / 128.0)  # processor.py:39-42
        def bn(v, p):
            s = torch.from_numpy(w[p + '_gamma'] / np.sqrt(w[p + '_var'] + np.float32(EPS)))
            t = torch.from_numpy(w[p + '_beta']) - torch.from_numpy(w[p + '_mean']) * s
            return v * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)
        def conv(v, p, stride):
            k = torch.from_numpy(np.ascontiguousarray(w[p + '_w'].transpose(3, 2, 0, 1)))  # [kh][kw][ci][co] -> [co][ci][kh][kw]
            return F.relu(F.conv2d(v, k, torch.from_numpy(w[p + '_b']), stride=stride))
        x = bn(conv(x, 'conv1', 3), 'bn1')
        x = bn(conv(x, 'conv2', 2), 'bn2')
        x = bn(conv(x, 'conv3', 2), 'bn3')
        x = x.reshape(x.shape[0], -1)  # Flatten, channels_first: index c*64 + y*8 + x
        x = F.relu(x @ torch.from_numpy(w['dense1_w']) + torch.from_numpy(w['dense1_b']))
        x = F.relu(x @ torch.from_numpy(w['dense2_w']) + torch.from_numpy(w['dense2_b']))
        return (x @ torch.from_numpy(w['out_w']) + torch.from_numpy(w['out_b'])).numpy()

    # ---- the same network with the engine's exact algebra: normalisation and BatchNorms folded into the NEXT
    # layer's weights/bias (float32 fold, then bfloat16 weights), bfloat16 activations, wide accumulation.
    q = lambda t: t.float().bfloat16().double()  # round to nearest even, like cvt.rn.bf16
    def affine(p):
        s = (w[p + '_gamma'] / np.sqrt(w[p + '_var'] + np.float32(EPS))).astype(np.float32)
        return s, (w[p + '_beta'] - w[p + '_mean'] * s).astype(np.float32)
    s1, t1 = affine('bn1'); s2, t2 = affine('bn2'); s3, t3 = affine('bn3')

    k1 = w['conv1_w'].astype(np.float32)
    b1 = (w['conv1_b'] - k1.astype(np.float64).sum(axis=(0, 1, 2)).astype(np.float32)).astype(np.float32)
    k1 = (k1 / np.float32(128.0)).astype(np.float32)
    k2 = (w['conv2_w'] * s1.reshape(1, 1, -1, 1)).astype(np.float32)
    b2 = (w['conv2_b'] + (w['conv2_w'].astype(np.float64) * t1.reshape(1, 1, -1, 1)).sum(axis=(0, 1, 2)).astype(np.float32)).astype(np.float32)
    k3 = (w['conv3_w'] * s2.reshape(1, 1, -1, 1)).astype(np.float32)
    b3 = (w['conv3_b'] + (w['conv3_w'].astype(np.float64) * t2.reshape(1, 1, -1, 1)).sum(axis=(0, 1, 2)).astype(np.float32)).astype(np.float32)
    s3f, t3f = np.repeat(s3, 64), np.repeat(t3, 64)  # flatten index c*64 + pixel
    d1 = (w['dense1_w'] * s3f[:, None]).astype(np.float32)
    bd1 = (w['dense1_b'] + (w['dense1_w'].astype(np.float64) * t3f[:, None]).sum(axis=0).astype(np.float32)).astype(np.float32)

    def conv(v, k, b, stride):
        kk = q(_t(k.transpose(3, 2, 0, 1)))
        return q(F.relu(F.conv2d(v, kk, _t(b), stride=stride).float().double()))
    import os
    if (conv1 or ('bf16' if os.environ.get('AHK_CONV1') == 'n144' else 'fp16')) == 'fp16':
        kk = _t(w['conv1_w'].astype(np.float16).astype(np.float32).transpose(3, 2, 0, 1))
        acc = F.conv2d(x, kk, None, stride=3).float()  # exact products, float32 accumulator
        x = q(F.relu(acc * np.float32(1.0 / 128.0) + torch.from_numpy(b1).view(1, -1, 1, 1)).double())
    else:
        x = conv(x, k1, b1, 3)
    x = conv(x, k2, b2, 2)
    x = conv(x, k3, b3, 2)
    x = x.reshape(x.shape[0], -1)
    x = q(F.relu((x @ q(_t(d1)) + _t(bd1)).float().double()))
    x = q(F.relu((x @ q(_t(w['dense2_w'])) + _t(w['dense2_b'])).float().double()))
    return (x @ q(_t(w['out_w'])) + _t(w['out_b'])).float().numpy()


def gray_diff_input(obs_u8):
    """DataProcessor's 'gray-diff' observation (processor.py:53-64) of a uint8 [n, 9, 128, 128] stack (newest first):
    uint8 wrap-around oldest - middle - newest per colour, 0.2989 R + 0.5870 G + 0.1140 B, (x - 128) / 128.
    float64 [n, 1, 128, 128], as the reference returns it."""
    o = np.ascontiguousarray(obs_u8, dtype=np.uint8)
    diff = o[:, 6:9] - o[:, 3:6] - o[:, 0:3]  # uint8 arithmetic wraps
    gray = 0.2989 * diff[:, 0] + 0.5870 * diff[:, 1] + 0.1140 * diff[:, 2]
    return ((gray - 128.0) / 128.0)[:, None]


def forward_gray(w, obs_u8, emulate_bf16=False):
    """The 'gray-diff' network (models.py:67-116; same Keras defaults as above): conv 16 k8 s4 -> 24 k4 s2 -> 32 k4 s1
    -> 48 k4 s1 (ReLU then BatchNorm each) -> flatten (C,H,W) 3072 -> dense 512 ReLU -> dense 10.  obs_u8: the uint8
    stack; w: gym_air_hockey_b200.policy.GRAY_SHAPES arrays.  float32 logits [n, 10]."""
    x = torch.from_numpy(gray_diff_input(obs_u8))
    strides = (4, 2, 1, 1)
    if not emulate_bf16:
        x = x.float()
        for i, st in enumerate(strides, start=1):
            k = torch.from_numpy(np.ascontiguousarray(w['conv%d_w' % i].transpose(3, 2, 0, 1)))
            x = F.relu(F.conv2d(x, k, torch.from_numpy(w['conv%d_b' % i]), stride=st))
            s = torch.from_numpy(w['bn%d_gamma' % i] / np.sqrt(w['bn%d_var' % i] + np.float32(EPS)))
            t = torch.from_numpy(w['bn%d_beta' % i]) - torch.from_numpy(w['bn%d_mean' % i]) * s
            x = x * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)
        x = x.reshape(x.shape[0], -1)
        x = F.relu(x @ torch.from_numpy(w['dense1_w']) + torch.from_numpy(w['dense1_b']))
        return (x @ torch.from_numpy(w['out_w']) + torch.from_numpy(w['out_b'])).numpy()

    # the engine's algebra: bf16 input image, each BatchNorm folded into the next layer, bf16 weights and activations
    q = lambda t: t.float().bfloat16().double()
    x = q(x.float())  # the input image is computed in float32 and stored as bfloat16
    s_prev = t_prev = None
    for i, st in enumerate(strides, start=1):
        k = w['conv%d_w' % i].astype(np.float32)
        b = w['conv%d_b' % i].astype(np.float32)
        if s_prev is not None:
            b = (b + (k.astype(np.float64) * t_prev.reshape(1, 1, -1, 1)).sum(axis=(0, 1, 2)).astype(np.float32)).astype(np.float32)
            k = (k * s_prev.reshape(1, 1, -1, 1)).astype(np.float32)
        x = q(F.relu(F.conv2d(x, q(_t(k.transpose(3, 2, 0, 1))), _t(b), stride=st).float().double()))
        s_prev = (w['bn%d_gamma' % i] / np.sqrt(w['bn%d_var' % i] + np.float32(EPS))).astype(np.float32)
        t_prev = (w['bn%d_beta' % i] - w['bn%d_mean' % i] * s_prev).astype(np.float32)
    sf, tf = np.repeat(s_prev, 64), np.repeat(t_prev, 64)  # flatten index c*64 + pixel
    d1 = (w['dense1_w'] * sf[:, None]).astype(np.float32)
    bd1 = (w['dense1_b'] + (w['dense1_w'].astype(np.float64) * tf[:, None]).sum(axis=0).astype(np.float32)).astype(np.float32)
    x = x.reshape(x.shape[0], -1)
    x = q(F.relu((x @ q(_t(d1)) + _t(bd1)).float().double()))
    return (x @ q(_t(w['out_w'])) + _t(w['out_b'])).float().numpy()
