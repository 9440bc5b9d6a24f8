"""`model.fit` on the device, hand-written (csrc/ahk_train.cu): ≙ the forward + backward + optimizer step Keras runs at
ai/reinforcement/rl_agents.py:189 (DDQN: mean-squared error, SGD lr 0.0075 momentum 0.5 decay 1e-6 clipnorm 2,
rl_agents.py:111,170) and :276 (PG: categorical cross-entropy, the pretrained model's Adam, models.py:119).

`NativeTrainer` owns ONE flat float32 parameter vector laid out as `model.get_weights()` (the order of
`ahk_policy_weights`, Keras layouts: conv [kh][kw][cin][cout], dense [in][out]; BatchNorm moving statistics included), its
gradient and optimizer-state twins, a policy handle for the training-mode forward pass and an `ahk_trainer` for the backward
pass.  Nothing here calls autograd, cuDNN or cuBLAS; PyTorch allocates the tensors and (with several ranks) all-reduces the
flat gradient over NCCL.  The float32 autograd restatement of the same step lives in oracle/train_ref.py as the checker.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .policy import SHAPES, PolicyNetwork, keras_default_init

L2_REGULARIZER = 1e-4  # models.py:17
BN_MOMENTUM = 0.99     # keras.layers.BatchNormalization default


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class NativeTrainer(object):
    def __init__(self, weights=None, max_batch=32, device=0, optimizer='sgd', lr=None, momentum=0.5, decay=1e-6, clipnorm=2.0,
                 beta1=0.9, beta2=0.999, eps=1e-7, l2=L2_REGULARIZER, l2_exclude=()):
        """optimizer 'sgd': keras SGD(lr 0.0075, momentum, decay, clipnorm) — rl_agents.py:111,170; 'adam': keras Adam(lr 1e-3)
        — models.py:119.  l2_exclude: kernels built without the regulariser (the DDQN agent's fresh linear 'out' layer,
        rl_agents.py:166)."""
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError('NativeTrainer needs a CUDA device; there is no CPU fallback')
        if optimizer not in ('sgd', 'adam'):
            raise ValueError('optimizer must be sgd or adam')
        self.device = torch.device('cuda', device)
        self.max_batch = int(max_batch)
        self.optimizer = optimizer
        self.lr = float(lr if lr is not None else (0.0075 if optimizer == 'sgd' else 1e-3))
        self.momentum, self.decay, self.clipnorm = float(momentum), float(decay), float(clipnorm)
        self.beta1, self.beta2, self.eps, self.l2 = float(beta1), float(beta2), float(eps), float(l2)
        self.iterations = 0
        weights = keras_default_init() if weights is None else weights
        self.offsets, at = {}, 0
        for name, shape in SHAPES:
            self.offsets[name] = (at, int(np.prod(shape)))
            at += int(np.prod(shape))
        self.count = at
        flat = np.concatenate([np.ascontiguousarray(weights[name], dtype=np.float32).reshape(-1) for name, _ in SHAPES])
        self.params = torch.from_numpy(flat).to(self.device)
        self.grads = torch.zeros_like(self.params)
        self.m1 = torch.zeros_like(self.params)
        self.m2 = torch.zeros_like(self.params) if optimizer == 'adam' else None
        self.batch_stats = torch.zeros(2 * _lib.BN_STATS, dtype=torch.float32, device=self.device)
        self.logits = torch.zeros((self.max_batch, 10), dtype=torch.float32, device=self.device)
        segs = (_lib.ParamSegment * len(SHAPES))()
        stat_at = {'bn1': 0, 'bn2': 8, 'bn3': 40}
        for i, (name, _) in enumerate(SHAPES):
            off, cnt = self.offsets[name]
            flags, stat = _lib.SEG_TRAINABLE, 0
            if name.endswith('_mean'):
                flags, stat = _lib.SEG_BN_MEAN, stat_at[name[:3]]
            elif name.endswith('_var'):
                flags, stat = _lib.SEG_BN_VAR, stat_at[name[:3]]
            elif name.endswith('_w') and name not in l2_exclude:
                flags |= _lib.SEG_L2
            segs[i] = _lib.ParamSegment(off, cnt, flags, stat)
        self.segments = segs
        self.w_struct = self._struct(self.params)
        self.g_struct = self._struct(self.grads)
        self.policy = PolicyNetwork(weights, max_batch=self.max_batch, device=device)  # the training-mode forward runs on its kernels
        h = C.c_void_p()
        _lib.check(self.lib.ahk_trainer_create(self.device.index, self.max_batch, C.byref(h)))
        self._h = h
        self._last = None

    def __del__(self):
        h = getattr(self, '_h', None)
        if h:
            self.lib.ahk_trainer_destroy(h)
            self._h = None

    def _struct(self, flat):
        st = _lib.PolicyWeights()
        for name, _ in SHAPES:
            setattr(st, name, C.c_void_p(flat.data_ptr() + 4 * self.offsets[name][0]))
        return st

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def tensor(self, name, flat=None):
        """View of one tensor of the flat vector (params by default) in its Keras shape."""
        off, cnt = self.offsets[name]
        return (self.params if flat is None else flat)[off:off + cnt].view(dict(SHAPES)[name])

    def keras_weights(self):
        """-> dict of float32 numpy arrays in Keras layouts: what model.get_weights() would list."""
        flat = self.params.cpu().numpy()
        return {name: flat[off:off + cnt].reshape(dict(SHAPES)[name]).copy() for name, (off, cnt) in self.offsets.items()}

    def forward(self, frames3, head=2):
        """Training-mode forward (BatchNorm on batch statistics) of uint8 frames [3, B, 3, 128, 128] in ring layout
        (ReplayStore.gather()[0:3], head 2) -> logits float32 [B, 10] (a view, overwritten by the next call)."""
        b = int(frames3.shape[1])
        if frames3.dtype != torch.uint8 or tuple(frames3.shape) != (3, b, 3, 128, 128) or not frames3.is_contiguous():
            raise ValueError('frames3 must be a contiguous uint8 [3, B, 3, 128, 128] tensor')
        if b > self.max_batch:
            raise ValueError('batch %d exceeds max_batch %d' % (b, self.max_batch))
        _lib.check(self.lib.ahk_policy_forward_train(self.policy._h, C.byref(self.w_struct), _ptr(frames3), b, int(head), _ptr(self.logits),
                                                     _ptr(self.batch_stats), self._stream()))
        self._last = (frames3, b, int(head))
        return self.logits[:b]

    def backward(self, dlogits):
        """d loss / d logits float32 [B, 10] -> self.grads (every trainable tensor, Keras layouts)."""
        frames3, b, head = self._last
        if dlogits.dtype != torch.float32 or tuple(dlogits.shape) != (b, 10) or not dlogits.is_contiguous():
            raise ValueError('dlogits must be a contiguous float32 [B, 10] tensor')
        _lib.check(self.lib.ahk_train_backward(self._h, self.policy._h, C.byref(self.w_struct), _ptr(frames3), b, head, _ptr(self.batch_stats),
                                               _ptr(dlogits), C.byref(self.g_struct), self._stream()))

    def activations(self):
        """Diagnostics: the layer outputs the last forward() left in the policy handle, as float32 tensors
        r1 [B,1681,8], r2 [B,361,32], r3 [B,64,64] (pixel, channel), h1 [B,512], h2 [B,512]."""
        _, b, _ = self._last
        out = []
        for which, shape in enumerate(((b, 1681, 8), (b, 361, 32), (b, 64, 64), (b, 512), (b, 512))):
            t = torch.empty(shape, dtype=torch.float32, device=self.device)
            _lib.check(self.lib.ahk_policy_copy_activation(self.policy._h, which, _ptr(t), t.numel(), self._stream()))
            out.append(t)
        return out

    def all_reduce(self, group=None):
        """Average the flat gradient and the batch statistics over the ranks of `group` (one NCCL all-reduce each): every
        rank then takes the same step and keeps the same BatchNorm moving statistics."""
        from .parallel import all_reduce_mean
        all_reduce_mean(self.grads, group)
        all_reduce_mean(self.batch_stats, group)

    def step(self, loss=None):
        """One optimizer step (and the BatchNorm moving-statistics update); loss: optional float32 [1] tensor that receives the
        regulariser's share l2 * sum ||W||^2."""
        if self.optimizer == 'sgd':
            lr_t = self.lr / (1.0 + self.decay * self.iterations)
            _lib.check(self.lib.ahk_train_sgd_step(self._h, _ptr(self.params), _ptr(self.grads), _ptr(self.m1), self.segments, len(SHAPES), lr_t,
                                                   self.momentum, self.clipnorm, self.l2, _ptr(self.batch_stats), BN_MOMENTUM, _ptr(loss), self._stream()))
        else:
            t = self.iterations + 1
            lr_t = self.lr * np.sqrt(1.0 - self.beta2 ** t) / (1.0 - self.beta1 ** t)
            _lib.check(self.lib.ahk_train_adam_step(self._h, _ptr(self.params), _ptr(self.grads), _ptr(self.m1), _ptr(self.m2), self.segments, len(SHAPES),
                                                    float(lr_t), self.beta1, self.beta2, self.eps, self.l2, _ptr(self.batch_stats), BN_MOMENTUM, _ptr(loss),
                                                    self._stream()))
        self.iterations += 1

    def push(self, policy):
        """The fitted network (moving statistics included) -> a PolicyNetwork's forward kernels, on the device."""
        _lib.check(self.lib.ahk_policy_set_weights_dev(policy._h, C.byref(self.w_struct), self._stream()))

    def fit_batch(self, frames3, backward_from, head=2, group=None, loss=None):
        """forward -> backward_from(logits) -> (dlogits, loss) -> backward -> all-reduce -> step.  Returns the logits view."""
        logits = self.forward(frames3, head)
        dlogits, loss_t = backward_from(logits)
        self.backward(dlogits)
        self.all_reduce(group)
        self.step(loss_t if loss is None else loss)
        return logits
