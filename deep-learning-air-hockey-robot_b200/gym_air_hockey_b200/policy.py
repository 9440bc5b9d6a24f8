"""The reference's 'rgb' CNN policy (ai/supervised/keras/models.py:20-65) as a forward pass on the frame ring.

`PolicyNetwork` owns an `ahk_policy` handle (include/ahk.h): six tcgen05 GEMM-tile kernels (csrc/ahk_policy.cu) that
read the rasteriser's uint8 ring directly — the float32 (9,128,128) observation the reference feeds to
`model.predict` (ai/reinforcement/rl_agents.py:78-98, ai/supervised/keras/test.py:42-43) is never materialised.
Weights are given in Keras layouts (conv kernels [kh][kw][cin][cout], dense [in][out], BatchNorm gamma/beta/moving
mean/moving variance), i.e. what `model.get_weights()` returns; a trained Keras `.h5` is read by `load_weights`
(h5lite.py: the package's own HDF5 reader).
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib

# name -> shape, in the order of ahk_policy_weights (include/ahk.h)
SHAPES = (
    ("conv1_w", (8, 8, 9, 5)), ("conv1_b", (5,)), ("bn1_gamma", (5,)), ("bn1_beta", (5,)), ("bn1_mean", (5,)), ("bn1_var", (5,)),
    ("conv2_w", (4, 4, 5, 32)), ("conv2_b", (32,)), ("bn2_gamma", (32,)), ("bn2_beta", (32,)), ("bn2_mean", (32,)), ("bn2_var", (32,)),
    ("conv3_w", (4, 4, 32, 64)), ("conv3_b", (64,)), ("bn3_gamma", (64,)), ("bn3_beta", (64,)), ("bn3_mean", (64,)), ("bn3_var", (64,)),
    ("dense1_w", (4096, 512)), ("dense1_b", (512,)), ("dense2_w", (512, 512)), ("dense2_b", (512,)),
    ("out_w", (512, 10)), ("out_b", (10,)),
)
FLOPS_PER_SAMPLE = 2 * 10227008
# MACs per sample by layer: conv1, conv2, conv3, dense1, dense2, out
LAYER_MACS = (4841280, 924160, 2097152, 2097152, 262144, 5120)


# the 'gray-diff' network (models.py:67-116), in the order of ahk_policy_gray_weights == model.get_weights()
GRAY_SHAPES = (
    ("conv1_w", (8, 8, 1, 16)), ("conv1_b", (16,)), ("bn1_gamma", (16,)), ("bn1_beta", (16,)), ("bn1_mean", (16,)), ("bn1_var", (16,)),
    ("conv2_w", (4, 4, 16, 24)), ("conv2_b", (24,)), ("bn2_gamma", (24,)), ("bn2_beta", (24,)), ("bn2_mean", (24,)), ("bn2_var", (24,)),
    ("conv3_w", (4, 4, 24, 32)), ("conv3_b", (32,)), ("bn3_gamma", (32,)), ("bn3_beta", (32,)), ("bn3_mean", (32,)), ("bn3_var", (32,)),
    ("conv4_w", (4, 4, 32, 48)), ("conv4_b", (48,)), ("bn4_gamma", (48,)), ("bn4_beta", (48,)), ("bn4_mean", (48,)), ("bn4_var", (48,)),
    ("dense1_w", (3072, 512)), ("dense1_b", (512,)), ("out_w", (512, 10)), ("out_b", (10,)),
)
GRAY_FLOPS_PER_SAMPLE = 2 * 6825984


def keras_default_init(seed=0, randomize_bn=False, shapes=None):
    """Random-init weights the way Keras builds models.py:20-65: Glorot-uniform kernels, zero biases, BatchNorm
    gamma 1 / beta 0 / mean 0 / variance 1.  randomize_bn=True draws non-trivial BatchNorm statistics instead (tests)."""
    rs = np.random.RandomState(seed)
    w = {}
    for name, shape in (shapes or SHAPES):
        if name.endswith("_w"):
            receptive = int(np.prod(shape[:-2])) if len(shape) == 4 else 1
            fan_in, fan_out = receptive * shape[-2], receptive * shape[-1]
            limit = np.sqrt(6.0 / (fan_in + fan_out))
            w[name] = rs.uniform(-limit, limit, size=shape).astype(np.float32)
        elif name.endswith(("_gamma", "_var")):
            w[name] = (rs.uniform(0.5, 1.5, size=shape) if randomize_bn else np.ones(shape)).astype(np.float32)
        elif name.endswith("_b") and randomize_bn:
            w[name] = rs.uniform(-0.1, 0.1, size=shape).astype(np.float32)
        elif randomize_bn:
            w[name] = rs.uniform(-0.2, 0.2, size=shape).astype(np.float32)
        else:
            w[name] = np.zeros(shape, dtype=np.float32)
    return w


def weights_from_keras(arrays, mode=None):
    """`model.get_weights()` of the reference's networks (models.py:20-65 'rgb', :67-116 'gray-diff') -> the dict
    PolicyNetwork takes.  Keras lists them layer by layer — Conv2D kernel, bias; BatchNormalization gamma, beta,
    moving_mean, moving_variance; ... Dense kernel, bias — which is exactly the order of SHAPES / GRAY_SHAPES (and of
    ahk_policy_weights).  Also accepts the DDQN variant whose last layer was swapped for a linear Dense(10)
    (rl_agents.py:160-172): same shapes.  mode=None picks the network by the number of arrays (24 / 28)."""
    arrays = list(arrays)
    if mode is None:
        mode = 'gray-diff' if len(arrays) == len(GRAY_SHAPES) else 'rgb'
    shapes = SHAPES if mode == 'rgb' else GRAY_SHAPES
    if len(arrays) != len(shapes):
        raise ValueError('expected %d weight arrays (models.py %s network), got %d' % (len(shapes), mode, len(arrays)))
    out = {}
    for (name, shape), a in zip(shapes, arrays):
        a = np.asarray(a, dtype=np.float32)
        if a.shape != shape:
            raise ValueError('%s: expected shape %s, got %s' % (name, shape, a.shape))
        out[name] = a
    return out


# layer names of models.py:20-116 (the `name=` arguments) and the weight names Keras 2.x gives their variables
_KERAS_SUFFIX = {'w': 'kernel:0', 'b': 'bias:0', 'gamma': 'gamma:0', 'beta': 'beta:0', 'mean': 'moving_mean:0',
                 'var': 'moving_variance:0'}


def _keras_layers(shapes, head='out'):
    """[(keras layer name, [(our name, keras weight name)])] in model order: conv1, batchnorm1, ..., dense1, ..., out."""
    layers = []
    for name, _ in shapes:
        ours, kind = name.rsplit('_', 1)
        layer = ours.replace('bn', 'batchnorm')
        if layer == 'out':
            layer = head
        if not layers or layers[-1][0] != layer:
            layers.append((layer, []))
        layers[-1][1].append((name, '%s/%s' % (layer, _KERAS_SUFFIX[kind])))
    return layers


def load_weights_h5(path):
    """A Keras 2.x HDF5 file of the reference — `model.save(...)` (keras/train.py:98-100, read back by `load_model`,
    rl_agents.py:153-154; the weights sit under `/model_weights`) or `model.save_weights(...)` (weights at the root):
    the group's `layer_names` attribute lists the layers in model order, each layer group's `weight_names` its
    variables in `get_weights()` order (Keras saving.py `save_weights_to_hdf5_group`).  Read with the package's own
    HDF5 reader (h5lite.py)."""
    from . import h5lite
    with h5lite.File(path, 'r') as f:
        g = f['model_weights'] if 'model_weights' in f else f
        if 'layer_names' not in g.attrs:
            raise ValueError('%s: no Keras layer_names attribute' % path)
        arrays = []
        for layer in np.atleast_1d(g.attrs['layer_names']):
            lg = g[layer.decode('utf-8')]
            for wname in np.atleast_1d(lg.attrs.get('weight_names', np.zeros(0, 'S1'))):
                arrays.append(np.asarray(lg[wname.decode('utf-8')], dtype=np.float32))
    return weights_from_keras(arrays)


def save_weights_h5(path, weights, mode='rgb', head='out'):
    """The layout `model.save_weights(path)` writes for models.py's networks, so that the reference side can
    `model.load_weights(path)` what was trained here (head='output' for the DDQN model, rl_agents.py:168)."""
    from . import h5lite
    shapes = SHAPES if mode == 'rgb' else GRAY_SHAPES
    layers = _keras_layers(shapes, head)
    with h5lite.File(path, 'w') as f:
        f.attrs['layer_names'] = np.array([l.encode() for l, _ in layers])
        f.attrs['backend'] = b'tensorflow'
        f.attrs['keras_version'] = b'2.1.5'
        for layer, names in layers:
            lg = f.create_group(layer)
            lg.attrs['weight_names'] = np.array([k.encode() for _, k in names])
            for ours, k in names:
                lg.create_dataset(k, data=np.asarray(weights[ours], dtype=np.float32))


def load_weights(path):
    """Weights saved from the reference: a Keras `.h5` (`load_weights_h5`), or an .npz written with
    np.savez(path, *model.get_weights()) (arr_0 .. arr_23) or with the SHAPES names as keys."""
    from . import h5lite
    if h5lite.is_hdf5(path):
        return load_weights_h5(path)
    with np.load(path) as f:
        for shapes in (SHAPES, GRAY_SHAPES):
            if all(name in f.files for name, _ in shapes) and len(f.files) == len(shapes):
                return weights_from_keras([f[name] for name, _ in shapes])
        return weights_from_keras([f['arr_%d' % i] for i in range(len(f.files))])


def save_weights_npz(path, weights, mode='rgb'):
    shapes = SHAPES if mode == 'rgb' else GRAY_SHAPES
    np.savez(path, **{name: np.asarray(weights[name], dtype=np.float32) for name, _ in shapes})


def softmax(logits):
    """The supervised model's output layer (models.py:62-65) on top of the linear logits the engine returns."""
    return torch.softmax(logits, dim=1)


class PolicyNetwork(object):
    """forward(env) -> (logits float32[N,10], action uint8[N]) as CUDA tensors, on the current stream.

    logits are pre-softmax (the DDQN head of rl_agents.py:160-172 is linear; softmax(logits) is the supervised
    model's output, models.py:62-65), action = argmax — what keras/test.py:42-43 feeds back into env.step."""

    def __init__(self, weights=None, max_batch=4096, device=0, mode='rgb'):
        """mode: 'rgb' (models.py:20-65) or 'gray-diff' (models.py:67-116) — ai/config.py:1-3."""
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError('PolicyNetwork needs a CUDA device; there is no CPU fallback')
        if mode not in ('rgb', 'gray-diff'):
            raise ValueError('Processing mode is not supported')
        self.mode = mode
        shapes = SHAPES if mode == 'rgb' else GRAY_SHAPES
        self.device = torch.device('cuda', device)
        self.max_batch = int(max_batch)
        weights = keras_default_init(shapes=shapes) if weights is None else weights
        self.weights = {}
        st = _lib.PolicyWeights() if mode == 'rgb' else _lib.GrayPolicyWeights()
        for name, shape in shapes:
            a = np.ascontiguousarray(weights[name], dtype=np.float32)
            if a.shape != shape:
                raise ValueError('%s: expected shape %s, got %s' % (name, shape, a.shape))
            self.weights[name] = a  # keeps the host memory alive during ahk_policy_create
            setattr(st, name, a.ctypes.data_as(C.c_void_p))
        h = C.c_void_p()
        create = self.lib.ahk_policy_create if mode == 'rgb' else self.lib.ahk_policy_gray_create
        _lib.check(create(self.device.index, self.max_batch, C.byref(st), C.byref(h)))
        self._h = h
        self.launches = 0
        # kernels per inference forward: conv2 -> conv3 are ONE kernel (conv23_kernel) unless an A/B switch selects other conv kernels
        ab = any(os.environ.get(k) for k in ('AHK_CONVTM', 'AHK_CONV2', 'AHK_CONV23', 'AHK_GEMM')) or os.environ.get('AHK_FUSE23') == '0'
        self._forward_launches = (6 if ab else 5) if mode == 'rgb' else 7

    def forward_ring(self, frames, head, logits=None, action=None):
        """frames: uint8 CUDA tensor [3, n, 3, 128, 128] (the env's ring), head: newest slot."""
        n = frames.shape[1]
        if frames.dtype != torch.uint8 or tuple(frames.shape) != (3, n, 3, 128, 128) or not frames.is_contiguous():
            raise ValueError('frames must be a contiguous uint8 [3, n, 3, 128, 128] ring')
        if logits is None:
            logits = torch.empty((n, 10), dtype=torch.float32, device=self.device)
        if action is None:
            action = torch.empty((n,), dtype=torch.uint8, device=self.device)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self.lib.ahk_policy_forward(self._h, C.c_void_p(frames.data_ptr()), n, int(head),
                                               C.c_void_p(logits.data_ptr()), C.c_void_p(action.data_ptr()), stream))
        self.launches += self._forward_launches
        return logits, action

    def forward_profiled(self, env, logits, action):
        """forward() with CUDA events around each layer kernel -> [conv1, conv2, conv3, dense1, dense2, out] in ms."""
        ms = (C.c_float * 6)()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        st = env.stack_view()
        _lib.check(self.lib.ahk_policy_forward_stack_profiled(self._h, C.byref(st), C.c_void_p(logits.data_ptr()),
                                                              C.c_void_p(action.data_ptr()), stream, ms))
        self.launches += 6 if self.mode == 'rgb' else 7  # one kernel per layer (conv2 and conv3 are not fused here)
        return list(ms)

    def forward(self, env, logits=None, action=None):
        """The network on the env's current observation — the 3-slot ring or a deep ring (ahk_stack_view)."""
        n = env.n_envs
        if logits is None:
            logits = torch.empty((n, 10), dtype=torch.float32, device=self.device)
        if action is None:
            action = torch.empty((n,), dtype=torch.uint8, device=self.device)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        st = env.stack_view()
        _lib.check(self.lib.ahk_policy_forward_stack(self._h, C.byref(st), C.c_void_p(logits.data_ptr()),
                                                     C.c_void_p(action.data_ptr()), stream))
        self.launches += self._forward_launches
        return logits, action

    def close(self):
        if getattr(self, '_h', None):
            self.lib.ahk_policy_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
