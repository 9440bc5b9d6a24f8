"""The package's own HDF5 reader / writer (gym_air_hockey_b200/h5lite.py): the container of the reference's dataset
(ai/supervised/utils/data_utils.py:3-13) and of its Keras model files (ai/reinforcement/rl_agents.py:153-154).
The reader is pinned against a file written by libhdf5 itself (a MATLAB 7.3 file from scipy's test data)."""
import os
import struct
import threading

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def test_reads_a_libhdf5_file():
    """tests/golden/libhdf5_matlab73_testdouble.mat = scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat: 512-byte user
    block, superblock 0, symbol-table root group, version-1 object header, contiguous float64 dataset (layout message
    version 2), fixed-length string attribute.  scipy documents its contents as 0 .. 2 pi in steps of pi / 4."""
    from gym_air_hockey_b200 import h5lite
    with h5lite.File(os.path.join(GOLDEN, 'libhdf5_matlab73_testdouble.mat')) as f:
        assert f.keys() == ['testdouble']
        d = f['testdouble']
        assert d.shape == (9, 1) and d.dtype == np.dtype('<f8')
        assert np.array_equal(d[:].ravel(), np.arange(9) * (np.pi / 4))
        assert d.attrs['MATLAB_class'] == b'double'
        assert 'nothing' not in f
        with pytest.raises(KeyError):
            f['nothing']


def test_dataset_container_round_trip(tmp_path):
    """What data_utils.save_data writes and load_data reads back: three root datasets, float32 / int8."""
    from gym_air_hockey_b200 import dataset, h5lite
    rs = np.random.RandomState(1)
    states = rs.uniform(-1, 1, size=(7, 9, 128, 128)).astype(np.float32)
    robot, human = rs.randint(0, 9, size=7).astype(np.int8), rs.randint(0, 9, size=7).astype(np.int8)
    path = str(tmp_path / dataset.data_file_name('rgb', 7, when=1500000000))
    assert path.endswith('_rgb_7.h5')
    dataset.save_data(path, states, robot, human)
    assert h5lite.is_hdf5(path)
    s, r, h = dataset.load_data(path)
    assert s.dtype == np.float32 and r.dtype == np.int8 and h.dtype == np.int8
    assert np.array_equal(s, states) and np.array_equal(r, robot) and np.array_equal(h, human)
    with h5lite.File(path) as f:
        assert f.keys() == ['human_actions', 'robot_actions', 'states'] and len(f['states']) == 7
        assert np.array_equal(f['states'][2:4, 3], states[2:4, 3])


def test_file_structure_is_what_libhdf5_expects(tmp_path):
    """Superblock 0 fields, end-of-file address, 8-byte aligned version-1 object headers whose messages are the ones the
    libhdf5-written golden file carries for a dataset (dataspace, datatype, fill value, layout)."""
    from gym_air_hockey_b200 import h5lite
    path = str(tmp_path / 'a.h5')
    with h5lite.File(path, 'w') as f:
        f.create_dataset('testdouble', data=(np.arange(9) * (np.pi / 4)).reshape(9, 1)).attrs['MATLAB_class'] = b'double'
        f.create_group('empty')
    raw = open(path, 'rb').read()
    assert raw[:8] == h5lite.SIGNATURE and raw[8:16] == bytes([0, 0, 0, 0, 0, 8, 8, 0])
    base, free, eof, driver = struct.unpack_from('<QQQQ', raw, 24)
    assert (base, free, eof, driver) == (0, h5lite.UNDEF, len(raw), h5lite.UNDEF)
    name_off, root, cache, _r, btree, heap = struct.unpack_from('<QQIIQQ', raw, 56)
    assert cache == 1 and root % 8 == 0 and raw[btree:btree + 4] == b'TREE' and raw[heap:heap + 4] == b'HEAP'
    with h5lite.File(path) as mine, h5lite.File(os.path.join(GOLDEN, 'libhdf5_matlab73_testdouble.mat')) as ref:
        a, b = mine['testdouble'], ref['testdouble']
        assert a._addr % 8 == 0
        assert (a.shape, a.dtype, a.attrs) == (b.shape, b.dtype, b.attrs) and np.array_equal(a[:], b[:])
        types = lambda d: [t for t, _ in d._msgs if t not in (0, 0x12)]  # but for NIL padding and the modification time
        assert sorted(types(a)) == sorted(types(b))
        body = lambda d, t: [m for k, m in d._msgs if k == t][0]
        assert body(a, 3) == body(b, 3)                    # the float64 datatype message, byte for byte
        assert body(a, 1)[:24] == body(b, 1)[:24]          # the dataspace message
        # the attribute message: identical but for the string padding flag (h5py stores numpy 'S' as null-padded = 1,
        # MATLAB as null-terminated = 0)
        mine_attr, ref_attr = body(a, 0xC), body(b, 0xC)
        assert mine_attr[:25] == ref_attr[:25] and (mine_attr[25], ref_attr[25]) == (1, 0) and mine_attr[26:48] == ref_attr[26:48]
        assert mine['empty'].keys() == []


def test_keras_weight_files(tmp_path):
    """N3: `model.save_weights` layout (layer_names / weight_names attributes, Keras 2.x saving.py) and the `model.save`
    layout (the same under /model_weights) -> PolicyNetwork's weight dict, both networks of models.py."""
    from gym_air_hockey_b200 import h5lite, policy
    for mode, shapes in (('rgb', policy.SHAPES), ('gray-diff', policy.GRAY_SHAPES)):
        w = policy.keras_default_init(seed=4, randomize_bn=True, shapes=shapes)
        path = str(tmp_path / ('w_%s.h5' % mode))
        policy.save_weights_h5(path, w, mode=mode, head='output' if mode == 'rgb' else 'out')
        with h5lite.File(path) as f:
            names = [n.decode() for n in f.attrs['layer_names']]
            assert names[:3] == ['conv1', 'batchnorm1', 'conv2'] and names[-1] == ('output' if mode == 'rgb' else 'out')
            assert [n.decode() for n in f['batchnorm1'].attrs['weight_names']] == [
                'batchnorm1/gamma:0', 'batchnorm1/beta:0', 'batchnorm1/moving_mean:0', 'batchnorm1/moving_variance:0']
            assert f['conv1/conv1/kernel:0'].shape == shapes[0][1]
        back = policy.load_weights(path)
        assert list(back) == [n for n, _ in shapes]
        assert all(np.array_equal(back[k], w[k]) for k in w)
    # model.save: weights under /model_weights, model_config / training_config as root attributes
    w = policy.keras_default_init(seed=5)
    path = str(tmp_path / 'robot_model.h5')
    with h5lite.File(path, 'w') as f:
        f.attrs['keras_version'] = b'2.1.5'
        f.attrs['model_config'] = b'{"class_name": "Sequential"}'
        g = f.create_group('model_weights')
        layers = policy._keras_layers(policy.SHAPES)
        g.attrs['layer_names'] = np.array([b'conv1', b'batchnorm1', b'conv2', b'batchnorm2', b'conv3', b'batchnorm3',
                                           b'flatten', b'dense1', b'dense2', b'out'])
        g.create_group('flatten').attrs['weight_names'] = np.zeros((0,), dtype='S1')
        for layer, names in layers:
            lg = g.create_group(layer)
            lg.attrs['weight_names'] = np.array([k.encode() for _, k in names])
            for ours, k in names:
                lg.create_dataset(k, data=w[ours])
        f.create_group('optimizer_weights')
    back = policy.load_weights(path)
    assert all(np.array_equal(back[k], w[k]) for k in w)


def test_streaming_writer_fills_the_file_in_place(tmp_path):
    """dataset._Writer (the consumer of the pinned D2H buffers) without a GPU: blocks of uneven size, one HDF5 file
    declared at its final shape, and shards."""
    from gym_air_hockey_b200 import dataset
    rs = np.random.RandomState(2)
    total = 37
    states = rs.randint(0, 256, size=(total, 9, 16, 16)).astype(np.uint8)
    robot, human = rs.randint(0, 9, size=total).astype(np.int8), rs.randint(0, 9, size=total).astype(np.int8)
    for shard, container in ((0, 'h5'), (10, 'h5'), (10, 'npz')):
        out = tmp_path / ('%s_%d' % (container, shard))
        out.mkdir()
        w = dataset._Writer(str(out), 'rgb', shard, 1500000000, container, total, savers=3)
        w.start()
        at = 0
        for k in (8, 8, 8, 8, 5):
            ev = threading.Event()
            w.q.put((states[at:at + k], robot[at:at + k], human[at:at + k], ev))
            assert ev.wait(30)
            at += k
        w.q.put(None)
        w.join()
        assert w.error is None and len(w.files) == (1 if shard == 0 else 4)
        assert all(f.endswith('.' + container) for f in w.files)
        s, r, h = dataset.load_data(os.path.join(str(out), '*'))
        assert np.array_equal(s, states) and np.array_equal(r, robot) and np.array_equal(h, human)
