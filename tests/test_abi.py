"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/ahk.h declares, refuses to
run without a GPU (no CPU fallback), and the host-side mirror of the reference interface behaves like the reference."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ahk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ahk_[a-z_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from gym_air_hockey_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), name
    assert set(names) == set(_lib.SYMBOLS), set(names) ^ set(_lib.SYMBOLS)   # the binding covers the whole header
    assert lib.ahk_version().decode().startswith("ahk")


def test_struct_layouts_match_header():
    from gym_air_hockey_b200 import _lib
    assert C.sizeof(_lib.Config) == 48            # int32 x2, uint64, uint32, int32 x3, int32 x4 (+ padding to 8)
    assert _lib.Config.ring_depth.offset == 28
    assert C.sizeof(_lib.Buffers) == 10 * 8
    assert C.sizeof(_lib.Stack) == 3 * 8 + 3 * 4 + 4
    assert C.sizeof(_lib.StepLog) == 4 * 8
    assert C.sizeof(_lib.PolicyWeights) == 24 * 8
    assert C.sizeof(_lib.GrayPolicyWeights) == 28 * 8


def test_no_cpu_fallback():
    """Without a CUDA device the engine must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this checks the CPU-only behaviour")
    from gym_air_hockey_b200 import _lib, BatchedAirHockeyEnv
    lib = _lib.load()
    cfg = _lib.Config(n_envs=4, device=0, seed=0, env_id_base=0, auto_reset=1, max_episode_steps=200)
    h = C.c_void_p()
    rc = lib.ahk_create(C.byref(cfg), C.byref(h))
    assert rc == -4 and b"no CPU fallback" in lib.ahk_last_error()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BatchedAirHockeyEnv(4)


def test_argument_errors():
    from gym_air_hockey_b200 import _lib
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.ahk_create(None, C.byref(h)) == -1
    cfg = _lib.Config(n_envs=0, device=0)
    assert lib.ahk_create(C.byref(cfg), C.byref(h)) == -1 and b"n_envs" in lib.ahk_last_error()
    assert lib.ahk_step(None, None, None, None) == -1
    with pytest.raises(ValueError):
        _lib.check(-1)
    with pytest.raises(RuntimeError):
        _lib.check(-2)


def test_data_processor_mirror():
    """Same table, label formula and error text as processor.py:7-18,66,74-76,84-85 (golden from the reference)."""
    from conftest import golden
    from gym_air_hockey_b200 import DataProcessor
    g = golden("processor.npz")
    p = DataProcessor()
    assert p.dim == 128 and p.mode == 'rgb' and p.actions.dtype == np.int8
    assert np.array_equal(p.actions, g["actions"])
    assert [p.action_to_label(p.process_action(l)) for l in range(10)] == g["labels"].tolist()
    assert p.process_action(None) is None
    assert p.process_reward(3.0) == 1.0 and p.process_reward(-7) == -1.0
    with pytest.raises(ValueError, match='Processing mode is not supported'):
        DataProcessor(mode='hsv').check_mode()
    batch = [np.zeros((1, 9, 128, 128), np.float32)] * 3
    assert p.process_state_batch(batch).shape == (3, 9, 128, 128)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "deep-learning-air-hockey-robot_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import orc" not in text and "ahk_oracle" not in text and "libahk_oracle" not in text, f


def test_bench_reads_ncu_traffic_from_the_newest_profile():
    """bench.py's roofline.traffic comes from the newest committed ncu summary, not from a constant."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    traffic, src = bench.ncu_traffic("step_kernel_*_ncu.txt")
    assert src is not None and src.startswith("step_kernel_") and 100e6 < traffic < 250e6
    traffic, src = bench.ncu_traffic("step_many_T*_ncu.txt")
    assert src.startswith("step_many_T16_") and 16 * 150e6 < traffic < 16 * 250e6
    assert bench.ncu_traffic("no_such_kernel_*.txt") == (None, None)
    assert bench.cpu_reference_wrapper()["B3_reference_wrapper_over_oracle_simulator"]["rgb"]["frames"] == 10000
