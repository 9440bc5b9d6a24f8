#!/usr/bin/env python
"""Time the REFERENCE'S OWN CPU code for the path (BASELINE.md §4 lines B2 and B3, BASELINE.json configs[0]) in the build
container — /root/reference does not exist on the GPU box, so the result is committed as profiles/cpu_reference_wrapper.json
and quoted by bench.py next to its cpu_baseline.

  B2  gym-air-hockey/gym_air_hockey/processor.py:31-66  DataProcessor.process_observation on oracle-drawn native frames
      (450x700 RGB), 'rgb' and 'gray-diff', Pillow filter NEAREST (the 2018 default the reference relied on) and BICUBIC
      (what today's Pillow picks for resize() with no filter), one process and one process per core.
  B3  the unmodified AirHockeyEnv + DataProcessor (air_hockey_env.py:8-155) over oracle/air_hockey (the stand-in for the
      missing simulator, backed by oracle/ahk_oracle.c) driven by the ai/supervised/generate_data.py:44-58 loop:
      10 000 labelled frames, one process; and one such process per core.

    python tools/cpu_reference.py            # writes profiles/cpu_reference_wrapper.json
"""
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "gym_shim"), "/root/reference/gym-air-hockey",
                os.path.join(ROOT, "deep-learning-air-hockey-robot_b200")]


def _pin_filter(which):
    from PIL import Image
    base = getattr(Image.Image, "_ahk_resize", None) or Image.Image.resize
    Image.Image._ahk_resize = base
    flt = {"nearest": Image.NEAREST, "bicubic": Image.BICUBIC}[which]

    def resize(self, size, resample=None, *a, **k):
        return base(self, size, flt if resample is None else resample, *a, **k)

    Image.Image.resize = resize


def _frames(k=64):
    import orc
    from gym_air_hockey_b200.assets import make_table_image
    table = make_table_image()
    sim = orc.Sim(seed=0, env_id=0, table=table)
    sim.reset()
    out = []
    for _ in range(k):
        sim.step(None, None, orc.USE_ALL)
        out.append(sim.draw(orc.USE_ALL))
    return out


def processor_worker(args):
    mode, flt, n = args
    _pin_filter(flt)
    from gym_air_hockey import DataProcessor  # THE REFERENCE (processor.py)
    proc = DataProcessor(mode=mode)
    frames = _frames()
    for f in frames[:8]:
        proc.process_observation(f)
    t0 = time.perf_counter()
    for i in range(n):
        proc.process_observation(frames[i % len(frames)])
    return n / (time.perf_counter() - t0)


def wrapper_worker(args):
    """generate_data.py:44-58 with the reference wrapper, unmodified."""
    mode, n_states, seed = args
    _pin_filter("nearest")
    import contextlib
    import io
    import air_hockey
    from gym_air_hockey_b200.assets import make_table_image
    air_hockey.SEED, air_hockey.ENV_ID, air_hockey.TABLE = 20181104, seed, make_table_image()
    import gym  # shim
    import gym_air_hockey  # noqa: F401  THE REFERENCE (registers AirHockey-v0)
    np.random.seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):  # the wrapper prints the episode type at every reset
        env = gym.make('AirHockey-v0')
        env.update(mode=mode)
        states, robot, human = [], [], []
        reset = True
        t0 = time.perf_counter()
        for i in range(n_states):
            if reset:
                reset = False
                env.reset()
            state, _, terminal, info = env.step()
            states.append(state)
            robot.append(info['robot_action'])
            human.append(info['human_action'])
            if terminal or i % 200 == 0:
                reset = True
            if len(states) >= 256:  # the script keeps all frames in a deque; memory is not what is being measured
                states.clear()
        dt = time.perf_counter() - t0
    return n_states / dt


def main():
    cores = os.cpu_count() or 1
    out = {"where": "build container (no GPU); /root/reference imported unmodified", "cores": cores,
           "python": sys.version.split()[0], "unit_processor": "frames/s", "unit_wrapper": "labelled env-steps/s"}
    import PIL
    out["pillow"] = PIL.__version__
    ctx = mp.get_context("spawn")
    proc = {}
    for mode in ("rgb", "gray-diff"):
        for flt in ("nearest", "bicubic"):
            one = processor_worker((mode, flt, 1500))
            with ctx.Pool(cores) as pool:
                t0 = time.perf_counter()
                rates = pool.map(processor_worker, [(mode, flt, 1500)] * cores)
            proc["%s/%s" % (mode, flt)] = {"one_process": one, "all_cores": float(sum(rates)), "processes": cores}
            print(mode, flt, proc["%s/%s" % (mode, flt)], file=sys.stderr)
    out["B2_processor_process_observation"] = proc
    wrap = {}
    for mode, n_states in (("rgb", 10000), ("gray-diff", 3000)):
        one = wrapper_worker((mode, n_states, 7))
        with ctx.Pool(cores) as pool:
            rates = pool.map(wrapper_worker, [(mode, max(n_states // 4, 1000), 100 + k) for k in range(cores)])
        wrap[mode] = {"frames": n_states, "one_process": one, "all_cores": float(sum(rates)), "processes": cores}
        print(mode, wrap[mode], file=sys.stderr)
    out["B3_reference_wrapper_over_oracle_simulator"] = wrap
    out["note"] = ("B3 'rgb' one_process is BASELINE.json configs[0] (single env, 10 000 labelled frames, the reference's "
                   "generate_data.py loop); the simulator under the wrapper is oracle/ahk_oracle.c through ctypes (the Pygame "
                   "original is a missing submodule), so the Python wrapper + DataProcessor cost dominates")
    path = os.path.join(ROOT, "profiles", "cpu_reference_wrapper.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
