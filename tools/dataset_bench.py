"""Dataset writer throughput: python tools/dataset_bench.py [--envs 2048] [--states 40960] [--dir /tmp/ahk_ds]"""
import argparse, json, os, shutil, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))
from gym_air_hockey_b200 import BatchedAirHockeyEnv, dataset  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=2048)
ap.add_argument("--states", type=int, default=40960)
ap.add_argument("--dir", default="/tmp/ahk_ds")
args = ap.parse_args()
res = {}
for fmt, threads, container in (("u8", 1, "h5"), ("u8", 4, "h5"), ("f32", 4, "h5"), ("u8", 4, "npz")):
    shutil.rmtree(args.dir, ignore_errors=True)
    env = BatchedAirHockeyEnv(args.envs, seed=0, max_episode_steps=200)
    env.reset()
    n = args.states if fmt == "u8" else args.states // 4
    # HDF5: ONE file like the reference's (data_utils.py:3-7), filled in place; .npz: shards (round 1's container)
    out = dataset.generate(env, n, args.dir, states=fmt, shard_states=0 if container == "h5" else 8192, label_random_as_undefined=True,
                           writer_threads=threads, container=container)
    size = sum(os.path.getsize(f) for f in out["files"])
    res["%s_%s_x%d" % (fmt, container, threads)] = {"samples": out["samples"], "seconds": out["seconds"], "samples_per_s": out["samples_per_s"],
                "host_GBs": out["bytes_to_host"] / out["seconds"] / 1e9, "files": len(out["files"]), "file_bytes": size}
    env.close()
shutil.rmtree(args.dir, ignore_errors=True)
print(json.dumps({"envs": args.envs, "results": res}))
