#!/usr/bin/env python
"""BASELINE.json configs[4] / SURVEY.md §8(d): collision-stress sweep — adversarial initial states (pucks at top speed
aimed at corner arcs, grazing walls, at goal posts, on mallet rims, on top of a mallet), a long horizon (10 000 steps),
the CUDA engine against the CPU oracle at EVERY step: first-divergence step and max |delta| of the float32 state
(expected: none / 0 — the parity bar is bit-exact), all outputs (reward, terminal, labels, flags), and the whole frame
ring at the end of every chunk.  Envs shard by global id over the ranks of a torchrun launch; the per-env trajectory
checksums must be identical at 1, 2, 4 and 8 GPUs.

    python tools/stress_sweep.py --envs 384 --horizon 10000 --out profiles/stress_sweep_N1.json
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \\
        tools/stress_sweep.py --envs 384 --horizon 10000 --out profiles/stress_sweep_N8.json

The oracle is the checker here (tools/ and tests/ only); the engine runs through the C ABI as in the tests.
"""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"), os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

FNV = np.uint64(0x100000001B3)


def fold(acc, words):
    """Running FNV-style checksum per env over uint32 words [k, n] (vectorised over envs, order-dependent)."""
    with np.errstate(over="ignore"):
        for w in words:
            acc = (acc ^ w.astype(np.uint64)) * FNV
    return acc


def sweep(total_envs=384, horizon=10000, chunk=500, seed=8, max_episode_steps=1000, rank=0, world=1, device=0, threads=None,
          verbose=False):
    """Runs this rank's shard.  -> dict with the divergence report and per-env checksums (global id order within the shard)."""
    import torch
    import orc
    import stress
    from gym_air_hockey_b200 import BatchedAirHockeyEnv
    from gym_air_hockey_b200.assets import make_table_image
    from gym_air_hockey_b200.parallel import shard_range

    orc.build()
    table = make_table_image()
    lo, hi = shard_range(total_envs, rank, world)
    n = hi - lo
    env = BatchedAirHockeyEnv(n, device=device, seed=seed, env_id_base=lo, auto_reset=True, max_episode_steps=max_episode_steps, table=table)
    env.reset(episode=np.zeros(n, dtype=np.uint8), use_object=np.full(n, 7, dtype=np.uint8))
    s0 = stress.stress_states(lo, hi)
    env.set_state(s0)
    oracles = [orc.Env(seed=seed, env_id=lo + i, table=table) for i in range(n)]
    for i, o in enumerate(oracles):
        o.reset(0, 7)
        _, m = o.get_state()
        o.set_state(s0[i], m)
    pool = orc.Pool(oracles, threads=threads or max(1, (os.cpu_count() or 1) // world), want_obs=False)

    dev = env.device
    state_log = torch.empty((chunk, 16, n), dtype=torch.int32, device=dev)
    out_log = torch.empty((chunk, 4, n), dtype=torch.int32, device=dev)
    acc = np.zeros(n, dtype=np.uint64)
    first_div, max_delta, bad_outputs, frames_checked, frame_mismatch = None, 0.0, 0, 0, 0
    goals = resets = 0
    t0 = time.time()
    done_steps = 0
    while done_steps < horizon:
        k = min(chunk, horizon - done_steps)
        for t in range(k):
            _, reward, done, info = env.step()
            state_log[t].copy_(env.state)
            out_log[t, 0].copy_((reward * 2 + 2).to(torch.int32))
            out_log[t, 1].copy_(env.done)
            out_log[t, 2].copy_(env.robot_label.to(torch.int32) | (env.human_label.to(torch.int32) << 4))
            out_log[t, 3].copy_(env.info_bits)
        got_state = state_log[:k].cpu().numpy().view(np.uint32)
        got_out = out_log[:k].cpu().numpy().view(np.uint32)
        want_state, want_out = pool.trace(k, max_episode_steps, auto_reset=True, draw_last=3)
        u = orc.unpack_out(want_out)
        want_packed = np.stack([(want_out & 7), u["terminal"], u["robot_label"] | u["human_label"] << 4, u["flags"] | u["scored"] << 5], axis=1).astype(np.uint32)
        if not np.array_equal(got_state, want_state):
            bad = np.nonzero((got_state != want_state).any(axis=(1, 2)))[0]
            if first_div is None:
                first_div = done_steps + int(bad[0])
            d = np.abs(got_state[:, :12].view(np.float32).astype(np.float64) - want_state[:, :12].view(np.float32).astype(np.float64))
            max_delta = max(max_delta, float(np.nanmax(d)))
        bad_outputs += int((got_out != want_packed).sum())
        goals += int(u["terminal"].sum())
        resets = int(want_state[-1, 13].astype(np.int64).sum())
        for t in range(k):
            acc = fold(acc, got_state[t])
            acc = fold(acc, got_out[t])
        ring = env.frames.cpu().numpy()
        h = env.head
        for i, o in enumerate(oracles):
            want = o.ring()
            for age in range(3):
                frames_checked += 1
                frame_mismatch += int(not np.array_equal(ring[(h + 3 - age) % 3, i], want[age]))
        done_steps += k
        if verbose:
            print("rank %d: %d / %d steps, divergence %s, %.1f s" % (rank, done_steps, horizon, first_div, time.time() - t0), file=sys.stderr)
    pool.close()
    speeds = np.hypot(want_state[:, 2].view(np.float32), want_state[:, 3].view(np.float32))
    return {"lo": lo, "hi": hi, "first_divergence_step": first_div, "max_abs_delta": max_delta, "output_mismatches": bad_outputs,
            "frames_checked": frames_checked, "frame_mismatches": frame_mismatch, "goals": goals, "env_resets": resets,
            "max_puck_speed_last_chunk": float(speeds.max()), "checksums": acc, "seconds": time.time() - t0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=384)
    ap.add_argument("--horizon", type=int, default=10000)
    ap.add_argument("--chunk", type=int, default=500)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = sweep(args.envs, args.horizon, args.chunk, rank=rank, world=world, device=local, verbose=True)
    sums = torch.from_numpy(r["checksums"].view(np.int64)).cuda()
    report = torch.tensor([-1 if r["first_divergence_step"] is None else r["first_divergence_step"], r["output_mismatches"],
                           r["frames_checked"], r["frame_mismatches"], r["goals"], r["env_resets"]], dtype=torch.int64, device="cuda")
    delta = torch.tensor([r["max_abs_delta"], r["max_puck_speed_last_chunk"], r["seconds"]], dtype=torch.float64, device="cuda")
    if world > 1:
        parts = [torch.empty_like(sums) for _ in range(world)]  # shards are equal-sized when world divides --envs
        dist.all_gather(parts, sums)
        sums = torch.cat(parts)
        reports = [torch.empty_like(report) for _ in range(world)]
        dist.all_gather(reports, report)
        dist.all_reduce(delta, op=dist.ReduceOp.MAX)
    else:
        reports = [report]
    if rank == 0:
        reports = torch.stack(reports).cpu().numpy()
        div = reports[:, 0][reports[:, 0] >= 0]
        per_env = sums.cpu().numpy().view(np.uint64)
        line = {
            "what": "configs[4] collision-stress sweep: CUDA engine vs CPU oracle at every step", "n_gpus": world, "envs": args.envs,
            "horizon_steps": args.horizon, "env_steps_compared": args.envs * args.horizon,
            "first_divergence_step": int(div.min()) if len(div) else None, "max_abs_delta_vs_oracle": float(delta[0]),
            "output_mismatches": int(reports[:, 1].sum()), "frames_compared": int(reports[:, 2].sum()),
            "frame_mismatches": int(reports[:, 3].sum()), "goals": int(reports[:, 4].sum()), "env_resets": int(reports[:, 5].sum()),
            "max_puck_speed_last_chunk": float(delta[1]), "seconds": float(delta[2]),
            "per_env_checksum_sha256": hashlib.sha256(per_env.tobytes()).hexdigest(),
            "per_env_checksums_first8": ["%016x" % int(v) for v in per_env[:8]],
            "initial_states": "oracle/stress.py (function of the global env id)", "auto_reset": True, "max_episode_steps": 1000,
        }
        text = json.dumps(line)
        print(text)
        if args.out:
            with open(args.out, "w") as f:
                f.write(text + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
