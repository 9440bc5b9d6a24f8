"""RL-step measurement: the batched train.py loop, with its parts timed by CUDA events.
    python tools/rl_bench.py [--envs 8192] [--steps 200] [--batch 32] [--interval 25] > profiles/rl_<round>.json
Under torchrun (one rank per GPU, NCCL) every rank runs its own shard of the envs, gradients and BatchNorm statistics are
all-reduced at every fit, rank 0 prints the line (times = max over ranks) and the ranks' parameters are compared at the end.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"))

import torch  # noqa: E402

from gym_air_hockey_b200 import BatchedAirHockeyEnv  # noqa: E402
from gym_air_hockey_b200.rl import DDQNAgent, PGAgent, run_training  # noqa: E402


def timed(fn, reps=1):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


def pg_main(args):
    """PGAgent (rl_agents.py:199-286) batched: `episode_length` rollout steps, then one pass of fits over every recorded
    sample (returns per episode by ahk_rl_pg_returns, targets and output gradient by ahk_rl_pg_targets)."""
    if os.environ.get("RL_CUDNN_BENCHMARK"):
        torch.backends.cudnn.benchmark = True
    if os.environ.get("RL_TF32"):
        torch.backends.cuda.matmul.allow_tf32 = True
    env = BatchedAirHockeyEnv(args.envs, seed=0, max_episode_steps=256)
    env.reset()
    agent = PGAgent(env, episode_length=args.episode_length, fit_batch=args.fit_batch, discount_rate=0.98, eps_min=0.05, eps_max=0.99,
                    eps_decay_steps=100000)
    run_training(agent, args.episode_length)   # warm-up: one rollout + one training pass
    steps = 2 * args.episode_length
    ms, losses = timed(lambda: run_training(agent, steps))
    print(json.dumps({"agent": "pg", "envs": args.envs, "episode_length": args.episode_length, "fit_batch": args.fit_batch,
                      "steps": steps, "ms": ms, "env_steps_per_s": args.envs * steps / ms * 1e3,
                      "samples_fitted_per_s": args.envs * steps / ms * 1e3, "losses": losses}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--batch", type=int, default=32)       # train.py:49
    ap.add_argument("--interval", type=int, default=25)    # train.py:47
    ap.add_argument("--buffer-steps", type=int, default=32)
    ap.add_argument("--agent", choices=["ddqn", "pg"], default="ddqn")
    ap.add_argument("--copy-store", action="store_true")       # 3-slot ring + a separate replay store filled by one frame copy per step (round 1)
    ap.add_argument("--episode-length", type=int, default=64)   # pg: steps between fits (train.py:45 uses 256)
    ap.add_argument("--fit-batch", type=int, default=4096)      # pg: samples per optimizer step
    args = ap.parse_args()
    if args.agent == "pg":
        return pg_main(args)
    import torch.distributed as dist
    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    env = BatchedAirHockeyEnv(args.envs, device=local, seed=0, env_id_base=rank * args.envs, max_episode_steps=256,
                              ring_depth=0 if args.copy_store else args.buffer_steps)  # the ring is the replay memory: no frame copies
    env.reset()
    agent = DDQNAgent(env, buffer_steps=args.buffer_steps, batch_size=args.batch, discount_rate=0.98, eps_min=0.05, eps_max=0.99,
                      eps_decay_steps=100000)
    agent.replay.record_observation(env)

    def rollout_step():
        action = agent.act()
        _, reward, done, _ = env.step(action)
        agent.store_experience(action, reward, done)

    for _ in range(8):
        rollout_step()
    agent.train()
    out = {"n_gpus": world, "envs": args.envs, "batch": args.batch, "interval": args.interval, "buffer_steps": args.buffer_steps}
    out["rollout_step_ms"], _ = timed(rollout_step, 50)                        # act (6 forward kernels) + step + store
    out["act_ms"], _ = timed(agent.act, 20)
    out["store_ms"], _ = timed(lambda: agent.store_experience(agent.greedy, env.reward, env.done), 20)
    batch = agent.replay.sample(args.batch, agent.gen)
    out["sample_ms"], _ = timed(lambda: agent.replay.sample(args.batch, agent.gen), 10)
    big = min(4096, args.envs)
    t_idx = torch.randint(2, args.buffer_steps - 2, (big,), device=env.device)
    e_idx = torch.randint(0, args.envs, (big,), device=env.device)
    ms, _ = timed(lambda: agent.replay.gather(t_idx, e_idx), 20)
    out["gather"] = {"batch": big, "ms": ms, "GBs_read_plus_write": 2 * 4 * big * 49152 / ms / 1e6}
    out["train_ms"], _ = timed(agent.train, 10)                                 # sample, 3 Q forwards, targets, fit, push
    t0 = time.perf_counter()
    agent.push_weights()
    torch.cuda.synchronize()
    out["push_weights_ms_host"] = (time.perf_counter() - t0) * 1e3

    def loop():
        for _ in range(args.steps):
            rollout_step()
            if agent.iteration % args.interval == 0:
                agent.train()
    if world > 1:
        dist.barrier()
    ms, _ = timed(loop)
    if world > 1:  # max over ranks; all ranks must still hold the same network
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
        flat = agent.trainer.params  # the flat float32 weight vector, BatchNorm moving statistics included
        lo, hi = flat.clone(), flat.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        out["max_parameter_spread_across_ranks"] = float((hi - lo).abs().max().item())
    out["loop"] = {"steps": args.steps, "ms": ms, "env_steps_per_s": world * args.envs * args.steps / ms * 1e3,
                   "trainings": args.steps // args.interval, "losses_tail": agent.loss_buffer[-3:]}
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
