"""Multi-GPU plumbing: one process per GPU, environments sharded by GLOBAL id, nothing exchanged on the step path.

The reference is single-process (SURVEY.md §2 "Parallelism strategies": none).  Here `torch.distributed` is used for
exactly two things: summing the episode statistics of all ranks, and all-reducing policy gradients in
the RL configuration (ai/reinforcement/rl_agents.py:177-191,255-282 train on one process; with envs on 8 GPUs each
rank computes gradients on its own rollouts).  Works with the `nccl` backend on GPUs and `gloo` on CPU (tests).
"""
import torch
import torch.distributed as dist


def shard_range(total_envs, rank, world):
    """Global env ids [lo, hi) owned by `rank`: contiguous blocks, the remainder spread over the first ranks.
    Because every random draw is keyed by the global id (SPEC.md §4), the union over ranks is the same set of
    trajectories whatever `world` is."""
    if not (0 <= rank < world):
        raise ValueError('rank %d not in [0, %d)' % (rank, world))
    base, extra = divmod(int(total_envs), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def reduce_statistics(stats, group=None):
    """Sum the engine's running statistics (int64[32], see ahk_buffers.stats) over all ranks, in place on a copy."""
    out = stats.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def statistics_dict(stats):
    s = stats.cpu().tolist()
    return {'steps': s[0], 'episodes': s[1], 'goals_top': s[2], 'goals_bottom': s[3], 'reward_sum': 0.5 * s[4],
            'robot_label_hist': s[5:15], 'human_label_hist': s[15:25]}


def all_reduce_mean(flat, group=None, async_op=False):
    """Average a flat tensor over the ranks of `group`, in place, with ONE collective: the trainer's whole gradient
    (2 403 961 float32 = 9.6 MB for the 'rgb' network — at this size the all-reduce is latency-bound, so one launch beats
    per-tensor calls) and the BatchNorm batch statistics.  A no-op on one rank.  async_op: returns the work handle and
    leaves the division to `finish_mean` (call it after handle.wait())."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None
    work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    if not async_op:
        flat /= dist.get_world_size(group)
    return work


def finish_mean(flat, group=None):
    flat /= dist.get_world_size(group)
