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


class GradientBucket(object):
    """One flat buffer for all policy gradients (2 403 961 float32 = 9.6 MB for the 'rgb' network), all-reduced with a
    single collective and averaged: at this size the all-reduce is latency-bound, so one launch beats per-tensor calls."""

    def __init__(self, params, dtype=None):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        first = self.params[0]
        self.flat = torch.zeros(n, dtype=dtype or first.dtype, device=first.device)

    def all_reduce(self, group=None, async_op=False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None  # one rank: .grad already holds the average
        offset = 0
        for p in self.params:
            k = p.numel()
            if p.grad is not None:
                self.flat[offset:offset + k].copy_(p.grad.reshape(-1))
            else:
                self.flat[offset:offset + k].zero_()
            offset += k
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        work = None
        if world > 1:
            work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if not async_op:
            self.scatter(world)
        return work

    def scatter(self, world):
        """Write the averaged gradients back into .grad (call after an async all_reduce has completed)."""
        offset = 0
        for p in self.params:
            k = p.numel()
            g = self.flat[offset:offset + k].reshape(p.shape) / world
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)
            offset += k
