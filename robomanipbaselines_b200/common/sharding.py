"""Multi-GPU plumbing of the batched rollout: one process per GPU, weights replicated, episodes sharded by
contiguous index block, and ONE all-gather of the per-episode actions (and of the per-episode result records at episode
end) so that every rank sees the full matrices in global episode order. The reference has no distributed code
(SURVEY F2); this is the new contract. Works with the `nccl` backend on GPUs and `gloo` on CPU tensors (tests).
"""
import torch
import torch.distributed as dist


def episode_block(total_episodes, rank, world_size):
    """[start, stop) of the episodes owned by `rank`: contiguous blocks, the first `total % world` ranks get one more."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(int(total_episodes), int(world_size))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def episode_seed(base_seed, global_episode_index):
    """Per-episode RNG seed, invariant to the number of GPUs."""
    return int(base_seed) + int(global_episode_index)


def gather_episodes(local, total_episodes, group=None):
    """All-gather a per-episode tensor (B_local, ...) into (total_episodes, ...) in global episode order."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = [episode_block(total_episodes, r, world) for r in range(world)]
    if local.shape[0] != sizes[rank][1] - sizes[rank][0]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} episodes, its block is {sizes[rank]}")
    if total_episodes % world == 0:
        out = torch.empty((total_episodes,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    # ragged blocks: pad every rank's block to the largest one (collectives want equal sizes), gather, drop the padding
    mx = max(b - a for a, b in sizes)
    padded = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * mx:r * mx + (b - a)] for r, (a, b) in enumerate(sizes)], dim=0)


def gather_results(success, reward, duration, total_episodes, group=None):
    """The yaml result lists of RolloutBase (common/base/RolloutBase.py:105-107) across ranks: (success u8, reward f32,
    duration f32) per episode, packed into one float32 all-gather."""
    packed = torch.stack([success.to(torch.float32), reward.to(torch.float32), duration.to(torch.float32)], dim=1)
    allp = gather_episodes(packed, total_episodes, group)
    return allp[:, 0].to(torch.uint8), allp[:, 1], allp[:, 2]
