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


class AsyncEpisodeGather:
    """The per-step action all-gather OFF the compute stream ("issue it on a side stream").

    submit(local) snapshots the rank-local (B_local, ...) tensor into a staging slot on the compute stream and enqueues
    the all-gather on a side stream behind an event; the compute stream goes on with the next policy step. Ranks may
    therefore drift apart by up to `depth` steps instead of meeting at a collective every step (every step otherwise
    costs the slowest rank's time). result() returns the gathered (total, ...) tensor of the OLDEST outstanding submit
    and makes the current stream wait for just that gather. On CPU tensors (gloo tests) it degrades to the synchronous
    gather_episodes with the same submit / result order."""

    def __init__(self, total_episodes, local_shape, dtype, device, depth=2, group=None):
        self.total, self.group, self.depth = int(total_episodes), group, int(depth)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.active = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        if self.active and self.total % dist.get_world_size(group) != 0:
            raise ValueError("AsyncEpisodeGather needs equal episode blocks (use gather_episodes for ragged ones)")
        self.stage = [torch.empty(tuple(local_shape), dtype=dtype, device=self.device) for _ in range(self.depth)]
        self.out = [torch.empty((self.total,) + tuple(local_shape[1:]), dtype=dtype, device=self.device)
                    for _ in range(self.depth)] if self.active else None
        self.side = torch.cuda.Stream(device=self.device) if (self.cuda and self.active) else None
        self.ready = [None] * self.depth          # event: staging slot filled (compute stream)
        self.done = [None] * self.depth           # event: gather finished (side stream)
        self.n_submit = 0
        self.n_result = 0

    def submit(self, local):
        if self.n_submit - self.n_result >= self.depth:
            raise RuntimeError("AsyncEpisodeGather: more than `depth` gathers outstanding; call result() first")
        slot = self.n_submit % self.depth
        self.n_submit += 1
        if not self.active:
            self.stage[slot].copy_(local)
            return
        if not self.cuda:
            self.stage[slot].copy_(local)
            dist.all_gather_into_tensor(self.out[slot], self.stage[slot], group=self.group)
            return
        cur = torch.cuda.current_stream(self.device)
        if self.done[slot] is not None:
            cur.wait_event(self.done[slot])       # the gather that last read this staging slot
        self.stage[slot].copy_(local)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.ready[slot] = ev
        with torch.cuda.stream(self.side):
            self.side.wait_event(ev)
            dist.all_gather_into_tensor(self.out[slot], self.stage[slot], group=self.group)
            dn = torch.cuda.Event()
            dn.record(self.side)
            self.done[slot] = dn

    def result(self):
        if self.n_result >= self.n_submit:
            raise RuntimeError("AsyncEpisodeGather: nothing outstanding")
        slot = self.n_result % self.depth
        self.n_result += 1
        if not self.active:
            return self.stage[slot]
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_event(self.done[slot])
        return self.out[slot]

    def drain(self):
        last = None
        while self.n_result < self.n_submit:
            last = self.result()
        return last
