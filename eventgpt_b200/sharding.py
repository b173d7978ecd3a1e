"""Multi-GPU plan for the visual-token stage: one process per GPU, windows sharded in
contiguous blocks, weights replicated, NO collective on the hot path -- only a final
gather of the tokens.  The reference has no multi-GPU path;
this is the data-parallel driver around its per-sample loop
(model/EventChatModel.py:332-343)."""
import torch
import torch.distributed as dist


def shard_bounds(n_windows, world_size, rank):
    """Contiguous block [lo, hi) of rank ``rank``; the first n % world ranks get one extra window."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world {world_size}")
    base, extra = divmod(n_windows, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_events(window_offsets, world_size, rank):
    """Event range and re-based offsets of this rank's windows (ragged windows split at window
    boundaries).  ``window_offsets``: 1-D integer tensor/array [B+1]."""
    lo, hi = shard_bounds(len(window_offsets) - 1, world_size, rank)
    local = window_offsets[lo:hi + 1] - window_offsets[lo]
    return (int(window_offsets[lo]), int(window_offsets[hi])), local, (lo, hi)


def gather_tokens(tokens, n_windows_total=None):
    """Final gather: every rank ends with [B_total, n+577, 4096].  Equal shards use one
    all_gather_into_tensor (NCCL over NVLink); ragged shards pad to the largest shard."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tokens
    world = dist.get_world_size()
    counts = torch.tensor([tokens.shape[0]], dtype=torch.int64, device=tokens.device)
    if n_windows_total is not None and n_windows_total % world == 0:
        out = torch.empty((tokens.shape[0] * world,) + tuple(tokens.shape[1:]), dtype=tokens.dtype, device=tokens.device)
        dist.all_gather_into_tensor(out, tokens.contiguous())
        return out
    all_counts = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(all_counts, counts)
    sizes = [int(c.item()) for c in all_counts]
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(tokens.shape[1:]), dtype=tokens.dtype, device=tokens.device)
    pad[: tokens.shape[0]] = tokens
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], dim=0)
