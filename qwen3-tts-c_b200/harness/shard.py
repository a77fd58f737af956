"""Request sharding across GPUs: utterances are independent, so a batch of B requests is split
contiguously over the ranks (utterance i -> rank floor(i * world / B)), every rank decodes its share on its own GPU with
replicated weights, and the only collectives are a barrier and the max-over-ranks of the device time.  No data-path
collective exists; this module is the whole multi-GPU "protocol" and runs unchanged on gloo (CPU tests) and NCCL."""
from typing import List, Sequence, Tuple


def partition(n_requests: int, world: int) -> List[Tuple[int, int]]:
    """[lo, hi) of every rank: contiguous, sizes differ by at most one, earlier ranks never get fewer than later ones"""
    if world < 1 or n_requests < 0:
        raise ValueError("world >= 1 and n_requests >= 0 required")
    out = []
    for r in range(world):
        lo = (n_requests * r + world - 1) // world
        hi = (n_requests * (r + 1) + world - 1) // world
        out.append((lo, hi))
    return out


def owner(i: int, n_requests: int, world: int) -> int:
    for r, (lo, hi) in enumerate(partition(n_requests, world)):
        if lo <= i < hi:
            return r
    raise IndexError(i)


def micro_batches(lo: int, hi: int, max_batch: int = 64) -> List[Tuple[int, int]]:
    """a rank's share cut into lock-step batches of at most max_batch utterances (qwen_tts_cuda_generate_batch limit)"""
    return [(s, min(s + max_batch, hi)) for s in range(lo, hi, max_batch)]


def reduce_times(dist, values: Sequence[float], device=None) -> List[float]:
    """max over ranks of each value (a multi-GPU number is the slowest rank's); identity without a process group"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(values)
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def gather_counts(dist, count: int, device=None) -> List[int]:
    """frames (or utterances) every rank produced, on every rank"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [count]
    import torch
    mine = torch.tensor([count], dtype=torch.int64, device=device)
    out = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [int(x.item()) for x in out]
