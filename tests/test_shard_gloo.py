"""The N > 1 path on CPU: request partition and the two collectives the bench uses, world_size 2 over gloo."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import shard  # noqa: E402


def test_partition_is_contiguous_and_balanced():
    for n in (0, 1, 2, 63, 64, 65, 256, 257):
        for w in (1, 2, 3, 4, 8):
            parts = shard.partition(n, w)
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1
            for i in range(n):
                lo, hi = parts[shard.owner(i, n, w)]
                assert lo <= i < hi
    assert shard.partition(256, 8) == [(32 * r, 32 * (r + 1)) for r in range(8)]        # BASELINE config 5
    assert shard.micro_batches(0, 128) == [(0, 64), (64, 128)]
    assert shard.micro_batches(10, 75) == [(10, 74), (74, 75)]
    with pytest.raises(ValueError):
        shard.partition(4, 0)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.partition(5, world)[rank]
    # every rank "decodes" its share: frames = 10 per utterance, device time = 1.5 s on rank 0, 2.5 s on rank 1
    frames = 10 * (hi - lo)
    dist.barrier()
    t = shard.reduce_times(dist, [1.5 + rank, 0.25 * (rank + 1)])
    counts = shard.gather_counts(dist, frames)
    q.put((rank, (lo, hi), t, counts))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_over_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][1] == (0, 3) and got[1][1] == (3, 5)          # contiguous shares of 5 requests
    for _, _, t, counts in got:
        assert t == [2.5, 0.5]                                    # max over ranks, identical everywhere
        assert counts == [30, 20] and sum(counts) == 50           # whole-job frames = sum over ranks
