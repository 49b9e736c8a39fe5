"""Host-side multi-rank logic on CPU: shard plan properties, and a world_size-2 gloo run that
checks every rank derives a consistent, complete partition and that mask tables keyed by global
utterance id are identical however the batch is split."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import logmel_oracle as O
from speech_recognition_b200.sharding import shard_ranges


def test_shard_ranges_tile_and_balance():
    rng = np.random.default_rng(1)
    lengths = np.clip(rng.normal(12.3, 4.5, 2930), 1.0, 20.48) * 16000   # BASELINE config 5 durations
    lengths = lengths.astype(np.int64).tolist()
    for world in (1, 2, 4, 8):
        ranges = shard_ranges(lengths, world)
        assert ranges[0][0] == 0 and ranges[-1][1] == len(lengths)
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
        loads = [sum(lengths[a:b]) for a, b in ranges]
        assert max(loads) - min(loads) <= 2 * max(lengths)
    assert shard_ranges([], 4) == [(0, 0)] * 4
    assert shard_ranges([0, 0, 0, 0], 2) == [(0, 2), (2, 4)]
    assert shard_ranges([5], 3)[0][0] == 0 and sum(b - a for a, b in shard_ranges([5], 3)) == 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, lengths, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a, b = shard_ranges(lengths, world)[rank]
    # each rank's mask tables for ITS utterances, keyed by global id
    tables = [O.specaug_draw_masks(O.frame_count(lengths[i], 320, 160), 80, 27, 2, 100, 1.0, 2, seed=43,
                                   utterance_id=i) for i in range(a, b)]
    flat = torch.tensor([v for t in tables for pair in t for v in pair], dtype=torch.int64)
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([flat.numel()]))
    bufs = [torch.zeros(int(s.item()), dtype=torch.int64) for s in sizes]
    # gloo all_gather needs equal sizes: pad to the max
    m = max(int(s.item()) for s in sizes)
    padded = torch.zeros(m, dtype=torch.int64)
    padded[:flat.numel()] = flat
    outs = [torch.zeros(m, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(outs, padded)
    gathered = torch.cat([o[:int(s.item())] for o, s in zip(outs, sizes)])
    # timing contract of bench.py: max over ranks
    t = torch.tensor([1.0 + rank])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ret["gathered"] = gathered.tolist()
        ret["max_t"] = float(t.item())
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_partition_is_consistent():
    lengths = [48000, 31000, 16500, 19000, 64000, 25000, 32000]   # all > 100 frames: no TF error case
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), lengths, ret), nprocs=world, join=True)
    single = [v for i in range(len(lengths))
              for pair in O.specaug_draw_masks(O.frame_count(lengths[i], 320, 160), 80, 27, 2, 100, 1.0, 2, seed=43,
                                               utterance_id=i) for v in pair]
    assert ret["gathered"] == single
    assert ret["max_t"] == 2.0
