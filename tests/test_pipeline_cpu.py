"""Host logic of the multi-GPU path on CPU: pair / sequence sharding and the exchange of match records
(padded all_gather and gather-to-consumer) with the gloo backend at world_size 2."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from featurematching_b200 import _lib, pipeline


def test_shard_pairs_partitions_everything():
    for n in (0, 1, 7, 32, 9999):
        for world in (1, 2, 3, 8):
            spans = [pipeline.shard_pairs(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1


def test_shard_sequence_has_one_frame_halo():
    frames, world = 10000, 8   # BASELINE config 4: 10k frames -> 9999 consecutive pairs
    total = 0
    for r in range(world):
        b, e, (f0, f1) = pipeline.shard_sequence(frames, r, world)
        assert (f0, f1) == (b, e + 1)      # pairs (k, k + 1), k in [b, e) need frames b .. e
        total += e - b
    assert total == frames - 1
    assert pipeline.shard_sequence(1, 0, 2)[:2] == (0, 0)


def _fake_records(rank, capacity, n):
    rec = np.zeros(capacity + 1, dtype=pipeline.RECORD_DTYPE)
    rec[0]["pair"] = n
    rec[0]["i"] = rank * 100
    for k in range(n):
        rec[1 + k] = (rank * 100 + k // 3, k, 2 * k, 0.5 + 0.01 * k, k, k + 1, 0.1, k + 2, k + 3, 0.2)
    return torch.from_numpy(rec.view(np.uint8).copy())


def _worker(rank, world, port, capacity, counts, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = _fake_records(rank, capacity, counts[rank])

        def complete(out):
            per_rank = pipeline.unpack_records(out, capacity, world)
            good = all(len(per_rank[r]) == counts[r] for r in range(world))
            for r in range(world):
                exp = _fake_records(r, capacity, counts[r]).numpy().view(pipeline.RECORD_DTYPE)[1:1 + counts[r]]
                good = good and bool((per_rank[r] == exp).all())
            return good

        ok = complete(pipeline.gather_matches(mine, capacity))            # every rank receives everything
        for dst in range(world):                                          # only the consumer rank receives
            out = pipeline.gather_matches(mine, capacity, dst=dst)
            ok = ok and ((out is None) if rank != dst else complete(out))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_gather_matches_gloo_world2():
    """all_gather form: every rank ends with every rank's records; gather-to-consumer form (dst): only that
    rank does.  A rank with zero matches still takes part (the reference's DDP-deadlock concern, config.py:35)."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world, capacity, counts = 2, 64, [17, 0]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, capacity, counts, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_record_layout_matches_header():
    assert pipeline.RECORD_DTYPE.itemsize == _lib.RECORD_BYTES == 40
    assert pipeline.RECORD_DTYPE.names == ("pair", "i", "j", "conf", "x0", "y0", "std0", "x1", "y1", "std1")
