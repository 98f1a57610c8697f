"""CPU suite: the N>1 host logic (batch sharding + result gather) on the gloo backend, world_size 2."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_total, ragged, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import hpe_b200 as hpe
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        K = 17
        g = torch.Generator().manual_seed(0)
        preds = torch.rand(n_total, K, 2, generator=g); maxvals = torch.rand(n_total, K, 1, generator=g)
        lo, hi = hpe.shard_range(n_total, rank, world)
        if ragged:
            out = hpe.gather_keypoints(preds[lo:hi], maxvals[lo:hi])
        elif ragged is None:       # packed rows straight from the decode kernel: nothing to concatenate
            out = hpe.gather_keypoints_equal(torch.cat([preds[lo:hi], maxvals[lo:hi]], 2).contiguous())
        else:
            out = hpe.gather_keypoints_equal(preds[lo:hi], maxvals[lo:hi])
        want = torch.cat([preds, maxvals], 2)
        q.put((rank, bool(torch.equal(out, want)), tuple(out.shape)))
    finally:
        dist.destroy_process_group()


def _run(n_total, ragged):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, ragged, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, shape in res:
        assert ok, f"rank {rank} gathered wrong data"
        assert shape == (n_total, 17, 3)


def test_shard_range_partitions_exactly():
    import hpe_b200 as hpe
    for n in (0, 1, 7, 256, 513):
        for w in (1, 2, 3, 8):
            spans = [hpe.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_gather_equal_shards_gloo_world2():
    _run(8, ragged=False)


def test_gather_packed_rows_gloo_world2():
    _run(8, ragged=None)


def test_gather_ragged_shards_gloo_world2():
    _run(7, ragged=True)


def test_gather_single_process_is_identity():
    import hpe_b200 as hpe
    p, m = torch.rand(3, 17, 2), torch.rand(3, 17, 1)
    assert torch.equal(hpe.gather_keypoints(p, m), torch.cat([p, m], 2))
