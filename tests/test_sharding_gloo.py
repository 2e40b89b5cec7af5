"""CPU, world_size 2 over gloo: the data-parallel host logic (batch sharding at image boundaries, CSR offsets of the
shards, max-over-ranks timing, optional depth gather).  No kernels run here — there is no CPU compute path."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from objcavit_b200 import sharding
from oracle import synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        feats = synth.make_features(B, 32, 48, seed=7)
        objs, boxes = synth.make_objects(B, 32, 48, 5, seed=7, ragged=True, none_at=1)
        f, o, b = sharding.shard_batch(feats, objs, boxes, rank, world)
        lo, hi = sharding.shard_bounds(B, rank, world)
        assert f.shape[0] == hi - lo == len(o) == len(b)
        assert all(torch.equal(x, y) for x, y in zip(o, objs[lo:hi]))
        assert o is not objs                                              # F11: the head may overwrite the shard's list
        # stand-in "depth": a per-image checksum image, so the gather order can be verified exactly
        depth = f.sum(dim=1, keepdim=True)
        full = sharding.gather_depth(depth, B)
        ms = sharding.max_over_ranks_ms(10.0 + rank)
        assert ms == 10.0 + world - 1
        counts = [int(x.shape[0]) for x in o]
        offs = sharding.csr_offsets(counts)
        assert offs[-1] == sum(counts) and len(offs) == len(o) + 1
        if rank == 0:
            assert torch.equal(full, feats.sum(dim=1, keepdim=True))
            out.put("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [4, 5])
def test_two_rank_sharding_gloo(B):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == "ok"


def test_shard_bounds_cover_batch():
    for B in range(1, 20):
        for world in (1, 2, 4, 8):
            spans = [sharding.shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1
