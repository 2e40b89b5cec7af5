"""CPU, world_size 2 over gloo: one flat all-reduce of the head gradients, parameters without a gradient
contributing zeros; matches per-parameter averaging."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from objcavit_b200.grad_sync import FlatGradBuffer


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3), torch.nn.Linear(3, 3))  # last one unused
        buf = FlatGradBuffer(net.parameters())
        assert buf.flat.numel() % 64 == 0 and buf.numel() == sum(p.numel() for p in net.parameters())
        x = torch.full((4, 7), float(rank + 1))
        net[1](net[0](x)).sum().backward()            # autograd accumulates INTO the flat views
        assert all(p.grad.data_ptr() == buf.flat[o:].data_ptr() for p, o in zip(buf.params, buf.offsets))
        local = [p.grad.clone() for p in net.parameters()]
        handle = buf.all_reduce_mean(async_op=True)
        handle.wait()
        gathered = [[torch.zeros_like(g) for _ in range(world)] for g in local]
        for g, lst in zip(local, gathered):
            dist.all_gather(lst, g)
        for p, lst in zip(net.parameters(), gathered):
            assert torch.allclose(p.grad, sum(lst) / world, atol=1e-6)
        assert float(net[2].weight.grad.abs().max()) == 0.0      # unused parameters: zeros, no special handling
        before = torch.linalg.vector_norm(buf.flat).item()
        total = buf.clip_grad_norm_(0.1)
        assert abs(total.item() - before) < 1e-5 and torch.linalg.vector_norm(buf.flat).item() <= 0.1 + 1e-5
        if rank == 0:
            out.put("ok")
    finally:
        dist.destroy_process_group()


def test_flat_grad_allreduce_two_ranks():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == "ok"
