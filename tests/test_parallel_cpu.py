"""World-size-2 (gloo, CPU) test of the data-parallel gradient exchange:
GradArena must leave every rank with the mean of the per-rank gradients, keep never-used
parameters out of the exchange, align every gradient slice to 16 bytes and issue no collective
while ``exchange`` is off."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(7, 5)          # odd sizes: exercises the 16-byte slice padding
        self.b = torch.nn.Linear(5, 3)
        self.unused = torch.nn.Linear(3, 3)     # never used in forward (SURVEY note U)

    def forward(self, x):
        return self.b(torch.tanh(self.a(x)))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import hybridnet  # noqa: F401  (package import must not need CUDA)
        from hybridnet.parallel import GradArena
        torch.manual_seed(0)
        net = _Net()
        used = {n for n, _ in net.named_parameters() if not n.startswith("unused")}
        arena = GradArena(net, used=used, buckets=2)
        assert all(s % 4 == 0 for s, _ in arena.slices), "gradient slices must start on 16-byte boundaries"
        g = torch.Generator().manual_seed(100 + rank)
        x = torch.randn(6, 7, generator=g)
        # local gradients (no exchange)
        arena.zero_grad()
        arena.exchange = False
        net(x).square().mean().backward()
        arena.finish()
        local = {n: p.grad.clone() for n, p in net.named_parameters() if n in used}
        # exchanged gradients
        arena.exchange = True
        arena.zero_grad()
        net(x).square().mean().backward()
        arena.finish()
        ok = True
        for n, p in net.named_parameters():
            if n not in used:
                ok &= p.grad is None
                continue
            gathered = [torch.zeros_like(local[n]) for _ in range(world)]
            dist.all_gather(gathered, local[n])
            mean = sum(gathered) / world
            ok &= bool(torch.allclose(p.grad, mean, atol=1e-6, rtol=1e-5))
            ok &= p.grad.data_ptr() % 16 == 0
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_grad_arena_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
