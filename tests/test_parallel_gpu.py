"""N-rank data-parallel parity on real GPUs over NCCL: the product model on 2 ranks, each
with its own shard of the batch, must end the step with the MEAN over ranks of the per-shard oracle gradients
(BatchNorm statistics per rank, as DDP gives -- the reference has no SyncBN), through GradArena's bucketed all-reduce
and through the fused AdamW's folded 1/world.  Skipped on boxes with fewer than 2 GPUs (run with `gpurun --gpus 2`)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import hybridnet
        from hybridnet import ops, parallel
        from oracle import nets
        from tests.golden import trained
        from tests.util import cosine, rel_l2
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        sd = nets.init_state_dict(name, seed=5, jitter=0.2)
        S, Bper = 64, 2
        x, y = trained.blobs(Bper * world, S, 77)
        xs, ys = x[rank * Bper:(rank + 1) * Bper], y[rank * Bper:(rank + 1) * Bper]
        # product on this rank's shard
        m = hybridnet.MODELS[name]()
        m.load_state_dict(sd, strict=True)
        m = m.to(dev).train()
        for mod in m.modules():
            if isinstance(mod, (torch.nn.Dropout, torch.nn.Dropout2d)):
                mod.p = 0.0
        dead = parallel.used_parameter_names(name)
        arena = parallel.GradArena(m, used={n for n, _ in m.named_parameters() if n not in dead}, buckets=3).direct()
        crit = hybridnet.DiceLoss()
        arena.zero_grad()
        ops.begin_step(dev, sum(p.numel() for p in m.parameters()) + (1 << 16))
        loss = crit(m(xs.to(dev)), ys.to(dev))
        loss.backward()
        arena.finish()                                   # all-reduce (sum) joined, scaled by 1/world
        torch.cuda.synchronize()
        got = {n: p.grad.detach().float().cpu().clone() for n, p in m.named_parameters() if n not in dead}
        ops.set_grad_sink(None, None)
        ops.end_zero_pool()
        # oracle: every shard independently (fp32, this GPU), gradients averaged
        want = None
        for r in range(world):
            params = {k: (v.to(dev).clone().requires_grad_(True) if v.is_floating_point() and "running" not in k else v.to(dev).clone())
                      for k, v in sd.items()}
            out = nets.FORWARDS[name](params, x[r * Bper:(r + 1) * Bper].to(dev), nets.Ctx(training=True, dropout="off"))
            trained.loss_of(name, out, None, y[r * Bper:(r + 1) * Bper].to(dev)).backward()
            g = {k: p.grad.detach().cpu() / world for k, p in params.items() if torch.is_tensor(p) and p.requires_grad and p.grad is not None}
            want = g if want is None else {k: want[k] + g[k] for k in g}
        assert set(got) == set(want)
        ga = torch.cat([got[k].flatten() for k in want])
        gb = torch.cat([want[k].flatten() for k in want])
        # and the same gradients on every rank (bit-identical after the all-reduce)
        chk = ga.to(dev).clone()
        dist.broadcast(chk, 0)
        same = bool(torch.equal(chk.cpu(), ga))
        q.put((rank, cosine(ga, gb), rel_l2(ga, gb), same))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["unet", "fct"])
def test_two_rank_nccl_gradients_equal_mean_of_shard_oracles(name):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, cos, rel, same in res:
        assert same, "ranks disagree after the all-reduce"
        # random-init weights with jitter at 64x64: the bound of tests/test_models_gpu.py::test_against_oracle_small
        assert cos >= (0.99 if name == "unet" else 0.999), (name, rank, cos, rel)
        assert rel <= (0.15 if name == "unet" else 3e-2), (name, rank, cos, rel)
