"""Debug aid: run one product step with the GradArena's world>1 bookkeeping forced on a single GPU and print the
stack of every notification a parameter receives beyond its first."""
import os
import sys
import traceback

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hybridnet
from hybridnet import ops, parallel
from tests.golden import trained


def main(name):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    m = hybridnet.MODELS[name]().to(dev).train()
    dead = parallel.used_parameter_names(name)
    arena = parallel.GradArena(m, used={n for n, _ in m.named_parameters() if n not in dead}, buckets=3)
    arena.world = 2
    seen = {}

    def wrap(i, h):
        def f(p):
            st = "".join(traceback.format_stack(limit=8))
            if i in seen:
                print(f"=== second notification for {arena.names[i]} (pending of its bucket {arena.pending[arena.bucket_of[i]]})")
                print("--- first:\n" + seen[i] + "--- second:\n" + st)
            seen[i] = st
            try:
                h(p)
            except RuntimeError as e:
                print("RAISED:", e)
        return f
    arena._hooks = [wrap(i, h) for i, h in enumerate(arena._hooks)]
    for i, p in enumerate(arena.params):
        p.register_post_accumulate_grad_hook(arena._hooks[i])
    arena.direct()
    x, y = trained.blobs(2, 64, 77)
    crit = hybridnet.DiceLoss()
    for it in range(2):
        seen.clear()
        arena.zero_grad()
        ops.begin_step(dev, sum(p.numel() for p in m.parameters()) + (1 << 16))
        loss = crit(m(x.to(dev)), y.to(dev))
        loss.backward()
        arena.finish()
        torch.cuda.synchronize()
        print(name, "iteration", it, "notified", len(seen), "of", len(arena.params), "pending", arena.pending)
    dist.destroy_process_group()


if __name__ == "__main__":
    for n in sys.argv[1:] or ["unet", "fct"]:
        main(n)
