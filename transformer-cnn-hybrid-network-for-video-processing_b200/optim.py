"""``hybridnet.optim.AdamW`` -- the optimizer step right after the hot path (the reference's
``torch.optim.AdamW(model.parameters(), lr)`` / ``optimizer.step()``, FCT.py:305,338, Model.py:153,59) as ONE
kernel launch over flat arenas instead of ~300 per-tensor launches:

* the live parameters are re-pointed into one flat fp32 buffer laid out exactly like the gradient arena
  (``parallel.GradArena``), so ``hn_adamw_step_dev`` walks four congruent arrays (param, grad, exp_avg,
  exp_avg_sq).  The ``nn.Parameter`` objects, their names, shapes and ``state_dict`` are untouched;
* parameters that never receive a gradient (SURVEY note U, frozen encoders) are not in the arena and are left
  bit-identical -- what torch's AdamW does for ``p.grad is None`` (no weight decay either);
* the 1/world_size of the data-parallel mean rides the same pass (``grad_scale``), so ``GradArena.finish``
  needs no extra pass over the gradients;
* the step count lives on the device, which makes the update capturable in the step's CUDA graph;
* ``state_dict()`` / ``load_state_dict()`` speak torch.optim.AdamW's format (``state[i] = {step, exp_avg,
  exp_avg_sq}`` indexed over ALL parameters handed in, one param group), because the reference stores it in its
  ``.tar`` checkpoints (FCT.py:371) -- either optimizer can resume from the other's file.
"""
import torch

from ._lib import call


class AdamW:
    def __init__(self, params, arena, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, fold_world_scale=True):
        self.all_params = list(params)
        self.arena = arena
        self.defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        template = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))], lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self._group = {k: v for k, v in template.state_dict()["param_groups"][0].items() if k != "params"}
        self.param_groups = [dict(self._group, params=self.all_params)]
        n = arena.arena.numel()
        dev = arena.arena.device
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(2, dtype=torch.int64, device=dev)        # [steps done, scratch counter]
        self._steps_host = 0
        with torch.no_grad():
            for p, (s, e) in zip(arena.params, arena.slices):
                self.flat[s:e].copy_(p.detach().reshape(-1))
                p.data = self.flat[s:e].view_as(p)
        arena.rebind()
        # the data-parallel mean: sum-all-reduce, then this scale inside the update
        self.grad_scale = 1.0 / arena.world if (fold_world_scale and arena.world > 1) else 1.0
        if self.grad_scale != 1.0:
            arena.scale_in_finish = False
        self._index = {id(p): i for i, p in enumerate(self.all_params)}

    # -- torch.optim.Optimizer surface used by the reference's trainers --------------------------------------
    def zero_grad(self, set_to_none=False):
        self.arena.zero_grad()

    def step(self):
        from . import ops
        g = self.param_groups[0]
        if not self.flat.is_cuda:
            raise RuntimeError("hybridnet.optim.AdamW.step: CUDA only (there is no CPU fallback)")
        b1, b2 = g["betas"]
        call("hn_adamw_step_dev", self.flat.data_ptr(), self.arena.arena.data_ptr(), self.exp_avg.data_ptr(),
             self.exp_avg_sq.data_ptr(), self.flat.numel(), float(g["lr"]), float(b1), float(b2), float(g["eps"]),
             float(g["weight_decay"]), self.step_dev.data_ptr(), float(self.grad_scale),
             torch.cuda.current_stream().cuda_stream)
        self._steps_host += 1
        # the kernel wrote through raw pointers: Parameter._version did not move, so refresh every derived bf16
        # weight copy here (one launch) instead of relying on the version check of the lazy path
        ops.repack_all()

    def steps_done(self):
        """Step count as the device holds it (reads the device: use outside captured regions)."""
        return int(self.step_dev[0].item())

    def state_dict(self):
        steps = self.steps_done() if self.flat.is_cuda else self._steps_host
        state = {}
        if steps > 0:
            for p, (s, e) in zip(self.arena.params, self.arena.slices):
                state[self._index[id(p)]] = {"step": torch.tensor(float(steps)),
                                             "exp_avg": self.exp_avg[s:e].view_as(p).clone(),
                                             "exp_avg_sq": self.exp_avg_sq[s:e].view_as(p).clone()}
        group = dict(self.param_groups[0])
        group["params"] = list(range(len(self.all_params)))
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        groups = sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(self.all_params):
            raise ValueError("hybridnet.optim.AdamW.load_state_dict: expected one param group over the same parameter list")
        if groups[0].get("amsgrad") or groups[0].get("maximize"):
            raise ValueError("hybridnet.optim.AdamW: amsgrad / maximize are not part of the reference path")
        for k, v in groups[0].items():
            if k != "params":
                self.param_groups[0][k] = v
        steps = set()
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        live = {self._index[id(p)]: (p, se) for p, se in zip(self.arena.params, self.arena.slices)}
        for idx, st in sd["state"].items():
            idx = int(idx)
            if idx not in live:
                raise ValueError(f"hybridnet.optim.AdamW.load_state_dict: parameter {idx} has optimizer state but never "
                                 "receives a gradient here")
            p, (s, e) = live[idx]
            self.exp_avg[s:e].copy_(st["exp_avg"].reshape(-1))
            self.exp_avg_sq[s:e].copy_(st["exp_avg_sq"].reshape(-1))
            steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError("hybridnet.optim.AdamW.load_state_dict: per-parameter step counts differ")
        n = steps.pop() if steps else 0
        self.step_dev.zero_()
        self.step_dev[0] = n
        self._steps_host = n
