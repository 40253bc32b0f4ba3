"""Adam over the Gaussian parameter groups in ONE kernel launch (b200splat_adam_step).

`FusedAdam` subclasses torch.optim.Adam so that everything the reference does to its optimizer keeps working
unchanged -- construction with named param groups (scene/gaussian_model.py:185-197), per-iteration
`param_group['lr'] = ...` (:207-214), the state surgery of densification (`optimizer.state[...]['exp_avg']`,
deleting / re-inserting state entries: replace_tensor_to_optimizer, _prune_optimizer, cat_tensors_to_optimizer),
state_dict()/load_state_dict() -- and only `step()` is replaced.  Same state layout as torch (state[p] =
{'step': tensor, 'exp_avg', 'exp_avg_sq'}), same formulas; CUDA fp32 parameters only, no fallback.

One addition: a group may carry `row_width`, `row_split` and `lr_tail`; its tensors are then treated as
[n / row_width, row_width] matrices whose columns >= row_split train with `lr_tail` instead of `lr`.  That lets
the SH coefficients live in ONE [P, M, 3] parameter (row_width = 3 M, row_split = 3: f_dc / f_rest learning rates
of gaussian_model.py:187-188), so the torch.cat of get_features (:119-122) and its backward split disappear."""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib


def expon_lr(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """The position learning-rate schedule the reference builds at scene/gaussian_model.py:200-205 with
    get_expon_lr_func (utils/general_utils.py:30-69): geometric interpolation lr_init -> lr_final over max_steps
    (lr_init^(1-t) * lr_final^t), times an optional sine-eased warm-up factor that rises from lr_delay_mult to 1
    over lr_delay_steps.  Returns step -> learning rate (a Python float)."""
    disabled = (lr_init == 0.0 and lr_final == 0.0)
    log_a, log_b = (0.0, 0.0) if disabled else (math.log(lr_init), math.log(lr_final))

    def schedule(step):
        if disabled or step < 0:
            return 0.0
        t = min(max(step / max_steps, 0.0), 1.0)
        lr = math.exp(log_a * (1.0 - t) + log_b * t)
        if lr_delay_steps > 0:
            ramp = min(max(step / lr_delay_steps, 0.0), 1.0)
            lr *= lr_delay_mult + (1.0 - lr_delay_mult) * math.sin(0.5 * math.pi * ramp)
        return lr

    return schedule


class FusedAdam(torch.optim.Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        todo = []          # (beta1, beta2, eps) -> list of AdamGroup, launched per distinct hyper-parameter set
        for group in self.param_groups:
            if group.get("weight_decay", 0) != 0 or group.get("amsgrad", False) or group.get("maximize", False):
                raise RuntimeError("FusedAdam implements plain Adam only (the reference's settings)")
            key = (float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]))
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("FusedAdam needs CUDA parameters (there is no CPU path)")
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam needs contiguous fp32 parameters")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                g = p.grad if (p.grad.dtype == torch.float32 and p.grad.is_contiguous()) else p.grad.float().contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if not torch.is_tensor(st["step"]):
                    st["step"] = torch.tensor(float(st["step"]), dtype=torch.float32)
                st["step"] += 1
                for k in ("exp_avg", "exp_avg_sq"):
                    if not (st[k].is_contiguous() and st[k].dtype == torch.float32 and st[k].device == p.device):
                        st[k] = st[k].to(device=p.device, dtype=torch.float32).contiguous()
                step = int(st["step"].item()) if st["step"].device.type == "cpu" else int(st["step"].cpu().item())
                rw, rs = int(group.get("row_width", 0)), int(group.get("row_split", 0))
                if rw and p.numel() % rw != 0:
                    raise RuntimeError(f"group {group.get('name')}: {p.numel()} elements are not rows of {rw}")
                todo.append((key, p.device, _lib.AdamGroup(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                                           st["exp_avg_sq"].data_ptr(), p.numel(), float(group["lr"]), step,
                                                           rw, rs, 0, float(group.get("lr_tail", group["lr"]))), g))
        L = _lib.lib()
        i = 0
        while i < len(todo):
            key, dev = todo[i][0], todo[i][1]
            batch = []
            while i < len(todo) and len(batch) < _lib.ADAM_MAX_GROUPS and todo[i][0] == key and todo[i][1] == dev:
                batch.append(todo[i][2])
                i += 1
            arr = (_lib.AdamGroup * len(batch))(*batch)
            with torch.cuda.device(dev):
                _lib.check(L.b200splat_adam_step(len(batch), ctypes.addressof(arr), key[0], key[1], key[2],
                                                 _lib.stream_ptr(dev)), "adam step")
        return loss
