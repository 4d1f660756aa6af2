"""TEST INFRASTRUCTURE ONLY -- the reference's update path spelled with the very
torch calls the reference makes, on per-tensor lists, for any device.

Why it exists next to dpk_oracle.c: the reference's arithmetic *is* torch's
(``torch.nn.utils.clip_grad_norm_``, ``torch.optim.AdamW``, eager ``mul_/add_``),
and /root/reference cannot travel to the GPU box.  This port is
  * the timed CPU baseline of bench.py (``cpu_baseline.kind = "port"`` and
    ``--impl reference``): same torch CPU kernels the reference would run;
  * the tight-tolerance oracle on the B200 box (torch-CUDA eager), SURVEY §8c T1.
It is checked against the real reference through tests/golden (CPU).

Follows /root/reference/utils/trainer.py:
  ctor  :91-113   params list, AdamW(lr, weight_decay), K deep-copied EMA lists
  optimize :237-244, grad_clip :246-255, _anneal_lr :257-263,
  _log_grad_norm :265-271, update_ema :360-370, _zero_grad :203-207
"""
import copy
import math

import torch


class ReferenceUpdatePort:
    def __init__(self, params, lr, ema_rates, weight_decay=0.0, learning_steps=0, gradient_clipping=-1.0,
                 fused=False):
        self.params = list(params)
        self.lr = float(lr)
        self.ema_rate = [float(r) for r in ema_rates]
        self.weight_decay = weight_decay
        self.learning_steps = learning_steps
        self.gradient_clipping = gradient_clipping
        self.step = 0
        self.resume_step = 0
        kw = {"fused": True} if fused else {}
        self.opt = torch.optim.AdamW(self.params, lr=self.lr, weight_decay=self.weight_decay, **kw)
        self.ema_params = [copy.deepcopy(self.params) for _ in self.ema_rate]
        self.last_total_norm = None
        self.last_logged_norm = None

    def zero_grad(self):
        for p in self.params:
            if p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()

    def optimize(self):
        if self.gradient_clipping > 0:
            self.last_total_norm = torch.nn.utils.clip_grad_norm_(self.params, self.gradient_clipping)
        sqsum = 0.0
        for p in self.params:
            if p.grad is not None:
                sqsum += (p.grad ** 2).sum().item()
        self.last_logged_norm = math.sqrt(sqsum)
        if self.learning_steps:
            frac_done = (self.step + self.resume_step) / self.learning_steps
            for group in self.opt.param_groups:
                group["lr"] = self.lr * (1 - frac_done)
        self.opt.step()
        for rate, targets in zip(self.ema_rate, self.ema_params):
            for trg, src in zip(targets, self.params):
                trg.detach().mul_(rate).add_(src, alpha=1 - rate)
        self.step += 1
