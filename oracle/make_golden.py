"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the REAL
reference ``TrainLoop`` (imported from /root/reference, unmodified) on CPU.

    python oracle/make_golden.py            # all fixtures (needs /root/reference)

Runs only in the build container: /root/reference does not exist on the GPU
box, which is why the outputs are committed.  Shims applied, none of which
touches the arithmetic:
  * ``blobfile`` stand-in (exists / join / dirname / BlobFile -> open)
  * wandb disabled
  * ``utils.trainer.DistributedDataParallel`` wrapped to drop ``device_ids`` /
    ``output_device`` on CPU (the reference's ctor call crashes there, F3)

Each fixture records, for N optimiser steps of the reference loop: the initial
parameters, the gradients exactly as ``optimize()`` saw them (after DDP's
all-reduce, before clipping), the total norm ``clip_grad_norm_`` returned, the
``grad_norm`` value the reference logged, the learning rate AdamW used, and the
parameters / AdamW moments / EMA copies after every step.
"""
import argparse
import os
import subprocess
import sys
import tempfile
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def install_shims():
    bf = types.ModuleType("blobfile")
    bf.exists = os.path.exists
    bf.join = os.path.join
    bf.dirname = os.path.dirname
    bf.BlobFile = lambda path, mode="rb": open(path, mode)
    sys.modules["blobfile"] = bf
    os.environ.setdefault("WANDB_MODE", "disabled")
    os.environ.setdefault("LOCAL_RANK", "0")
    if REF not in sys.path:
        sys.path.insert(0, REF)


def build_model(torch, seed):
    """Small stand-in with the tensor kinds of a DiffuSeq block: embedding,
    ragged-size linears with bias, LayerNorm.  5 307 parameters, 10 tensors."""
    torch.manual_seed(seed)
    nn = torch.nn

    class Toy(nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = nn.Embedding(37, 19)
            self.up = nn.Linear(19, 64)
            self.ln = nn.LayerNorm(64)
            self.mid = nn.Linear(64, 45)
            self.down = nn.Linear(45, 7)
            self.scale = nn.Parameter(torch.ones(3))

        def forward(self, ids):
            h = torch.tanh(self.up(self.emb(ids)))
            h = self.ln(h)
            h = torch.nn.functional.gelu(self.mid(h))
            return self.down(h) * self.scale.sum()

    return Toy()


def run(cfg, out_path):
    install_shims()
    import torch
    import wandb
    wandb.init(mode="disabled")
    from basic_utils import dist_util, logger
    import utils.trainer as ref_trainer

    torch.set_num_threads(1)
    distributed = "WORLD_SIZE" in os.environ and int(os.environ["WORLD_SIZE"]) > 1
    if distributed:
        dist_util.setup_dist(silent=True)
        real_ddp = ref_trainer.DistributedDataParallel

        def cpu_ddp(module, device_ids=None, output_device=None, **kw):  # F3
            return real_ddp(module, **kw)
        ref_trainer.DistributedDataParallel = cpu_ddp
    rank = dist_util.get_rank()
    tmp = tempfile.mkdtemp(prefix="golden-log-")
    logger.configure(dir=tmp, format_strs=[])

    model = build_model(torch, seed=102)  # same seed on every rank (sidesteps F4)
    if cfg.get("dtype") == "bf16":
        # "DiffuSeq-base bf16" of BASELINE configs[2]: the reference keeps no fp32 master copy
        # (utils/trainer.py:91-92), so AdamW's moments (zeros_like) and the EMA deep copies (:111-113)
        # are bf16 too and every torch op rounds to bf16
        model = model.to(torch.bfloat16)
    rec = {"grads": [], "total_norm": [], "logged": [], "lr": [], "p": [], "m": [], "v": [],
           "ema0": [], "ema1": []}

    def flat(ts):
        # bf16 values are stored widened to float32 (exact); the fixture's `dtype` field says what they are
        return np.concatenate([t.detach().reshape(-1).float().numpy().astype(np.float32) for t in ts])

    real_clip = torch.nn.utils.clip_grad_norm_

    def recording_clip(params, max_norm, *a, **kw):
        tn = real_clip(params, max_norm, *a, **kw)
        rec["total_norm"].append(float(tn))
        return tn
    torch.nn.utils.clip_grad_norm_ = recording_clip

    class Loop(ref_trainer.TrainLoop):
        def compute_losses(self, micro_batch):
            out = self.ddp_model(micro_batch["ids"])
            return {"loss": ((out - micro_batch["target"]) ** 2).mean(dim=(1, 2))}

        @staticmethod
        def backward_from_losses(losses):
            losses["loss"].mean().backward()

        def log_loss_dict(self, mode, losses, *args, **kwargs):
            pass

        def optimize(self):
            rec["grads"].append(flat([p.grad for p in self.master_params]))
            super().optimize()
            rec["logged"].append(float(logger.getkvs()["grad_norm"]))
            logger.getkvs().clear()
            logger.get_current().name2cnt.clear()
            rec["lr"].append(float(self.opt.param_groups[0]["lr"]))
            rec["p"].append(flat(self.master_params))
            rec["m"].append(flat([self.opt.state[p]["exp_avg"] for p in self.master_params]))
            rec["v"].append(flat([self.opt.state[p]["exp_avg_sq"] for p in self.master_params]))
            rec["ema0"].append(flat(self.ema_params[0]))
            rec["ema1"].append(flat(self.ema_params[1]))

    gen = torch.Generator().manual_seed(1234 + rank)
    B, L = cfg["batch"], 11

    def data():
        while True:
            yield {"ids": torch.randint(0, 37, (B, L), generator=gen),
                   "target": torch.randn(B, L, 7, generator=gen) * cfg["target_scale"]}

    names = [n for n, _ in model.named_parameters()]
    shapes = [tuple(p.shape) for p in model.parameters()]
    init = flat(list(model.parameters()))
    loop = Loop(model=model, data=data(), batch_size=B, microbatch=cfg["microbatch"], lr=cfg["lr"],
                ema_rate=cfg["ema_rate"], log_interval=10 ** 9, save_interval=10 ** 9, resume_checkpoint="",
                weight_decay=cfg["wd"], learning_steps=cfg["learning_steps"], checkpoint_path=tmp,
                gradient_clipping=cfg["clip"])
    it = data()
    for _ in range(cfg["steps"]):
        loop.run_step(next(it))
        loop.step += 1
    if cfg["clip"] <= 0:
        rec["total_norm"] = [float("nan")] * cfg["steps"]
    if rank == 0:
        np.savez_compressed(
            out_path, names=np.array(names), shapes=np.array([str(s) for s in shapes]),
            numels=np.array([int(np.prod(s)) for s in shapes], dtype=np.int64), init=init,
            grads=np.stack(rec["grads"]), total_norm=np.array(rec["total_norm"]),
            logged=np.array(rec["logged"]), lr=np.array(rec["lr"]), p=np.stack(rec["p"]),
            m=np.stack(rec["m"]), v=np.stack(rec["v"]), ema0=np.stack(rec["ema0"]), ema1=np.stack(rec["ema1"]),
            ema_rates=np.array([float(x) for x in cfg["ema_rate"].split(",")]),
            wd=cfg["wd"], clip=cfg["clip"], lr0=cfg["lr"], learning_steps=cfg["learning_steps"],
            world_size=int(os.environ.get("WORLD_SIZE", "1")), dtype=cfg.get("dtype", "f32"),
            torch_version=str(torch.__version__))
        print("wrote", out_path)


CONFIGS = {
    # clip active (large targets -> norm >> 1), decay, annealed lr, two EMA rates
    "ref_clip_wd_ws1": dict(batch=8, microbatch=4, lr=1e-3, ema_rate="0.9999,0.999", wd=0.01,
                            learning_steps=50, clip=1.0, steps=6, target_scale=30.0, world=1),
    # clip off, no decay (AdamW skips the multiply), constant lr
    "ref_noclip_ws1": dict(batch=8, microbatch=8, lr=3e-4, ema_rate="0.9999,0.5", wd=0.0,
                           learning_steps=0, clip=-1.0, steps=6, target_scale=1.0, world=1),
    # the reference's CPU/gloo DDP path, two ranks, microbatching with no_sync
    "ref_clip_wd_ws2": dict(batch=8, microbatch=4, lr=1e-3, ema_rate="0.9999,0.999", wd=0.01,
                            learning_steps=50, clip=1.0, steps=6, target_scale=30.0, world=2),
    # every state tensor bf16 (F2/F7): bf16 per-tensor norms, bf16 clip coefficient, bf16 moments and EMA
    "ref_bf16_clip_ema2_ws1": dict(batch=8, microbatch=4, lr=1e-3, ema_rate="0.9999,0.999", wd=0.01,
                                   learning_steps=50, clip=1.0, steps=6, target_scale=30.0, world=1, dtype="bf16"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worker", default=None)
    ap.add_argument("--only", default=None, help="generate just this fixture")
    args = ap.parse_args()
    if args.worker:
        run(CONFIGS[args.worker], os.path.join(OUT, args.worker + ".npz"))
        return
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    os.makedirs(OUT, exist_ok=True)
    for name, cfg in CONFIGS.items():
        if args.only and name != args.only:
            continue
        env = dict(os.environ, WANDB_MODE="disabled")
        if cfg["world"] == 1:
            env.pop("WORLD_SIZE", None)
            cmd = [sys.executable, __file__, "--worker", name]
        else:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={cfg['world']}",
                   "--master-addr", "127.0.0.1", "--master-port", "29731", __file__, "--worker", name]
            env["CUDA_VISIBLE_DEVICES"] = ""
        subprocess.run(cmd, check=True, env=env)


if __name__ == "__main__":
    main()
