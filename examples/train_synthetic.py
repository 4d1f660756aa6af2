"""Minimal use of the drop-in TrainLoop, mirroring the reference's run/train.py:27-121 with a synthetic
DiffuSeq-shaped model and random token batches.

    python examples/train_synthetic.py                       # one GPU
    torchrun --standalone --nproc-per-node 2 examples/train_synthetic.py   # data parallel

Only the three user hooks are written here; everything after loss.backward() -- gradient all-reduce, grad-norm,
clipping, AdamW, EMA -- is libdpk.
"""
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pipeline_b200 import TrainLoop, dist_util, logger  # noqa: E402
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq  # noqa: E402


class Loop(TrainLoop):
    def compute_losses(self, micro_batch):
        out = self.ddp_model(micro_batch["ids"], micro_batch["t"])
        return {"loss": ((out - micro_batch["target"]) ** 2).mean(dim=(1, 2))}

    @staticmethod
    def backward_from_losses(losses):
        losses["loss"].mean().backward()

    def log_loss_dict(self, mode, losses, *args, **kwargs):
        self.logger.logkv_mean(f"{mode}_loss", float(losses["loss"].mean()))


def batches(batch_size, seq, vocab, emb, seed):
    gen = torch.Generator().manual_seed(seed)
    while True:
        yield {"ids": torch.randint(0, vocab, (batch_size, seq), generator=gen),
               "t": torch.randint(0, 2000, (batch_size,), generator=gen),
               "target": torch.randn(batch_size, seq, emb, generator=gen)}


def main():
    dist_util.setup_dist()
    out_dir = os.environ.get("OPENAI_LOGDIR") or tempfile.mkdtemp(prefix="dpk-example-")
    logger.configure(dir=out_dir, stdout=True)
    torch.manual_seed(102 + dist_util.get_rank())          # the reference seeds seed + rank
    model = SyntheticDiffuSeq(n_layers=2, d=128, ffn=512, vocab=4096).to(dist_util.dev())
    Loop(model=model, data=batches(16, 32, 4096, 128, 1234 + dist_util.get_rank()), batch_size=16, microbatch=8,
         lr=1e-3, ema_rate="0.9999,0.999", log_interval=5, save_interval=10, resume_checkpoint="",
         weight_decay=0.01, learning_steps=20, checkpoint_path=out_dir, gradient_clipping=1.0).run_loop()
    if dist_util.get_rank() == 0:
        print("checkpoints:", sorted(f for f in os.listdir(out_dir) if f.endswith(".pt")))


if __name__ == "__main__":
    main()
