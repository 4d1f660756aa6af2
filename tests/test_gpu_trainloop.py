"""The drop-in TrainLoop end to end on one GPU against the reference's loop
restated with torch (oracle/torch_port.py): same model, same batches, N steps."""
import copy
import os

import numpy as np
import pytest
import torch

from distributed_pipeline_b200 import TrainLoop, logger, update_ema
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq
from oracle.torch_port import ReferenceUpdatePort

pytestmark = pytest.mark.gpu

TOY = dict(n_layers=2, d=128, ffn=512, vocab=4096)


class Loop(TrainLoop):
    def compute_losses(self, micro_batch):
        out = self.ddp_model(micro_batch["ids"], micro_batch["t"])
        return {"loss": ((out - micro_batch["target"]) ** 2).mean(dim=(1, 2))}

    @staticmethod
    def backward_from_losses(losses):
        losses["loss"].mean().backward()

    def log_loss_dict(self, mode, losses, *args, **kwargs):
        pass


def batches(seed, B, L=16, n=100, scale=1.0):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        out.append({"ids": torch.randint(0, TOY["vocab"], (B, L), generator=gen),
                    "t": torch.randint(0, 2000, (B,), generator=gen),
                    "target": torch.randn(B, L, 128, generator=gen) * scale})
    return out


def reference_steps(model, data, n_steps, microbatch, **kw):
    """forward_backward + optimize of the reference (utils/trainer.py:230-244), torch only."""
    port = ReferenceUpdatePort(list(model.parameters()), **kw)
    logged = []
    for batch in data[:n_steps]:
        port.zero_grad()
        B = batch["ids"].shape[0]
        for i in range(0, B, microbatch):
            mb = {k: v[i:i + microbatch].cuda() for k, v in batch.items()}
            out = model(mb["ids"], mb["t"])
            ((out - mb["target"]) ** 2).mean(dim=(1, 2)).mean().backward()
        port.optimize()
        logged.append(port.last_logged_norm)
    return port, logged


def make_loop(model, data, tmp_path, **kw):
    logger.configure(dir=str(tmp_path))
    args = dict(model=model, data=iter(data), batch_size=8, microbatch=4, lr=1e-3, ema_rate="0.9999,0.999",
                log_interval=10 ** 9, save_interval=10 ** 9, resume_checkpoint="", weight_decay=0.01,
                learning_steps=0, checkpoint_path=str(tmp_path), gradient_clipping=-1.0)
    args.update(kw)
    return Loop(**args)


def close(a, b, rtol=1e-6):
    a, b = a.double(), b.double()
    rms = b.pow(2).mean().sqrt()
    return bool(((a - b).abs() <= rtol * (b.abs() + rms)).all())


@pytest.mark.parametrize("clip", [-1.0, 0.5])
def test_trainloop_matches_reference_loop(tmp_path, clip):
    torch.manual_seed(102)
    model = SyntheticDiffuSeq(**TOY).cuda()
    ref_model = copy.deepcopy(model)
    data = batches(1, 8, scale=3.0 if clip > 0 else 1.0)
    n_steps = 6
    loop = make_loop(model, data, tmp_path, gradient_clipping=clip, learning_steps=50)
    for batch in data[:n_steps]:
        loop.run_step(batch)
        loop.step += 1
    loop.flush_logs()
    port, logged = reference_steps(ref_model, data, n_steps, 4, lr=1e-3, ema_rates=[0.9999, 0.999],
                                   weight_decay=0.01, learning_steps=50, gradient_clipping=clip)
    got_norm = logger.getkvs()["grad_norm"]
    assert abs(got_norm - np.mean(logged)) <= 1e-5 * np.mean(logged)
    for p, q in zip(loop.master_params, port.params):
        assert close(p, q)
    for k in range(2):
        for e, f in zip(loop.ema_params[k], port.ema_params[k]):
            assert close(e, f)
    sd = loop.opt.state_dict()
    for i, q in enumerate(port.params):
        assert close(sd["state"][i]["exp_avg"], port.opt.state[q]["exp_avg"])
        assert close(sd["state"][i]["exp_avg_sq"], port.opt.state[q]["exp_avg_sq"])
    # attributes the reference exposes
    assert loop.opt.param_groups[0]["lr"] == pytest.approx(1e-3 * (1 - (n_steps - 1) / 50))
    assert loop.global_batch == 8 and loop.use_ddp is False and loop.ddp_model is loop.model
    assert loop.master_params is loop.model_params and len(loop.ema_params) == 2
    if clip > 0:
        # "clipped" mode leaves in .grad what clip_grad_norm_ leaves there
        for p, q in zip(loop.master_params, port.params):
            assert close(p.grad, q.grad, 1e-5)


def test_checkpoints_are_torch_compatible_and_resume(tmp_path):
    torch.manual_seed(102)
    model = SyntheticDiffuSeq(**TOY).cuda()
    data = batches(2, 8)
    loop = make_loop(model, data, tmp_path, gradient_clipping=1.0)
    for batch in data[:3]:
        loop.run_step(batch)
        loop.step += 1
    loop.save()
    names = sorted(os.listdir(tmp_path))
    assert "model_000003.pt" in names and "opt_000003.pt" in names
    assert "ema_0.9999_000003.pt" in names and "ema_0.999_000003.pt" in names
    # the optimizer file is a stock torch.optim.AdamW state dict
    stock_model = SyntheticDiffuSeq(**TOY).cuda()
    stock_model.load_state_dict(torch.load(tmp_path / "model_000003.pt"))
    stock = torch.optim.AdamW(stock_model.parameters(), lr=1e-3, weight_decay=0.01)
    stock.load_state_dict(torch.load(tmp_path / "opt_000003.pt"))
    st = stock.state[next(iter(stock_model.parameters()))]
    assert float(st["step"]) == 3.0 and st["exp_avg"].shape == loop.master_params[0].shape
    saved = torch.load(tmp_path / "model_000003.pt")
    assert all(v._base is None for v in saved.values())  # no flat buffer dragged into the file

    # resume into a fresh loop: discovery is by newest model*.pt in the logger dir (reference :319-335)
    model2 = SyntheticDiffuSeq(**TOY).cuda()
    loop2 = make_loop(model2, data, tmp_path, gradient_clipping=1.0)
    assert loop2.resume_step == 3 and loop2.opt.step_count == 3
    for a, b in zip(loop.master_params, loop2.master_params):
        assert torch.equal(a, b)
    for k in range(2):
        for a, b in zip(loop.ema_params[k], loop2.ema_params[k]):
            assert torch.equal(a, b)
    assert torch.equal(loop.opt.flat.exp_avg, loop2.opt.flat.exp_avg)
    # and both continue identically
    loop.run_step(data[3])
    loop2.run_step(data[3])
    for a, b in zip(loop.master_params, loop2.master_params):
        assert torch.equal(a, b)
    # a stock AdamW checkpoint loads back into the fused optimizer
    loop2.opt.load_state_dict(stock.state_dict())
    assert loop2.opt.step_count == 3


def test_run_loop_logs_and_saves_like_the_reference(tmp_path):
    """run_loop (reference :175-196): dumpkvs every log_interval with step / samples /
    mean grad_norm, final save; more steps than the read-back ring has slots."""
    torch.manual_seed(102)
    model = SyntheticDiffuSeq(**TOY).cuda()
    data = batches(5, 8, n=40)
    dumps = []
    real_dump = logger.dumpkvs

    class Spy:
        def __getattr__(self, name):
            return getattr(logger, name)

        def dumpkvs(self):
            out = real_dump()
            dumps.append(out)
            return out

    loop = make_loop(model, data, tmp_path, learning_steps=36, log_interval=18, save_interval=1000, logger=Spy(),
                     gradient_clipping=1.0)
    loop.run_loop()
    assert loop.step == 36
    assert len(dumps) == 2                       # steps 0 and 18
    assert dumps[0]["step"] == 0 and dumps[0]["samples"] == 8 and dumps[0]["grad_norm"] > 0
    assert dumps[1]["step"] == 18 and dumps[1]["samples"] == 19 * 8
    # clipped to 1.0 every step -> the logged post-clip norm is <= 1 (+ rounding)
    assert 0 < dumps[1]["grad_norm"] <= 1.0 + 1e-5
    assert (tmp_path / "model_000036.pt").exists() and (tmp_path / "opt_000036.pt").exists()
    assert not loop._norm_inflight


def test_update_ema_standalone_and_fused_flag(tmp_path):
    a = [torch.randn(100, device="cuda"), torch.randn(7, 3, device="cuda")]
    b = [torch.randn(100, device="cuda"), torch.randn(7, 3, device="cuda")]
    want = [x.clone().mul_(0.99).add_(y, alpha=1 - 0.99) for x, y in zip(a, b)]
    update_ema(a, b, rate=0.99)
    for x, w in zip(a, want):
        assert torch.equal(x, w)
    with pytest.raises(RuntimeError):
        update_ema([torch.zeros(3)], [torch.zeros(3)], 0.9)  # no CPU implementation


def test_loss_scale_skips_on_overflow(tmp_path):
    torch.manual_seed(102)
    model = SyntheticDiffuSeq(**TOY).cuda()
    data = batches(3, 8)
    loop = make_loop(model, data, tmp_path, use_fp16=True)
    before = [p.detach().clone() for p in loop.master_params]
    bad = dict(data[0])
    bad["target"] = data[0]["target"].clone()
    bad["target"][0, 0, 0] = float("inf")
    loop.run_step(bad)
    assert loop.lg_loss_scale == 19.0 and loop.opt.step_count == 0
    for p, q in zip(loop.master_params, before):
        assert torch.equal(p, q)
    loop.run_step(data[1])
    assert loop.opt.step_count == 1 and loop.lg_loss_scale == pytest.approx(19.0 + 1e-3)
    assert any(not torch.equal(p, q) for p, q in zip(loop.master_params, before))


def test_cpu_model_is_refused(tmp_path):
    logger.configure(dir=str(tmp_path))
    with pytest.raises(RuntimeError, match="CUDA"):
        make_loop(SyntheticDiffuSeq(**TOY), batches(0, 8, n=1), tmp_path)
