"""Round-2 additions on one GPU: frozen tensors, the schedule sampler wired into TrainLoop, the deferred loss
log, the peer-timeout guard of the update kernels."""
import copy

import numpy as np
import pytest
import torch

from distributed_pipeline_b200 import FusedAdamW, _lib, logger
from distributed_pipeline_b200.sampler import LossSecondMomentResampler, UniformSampler
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq
from oracle.sampler_oracle import LossSecondMomentResamplerOracle
from test_gpu_trainloop import TOY, Loop, batches, make_loop

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("clip", [-1.0, 0.5])
def test_frozen_tensor_matches_torch_adamw(dtype, clip):
    """A requires_grad=False tensor: torch.optim.AdamW never touches it (no weight decay either) and
    clip_grad_norm_ leaves it out of the norm; the reference's update_ema still walks it."""
    shapes = [(300, 7), (5,), (4099,), (64, 33), (9,)]
    frozen = [False, True, False, True, False]
    torch.manual_seed(3)
    ps = [torch.nn.Parameter((torch.randn(s, device="cuda") * 0.05).to(dtype), requires_grad=not f)
          for s, f in zip(shapes, frozen)]
    ref = [torch.nn.Parameter(p.detach().clone(), requires_grad=p.requires_grad) for p in ps]
    opt = FusedAdamW(ps, lr=1e-2, weight_decay=0.1, ema_rates=[0.9], grad_mode=_lib.GRAD_CLIPPED)
    ref_opt = torch.optim.AdamW(ref, lr=1e-2, weight_decay=0.1, foreach=True)
    ref_ema = [p.detach().clone() for p in ref]
    gen = torch.Generator(device="cuda").manual_seed(5)
    for _ in range(4):
        for p, r in zip(ps, ref):
            if not p.requires_grad:
                assert p.grad is None
                continue
            g = (torch.randn(p.shape, device="cuda", generator=gen) * 0.3).to(dtype)
            p.grad.copy_(g)
            r.grad = g.clone()
        # poison what the kernels must not read as data: the gradient bytes of the frozen tensors
        for view, f in zip(opt.flat.grad_views, frozen):
            if f:
                view.fill_(float("nan"))
        if clip > 0:
            opt.clip_grad_norm(clip)
            torch.nn.utils.clip_grad_norm_(ref, clip)
        opt.step()
        ref_opt.step()
        for e, r in zip(ref_ema, ref):
            e.mul_(0.9).add_(r.detach(), alpha=1 - 0.9)
    torch.cuda.synchronize()
    res = opt.engine.read_results_sync()
    assert res[_lib.R_NONFINITE] == 0.0
    for p, r, f, e, re_ in zip(ps, ref, frozen, opt.flat.ema_views[0], ref_ema):
        if f:
            assert torch.equal(p, r), "a frozen tensor moved"
        tol = 1e-6 if dtype == torch.float32 else 2 ** -7
        assert torch.allclose(p.float(), r.float(), rtol=tol, atol=tol * 0.05)
        assert torch.allclose(e.float(), re_.float(), rtol=tol, atol=tol * 0.05)
    # checkpoints: no state entry for the frozen tensors, as with torch; and they load back
    sd, ref_sd = opt.state_dict(), ref_opt.state_dict()
    assert sorted(sd["state"].keys()) == sorted(ref_sd["state"].keys()) == [0, 2, 4]
    opt.load_state_dict(ref_sd)
    assert opt.step_count == 4


def test_schedule_sampler_is_drawn_fed_and_applied(tmp_path):
    T = 40
    torch.manual_seed(102)
    model = SyntheticDiffuSeq(**TOY).cuda()
    data = batches(3, 8)

    seen = []

    class Recording(Loop):
        def compute_losses(self, micro_batch):
            seen.append(micro_batch["t"].clone())
            out = self.ddp_model(micro_batch["ids"], micro_batch["t"])
            return {"loss": ((out - micro_batch["target"]) ** 2).mean(dim=(1, 2))}

        backward_from_losses = staticmethod(lambda losses: losses["loss"].mean().backward())

    logger.configure(dir=str(tmp_path))
    loop = Recording(model=model, data=iter(data), batch_size=8, microbatch=4, lr=1e-3, ema_rate="0.999",
                     log_interval=10 ** 9, save_interval=10 ** 9, resume_checkpoint="", schedule_sampler="lossaware",
                     diffusion_steps=T)
    assert isinstance(loop.schedule_sampler, LossSecondMomentResampler)
    ref = LossSecondMomentResamplerOracle(T, 10)
    np.random.seed(11)
    for step in range(3):
        loop.run_step(data[step])
        loop.step += 1
    # the same draws from the same RNG state, in the same order: 2 microbatches per step
    np.random.seed(11)
    for i, t in enumerate(seen):
        want_t, _w = ref.sample(4)
        assert np.array_equal(t.cpu().numpy(), want_t), "sampled timesteps differ for the same RNG state"
    hist, counts = loop.schedule_sampler.state()
    assert int(counts.sum().item()) == 3 * 8          # every (t, loss) record of every microbatch arrived
    # forward_only (eval) draws but does not feed the history
    loop.forward_only(data[4])
    assert int(loop.schedule_sampler.state()[1].sum().item()) == 3 * 8
    # by object, uniform: weights are all one, nothing to feed
    loop2 = make_loop(SyntheticDiffuSeq(**TOY).cuda(), data, tmp_path, schedule_sampler=UniformSampler(T))
    loop2.run_step(data[0])
    with pytest.raises(ValueError):
        make_loop(SyntheticDiffuSeq(**TOY).cuda(), data, tmp_path, schedule_sampler="uniform")  # no diffusion_steps


def test_deferred_loss_log_equals_per_sample_logging(tmp_path):
    """log_loss_dict_deferred logs what upstream's log_loss_dict logs (mean per key, running mean per
    timestep quartile) without a host sync per value."""
    torch.manual_seed(0)
    loop = make_loop(SyntheticDiffuSeq(**TOY).cuda(), batches(1, 8), tmp_path)
    T = 100
    gen = torch.Generator().manual_seed(4)
    want = logger.Logger(None)
    for _ in range(5):
        losses = {"loss": torch.rand(6, generator=gen).cuda(), "mse": torch.rand(6, generator=gen).cuda()}
        ts = torch.randint(0, T, (6,), generator=gen).cuda()
        loop.log_loss_dict_deferred(losses, ts, T)
        for key, values in losses.items():                      # upstream log_loss_dict, verbatim semantics
            want.logkv_mean(key, values.mean().item())
            for sub_t, sub_loss in zip(ts.cpu().numpy(), values.detach().cpu().numpy()):
                want.logkv_mean(f"{key}_q{int(4 * sub_t / T)}", sub_loss)
    loop.flush_logs()
    got = logger.get_current().name2val
    for k, v in want.name2val.items():
        assert abs(got[k] - v) <= 1e-6 * max(abs(v), 1e-6), (k, got[k], v)


def test_update_is_a_noop_after_a_peer_timeout():
    """Once DPK_R_COMM_TIMEOUT is raised the gradients are unreduced: the fused pass must not touch the state."""
    ps = [torch.nn.Parameter(torch.randn(5000, device="cuda"))]
    opt = FusedAdamW(ps, lr=1e-2, weight_decay=0.1, ema_rates=[0.9])
    ps[0].grad.normal_()
    before = [t.clone() for t in (opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.flat.emas[0])]
    opt.engine.results[_lib.R_COMM_TIMEOUT] = 1.0
    opt.step()
    res = opt.engine.read_results_sync()
    assert res[_lib.R_COMM_TIMEOUT] == 1.0 and res[_lib.R_SKIPPED] == 1.0
    for a, b in zip(before, (opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.flat.emas[0])):
        assert torch.equal(a, b)
    opt.engine.results[_lib.R_COMM_TIMEOUT] = 0.0
    opt.step()
    torch.cuda.synchronize()
    assert not torch.equal(before[0], opt.flat.params)
