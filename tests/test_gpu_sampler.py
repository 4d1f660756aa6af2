"""Loss-aware sampler on the device vs the numpy restatement of the upstream
algorithm (oracle/sampler_oracle.py; parity unpinned by the reference, F1):
history, counts and weights bit for bit, hence identical np.random draws."""
import numpy as np
import pytest
import torch

from distributed_pipeline_b200.sampler import (LossSecondMomentResampler, UniformSampler,
                                                create_named_schedule_sampler)
from oracle.sampler_oracle import LossSecondMomentResamplerOracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("T,H", [(2000, 10), (37, 10), (300, 3), (130, 17)])
def test_history_weights_and_draws_bit_exact(T, H):
    rng = np.random.RandomState(5)
    dev = LossSecondMomentResampler(T, history_per_term=H, uniform_prob=0.001, max_local_batch=256)
    ref = LossSecondMomentResamplerOracle(T, H, 0.001)
    n_rounds = (T * (H + 2)) // 200 + 3
    for r in range(n_rounds):
        B = int(rng.randint(1, 257))
        if r % 2 == 0:  # sweep so that every timestep warms up; repeats inside a batch on purpose
            ts = (np.arange(B) * 7 + r * 131) % T
            ts[: B // 4] = ts[0]
        else:
            ts = rng.randint(0, T, size=B)
        losses = (rng.rand(B).astype(np.float32) * 3 + 1e-3).astype(np.float32)
        dev.update_with_local_losses(torch.from_numpy(ts).cuda(), torch.from_numpy(losses).cuda())
        ref.update_with_gathered([ts], [losses])
        if r % 5 == 0 or r == n_rounds - 1:
            hist, counts = dev.state()
            assert np.array_equal(counts.cpu().numpy().astype(np.int64), ref._loss_counts)
            assert np.array_equal(hist.cpu().numpy().view(np.uint64), ref._loss_history.view(np.uint64))
            w = dev.weights()
            assert np.array_equal(w.view(np.uint64), ref.weights().view(np.uint64)), f"weights differ at round {r}"
    # keep feeding until fully warm, then compare the draws
    full = np.arange(T)
    for _ in range(H):
        for s in range(0, T, 256):
            ts = full[s:s + 256]
            losses = (rng.rand(len(ts)).astype(np.float32) + 0.5).astype(np.float32)
            dev.update_with_local_losses(torch.from_numpy(ts).cuda(), torch.from_numpy(losses).cuda())
            ref.update_with_gathered([ts], [losses])
    assert ref._warmed_up()
    w = dev.weights()
    assert np.array_equal(w.view(np.uint64), ref.weights().view(np.uint64))
    assert not np.all(w == 1.0)
    np.random.seed(1234)
    idx_ref, wt_ref = ref.sample(64)
    np.random.seed(1234)
    idx, wt = dev.sample(64, "cuda")
    assert np.array_equal(idx.cpu().numpy(), idx_ref)          # bit-exact sampled timesteps
    assert np.array_equal(wt.cpu().numpy().view(np.uint32), wt_ref.view(np.uint32))


def test_factory_and_limits():
    assert isinstance(create_named_schedule_sampler("uniform", 10), UniformSampler)
    s = create_named_schedule_sampler("lossaware", 50, max_local_batch=8)
    from distributed_pipeline_b200 import DpkError, _lib
    with pytest.raises(DpkError) as e:
        s.update_with_local_losses(torch.zeros(9, dtype=torch.long).cuda(), torch.zeros(9).cuda())
    assert e.value.code == _lib.E_LIMIT
    s.update_with_local_losses(torch.zeros(0, dtype=torch.long).cuda(), torch.zeros(0).cuda())  # empty batch
    assert np.all(s.weights() == 1.0)
    with pytest.raises(NotImplementedError):
        create_named_schedule_sampler("nope", 10)
