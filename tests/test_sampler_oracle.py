"""Properties of the numpy restatement of the upstream loss-aware sampler
(parity unpinned by the reference: it has no sampler) and of the
numpy facts the device kernels rely on (summation order, SURVEY a-15)."""
import numpy as np

from oracle.sampler_oracle import LossSecondMomentResamplerOracle


def np_pairwise(a):
    """numpy's pairwise_sum, restated: what dpk_sampler.cu implements on the device."""
    n = len(a)
    if n < 8:
        r = 0.0
        for x in a:
            r += x
        return r
    if n <= 128:
        r = [a[j] for j in range(8)]
        i = 8
        while i < n - (n % 8):
            for j in range(8):
                r[j] += a[i + j]
            i += 8
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
        while i < n:
            res += a[i]
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return np_pairwise(a[:n2]) + np_pairwise(a[n2:])


def test_device_summation_order_is_numpys():
    rng = np.random.RandomState(0)
    for n in (1, 3, 7, 8, 9, 10, 17, 120, 128, 129, 300, 1000, 2000, 4097):
        a = rng.rand(n) * rng.choice([1e-3, 1.0, 1e3], size=n)
        assert np_pairwise(list(a)) == float(np.sum(a)), n
    h = rng.rand(50, 10)
    want = np.mean(h ** 2, axis=-1)
    got = np.array([np_pairwise(list(r * r)) / 10 for r in h])
    assert np.array_equal(want, got)
    p = rng.rand(2000)
    p /= p.sum()
    np.random.seed(3)
    a = np.random.choice(2000, size=64, p=p)
    np.random.seed(3)
    cdf = p.cumsum()
    cdf /= cdf[-1]
    b = cdf.searchsorted(np.random.random_sample(64), side="right")
    assert np.array_equal(a, b)


def test_warmup_shift_and_weights():
    s = LossSecondMomentResamplerOracle(5, history_per_term=3, uniform_prob=0.1)
    assert np.array_equal(s.weights(), np.ones(5))
    s.update_with_all_losses([0, 0, 0, 0], [1.0, 2.0, 3.0, 4.0])   # 4th shifts the row
    assert list(s._loss_history[0]) == [2.0, 3.0, 4.0] and s._loss_counts[0] == 3
    assert np.array_equal(s.weights(), np.ones(5))                 # others still cold
    for t in range(1, 5):
        s.update_with_all_losses([t] * 3, [1.0] * 3)
    w = s.weights()
    assert abs(w.sum() - 1.0) < 1e-12 and w[0] > w[1] and np.allclose(w[1:], w[1])
    assert w.min() >= 0.1 / 5
    # gathered order is rank-major: later ranks overwrite earlier ones on a full row
    s.update_with_gathered([[1], [1]], [[5.0], [7.0]])
    assert list(s._loss_history[1]) == [1.0, 5.0, 7.0]
    np.random.seed(0)
    idx, wt = s.sample(1000)
    assert idx.dtype == np.int64 and wt.dtype == np.float32 and idx.min() >= 0 and idx.max() < 5
    p = s.weights() / s.weights().sum()
    assert np.allclose(wt, (1 / (5 * p[idx])).astype(np.float32))
