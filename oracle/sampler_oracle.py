"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the loss-aware timestep
sampler that ``north_star`` names.

PARITY UNPINNED: the reference repository contains no sampler and no
copy of the upstream (improved-diffusion ``resample.py`` /
DiffuSeq ``step_sample.py``) exists on this machine; the semantics below follow
SURVEY Appendix B2, written from the published algorithm:

  LossSecondMomentResampler(T, history_per_term=10, uniform_prob=0.001)
    _loss_history float64[T, H], _loss_counts int[T]
    update_with_all_losses(ts, losses): sequential; full row -> shift left, append
    weights(): ones until every count == H, then sqrt(mean(hist**2, -1)),
               /= sum, *= 1-uniform_prob, += uniform_prob/T
  ScheduleSampler.sample(batch): p = w/sum(w); idx = np.random.choice(T, batch, p=p);
               weights = 1/(T*p[idx])
  LossAwareSampler.update_with_local_losses: all_gather (rank-major, batch order)
"""
import numpy as np


class LossSecondMomentResamplerOracle:
    def __init__(self, n_timesteps, history_per_term=10, uniform_prob=0.001):
        self.T = n_timesteps
        self.history_per_term = history_per_term
        self.uniform_prob = uniform_prob
        self._loss_history = np.zeros([n_timesteps, history_per_term], dtype=np.float64)
        self._loss_counts = np.zeros([n_timesteps], dtype=np.int64)

    def _warmed_up(self):
        return (self._loss_counts == self.history_per_term).all()

    def weights(self):
        if not self._warmed_up():
            return np.ones([self.T], dtype=np.float64)
        weights = np.sqrt(np.mean(self._loss_history ** 2, axis=-1))
        weights /= np.sum(weights)
        weights *= 1 - self.uniform_prob
        weights += self.uniform_prob / len(weights)
        return weights

    def update_with_all_losses(self, ts, losses):
        for t, loss in zip(ts, losses):
            t = int(t)
            if self._loss_counts[t] == self.history_per_term:
                self._loss_history[t, :-1] = self._loss_history[t, 1:]
                self._loss_history[t, -1] = loss
            else:
                self._loss_history[t, self._loss_counts[t]] = loss
                self._loss_counts[t] += 1

    def update_with_gathered(self, per_rank_ts, per_rank_losses):
        """per-rank lists in rank order; losses are fp32 values (``.item()``)."""
        ts = [int(t) for r in per_rank_ts for t in r]
        losses = [float(np.float32(x)) for r in per_rank_losses for x in r]
        self.update_with_all_losses(ts, losses)

    def sample(self, batch_size):
        """Uses numpy's global legacy RandomState, like upstream."""
        w = self.weights()
        p = w / np.sum(w)
        indices = np.random.choice(len(p), size=(batch_size,), p=p)
        weights = (1 / (len(p) * p[indices])).astype(np.float32)
        return indices.astype(np.int64), weights
