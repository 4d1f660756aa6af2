"""Synthetic DiffuSeq-shaped workloads: the shape table used by bench.py and a
small torch module with exactly those parameters, used only as a *gradient
generator* for end-to-end tests of the update path.

The reference ships no model (``models/__init__.py`` is empty,
``utils/initialization.py:18-27`` returns ``...``); shapes follow the upstream
DiffuSeq ``TransformerNetModel`` with a bert-base config:
209 tensors / 91 225 274 parameters for the defaults.
"""
import math

import torch
import torch.nn.functional as F
from torch import nn


def diffuseq_shapes(n_layers=12, d=768, ffn=3072, vocab=30522, emb=128, t_hidden=512, max_pos=512):
    """Parameter shapes in ``named_parameters()`` order."""
    shapes = [(vocab, emb), (vocab,)]                                   # word_embedding (tied), lm_head.bias
    shapes += [(t_hidden, emb), (t_hidden,), (d, t_hidden), (d,)]        # time_embed
    shapes += [(d, emb), (d,), (d, d), (d,)]                             # input_up_proj
    for _ in range(n_layers):
        shapes += [(d, d), (d,)] * 4                                     # q, k, v, attention output
        shapes += [(d,), (d,)]                                           # LayerNorm
        shapes += [(ffn, d), (ffn,), (d, ffn), (d,)]                     # intermediate, output
        shapes += [(d,), (d,)]                                           # LayerNorm
    shapes += [(max_pos, d)]                                             # position_embeddings
    shapes += [(d,), (d,)]                                               # LayerNorm
    shapes += [(d, d), (d,), (emb, d), (emb,)]                           # output_down_proj
    return shapes


def numel(shape):
    n = 1
    for s in shape:
        n *= int(s)
    return n


class _Layer(nn.Module):
    def __init__(self, d, ffn, heads):
        super().__init__()
        self.q, self.k, self.v, self.o = (nn.Linear(d, d) for _ in range(4))
        self.ln1 = nn.LayerNorm(d)
        self.up = nn.Linear(d, ffn)
        self.down = nn.Linear(ffn, d)
        self.ln2 = nn.LayerNorm(d)
        self.heads = heads

    def forward(self, x):
        b, l, d = x.shape
        split = lambda t: t.view(b, l, self.heads, d // self.heads).transpose(1, 2)
        a = F.scaled_dot_product_attention(split(self.q(x)), split(self.k(x)), split(self.v(x)))
        x = self.ln1(x + self.o(a.transpose(1, 2).reshape(b, l, d)))
        return self.ln2(x + self.down(F.gelu(self.up(x))))


class _Bias(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(n))


class SyntheticDiffuSeq(nn.Module):
    """forward(ids[B,L] int64, t[B] int64) -> [B,L,emb]; parameters match diffuseq_shapes()."""

    def __init__(self, n_layers=12, d=768, ffn=3072, vocab=30522, emb=128, t_hidden=512, max_pos=512, heads=None):
        super().__init__()
        heads = heads or max(1, d // 64)
        self.emb_dim = emb
        self.word_embedding = nn.Embedding(vocab, emb)
        self.lm_head = _Bias(vocab)  # weight tied to word_embedding, as upstream
        self.time_embed = nn.Sequential(nn.Linear(emb, t_hidden), nn.SiLU(), nn.Linear(t_hidden, d))
        self.input_up_proj = nn.Sequential(nn.Linear(emb, d), nn.Tanh(), nn.Linear(d, d))
        self.layers = nn.ModuleList(_Layer(d, ffn, heads) for _ in range(n_layers))
        self.position_embeddings = nn.Embedding(max_pos, d)
        self.LayerNorm = nn.LayerNorm(d)
        self.output_down_proj = nn.Sequential(nn.Linear(d, d), nn.Tanh(), nn.Linear(d, emb))
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.normal_(p, std=0.02)

    def _timestep_embedding(self, t):
        half = self.emb_dim // 2
        freqs = torch.exp(-math.log(10000) * torch.arange(half, device=t.device, dtype=torch.float32) / half)
        args = t[:, None].float() * freqs[None]
        return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)

    def forward(self, ids, t):
        x = self.word_embedding(ids)
        dt = x.dtype
        h = self.input_up_proj(x) + self.time_embed(self._timestep_embedding(t).to(dt))[:, None, :]
        h = h + self.position_embeddings(torch.arange(ids.shape[1], device=ids.device))[None]
        h = self.LayerNorm(h)
        for layer in self.layers:
            h = layer(h)
        out = self.output_down_proj(h)
        # tied lm_head: logits against the embedding matrix keep its bias in the graph
        logits_mean = (out @ self.word_embedding.weight[:64].t()).mean() + self.lm_head.bias.mean()
        return out + 0.0 * logits_mean
