"""Runs a few launches of one kernel configuration (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pipeline_b200.engine import Engine
from distributed_pipeline_b200.flat import FlatLayout
from distributed_pipeline_b200.synthetic import diffuseq_shapes, numel
dtype = torch.bfloat16 if sys.argv[1] == "bf16" else torch.float32
K = int(sys.argv[2]); gm = int(sys.argv[3]); iters = int(sys.argv[4]) if len(sys.argv) > 4 else 6
s = 2 if dtype == torch.bfloat16 else 4
numels = [numel(x) for x in diffuseq_shapes()]
lay = FlatLayout.from_numels(numels, s)
dev = torch.device("cuda", 0)
bufs = [torch.empty(lay.n_elems, dtype=dtype, device=dev).normal_(0, 0.01) for _ in range(4 + K)]
bufs[2].zero_(); bufs[3].zero_()
eng = Engine(lay, dtype, dev, [0.999] * K)
eng.bind(bufs[0], bufs[1], bufs[2], bufs[3], bufs[4:4 + K])
for i in range(iters):
    eng.grad_norm(1.0, 1.0)
    eng.fused_update(1e-4, 0.0, i + 1, use_clip_coef=True, grad_mode=gm)
torch.cuda.synchronize()
print("ok", eng.read_results_sync()[:3])
