"""Times the fused-update and sum-of-squares kernels alone (CUDA events inside
the library) over dtype x K x grad mode on DiffuSeq-base shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pipeline_b200 import _lib
from distributed_pipeline_b200.engine import Engine
from distributed_pipeline_b200.flat import FlatLayout
from distributed_pipeline_b200.synthetic import diffuseq_shapes, numel

PEAK = 6538.3
numels = [numel(s) for s in diffuseq_shapes()]
n = sum(numels)
dev = torch.device("cuda", 0)
tile = int(sys.argv[1]) if len(sys.argv) > 1 else 0
for dtype, s in ((torch.float32, 4), (torch.bfloat16, 2)):
    lay = FlatLayout.from_numels(numels, s)
    bufs = [torch.empty(lay.n_elems, dtype=dtype, device=dev).normal_(0, 0.01) for _ in range(8)]
    bufs[2].zero_(); bufs[3].zero_()
    for K in (0, 1, 2, 4):
        eng = Engine(lay, dtype, dev, [0.999] * K, tile_elems=tile)
        eng.bind(bufs[0], bufs[1], bufs[2], bufs[3], bufs[4:4 + K])
        for gm, gname in ((0, "keep"), (2, "zero")):
            for _ in range(3):
                eng.fused_update(1e-4, 0.0, 1, grad_mode=gm)
            eng.timing(True)
            for i in range(20):
                eng.fused_update(1e-4, 0.0, i + 1, grad_mode=gm)
            torch.cuda.synchronize()
            ms = eng.read_timings_sync(0)
            eng.timing(False)
            bpp = (7 + 2 * K) * s + (s if gm else 0)
            t = sum(ms) / len(ms)
            print(f"{str(dtype):15s} K={K} grad={gname:5s} {t*1e3:7.1f} us  {n*bpp/t/1e6:7.0f} GB/s  frac={n*bpp/t/1e6/PEAK:.3f}  "
                  f"{n/t/1e6:6.1f} Gelem/s", flush=True)
            bufs[1].normal_(0, 0.01)
        eng.timing(True)
        for _ in range(10):
            eng.grad_norm(1.0, 1.0)
        torch.cuda.synchronize()
        ms = eng.read_timings_sync(1)[3:]
        t = sum(ms) / len(ms)
        if K == 0:
            print(f"{str(dtype):15s} sumsq            {t*1e3:7.1f} us  {n*s/t/1e6:7.0f} GB/s  frac={n*s/t/1e6/PEAK:.3f}", flush=True)
            eng.set_option(_lib.OPT_NORM_STAGING, 1)
            for _ in range(10):
                eng.grad_norm(1.0, 1.0)
            torch.cuda.synchronize()
            ms = eng.read_timings_sync(1)[3:]
            t = sum(ms) / len(ms)
            print(f"{str(dtype):15s} sumsq (TMA)      {t*1e3:7.1f} us  {n*s/t/1e6:7.0f} GB/s  frac={n*s/t/1e6/PEAK:.3f}", flush=True)
            eng.set_option(_lib.OPT_NORM_STAGING, 0)
        eng.close()
