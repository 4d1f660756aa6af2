import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pipeline_b200.engine import Engine
from distributed_pipeline_b200.flat import FlatLayout
from distributed_pipeline_b200.synthetic import diffuseq_shapes, numel
PEAK = 6538.3
numels = [numel(s) for s in diffuseq_shapes()]
n = sum(numels)
dev = torch.device("cuda", 0)
for dtype, s, K in ((torch.float32, 4, 1), (torch.bfloat16, 2, 2)):
    lay = FlatLayout.from_numels(numels, s)
    bufs = [torch.empty(lay.n_elems, dtype=dtype, device=dev).normal_(0, 0.01) for _ in range(4 + K)]
    bufs[2].zero_(); bufs[3].zero_()
    for tile in (1024, 2048, 4096, 8192, 16384, 32768, 65536):
        eng = Engine(lay, dtype, dev, [0.999] * K, tile_elems=tile)
        eng.bind(bufs[0], bufs[1], bufs[2], bufs[3], bufs[4:4 + K])
        for _ in range(3):
            eng.fused_update(1e-4, 0.0, 1)
        eng.timing(True)
        for i in range(30):
            eng.fused_update(1e-4, 0.0, i + 1)
        torch.cuda.synchronize()
        ms = eng.read_timings_sync(0)
        bpp = (7 + 2 * K) * s
        t = sum(ms) / len(ms)
        print(f"{str(dtype):15s} K={K} tile={tile:6d} {t*1e3:7.1f} us (min {min(ms)*1e3:7.1f})  {n*bpp/t/1e6:7.0f} GB/s  frac={n*bpp/t/1e6/PEAK:.3f}", flush=True)
        eng.close()
