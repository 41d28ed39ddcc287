"""Exploration: K1 variants (B200HNSW_K1_VARIANT) with 1/2/3 ten-thousand-query launches in flight on a 1M x 128 graph."""
import os, sys, time
import numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from bench import synth_rows, draw_levels
N, d, B, k, ef = 1_000_000, 128, 10_000, 10, 64
variants = [int(x) for x in sys.argv[1].split(',')] if len(sys.argv) > 1 else [7, 9, 12]
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=16, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
nb = 16
dev = torch.device('cuda', 0)
Qd = torch.from_numpy(synth_rows(nb * B, d, 43).reshape(nb, B, d)).to(dev)
steps = 60
ref = None
for v in variants:
    if v < 0: os.environ.pop('B200HNSW_K1_VARIANT', None)
    else: os.environ['B200HNSW_K1_VARIANT'] = str(v)
    line = f"variant {v}:"
    for S in (1, 2, 3):
        st = [torch.cuda.Stream(dev) for _ in range(S)]
        bufs = [(torch.empty((B, k), dtype=torch.int32, device=dev), torch.empty((B, k), dtype=torch.float32, device=dev),
                 torch.empty((B,), dtype=torch.int32, device=dev)) for _ in range(S)]
        def run(n):
            for i in range(n):
                s = i % S; r, dd, c = bufs[s]
                g.search_device(Qd[i % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr(), 0, st[s].cuda_stream)
        run(6); torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            t0 = time.perf_counter(); run(steps); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
        line += f"  S={S}: {best / steps * 1e3:.4f} ms/step {steps * B / best / 1e6:.1f} Mq/s"
        if S == 1:
            r = bufs[0][0].cpu().numpy().copy()
            if ref is None: ref = r
            line += "  rows==first " + str(bool(np.array_equal(r, ref)))
    print(line, flush=True)
