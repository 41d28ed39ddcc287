"""Exploration: K1 kernel time against batch size and group width (how long one wave of traversals takes)."""
import os, sys, numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from bench import synth_rows, draw_levels
N = int(sys.argv[1]); d = 128; M = 16; ef = 64; k = 10
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
dev = torch.device("cuda", 0); stream = torch.cuda.current_stream(dev)
Qd = torch.from_numpy(synth_rows(4 * 10000, d, 43).reshape(4, 10000, d)).to(dev)
rows = torch.empty((10000, k), dtype=torch.int32, device=dev); dists = torch.empty((10000, k), dtype=torch.float32, device=dev)
counts = torch.empty((10000,), dtype=torch.int32, device=dev)
for G in ("8", "16", "32"):
    os.environ["B200HNSW_K1_G"] = G
    for B in (312, 625, 1250, 2500, 5000, 10000):
        ms = []
        for it in range(8):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            g.search_device(Qd[it % 4].data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(), 0, stream.cuda_stream)
            b.record(stream); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
        print(f"G={G} B={B} kernel_ms best={min(ms[1:]):.3f} med={np.median(ms[1:]):.3f}", flush=True)
