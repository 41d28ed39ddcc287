"""Exploration: K1 variants (B200HNSW_K1_VARIANT / B200HNSW_K1_G) on a wave-built graph; device-resident queries,
CUDA events around the traversal kernel alone."""
import os, sys, numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from bench import synth_rows, draw_levels
N = int(sys.argv[1]); d = 128; M = 16; ef = 64; k = 10
variants = sys.argv[2].split(',')
Bs = [int(x) for x in sys.argv[3].split(',')]
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
print("built", g.Topography(), "build launches", g.last_stats["launches"], flush=True)
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream(dev)
torch.cuda.cudart().cudaProfilerStart()
ref = {}
for v in variants:
    os.environ.pop("B200HNSW_K1_VARIANT", None); os.environ.pop("B200HNSW_K1_G", None)
    if v.startswith("G"): os.environ["B200HNSW_K1_G"] = v[1:]
    elif v != "0": os.environ["B200HNSW_K1_VARIANT"] = v
    for B in Bs:
        nb = 4
        Qd = torch.from_numpy(synth_rows(nb * B, d, 43).reshape(nb, B, d)).to(dev)
        rows = torch.empty((B, k), dtype=torch.int32, device=dev)
        dists = torch.empty((B, k), dtype=torch.float32, device=dev)
        counts = torch.empty((B,), dtype=torch.int32, device=dev)
        ctr = torch.zeros(8, dtype=torch.int64, device=dev)
        ms = []
        for it in range(9):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            g.search_device(Qd[it % nb].data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(),
                            ctr.data_ptr(), stream.cuda_stream)
            b.record(stream)
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
            if it == 0:
                r0 = rows.cpu().numpy().copy()
        if B not in ref: ref[B] = r0
        ok = np.array_equal(ref[B], r0)
        c = ctr.cpu().numpy()
        best = min(ms[1:])
        print(f"variant={v} B={B} kernel_ms best={best:.3f} med={np.median(ms[1:]):.3f} qps={B/best*1e3:.3e} same={ok} "
              f"E/q={c[0]/(9*B):.2f} H/q={c[1]/(9*B):.2f}", flush=True)
