"""Exploration: K1x (extended mode, traverse_ext.cuh) on a wave-built graph; device-resident queries, CUDA events around
the kernel alone.  usage: k1x_probe.py N dist ef[,ef..] B [build_mode]"""
import sys, numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from hnsw_b200 import _lib
from benchlib.data import gen, draw_levels
N = int(sys.argv[1]); dist = sys.argv[2]; efs = [int(x) for x in sys.argv[3].split(',')]; B = int(sys.argv[4])
bmode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
d, M, k = 128, 16, 10
X = gen(dist, N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=64, Distance=hnsw_b200.EuclideanDistance, capacity=N)
if bmode:
    g.set_build_mode(bmode)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=16384)
print("built", g.Topography(), "kernel_ms", g.last_stats["kernel_ms"], flush=True)
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream(dev)
Q = gen(dist, 2 * B, d, 43)
Qd = torch.from_numpy(Q.reshape(2, B, d)).to(dev)
xd = torch.from_numpy(X).to(dev)
qq = Qd[0][:1000]
d2 = (xd * xd).sum(1)[None, :] - 2.0 * (qq @ xd.T)
truth = d2.topk(k, dim=1, largest=False).indices.cpu().numpy()
del d2
rows = torch.empty((B, k), dtype=torch.int32, device=dev)
dists = torch.empty((B, k), dtype=torch.float32, device=dev)
counts = torch.empty((B,), dtype=torch.int32, device=dev)
torch.cuda.cudart().cudaProfilerStart()
for ef in efs:
    ctr = torch.zeros(8, dtype=torch.int64, device=dev)
    ms = []
    for it in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        g.search_device(Qd[it % 2].data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(),
                        ctr.data_ptr(), stream.cuda_stream, ef=ef, mode=_lib.MODE_EXTENDED)
        b.record(stream)
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
        if it == 0:
            r0 = rows.cpu().numpy()
    c = ctr.cpu().numpy()
    rec = np.mean([len(set(r0[i]) & set(truth[i])) / k for i in range(1000)])
    best = min(ms[1:])
    Eq = c[0] / (5 * B)
    print(f"ef={ef} B={B} kernel_ms best={best:.3f} qps={B/best*1e3:.3e} E/q={Eq:.1f} H/q={c[1]/(5*B):.1f} recall@10={rec:.4f} "
          f"alg GB/s={Eq*B*512/best/1e6:.0f} err={c[4]}", flush=True)
