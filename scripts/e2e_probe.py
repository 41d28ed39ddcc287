"""Exploration: where the host-pointer search call spends its time (wall vs device pipeline time)."""
import os, sys, time, numpy as np, torch
sys.path.insert(0, '.')
import hnsw_b200
from bench import synth_rows, draw_levels
N = int(sys.argv[1]); B = int(sys.argv[2]); d = 128; k = 10
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=16, Ml=0.25, EfSearch=64, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
Qh = torch.from_numpy(synth_rows(8 * B, d, 43).reshape(8, B, d)).pin_memory()
hr = torch.empty((B, k), dtype=torch.int32).pin_memory(); hd = torch.empty((B, k), dtype=torch.float32).pin_memory()
hc = torch.empty((B,), dtype=torch.int32).pin_memory()
for chunks in sys.argv[3].split(','):
    os.environ['B200HNSW_E2E_CHUNKS'] = chunks
    wall, dev = [], []
    for i in range(40):
        t = time.perf_counter()
        st = g.search_host_ptrs(Qh[i % 8].data_ptr(), B, k, hr.data_ptr(), hd.data_ptr(), hc.data_ptr())
        wall.append(time.perf_counter() - t); dev.append(st['kernel_ms'])
    print(f"chunks={chunks} wall_ms med={np.median(wall[5:])*1e3:.3f} min={min(wall[5:])*1e3:.3f} device_pipeline_ms med={np.median(dev[5:]):.3f} launches={st['launches']}", flush=True)
# raw copy timings
dq = torch.empty((B, d), device='cuda')
for n in (B, B // 4):
    torch.cuda.synchronize(); t = time.perf_counter()
    for i in range(20): dq[:n].copy_(Qh[i % 8][:n], non_blocking=True)
    torch.cuda.synchronize(); print(f"H2D {n*d*4/1e6:.2f} MB: {(time.perf_counter()-t)/20*1e6:.1f} us each")
