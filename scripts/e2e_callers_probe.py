"""Exploration: end-to-end rate of the synchronous host-pointer call (b200hnsw_search_batch) with 1..4 concurrent host
callers and 1/2/4 in-call chunks, plus the device-pointer call from 2 streams.  1M x 128 L2 graph, 10k-query batches."""
import os, sys, threading, time
import numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from bench import synth_rows, draw_levels
N, d, B, k, ef = 1_000_000, 128, 10_000, 10, 64
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=16, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
nb = 16
Qh = torch.from_numpy(synth_rows(nb * B, d, 43).reshape(nb, B, d)).pin_memory()
print("CUDA_DEVICE_MAX_CONNECTIONS =", os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), flush=True)
steps = 60
for chunks, nodirect in ((1, 0), (2, 0), (2, 1), (4, 0)):
    os.environ['B200HNSW_E2E_CHUNKS'] = str(chunks)
    if nodirect: os.environ['B200HNSW_NO_DIRECT_OUT'] = '1'
    else: os.environ.pop('B200HNSW_NO_DIRECT_OUT', None)
    for callers in (1, 2, 3, 4, 6):
        outs = [(torch.empty((B, k), dtype=torch.int32).pin_memory(), torch.empty((B, k), dtype=torch.float32).pin_memory(),
                 torch.empty((B,), dtype=torch.int32).pin_memory()) for _ in range(callers)]
        gate = threading.Barrier(callers + 1)
        def caller(t):
            r, dd, c = outs[t]
            for i in range(3):
                g.search_host_ptrs(Qh[(t + i) % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr())
            gate.wait()
            for i in range(t, steps, callers):
                g.search_host_ptrs(Qh[i % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr())
            gate.wait()
        th = [threading.Thread(target=caller, args=(t,)) for t in range(callers)]
        for x in th: x.start()
        gate.wait(); t0 = time.perf_counter(); gate.wait(); el = time.perf_counter() - t0
        for x in th: x.join()
        print(f"chunks={chunks} direct_out={1-nodirect} callers={callers}: {el / steps * 1e3:.3f} ms/step  {steps * B / el / 1e6:.1f} M q/s", flush=True)
# device-pointer call from S streams (queries resident): what the kernels alone sustain when calls overlap
dev = torch.device('cuda', 0)
Qd = Qh.to(dev)
for S in (1, 2, 3):
    st = [torch.cuda.Stream(dev) for _ in range(S)]
    bufs = [(torch.empty((B, k), dtype=torch.int32, device=dev), torch.empty((B, k), dtype=torch.float32, device=dev),
             torch.empty((B,), dtype=torch.int32, device=dev)) for _ in range(S)]
    def run(n):
        for i in range(n):
            s = i % S; r, dd, c = bufs[s]
            g.search_device(Qd[i % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr(), 0, st[s].cuda_stream)
    run(6); torch.cuda.synchronize()
    t0 = time.perf_counter(); run(steps); torch.cuda.synchronize(); el = time.perf_counter() - t0
    print(f"device-pointer, {S} streams: {el / steps * 1e3:.3f} ms/step  {steps * B / el / 1e6:.1f} M q/s", flush=True)
