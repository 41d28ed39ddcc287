"""Exploration: run-to-run spread of the 4-caller end-to-end measurement inside one process (with / without the NVML sampler)."""
import os, sys, threading, time
import numpy as np
sys.path.insert(0, '.')
import torch
import hnsw_b200
from bench import synth_rows, draw_levels, ClockSampler
N, d, B, k, ef = 1_000_000, 128, 10_000, 10, 64
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=16, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
nb = 16
Qh = torch.from_numpy(synth_rows(nb * B, d, 43).reshape(nb, B, d)).pin_memory()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
callers = 4
def measure(tag):
    outs = [(torch.empty((B, k), dtype=torch.int32).pin_memory(), torch.empty((B, k), dtype=torch.float32).pin_memory(),
             torch.empty((B,), dtype=torch.int32).pin_memory()) for _ in range(callers)]
    gate = threading.Barrier(callers + 1)
    lat = [[] for _ in range(callers)]
    def caller(t):
        r, dd, c = outs[t]
        for i in range(3):
            g.search_host_ptrs(Qh[(t + i) % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr())
        gate.wait()
        for i in range(t, steps, callers):
            t0 = time.perf_counter()
            g.search_host_ptrs(Qh[i % nb].data_ptr(), B, k, r.data_ptr(), dd.data_ptr(), c.data_ptr())
            lat[t].append(time.perf_counter() - t0)
        gate.wait()
    th = [threading.Thread(target=caller, args=(t,)) for t in range(callers)]
    for x in th: x.start()
    gate.wait(); t0 = time.perf_counter(); gate.wait(); el = time.perf_counter() - t0
    for x in th: x.join()
    al = np.concatenate([np.array(l) for l in lat]) * 1e3
    print(f"{tag}: {el / steps * 1e3:.3f} ms/step {steps * B / el / 1e6:.1f} Mq/s   call ms: med {np.median(al):.3f} max {al.max():.3f} first {[round(l[0]*1e3,2) for l in lat]}", flush=True)
for rep in range(6): measure(f"no sampler rep{rep}")
s = ClockSampler(0); s.start()
for rep in range(6): measure(f"sampler 2ms rep{rep}")
s.stop()
