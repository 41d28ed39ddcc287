"""Exploration: wave-batched Add rate on the device and K1 timing on the resulting graph."""
import sys, time, numpy as np
sys.path.insert(0, '.')
import os
os.environ.setdefault('B200HNSW_E2E_CHUNKS', '1')  # kernel_ms = the traversal kernel alone
import hnsw_b200
from bench import synth_rows, draw_levels
N = int(sys.argv[1]); d = int(sys.argv[2]); M = int(sys.argv[3]); ef = int(sys.argv[4]); wm = int(sys.argv[5])
metric = sys.argv[7] if len(sys.argv) > 7 else 'l2'
Bs = [int(x) for x in sys.argv[6].split(',')]
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance if metric == 'l2' else hnsw_b200.CosineDistance, capacity=N)
t = time.time(); g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=wm); dt = time.time() - t
st = g.last_stats
print(f"wave add N={N} wall {dt:.2f}s (C call {st.get('call_s', 0):.2f}s) kernel {st['kernel_ms']/1e3:.2f}s -> {N/(st['kernel_ms']/1e3):.0f} vec/s launches={st['launches']} "
      f"E={st['dist_evals']/N:.0f} evict={st['evict_evals']/N:.0f} repl={st['replenish_evals']/N:.0f} evictions={st['evictions']/N:.1f}", flush=True)
print(g.Topography())
for B in Bs:
    Q = synth_rows(B, d, 43)
    for it in range(3):
        rows, dists, cnt = g.search_rows(Q, 10, ef)
        st = g.last_stats
        byts = st['dist_evals'] * 4 * d + st['nbr_reads'] * 4
        print(f"B={B} it={it} kernel_ms={st['kernel_ms']:.3f} qps={B/st['kernel_ms']*1e3:.3e} E/q={st['dist_evals']/B:.1f} H/q={st['expansions']/B:.1f} GB/s={byts/st['kernel_ms']/1e6:.1f}", flush=True)
