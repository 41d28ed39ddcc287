import sys, time, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from oracle import binding as ob  # CPU study with the oracle only; not a product path
N=int(sys.argv[1]); d=128; nq=500
kind=sys.argv[2]
rng=np.random.default_rng(42)
if kind=='clustered':
    cent=rng.uniform(-1,1,(100,d)).astype(np.float32)
    X=(cent[rng.integers(0,100,N)]+rng.normal(0,0.1,(N,d))).astype(np.float32)
    rq=np.random.default_rng(43)
    Q=(cent[rq.integers(0,100,nq)]+rq.normal(0,0.1,(nq,d))).astype(np.float32)
else:
    X=rng.uniform(-1,1,(N,d)).astype(np.float32)
    Q=np.random.default_rng(43).uniform(-1,1,(nq,d)).astype(np.float32)
lv=np.random.default_rng(44)
nl=ob.max_level(0.25,N)
levels=np.minimum(np.floor(-np.log(1-lv.random(N))/np.log(4)).astype(np.int32), nl-1)
g=ob.Graph(d,16,ob.EUCLIDEAN)
if len(sys.argv)>3 and sys.argv[3]=='extbuild':
    g.set_build_mode(ob.MODE_EXTENDED)  # NOT the reference: textbook search inside Add
t=time.time()
g.add_waves(np.arange(N,dtype=np.int64),X,levels,64,wave_max=4096,n_threads=8)
print('build',time.time()-t,'layers',g.topography())
# ground truth
d2=(Q**2).sum(1)[:,None]-2*Q@X.T+(X**2).sum(1)[None]
gt=np.argsort(d2,1)[:,:10]
for mode,efs in ((0,[64]),(1,[64,128,256,512,1024])):
    for ef in efs:
        g.reset_counters()
        t=time.time()
        r=g.batch_search(Q,10,ef,mode=mode,n_threads=8)
        dt=time.time()-t
        rows=r[0]
        rec=np.mean([len(set(rows[i][:10])&set(gt[i]))/10 for i in range(nq)])
        print('mode',mode,'ef',ef,'recall',rec,'evals/q',g.counters().as_dict() if False else '', 'qps',nq/dt)
