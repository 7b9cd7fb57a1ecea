import sys, random, time
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
from vectorscan_b200 import litdb, gpu
import refrt, litgen
R = refrt.RefRuntime()
print(gpu.lib().hs_gpu_version())
rng = random.Random(7)
fails=0; total=0
for n in (1,3,20,48,100,1200):
    for alpha in (b"abcdefgh", bytes(range(32,127)), bytes(range(256))):
        for tl in (0, 5, 17, 600, 40000):
            lits,flags,ids,text = litgen.random_case(rng, None, n, alpha, tl)
            engines=[(litdb.ENGINE_AUTO,{})]
            if len(lits)>1:
                ml=min(len(l) for l in lits)
                for d,s_ in ((9,1),(13,2),(15,1),(12,4)):
                    if s_<=min(ml,8): engines.append((litdb.ENGINE_FDR,dict(fdr_domain=d,fdr_stride=s_)))
            if 1<len(lits)<=48:
                for m in (1,2,3,4): engines.append((litdb.ENGINE_TEDDY,dict(teddy_masks=m)))
            for eng,kw in engines:
                try: db = litdb.compile_lit_multi(lits, flags, ids, engine=eng, **kw)
                except litdb.CompileError: continue
                rdb = R.deserialize(db.serialize()); sc = R.alloc_scratch(rdb)
                rv,exp = R.scan(rdb, sc, text)
                exp = sorted(exp)
                gs = gpu.alloc_scratch(db)
                got=[]
                rv2 = gpu.scan(db, text, gs, lambda i,f,t,fl,c: got.append((i,t)) or 0)
                total+=1
                if rv2!=0 or sorted(got)!=exp:
                    fails+=1
                    if fails<8:
                        g=set(got); e=set(exp)
                        print("MISMATCH n=%d alpha=%d eng=%s %s tl=%d rv=%d got=%d exp=%d"%(len(lits),len(alpha),eng,kw,tl,rv2,len(got),len(exp)),"extra",sorted(g-e)[:3],"missing",sorted(e-g)[:3])
                gs.close(); R.L.hs_free_scratch(sc); R.L.hs_free_database(rdb)
print("total",total,"fails",fails)
# batch test
lits,flags,ids,_ = litgen.random_case(rng, None, 500, bytes(range(32,127)), 10)
db = litdb.compile_lit_multi(lits, flags, ids)
bufs=[]
for k in range(3000):
    _,_,_,t = litgen.random_case(random.Random(k), lits, 1, bytes(range(32,127)), rng.choice([0,1,7,64,300,1500]))
    t=bytearray(t)
    if len(t)>20:
        l=rng.choice(lits); p=rng.randrange(0,len(t)-len(l)+1) if len(l)<=len(t) else 0
        if len(l)<=len(t): t[p:p+len(l)]=l
    bufs.append(bytes(t))
rdb = R.deserialize(db.serialize()); sc = R.alloc_scratch(rdb)
exp=[]
for bi,b in enumerate(bufs):
    rv,m = R.scan(rdb, sc, b); exp += [(bi,i,t) for (i,t) in m]
exp=sorted(exp, key=lambda x:(x[0],x[2],x[1]))
gs = gpu.alloc_scratch(db)
t0=time.time(); m = gpu.scan_vector_ex(db, bufs, gs); t1=time.time()
got=[(int(a),int(b),int(c)) for a,b,c in m]
print("batch: exp",len(exp),"got",len(got),"equal", got==[(a,b,c) for a,b,c in exp], gs.stats())
