import sys, json
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
from vectorscan_b200 import gpu, litdb
cases = json.load(open("tests/golden/ref_limex_cases.json"))
for c in cases:
    db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"])
    bufs = [bytes.fromhex(t) for t in c["texts"]]
    exp = [(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m]
    sc = gpu.alloc_scratch(db)
    try:
        got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
        print(c["seed"], c["bits"], "ok" if got == exp else f"MISMATCH got {len(got)} exp {len(exp)}", flush=True)
        if got != exp:
            print(" missing", sorted(set(exp) - set(got))[:6], "extra", sorted(set(got) - set(exp))[:6], [len(b) for b in bufs])
    except Exception as e:
        print(c["seed"], c["bits"], "ERR", e, [len(b) for b in bufs], flush=True)
    sc.close()
