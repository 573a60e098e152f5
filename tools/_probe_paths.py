import sys, json, numpy as np, torch
sys.path.insert(0, '.')
from rag_application_b200 import Engine, synth
dev = torch.device('cuda', 0)
eng = Engine(0); eng.profile_enable(True)
def stage(fn, it=10):
    fn(); fn(); eng.profile_reset()
    for _ in range(it): fn()
    ms, calls, _ = eng.profile_read(); return ms['dense_scan'] / calls['dense_scan'], ms['dense_select'] / calls['dense_select']
for (n, d, dt) in ((10_000_000, 1024, torch.bfloat16), (1 << 20, 768, torch.float16), (12_500_000, 384, torch.float16)):
    x = synth.dense_corpus_torch(n, d, dt, dev); q = synth.dense_queries_torch(x, 128)
    eng.dense_set('c', x); del x; torch.cuda.empty_cache()
    gb = n * d * 2 / 1e9
    for path in ('gemv', 'tcgen05'):
        eng.force_dense_path(path)
        for B in (1, 8) if path == 'gemv' else (1, 8, 128):
            qb = q[:B].contiguous()
            a, b = stage(lambda: eng.dense_topk('c', qb, 100))
            print(n, d, path, B, round(a, 4), 'ms scan', round(gb / a, 1), 'TB/s;', round(b, 4), 'ms select', flush=True)
    eng.dense_drop('c')
