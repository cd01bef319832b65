import sys, time
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import numpy as np
import interplib_b200 as ib
from interplib_b200 import synthetic as syn
ctx = ib.Context(0)
cfg = syn.CONFIGS["c2"]
E = 1 << 20
basis, rules = list(cfg.basis), list(cfg.rules)
h_in = ctx.pinned((E, cfg.N)); h_out = ctx.pinned((E, 3, cfg.Q))
h_in.array[...] = np.random.default_rng(0).random((E, cfg.N))
for chunk in [0, 1 << 13, 1 << 14, 1 << 15, 1 << 16, 1 << 17, 1 << 18]:
    ctx.set_chunk_elements(chunk)
    best = 1e9
    for i in range(6):
        t0 = time.perf_counter()
        ctx.reconstruct(basis, rules, h_in.array, [0, 1, 2], out=h_out.array)
        best = min(best, time.perf_counter() - t0)
    print(chunk, f"{best*1e3:.2f} ms  {E/best:.3e} el/s  D2H {E*3*cfg.Q*8/best/1e9:.1f} GB/s")
