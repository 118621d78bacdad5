import sys, time
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
for cfg in sys.argv[1:]:
    mode, flags, rb, bt = cfg.split()
    t=time.time()
    try:
        k = tb.Kernel(P.kepler_tape(), 28, L.KIND_PLAIN, L.MODE_PARITY if mode=='parity' else L.MODE_FAST, flags=int(flags), reg_budget=int(rb), block_threads=int(bt))
        print(cfg, round(time.time()-t,1), k.info(), flush=True)
    except Exception as e:
        print(cfg, 'FAILED', str(e)[:300], flush=True)
