"""How long the CUDA context takes on this box vs. sct_create itself (the CLI's fixed cost)."""
import sys, os, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
t0 = time.perf_counter()
import numpy as np
from sctagger_b200 import _lib, synth
t1 = time.perf_counter()
_lib.device_count()
t2 = time.perf_counter()
p = _lib.Pinned(64, np.uint8)          # first runtime call that needs a context
t3 = time.perf_counter()
bc = synth.make_barcodes(25_000, 1002)
t4 = time.perf_counter()
m = _lib.Matcher(bc, 2, 16)
t5 = time.perf_counter()
m.close()
print({"import_s": round(t1 - t0, 3), "device_count_s": round(t2 - t1, 3), "context_s": round(t3 - t2, 3),
       "sct_create_s": round(t5 - t4, 3), "so_MB": round(os.path.getsize(_lib.SO_PATH) / 1e6, 1),
       "CUDA_MODULE_LOADING": os.environ.get("CUDA_MODULE_LOADING")})
