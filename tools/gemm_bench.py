import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200"); E = p.engine
from gguf import GGMLQuantizationType as T
rng = np.random.default_rng(0)
for name, rows, k in [("qkv", 12288, 4096), ("gate_up", 22016, 4096), ("down", 4096, 11008)]:
    W = p.synth.random_blocks(T.Q4_0, rows, k, 0.02, rng)
    for B in (8, 32):
        x = rng.standard_normal((B, k)).astype(np.float32)
        _, ms = E.gemm(E.TYPE_Q4_0, W, rows, k, x, iters=20)
        print(f"{name} {rows}x{k} B={B}: {ms*1e3:.1f} us  weights {rows*k*0.5625/ms/1e6:.0f} GB/s  {2*rows*k*B/ms/1e9:.1f} TOPS", flush=True)
