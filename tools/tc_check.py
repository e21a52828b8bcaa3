"""Bring-up check of the tcgen05 batched kernel (b2_gemm_tc): bits against the oracle, then timing on the 7B shapes."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
p = importlib.import_module("ollama-operator_b200"); import oracle
from gguf import GGMLQuantizationType as T
E = p.engine
ok = True
if "--one" in sys.argv:          # for ncu: one 7B qkv-shaped launch per accumulation mode, 512 columns
    rng = np.random.default_rng(1)
    W = p.synth.random_blocks(T.Q4_0, 12288, 4096, 4096 ** -0.5, rng)
    x = rng.standard_normal((512, 4096)).astype(np.float32)
    for acc32 in (False, True):
        _, ms = E.gemm_tc(E.TYPE_Q4_0, W, 12288, 4096, x, acc32=acc32, iters=2)
        print(f"qkv 12288x4096 x 512 columns acc32 {int(acc32)}: {ms * 1e3:.1f} us")
    sys.exit(0)
for rows, k, n in [(128, 256, 1), (128, 256, 32), (256, 512, 7), (300, 1024, 33), (1024, 4096, 64), (258, 11008, 17), (512, 4096, 128)]:
    rng = np.random.default_rng(rows + k + n)
    W = p.synth.random_blocks(T.Q4_0, rows, k, k ** -0.5, rng)
    x = rng.standard_normal((n, k)).astype(np.float32)
    for acc32 in (False, True):
        y, _ = E.gemm_tc(E.TYPE_Q4_0, W, rows, k, x, acc32=acc32)
        bad = 0; worst = 0.0
        for c in range(n):
            want = oracle.mul_mat_vec(2, W, rows, k, x[c], faithful=1 if acc32 else 2)
            if not np.array_equal(y[c].view(np.uint32), want.view(np.uint32)):
                bad += 1; worst = max(worst, float(np.abs(y[c] - want).max()))
        print(f"rows {rows} k {k} cols {n} acc32 {acc32}: {'OK' if bad == 0 else f'{bad} columns differ, max abs {worst}'}", flush=True)
        ok &= bad == 0
if "--time" in sys.argv:
    for name, rows, k in [("qkv", 12288, 4096), ("wo", 4096, 4096), ("gate|up", 22016, 4096), ("down", 4096, 11008)]:
        rng = np.random.default_rng(1)
        W = p.synth.random_blocks(T.Q4_0, rows, k, k ** -0.5, rng)
        for n in (32, 128, 512):
            x = rng.standard_normal((n, k)).astype(np.float32)
            for acc32 in (False, True):
                _, ms = E.gemm_tc(E.TYPE_Q4_0, W, rows, k, x, acc32=acc32, iters=20)
                print(f"{name:8s} {rows}x{k} cols {n:3d} acc32 {int(acc32)}: {ms * 1e3:8.1f} us  {2.0 * rows * k * n / ms / 1e9:7.1f} TOP/s", flush=True)
            if n == 32:
                _, ms = E.gemm(E.TYPE_Q4_0, W, rows, k, x, iters=20)
                print(f"{name:8s} {rows}x{k} cols {n:3d} mma.sync  : {ms * 1e3:8.1f} us", flush=True)
sys.exit(0 if ok else 1)
