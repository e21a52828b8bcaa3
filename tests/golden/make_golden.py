"""Regenerates tests/golden/quant_kat.npz from the in-container `gguf` python package
(gguf 0.19.0, gguf/quants.py) -- the only restatement of ggml's block formats that can be
executed here (the reference itself holds no vectors).

    python tests/golden/make_golden.py

Contents (all seeded): for each of Q4_0, Q8_0, Q4_K, Q6_K a set of block bytes (edge
cases first: all-zero, extreme f16 scales, every nibble/sextet value, then random) and
their `gguf.quants.dequantize` output; for Q4_0 and Q8_0 additionally f32 inputs and
their `gguf.quants.quantize` bytes.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import gguf
from gguf import GGMLQuantizationType as T

SIZES = {"Q4_0": (32, 18), "Q8_0": (32, 34), "Q4_K": (256, 144), "Q6_K": (256, 210)}
D_OFF = {"Q4_0": 0, "Q8_0": 0, "Q4_K": 0, "Q6_K": 208}


def f16b(v):
    return np.array([v], dtype=np.float16).view(np.uint8)


def blocks_for(name, rng, n_rand=48):
    bs, ts = SIZES[name]
    out = [np.zeros(ts, np.uint8)]
    # extreme / special scales with a counting payload
    for d in (65504.0, -65504.0, 6e-8, -6e-8, 1.0, -1.0, 0.333251953125, 0.0):
        b = (np.arange(ts) * 37 % 256).astype(np.uint8)
        b[D_OFF[name]:D_OFF[name] + 2] = f16b(d)
        if name == "Q4_K":
            b[2:4] = f16b(d / 2)
        out.append(b)
    # every payload byte value at least once
    for v in range(0, 256, 5):
        b = np.full(ts, v, np.uint8)
        b[D_OFF[name]:D_OFF[name] + 2] = f16b(0.0078125)
        if name == "Q4_K":
            b[2:4] = f16b(0.015625)
        out.append(b)
    for _ in range(n_rand):
        b = rng.integers(0, 256, ts, dtype=np.uint8)
        b[D_OFF[name]:D_OFF[name] + 2] = f16b(rng.uniform(-0.1, 0.1))
        if name == "Q4_K":
            b[0:2] = f16b(rng.uniform(0, 0.05)); b[2:4] = f16b(rng.uniform(0, 0.05))
        out.append(b)
    return np.stack(out)


def main():
    rng = np.random.default_rng(20261106)
    z = {}
    for name in SIZES:
        t = getattr(T, name)
        b = blocks_for(name, rng)
        z[f"{name}_blocks"] = b
        z[f"{name}_deq"] = gguf.quants.dequantize(b, t).astype(np.float32)
    for name in ("Q4_0", "Q8_0"):
        t = getattr(T, name)
        x = rng.standard_normal((24, 64)).astype(np.float32) * rng.uniform(1e-3, 30, (24, 1)).astype(np.float32)
        x[0] = 0.0                       # all-zero row: d == 0 branch
        x[1, :] = 1.0                    # constant row
        x[2, 5] = 1e4                    # one outlier
        x[3] = np.round(x[3] * 2) / 2    # ties for the rounding rule
        z[f"{name}_x"] = x
        z[f"{name}_quant"] = gguf.quants.quantize(x, t)
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "quant_kat.npz"), **z)
    print({k: v.shape for k, v in z.items()})


if __name__ == "__main__":
    main()
