"""Pins the CPU oracle (oracle/ggml_ref.c) to the golden vectors generated from the
in-container `gguf` package (tests/golden/make_golden.py): bit-exact for block
dequantisation and for the Q4_0 / Q8_0 reference quantisers."""
import numpy as np
import pytest

TYPES = {"Q4_0": (2, 32), "Q8_0": (8, 32), "Q4_K": (12, 256), "Q6_K": (14, 256)}


@pytest.mark.parametrize("name", list(TYPES))
def test_dequant_bit_exact_vs_golden(oracle, golden, name):
    t, bs = TYPES[name]
    blocks, want = golden[f"{name}_blocks"], golden[f"{name}_deq"]
    got = oracle.dequantize(t, blocks, bs)
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name", ["Q4_0", "Q8_0"])
def test_quantize_bit_exact_vs_golden(oracle, golden, name):
    t, _ = TYPES[name]
    got = oracle.quantize(t, golden[f"{name}_x"])
    assert np.array_equal(got, golden[f"{name}_quant"])


def test_q8_K_roundtrip_and_bsums(oracle):
    rng = np.random.default_rng(1)
    x = rng.standard_normal((3, 512)).astype(np.float32)
    x[2] = 0
    q = oracle.quantize(oracle.Q8_K, x).reshape(3, 2, 292)
    d = q[:, :, :4].copy().view(np.float32).reshape(3, 2)
    qs = q[:, :, 4:260].view(np.int8).astype(np.int32)
    bs = q[:, :, 260:].copy().view(np.int16).reshape(3, 2, 16)
    assert np.array_equal(qs.reshape(3, 2, 16, 16).sum(-1), bs)
    assert np.abs(qs).max() <= 127
    deq = (qs * d[..., None]).reshape(3, 512)
    assert np.max(np.abs(deq - x)) <= np.abs(x).max() / 127 * 0.51 + 1e-7
    assert np.all(d[2] == 0) and np.all(qs[2] == 0)


@pytest.mark.parametrize("name", ["Q4_K", "Q6_K"])
def test_kquant_makers_emit_valid_blocks(oracle, name):
    """The synthetic K-quant makers are not ggml's quantisers, but what they emit must
    decode (with the pinned dequantiser) to something close to the input."""
    import gguf
    t, _ = TYPES[name]
    rng = np.random.default_rng(2)
    x = rng.standard_normal((4, 512)).astype(np.float32) * 0.05
    q = oracle.quantize(t, x)
    deq = gguf.quants.dequantize(q, gguf.GGMLQuantizationType(t))
    assert np.array_equal(deq, oracle.dequantize(t, q, 512))
    err = np.sqrt(np.mean((deq - x) ** 2)) / np.sqrt(np.mean(x ** 2))
    assert err < (0.12 if name == "Q4_K" else 0.04), err


@pytest.mark.parametrize("name,act", [("Q4_0", "Q8_0"), ("Q8_0", "Q8_0"), ("Q4_K", "Q8_K"), ("Q6_K", "Q8_K")])
def test_vec_dot_matches_dequantised_product(oracle, pkg, name, act):
    """Integer block dot == dot of the dequantised weight row with the dequantised
    activation row (f64), to float rounding."""
    import gguf
    from gguf import GGMLQuantizationType as T
    t, _ = TYPES[name]
    rng = np.random.default_rng(3)
    k, rows = 1024, 8
    W = pkg.synth.random_blocks(T(t), rows, k, 0.03, rng)
    x = rng.standard_normal(k).astype(np.float32)
    y = oracle.mul_mat_vec(t, W, rows, k, x, faithful=True)
    wd = gguf.quants.dequantize(W, T(t)).astype(np.float64)
    at = getattr(oracle, act)
    xq = oracle.quantize(at, x[None])
    if act == "Q8_0":
        xd = gguf.quants.dequantize(xq, T.Q8_0)[0].astype(np.float64)
    else:
        b = xq.reshape(-1, 292)
        xd = (b[:, 4:260].view(np.int8).astype(np.float64) * b[:, :4].copy().view(np.float32).astype(np.float64)).reshape(-1)
    want = wd @ xd
    assert np.allclose(y, want, rtol=2e-5, atol=2e-5 * np.abs(want).max())
    # ideal path: against the un-quantised activations
    yi = oracle.mul_mat_vec(t, W, rows, k, x, faithful=False)
    assert np.allclose(yi, wd @ x.astype(np.float64), rtol=1e-6, atol=1e-6)


def test_avx2_and_scalar_q4_0_dot_agree_bitwise(oracle):
    """The SIMD helper only speeds up the exact integer part; results must equal a pure
    python evaluation in ggml's order."""
    rng = np.random.default_rng(4)
    k = 256
    x = rng.standard_normal((1, k)).astype(np.float32)
    w = rng.standard_normal((1, k)).astype(np.float32)
    wq, xq = oracle.quantize(oracle.Q4_0, w), oracle.quantize(oracle.Q8_0, x)
    got = oracle.mul_mat_vec(oracle.Q4_0, wq, 1, k, x[0])[0]
    wb, xb = wq.reshape(-1, 18), xq.reshape(-1, 34)
    s = np.float32(0)
    for i in range(k // 32):
        dw = wb[i, :2].copy().view(np.float16)[0].astype(np.float32)
        dx = xb[i, :2].copy().view(np.float16)[0].astype(np.float32)
        q = wb[i, 2:].astype(np.int32)
        a = xb[i, 2:].view(np.int8).astype(np.int32)
        sumi = int(((q & 15) - 8) @ a[:16] + ((q >> 4) - 8) @ a[16:])
        s = np.float32(s + np.float32(np.float32(np.float32(sumi) * dw) * dx))
    assert np.float32(got).view(np.uint32) == s.view(np.uint32)
