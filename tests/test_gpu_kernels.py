"""Parity of the individual CUDA kernels with the CPU oracle, called through the C ABI.
Bit-exact for dequantisation (integer/byte work) AND for the W4A8 GEMV against the oracle's
canonical mode (exact integer block dots, per-block float terms as in ggml's generic path,
terms added in f64 so the order is immaterial); within a few ulps of the row's magnitude
against ggml's generic f32 accumulation order."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TYPES = {"Q4_0": (2, 32, 18), "Q8_0": (8, 32, 34), "Q4_K": (12, 256, 144), "Q6_K": (14, 256, 210)}


@pytest.mark.parametrize("name", list(TYPES))
def test_dequant_bit_exact_vs_golden(pkg, golden, name):
    t, bs, _ = TYPES[name]
    blocks, want = golden[f"{name}_blocks"], golden[f"{name}_deq"]
    got = pkg.engine.dequantize(t, blocks, want.size).reshape(want.shape)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name", list(TYPES))
def test_dequant_bit_exact_vs_oracle_random(pkg, oracle, name):
    from gguf import GGMLQuantizationType as T
    t, bs, ts = TYPES[name]
    rng = np.random.default_rng(7)
    blocks = pkg.synth.random_blocks(T(t), 33, 4 * 256, 0.05, rng)
    want = oracle.dequantize(t, blocks, 1024)
    got = pkg.engine.dequantize(t, blocks, want.size).reshape(want.shape)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_dequant_empty_input_is_a_no_op(pkg):
    assert pkg.engine.dequantize(2, np.zeros(0, np.uint8), 0).size == 0


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("rows,k", [(2, 256), (6, 1024), (10, 2560 if True else 0), (4, 11008 // 256 * 256)])
def test_repack_then_dequant_equals_dequant(pkg, oracle, name, rows, k):
    """The load-time split-plane layout must decode to exactly what the file blocks decode to."""
    from gguf import GGMLQuantizationType as T
    t, bs, ts = TYPES[name]
    if k % bs:
        pytest.skip("k not a multiple of the block size")
    rng = np.random.default_rng(rows * 1000 + k)
    blocks = pkg.synth.random_blocks(T(t), rows, k, 0.05, rng)
    want = oracle.dequantize(t, blocks, k)
    got = pkg.engine.dequantize_repacked(t, blocks, rows, k)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


GEMV_SHAPES = [(2, 256), (64, 4096), (130, 1024), (32, 11008), (48, 2560), (20, 5120), (16, 8192), (8, 14336), (8, 28672)]


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("rows,k", GEMV_SHAPES)
def test_gemv_matches_oracle(pkg, oracle, name, rows, k):
    from gguf import GGMLQuantizationType as T
    t, bs, ts = TYPES[name]
    if k % bs:
        pytest.skip("k not a multiple of the block size")
    rng = np.random.default_rng(rows + k)
    W = pkg.synth.random_blocks(T(t), rows, k, k ** -0.5, rng)
    x = rng.standard_normal(k).astype(np.float32)
    got = pkg.engine.gemv(t, W, rows, k, x)
    canon = oracle.mul_mat_vec(t, W, rows, k, x, faithful=2)
    assert np.array_equal(got.view(np.uint32), canon.view(np.uint32)), np.max(np.abs(got - canon))
    generic = oracle.mul_mat_vec(t, W, rows, k, x, faithful=1)      # ggml's f32 block-after-block order
    assert np.max(np.abs(got - generic)) <= 3e-5 * np.abs(generic).max() + 1e-6


@pytest.mark.parametrize("name", ["Q4_0", "Q6_K"])
def test_gemv_with_fused_rmsnorm_matches_oracle(pkg, oracle, name):
    from gguf import GGMLQuantizationType as T
    t, bs, ts = TYPES[name]
    rows, k = 96, 4096
    rng = np.random.default_rng(11)
    W = pkg.synth.random_blocks(T(t), rows, k, k ** -0.5, rng)
    x = (rng.standard_normal(k) * 3).astype(np.float32)
    g = (1 + 0.02 * rng.standard_normal(k)).astype(np.float32)
    xn = oracle.rms_norm(x, g, 1e-5)
    want = oracle.mul_mat_vec(t, W, rows, k, xn, faithful=2)
    got = pkg.engine.gemv(t, W, rows, k, x, gain=g, eps=1e-5)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), np.max(np.abs(got - want))


def test_gemv_edge_inputs(pkg, oracle):
    """All-zero activations (d == 0 branch of the quantiser), a single spike, constant rows."""
    from gguf import GGMLQuantizationType as T
    rng = np.random.default_rng(5)
    rows, k = 8, 512
    for t in (2, 12, 14):
        W = pkg.synth.random_blocks(T(t), rows, k, 0.05, rng)
        for x in (np.zeros(k, np.float32), np.eye(1, k, 37, dtype=np.float32)[0] * 100, np.full(k, 0.25, np.float32)):
            want = oracle.mul_mat_vec(t, W, rows, k, x, faithful=2)
            got = pkg.engine.gemv(t, W, rows, k, x)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (t, np.abs(got - want).max())


@pytest.mark.parametrize("n_ctx", [1, 2, 63, 64, 65, 300, 1024, 2500])
@pytest.mark.parametrize("n_head,n_head_kv,hd", [(4, 4, 128), (8, 2, 64), (8, 1, 128), (4, 2, 80)])
@pytest.mark.parametrize("cluster", [1, 4, 8])
def test_attention_matches_reference(pkg, oracle, n_ctx, n_head, n_head_kv, hd, cluster):
    if cluster == 8 and n_ctx not in (65, 2500):
        pytest.skip("covered by the other cluster sizes")
    rng = np.random.default_rng(n_ctx * 7 + n_head)
    q = rng.standard_normal(n_head * hd).astype(np.float32)
    k = rng.standard_normal((n_ctx, n_head_kv * hd)).astype(np.float16)
    v = rng.standard_normal((n_ctx, n_head_kv * hd)).astype(np.float16)
    scale = hd ** -0.5
    # oracle canonical mode: llama.cpp's non-flash path with both dots in ggml's 32-lane SIMD order
    want = oracle.attention_canon(q, k, v, n_head, n_head_kv, hd, scale)
    got = pkg.engine.attn_decode(q, k, v, n_head, n_head_kv, hd, scale, cluster)
    # a warp reproduces the SIMD lane order exactly, so the kernel must agree bit for bit for any
    # split of the work over the cluster
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), np.max(np.abs(got - want))


@pytest.mark.parametrize("rows,k,n_cols", [(2, 256, 1), (300, 512, 5), (1024, 4096, 32), (258, 11008, 17), (64, 2560 // 256 * 256, 8)])
def test_batched_gemm_equals_gemv_per_column(pkg, oracle, rows, k, n_cols):
    """The tensor-core batched kernel must give, column by column, the bits of the GEMV (and of the oracle)."""
    from gguf import GGMLQuantizationType as T
    E = pkg.engine
    rng = np.random.default_rng(rows + k + n_cols)
    W = pkg.synth.random_blocks(T.Q4_0, rows, k, k ** -0.5, rng)
    x = rng.standard_normal((n_cols, k)).astype(np.float32)
    x[0, : k // 2] = 0
    g = (1 + 0.02 * rng.standard_normal(k)).astype(np.float32)
    y, _ = E.gemm(E.TYPE_Q4_0, W, rows, k, x)
    yn, _ = E.gemm(E.TYPE_Q4_0, W, rows, k, x, gain=g, eps=1e-5)
    for c in range(n_cols):
        want = oracle.mul_mat_vec(2, W, rows, k, x[c], faithful=2)
        assert np.array_equal(y[c].view(np.uint32), want.view(np.uint32)), (c, np.abs(y[c] - want).max())
        wn = oracle.mul_mat_vec(2, W, rows, k, oracle.rms_norm(x[c], g, 1e-5), faithful=2)
        assert np.array_equal(yn[c].view(np.uint32), wn.view(np.uint32)), c


@pytest.mark.parametrize("rows,k,n_cols", [(128, 256, 1), (300, 512, 33), (1024, 4096, 64), (258, 11008, 17), (512, 4096, 128), (256, 1024, 512)])
def test_tcgen05_gemm_matches_the_oracle_bit_for_bit(pkg, oracle, rows, k, n_cols):
    """The tcgen05 / TMEM batched kernel (gemm_tc.cu): with f64 accumulation every column carries the bits of the
    canonical oracle (and therefore of the GEMV); with f32 accumulation the bits of ggml's generic C order (oracle
    mode 1: ggml_vec_dot_q4_0_q8_0, scalar branch) -- exact integer block sums either way."""
    from gguf import GGMLQuantizationType as T
    E = pkg.engine
    rng = np.random.default_rng(rows + k + n_cols)
    W = pkg.synth.random_blocks(T.Q4_0, rows, k, k ** -0.5, rng)
    x = rng.standard_normal((n_cols, k)).astype(np.float32)
    x[0, : k // 2] = 0
    g = (1 + 0.02 * rng.standard_normal(k)).astype(np.float32)
    y64, _ = E.gemm_tc(E.TYPE_Q4_0, W, rows, k, x)
    y32, _ = E.gemm_tc(E.TYPE_Q4_0, W, rows, k, x, acc32=True)
    yn, _ = E.gemm_tc(E.TYPE_Q4_0, W, rows, k, x, gain=g, eps=1e-5)
    for c in range(0, n_cols, max(1, n_cols // 24)):
        want = oracle.mul_mat_vec(2, W, rows, k, x[c], faithful=2)
        assert np.array_equal(y64[c].view(np.uint32), want.view(np.uint32)), (c, np.abs(y64[c] - want).max())
        w1 = oracle.mul_mat_vec(2, W, rows, k, x[c], faithful=1)
        assert np.array_equal(y32[c].view(np.uint32), w1.view(np.uint32)), (c, np.abs(y32[c] - w1).max())
        wn = oracle.mul_mat_vec(2, W, rows, k, oracle.rms_norm(x[c], g, 1e-5), faithful=2)
        assert np.array_equal(yn[c].view(np.uint32), wn.view(np.uint32)), c


def test_tcgen05_gemm_rejects_what_it_does_not_serve(pkg):
    E = pkg.engine
    W = np.zeros(64 * 320 // 32 * 18, np.uint8)
    with pytest.raises(E.B2Error):
        E.gemm_tc(E.TYPE_Q4_0, W, 64, 320, np.zeros((2, 320), np.float32))        # k not a multiple of 256
