"""End-to-end parity of the decode engine with the CPU oracle on seeded synthetic GGUF
models, through the C ABI: logits within max-abs 2e-2 (BASELINE.json north_star), greedy
tokens identical for 64 steps.

Which oracle.  W4A8 inference is chaotic at the 1e-7 level: two evaluations whose f32
summation ORDER differs (ggml's own generic-C, AVX2 and NEON paths already do) sooner or later
round one activation to a different int8, after which their logits differ by ~1e-1.  The
engine therefore implements, and these tests pin it to, the oracle's order-free "canonical"
mode (oracle/ggml_ref.h: same integer block dots and float terms as ggml's generic path,
terms added in f64).  Against ggml's generic f32 order the comparison is statistical."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2      # max-abs on logits vs the ggml-faithful oracle (north_star)


@pytest.fixture(scope="module")
def eng(pkg):
    e = pkg.engine.Engine(0)
    yield e
    e.close()


def _teacher_forced_compare(pkg, oracle, eng, path, n_prompt, n_steps, flags=0, mode=2):
    om = oracle.OracleModel(path, faithful=mode)
    m = eng.load(path, max_batch=2, max_ctx=256, flags=flags)
    toks = pkg.synth.prompt_tokens(n_prompt, om.m.n_vocab)
    seq = m.seq_new()
    lg = m.prefill(seq, toks)
    for t in toks:
        want = om.eval(int(t))
    worst = float(np.max(np.abs(lg - want)))
    n_same = 0
    tok = int(np.argmax(want))
    for i in range(n_steps):
        g_tok = int(np.argmax(lg))
        n_same += int(g_tok == tok)
        nxt, lgs = m.decode([seq], [tok], want_logits=True)      # teacher-forced with the oracle's token
        lg = lgs[0]
        want = om.eval(tok)
        worst = max(worst, float(np.max(np.abs(lg - want))))
        assert int(nxt[0]) == int(np.argmax(lg))
        tok = int(np.argmax(want))
    m.close()
    return worst, n_same


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2", "tiny-gqa8", "tiny-phi2-80", "tiny-rope31"])
def test_logits_match_oracle(pkg, oracle, eng, model_dir, preset):
    path = pkg.synth.cached_model(preset, model_dir)
    worst, n_same = _teacher_forced_compare(pkg, oracle, eng, path, 9, 40)
    assert worst <= LOGIT_TOL, worst
    assert n_same == 40


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2"])
def test_logits_vs_ggml_generic_order(pkg, oracle, eng, model_dir, preset):
    """Against ggml's generic f32 accumulation order the engine is as close as two ggml builds
    are to each other: bounded by the quantisation-noise floor, tokens equal wherever the
    oracle's own top-2 margin is above it."""
    path = pkg.synth.cached_model(preset, model_dir)
    worst, n_same = _teacher_forced_compare(pkg, oracle, eng, path, 9, 40, mode=1)
    assert worst <= 0.5, worst
    assert n_same >= 36, n_same


@pytest.mark.parametrize("flags", [1, 2, 3, 4, 6, 7, 8, 15])
def test_logits_match_without_pdl_or_graphs(pkg, oracle, eng, model_dir, flags):
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    """flags: 1 no PDL, 2 no CUDA graphs, 4 no persistent kernel (one kernel per matrix), 8 no batched prompt kernels, and combinations."""
    worst, _ = _teacher_forced_compare(pkg, oracle, eng, path, 5, 6, flags=flags)
    assert worst == 0.0, worst          # every path computes the same bits


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2"])
def test_greedy_64_tokens_identical(pkg, oracle, eng, model_dir, preset):
    path = pkg.synth.cached_model(preset, model_dir)
    om = oracle.OracleModel(path, faithful=2)
    toks = pkg.synth.prompt_tokens(16, om.m.n_vocab)
    want, want_logits = om.generate(toks, 64)
    m = eng.load(path, max_batch=1, max_ctx=256)
    seq = m.seq_new()
    lg = m.prefill(seq, toks)
    first = int(np.argmax(lg))
    rest = m.decode_greedy(seq, 63)            # device-resident loop continues from the arg-max of the prefill
    got = [first] + [int(t) for t in rest]
    if got != want:
        i = next(j for j in range(64) if got[j] != want[j])
        top2 = np.sort(want_logits[i])[-2:]
        pytest.fail(f"first divergence at step {i}: gpu {got[i]} vs oracle {want[i]} (oracle top-2 margin {top2[1] - top2[0]:.2e})")
    assert m.seq_len(seq) == 16 + 63
    m.close()


def test_kv_cache_rows_match_oracle(pkg, oracle, eng, model_dir):
    """f16 K/V rows written by the fused QKV+RoPE epilogue == the oracle's cache, bit for bit."""
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    m = eng.load(path, max_batch=1, max_ctx=256)
    toks = pkg.synth.prompt_tokens(70, m.info.n_vocab)          # crosses a 64-token page boundary
    seq = m.seq_new()
    m.prefill(seq, toks, want_logits=False)
    k, v = m.kv_read(seq, 0, 70)
    # layer-0 K/V depend only on the embedding row, the norm and the projections
    om = oracle.OracleModel(path, faithful=2)
    nkv = m.info.n_head_kv * m.info.head_dim
    # the oracle keeps its cache private; recompute layer-0 K/V rows through its public pieces
    import gguf
    tens = om.tensors
    E = m.info.n_embd
    emb = gguf.quants.dequantize(np.ascontiguousarray(tens["token_embd.weight"].data), tens["token_embd.weight"].tensor_type)
    g = np.array(tens["blk.0.attn_norm.weight"].data, dtype=np.float32)
    wk, wv = tens["blk.0.attn_k.weight"], tens["blk.0.attn_v.weight"]
    for pos in (0, 1, 63, 64, 69):
        xn = oracle.rms_norm(emb[toks[pos]], g, om.m.norm_eps)
        kk = oracle.mul_mat_vec(int(wk.tensor_type), np.ascontiguousarray(wk.data), nkv, E, xn, faithful=2)
        vv = oracle.mul_mat_vec(int(wv.tensor_type), np.ascontiguousarray(wv.data), nkv, E, xn, faithful=2)
        kk = oracle.rope(kk, m.info.n_head_kv, m.info.head_dim, m.info.n_rot, pos, om.m.rope_freq_base, 0)
        assert np.array_equal(k[pos], kk.astype(np.float16))
        assert np.array_equal(v[pos], vv.astype(np.float16))
    m.close()


def test_sequences_are_independent_and_reusable(pkg, eng, model_dir):
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    m = eng.load(path, max_batch=2, max_ctx=128)
    a, b = m.seq_new(), m.seq_new()
    ta, tb = pkg.synth.prompt_tokens(10, 512, seed=1), pkg.synth.prompt_tokens(7, 512, seed=2)
    la, lb = m.prefill(a, ta), m.prefill(b, tb)
    # same prompts in the other order, fresh slots
    m.seq_free(a); m.seq_free(b)
    b2, a2 = m.seq_new(), m.seq_new()
    lb2, la2 = m.prefill(b2, tb), m.prefill(a2, ta)
    assert np.array_equal(la, la2) and np.array_equal(lb, lb2)
    # batched b2_decode == one by one
    nxt = m.decode([a2, b2], [int(np.argmax(la)), int(np.argmax(lb))])
    m.seq_truncate(a2, 10)
    one = m.decode([a2], [int(np.argmax(la))])
    assert int(one[0]) == int(nxt[0])
    with pytest.raises(pkg.engine.B2Error):
        m.seq_new()                      # both slots taken
    with pytest.raises(pkg.engine.B2Error):
        m.decode([5], [0])               # bad sequence id
    with pytest.raises(pkg.engine.B2Error):
        m.prefill(a2, [999999])          # token outside the vocabulary
    with pytest.raises(pkg.engine.B2Error):
        m.prefill(a2, list(range(200)))  # beyond max_ctx
    m.close()


def test_load_errors(pkg, eng, tmp_path):
    with pytest.raises(pkg.engine.B2Error) as e:
        eng.load(str(tmp_path / "missing.gguf"))
    assert e.value.code == -3
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"NOPE" + bytes(64))
    with pytest.raises(pkg.engine.B2Error) as e:
        eng.load(str(bad))
    assert e.value.code == -4


def test_tensor_parallel_and_tooling_argument_errors(pkg, eng, model_dir):
    E = pkg.engine
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    with pytest.raises(E.B2Error) as e:          # one device bound, two asked for
        eng.load(path, tp_size=2)
    assert e.value.code == -1 and "tp_size" in str(e.value)
    m = eng.load(path, max_batch=1, max_ctx=128)
    s = m.seq_new()
    m.prefill(s, pkg.synth.prompt_tokens(10, m.info.n_vocab))
    with pytest.raises(E.B2Error):               # cannot "assume" fewer rows than the sequence holds
        m.seq_assume(s, 5)
    with pytest.raises(E.B2Error) as e:          # nor more than the cache can hold
        m.seq_assume(s, 1000)
    assert e.value.code == -6
    m.seq_assume(s, 100)
    assert m.seq_len(s) == 100
    nxt = m.decode([s], [3])
    assert 0 <= int(nxt[0]) < m.info.n_vocab and m.seq_len(s) == 101
    m.close()


def test_cache_edges_and_prefix_reuse(pkg, eng, model_dir):
    """Empty prompts, a cache filled to its last row, truncation and re-evaluation of a suffix (prefix reuse)."""
    E = pkg.engine
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    m = eng.load(path, max_batch=1, max_ctx=64)
    V = m.info.n_vocab
    s = m.seq_new()
    with pytest.raises(E.B2Error):
        m.prefill(s, [])                                 # nothing to evaluate
    toks = pkg.synth.prompt_tokens(64, V, seed=11)
    full = m.prefill(s, toks)                            # exactly max_ctx tokens: the last cache row is used
    assert m.seq_len(s) == 64
    with pytest.raises(E.B2Error) as e:
        m.decode([s], [1])                               # one more does not fit
    assert e.value.code == -6
    with pytest.raises(E.B2Error):
        m.decode_greedy(s, 1, first_tok=1)
    # drop the last 24 tokens and evaluate them again: same logits as the first time, bit for bit
    m.seq_truncate(s, 40)
    assert m.seq_len(s) == 40
    again = m.prefill(s, toks[40:])
    assert np.array_equal(again.view(np.uint32), full.view(np.uint32))
    # and token by token through b2_decode
    m.seq_truncate(s, 60)
    lg = None
    for t in toks[60:]:
        _, lgs = m.decode([s], [int(t)], want_logits=True)
        lg = lgs[0]
    assert np.array_equal(lg.view(np.uint32), full.view(np.uint32))
    with pytest.raises(E.B2Error):
        m.seq_truncate(s, 65)
    m.seq_free(s)
    with pytest.raises(E.B2Error):
        m.seq_len(s)
    m.close()


def test_wide_model_long_prompt_through_the_tensor_core_path(pkg, oracle, eng, model_dir):
    """SURVEY 8c (iv)/(v) at 7B width: a two-layer model with llama2-7B's matrix shapes.  A 200-token prompt goes
    through the tcgen05 batched kernel in one chunk; its logits and the 64 greedy tokens that follow must be the
    canonical oracle's, bit for bit -- and the same prompt evaluated token by token gives the same bits."""
    path = pkg.synth.cached_model("wide-llama-2l", model_dir, source="blocks")
    om = oracle.OracleModel(path, faithful=2)
    m = eng.load(path, max_batch=1, max_ctx=300)
    toks = pkg.synth.prompt_tokens(200, m.info.n_vocab)
    seq = m.seq_new()
    lg = m.prefill(seq, toks)
    for t in toks:
        want = om.eval(int(t))
    assert np.array_equal(lg.view(np.uint32), want.view(np.uint32)), float(np.max(np.abs(lg - want)))
    got = [int(np.argmax(lg))] + [int(t) for t in m.decode_greedy(seq, 63)]      # the loop continues from the prefill's arg-max
    tok, ref = int(np.argmax(want)), []
    for _ in range(64):
        ref.append(tok)
        tok = int(np.argmax(om.eval(tok)))
    assert got == ref
    m.close()


def test_13b_width_tiles_smaller_than_a_worker_round(pkg, oracle, eng, model_dir):
    """llama2-13B's matrix widths (K = 5120 / 13824) make the persistent kernel's weight tiles hold fewer row pairs than
    there are row workers, so a worker owns a pair only in some tiles.  The greedy tokens must still be the oracle's and
    those of the per-matrix kernels (a worker that skipped a tile's barrier lap once read a tile before it had landed)."""
    path = pkg.synth.cached_model("wide13-llama-1l", model_dir, source="blocks")
    om = oracle.OracleModel(path, faithful=2)
    toks = pkg.synth.prompt_tokens(24, om.m.n_vocab)
    want, _ = om.generate(toks, 48)
    for flags in (0, pkg.engine.LOAD_NO_MEGA):
        m = eng.load(path, max_batch=1, max_ctx=128, flags=flags)
        seq = m.seq_new()
        lg = m.prefill(seq, toks)
        got = [int(np.argmax(lg))] + [int(t) for t in m.decode_greedy(seq, 47)]
        assert got == want, (flags, next(i for i in range(48) if got[i] != want[i]))
        m.close()


def test_stage_handover_without_grid_barriers_keeps_the_bits(pkg, oracle, eng, model_dir, monkeypatch):
    """The persistent kernel hands x and the attention output from stage to stage through poisoned buffers that the
    next stage polls (no grid barrier at attention->wo, wo->up, down->qkv).  A race there would show as a token or a
    logit bit that differs from the same kernel with every grid barrier in place (B2_MEGA_NO_LL) -- over a few hundred
    steps and two sequences that reuse the buffers of one slot after another -- or from the canonical oracle."""
    path = pkg.synth.cached_model("wide-llama-2l", model_dir, source="blocks")
    om = oracle.OracleModel(path, faithful=2)
    toks = pkg.synth.prompt_tokens(40, om.m.n_vocab, seed=5)
    want, _ = om.generate(toks, 32)
    runs = {}
    for name, env in (("handover", None), ("barriers", "1"), ("handover_h", "h")):
        monkeypatch.delenv("B2_MEGA_NO_LL", raising=False)
        monkeypatch.delenv("B2_MEGA_LL_H", raising=False)
        if env == "1":
            monkeypatch.setenv("B2_MEGA_NO_LL", "1")
        if env == "h":
            monkeypatch.setenv("B2_MEGA_LL_H", "1")          # also hand h over (off by default: slower, same bits)
        m = eng.load(path, max_batch=2, max_ctx=384)
        out = []
        for rep in range(2):                                   # the second sequence reuses slot state the first one left
            seq = m.seq_new()
            lg = m.prefill(seq, toks)
            got = [int(np.argmax(lg))] + [int(t) for t in m.decode_greedy(seq, 300)]
            _, lgs = m.decode([seq], [got[-1]], want_logits=True)
            out.append((got, lgs[0].view(np.uint32).copy()))
            m.seq_free(seq)
        m.close()
        assert out[0][0] == out[1][0] and np.array_equal(out[0][1], out[1][1]), name
        runs[name] = out[0]
    assert runs["handover"][0][:32] == want
    for name in ("barriers", "handover_h"):
        assert runs["handover"][0] == runs[name][0], name
        assert np.array_equal(runs["handover"][1], runs[name][1]), name


def test_long_context_decode_does_not_lap_the_ring(pkg, eng, model_dir):
    """Beyond ~3800 positions the KV prefetch issued before the QKV row loop staggers the warps of a CTA by more than one
    weight tile.  With tiles of fewer row pairs than row workers a worker without a pair in a tile used to hold nothing,
    the producer could refill the slot before that worker had looked at it, and its parity wait then matched the wrong
    phase: the kernel deadlocked (reported as B2_ERR_CUDA by the spin guards after 4 s).  Decode across that range, and
    compare with the per-matrix kernels."""
    path = pkg.synth.cached_model("wide-llama-2l-long", model_dir, source="blocks")
    toks = pkg.synth.prompt_tokens(3790, 2048, seed=9)
    got = {}
    for flags in (0, pkg.engine.LOAD_NO_MEGA):
        m = eng.load(path, max_batch=1, max_ctx=3900, flags=flags)
        seq = m.seq_new()
        m.prefill(seq, toks, want_logits=False)
        got[flags] = [int(t) for t in m.decode_greedy(seq, 60, first_tok=1)]
        m.close()
    assert got[0] == got[pkg.engine.LOAD_NO_MEGA]


def test_single_sequence_decode_graph_matches_the_general_path(pkg, eng, model_dir):
    """b2_decode for one sequence without logits is one graph launch whose step kernels read the token from and write
    the new state to pinned host memory; it must produce the tokens of the general path (and of decode_greedy)."""
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    m = eng.load(path, max_batch=2, max_ctx=128)
    toks = pkg.synth.prompt_tokens(12, m.info.n_vocab, seed=3)
    a, b, c = m.seq_new(), m.seq_new(), None
    la, lb = m.prefill(a, toks), m.prefill(b, toks)
    ta = tb = int(np.argmax(la))
    out_a, out_b = [], []
    for _ in range(40):
        ta = m.decode_one(a, ta); out_a.append(ta)
        nb, _ = m.decode([b], [tb], want_logits=True); tb = int(nb[0]); out_b.append(tb)
    assert out_a == out_b
    m.seq_truncate(a, len(toks))
    assert [int(t) for t in m.decode_greedy(a, 40, first_tok=int(np.argmax(la)))] == out_a
    m.close()
