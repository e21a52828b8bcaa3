"""Batched steps (several sequences per b2_decode call, prompts in chunks of 32 tokens) must give
the bits of the one-token-at-a-time path: llama.cpp's batch evaluation quantises and multiplies
every activation row on its own, so batching may change the schedule but not one result."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(pkg):
    e = pkg.engine.Engine(0)
    yield e
    e.close()


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.mark.parametrize("preset,n_prompt", [("tiny-llama", 3), ("tiny-llama", 32), ("tiny-llama", 33), ("tiny-llama", 101), ("tiny-phi2", 70)])
def test_chunked_prefill_equals_token_by_token(pkg, eng, model_dir, preset, n_prompt):
    E = pkg.engine
    path = pkg.synth.cached_model(preset, model_dir)
    ma = eng.load(path, max_batch=1, max_ctx=256)
    mb = eng.load(path, max_batch=1, max_ctx=256, flags=E.LOAD_NO_BATCH)
    toks = pkg.synth.prompt_tokens(n_prompt, ma.info.n_vocab, seed=n_prompt)
    sa, sb = ma.seq_new(), mb.seq_new()
    la, lb = ma.prefill(sa, toks), mb.prefill(sb, toks)
    assert ma.stats().last_steps < n_prompt or n_prompt < 3        # the chunked path really ran
    assert mb.stats().last_steps == n_prompt
    assert np.array_equal(_bits(la), _bits(lb)), np.abs(la - lb).max()
    for layer in range(ma.info.n_layer):
        ka, va = ma.kv_read(sa, layer, n_prompt)
        kb, vb = mb.kv_read(sb, layer, n_prompt)
        assert np.array_equal(ka.view(np.uint16), kb.view(np.uint16)) and np.array_equal(va.view(np.uint16), vb.view(np.uint16)), layer
    # a second prompt appended to the same sequence (positions continue mid-page), then greedy decoding
    more = pkg.synth.prompt_tokens(37, ma.info.n_vocab, seed=99)
    assert np.array_equal(_bits(ma.prefill(sa, more)), _bits(mb.prefill(sb, more)))
    assert np.array_equal(ma.decode_greedy(sa, 12), mb.decode_greedy(sb, 12))
    ma.close(); mb.close()


@pytest.mark.parametrize("preset,n_seq", [("tiny-llama", 3), ("tiny-llama", 8), ("tiny-llama", 32), ("tiny-llama", 37), ("tiny-phi2", 5)])
def test_batched_decode_equals_sequential(pkg, eng, model_dir, preset, n_seq):
    E = pkg.engine
    path = pkg.synth.cached_model(preset, model_dir)
    ma = eng.load(path, max_batch=n_seq, max_ctx=128)
    mb = eng.load(path, max_batch=n_seq, max_ctx=128, flags=E.LOAD_NO_BATCH)
    V = ma.info.n_vocab
    seqs_a, seqs_b, tok = [], [], []
    for i in range(n_seq):                      # ragged histories: 1 .. 70 cached tokens, crossing a page boundary
        n = 1 + (i * 23) % 70
        p = pkg.synth.prompt_tokens(n, V, seed=100 + i)
        a, b = ma.seq_new(), mb.seq_new()
        la, lb = ma.prefill(a, p), mb.prefill(b, p)
        assert np.array_equal(_bits(la), _bits(lb))
        seqs_a.append(a); seqs_b.append(b); tok.append(int(np.argmax(la)))
    for step in range(6):
        na, lga = ma.decode(seqs_a, tok, want_logits=True)
        assert ma.stats().last_steps == (n_seq + 31) // 32
        nb, lgb = mb.decode(seqs_b, tok, want_logits=True)
        assert mb.stats().last_steps == n_seq
        assert np.array_equal(_bits(lga), _bits(lgb)), (step, np.abs(lga - lgb).max())
        assert np.array_equal(na, nb)
        assert np.array_equal(na, np.argmax(lga, axis=1))
        tok = [int(t) for t in na]
    # a subset of the sequences, in another order, still decodes alone
    sub = [2, 0]
    na, _ = ma.decode([seqs_a[i] for i in sub], [tok[i] for i in sub], want_logits=True)
    nb, _ = mb.decode([seqs_b[i] for i in sub], [tok[i] for i in sub], want_logits=True)
    assert np.array_equal(na, nb)
    ma.close(); mb.close()


def test_batched_logits_match_oracle(pkg, oracle, eng, model_dir):
    """The batched path against the CPU oracle directly (canonical mode): bit-equal logits."""
    path = pkg.synth.cached_model("tiny-llama", model_dir)
    m = eng.load(path, max_batch=4, max_ctx=128)
    V = m.info.n_vocab
    oms, seqs, tok = [], [], []
    for i in range(4):
        om = oracle.OracleModel(path, faithful=2)
        p = pkg.synth.prompt_tokens(40 + 5 * i, V, seed=7 + i)
        s = m.seq_new()
        lg = m.prefill(s, p)
        for t in p:
            want = om.eval(int(t))
        assert np.array_equal(_bits(lg), _bits(want)), np.abs(lg - want).max()
        oms.append(om); seqs.append(s); tok.append(int(np.argmax(want)))
    for step in range(5):
        nxt, lgs = m.decode(seqs, tok, want_logits=True)
        for i, om in enumerate(oms):
            want = om.eval(tok[i])
            assert np.array_equal(_bits(lgs[i]), _bits(want)), (step, i)
        tok = [int(t) for t in nxt]
    m.close()
