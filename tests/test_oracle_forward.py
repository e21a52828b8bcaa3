"""Graph-wiring cross-check of the oracle's forward pass against an independent
implementation: HuggingFace `LlamaForCausalLM` / `PhiForCausalLM` (transformers is in the
image) fed the same dequantised weights.  This pins the ORDER of operations of
llama.cpp's build_llama / build_phi2 as restated in oracle/ggml_ref.c, not the reference
(which holds no vectors: "parity unpinned")."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


def _deq(om, name):
    import gguf
    t = om.tensors[name]
    a = gguf.quants.dequantize(np.ascontiguousarray(t.data), t.tensor_type)
    return torch.from_numpy(np.array(a, dtype=np.float32))


def _unpermute(w, n_head):
    # GGUF llama stores q/k rows so that adjacent-pair rope == HF's rotate-half rope
    rows, cols = w.shape
    hd = rows // n_head
    return w.reshape(n_head, hd // 2, 2, cols).transpose(1, 2).reshape(rows, cols)


def _hf_llama(om):
    from transformers import LlamaConfig, LlamaForCausalLM
    m = om.m
    cfg = LlamaConfig(hidden_size=m.n_embd, intermediate_size=m.n_ff, num_hidden_layers=m.n_layer,
                      num_attention_heads=m.n_head, num_key_value_heads=m.n_head_kv, vocab_size=m.n_vocab,
                      rms_norm_eps=m.norm_eps, max_position_embeddings=m.n_ctx, tie_word_embeddings=False,
                      attention_bias=False, mlp_bias=False)
    try:
        cfg.rope_theta = m.rope_freq_base
    except Exception:
        pass
    hf = LlamaForCausalLM(cfg).float().eval()
    sd = {"model.embed_tokens.weight": _deq(om, "token_embd.weight"), "model.norm.weight": _deq(om, "output_norm.weight"),
          "lm_head.weight": _deq(om, "output.weight")}
    for i in range(m.n_layer):
        p, q = f"blk.{i}.", f"model.layers.{i}."
        sd[q + "input_layernorm.weight"] = _deq(om, p + "attn_norm.weight")
        sd[q + "post_attention_layernorm.weight"] = _deq(om, p + "ffn_norm.weight")
        sd[q + "self_attn.q_proj.weight"] = _unpermute(_deq(om, p + "attn_q.weight"), m.n_head)
        sd[q + "self_attn.k_proj.weight"] = _unpermute(_deq(om, p + "attn_k.weight"), m.n_head_kv)
        sd[q + "self_attn.v_proj.weight"] = _deq(om, p + "attn_v.weight")
        sd[q + "self_attn.o_proj.weight"] = _deq(om, p + "attn_output.weight")
        sd[q + "mlp.gate_proj.weight"] = _deq(om, p + "ffn_gate.weight")
        sd[q + "mlp.up_proj.weight"] = _deq(om, p + "ffn_up.weight")
        sd[q + "mlp.down_proj.weight"] = _deq(om, p + "ffn_down.weight")
    missing = hf.load_state_dict(sd, strict=False)
    assert not [k for k in missing.missing_keys if "rotary" not in k], missing
    return hf


def _hf_phi(om):
    from transformers import PhiConfig, PhiForCausalLM
    m = om.m
    cfg = PhiConfig(hidden_size=m.n_embd, intermediate_size=m.n_ff, num_hidden_layers=m.n_layer,
                    num_attention_heads=m.n_head, num_key_value_heads=m.n_head_kv, vocab_size=m.n_vocab,
                    layer_norm_eps=m.norm_eps, max_position_embeddings=m.n_ctx, tie_word_embeddings=False,
                    partial_rotary_factor=m.n_rot / m.head_dim, hidden_act="gelu_new", qk_layernorm=False,
                    resid_pdrop=0.0, embd_pdrop=0.0, attention_dropout=0.0)
    hf = PhiForCausalLM(cfg).float().eval()
    E, kv = m.n_embd, m.n_head_kv * m.head_dim
    sd = {"model.embed_tokens.weight": _deq(om, "token_embd.weight"),
          "model.final_layernorm.weight": _deq(om, "output_norm.weight"), "model.final_layernorm.bias": _deq(om, "output_norm.bias"),
          "lm_head.weight": _deq(om, "output.weight"), "lm_head.bias": _deq(om, "output.bias")}
    for i in range(m.n_layer):
        p, q = f"blk.{i}.", f"model.layers.{i}."
        sd[q + "input_layernorm.weight"] = _deq(om, p + "attn_norm.weight")
        sd[q + "input_layernorm.bias"] = _deq(om, p + "attn_norm.bias")
        w, b = _deq(om, p + "attn_qkv.weight"), _deq(om, p + "attn_qkv.bias")
        for n, lo, hi in (("q", 0, E), ("k", E, E + kv), ("v", E + kv, E + 2 * kv)):
            sd[q + f"self_attn.{n}_proj.weight"] = w[lo:hi].clone()
            sd[q + f"self_attn.{n}_proj.bias"] = b[lo:hi].clone()
        sd[q + "self_attn.dense.weight"] = _deq(om, p + "attn_output.weight")
        sd[q + "self_attn.dense.bias"] = _deq(om, p + "attn_output.bias")
        sd[q + "mlp.fc1.weight"] = _deq(om, p + "ffn_up.weight")
        sd[q + "mlp.fc1.bias"] = _deq(om, p + "ffn_up.bias")
        sd[q + "mlp.fc2.weight"] = _deq(om, p + "ffn_down.weight")
        sd[q + "mlp.fc2.bias"] = _deq(om, p + "ffn_down.bias")
    missing = hf.load_state_dict(sd, strict=False)
    assert not [k for k in missing.missing_keys if "rotary" not in k], missing
    return hf


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2"])
def test_ideal_forward_matches_huggingface(oracle, pkg, model_dir, preset):
    path = pkg.synth.cached_model(preset, model_dir)
    om = oracle.OracleModel(path, faithful=False)
    hf = _hf_phi(om) if om.arch == "phi2" else _hf_llama(om)
    toks = pkg.synth.prompt_tokens(12, om.m.n_vocab)
    with torch.no_grad():
        want = hf(torch.from_numpy(toks[None].astype(np.int64))).logits[0].numpy()
    for i, t in enumerate(toks):
        got = om.eval(int(t))
        assert np.max(np.abs(got - want[i])) < 2e-3, (i, np.max(np.abs(got - want[i])))


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2"])
def test_faithful_close_to_ideal(oracle, pkg, model_dir, preset):
    """W4A8 (ggml-faithful) vs f32-ideal: the gap is the quantisation noise floor that the
    2e-2 logit tolerance of the GPU parity tests has to be read against."""
    path = pkg.synth.cached_model(preset, model_dir)
    a, b = oracle.OracleModel(path, faithful=True), oracle.OracleModel(path, faithful=False)
    toks = pkg.synth.prompt_tokens(6, a.m.n_vocab)
    for t in toks:
        la, lb = a.eval(int(t)), b.eval(int(t))
    gap = float(np.max(np.abs(la - lb)))
    assert 0 < gap < 0.5, gap


def test_eval_rejects_bad_token_and_position(oracle, pkg, model_dir):
    om = oracle.OracleModel(pkg.synth.cached_model("tiny-llama", model_dir))
    with pytest.raises(ValueError):
        om.eval(om.m.n_vocab)
    with pytest.raises(ValueError):
        om.eval(0, pos=om.m.n_ctx)


def test_rope_factors_and_linear_scaling(oracle, pkg, model_dir):
    """ggml's rope with `rope_freqs.weight` and `rope.scaling.type = linear` (Llama-3.1 files): the angle of pair p at
    position pos is freq_scale * pos * base^(-2p/n_rot) / factor[p]; checked against a float64 restatement, and the
    synthetic Llama-3.1-style model must differ from the same weights read without the two settings."""
    rng = np.random.default_rng(5)
    hd, n_rot, pos, base = 64, 64, 37, 500000.0
    x = rng.standard_normal((2, hd)).astype(np.float32)
    fac = np.linspace(1.0, 8.0, n_rot // 2).astype(np.float32)
    got = oracle.rope_ext(x, 2, hd, n_rot, pos, base, 0, 0.25, fac)
    p = np.arange(n_rot // 2)
    th = 0.25 * pos * base ** (-2.0 * p / n_rot) / fac.astype(np.float64)
    want = x.astype(np.float64).copy()
    want[:, 0::2] = x[:, 0::2] * np.cos(th) - x[:, 1::2] * np.sin(th)
    want[:, 1::2] = x[:, 0::2] * np.sin(th) + x[:, 1::2] * np.cos(th)
    assert np.max(np.abs(got - want)) < 2e-4
    plain = oracle.rope_ext(x, 2, hd, n_rot, pos, base, 0)
    assert np.max(np.abs(plain - got)) > 1e-2
    om = oracle.OracleModel(pkg.synth.cached_model("tiny-rope31", model_dir))
    assert abs(om.m.rope_freq_scale - 0.25) < 1e-7 and om.m.rope_factors
