"""Stage-by-stage comparison of a 1-layer llama model with the oracle's pieces (debug aid)."""
import dataclasses
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle
import gguf

p = importlib.import_module("ollama-operator_b200")
cfg = dataclasses.replace(p.synth.PRESETS["tiny-llama"], n_layer=1)
path = "/tmp/dbg1.gguf"
p.synth.write_gguf(path, cfg)
om = oracle.OracleModel(path, faithful=2)
T = om.tensors


def W(name):
    t = T[name]
    return int(t.tensor_type), np.ascontiguousarray(t.data)


def mv(name, rows, k, x):
    t, w = W(name)
    return oracle.mul_mat_vec(t, w, rows, k, x, faithful=2)


E, H, HK, hd, F, V = cfg.n_embd, cfg.n_head, cfg.n_head_kv, cfg.head_dim, cfg.n_ff, cfg.n_vocab
emb = gguf.quants.dequantize(np.ascontiguousarray(T["token_embd.weight"].data), T["token_embd.weight"].tensor_type)
f32 = lambda n: np.array(T[n].data, dtype=np.float32)
eng = p.engine.Engine(0)
m = eng.load(path, max_batch=1, max_ctx=128)
seq = m.seq_new()
toks = [5, 77, 300]
ks, vs = [], []
for pos, tok in enumerate(toks):
    lg = m.prefill(seq, [tok])
    got = {k: m.debug_read(seq, k) for k in ("x", "q", "att", "h", "logits")}
    x0 = emb[tok]
    xn = oracle.rms_norm(x0, f32("blk.0.attn_norm.weight"), cfg.norm_eps)
    q = oracle.rope(mv("blk.0.attn_q.weight", H * hd, E, xn), H, hd, hd, pos, cfg.rope_freq_base, 0)
    k = oracle.rope(mv("blk.0.attn_k.weight", HK * hd, E, xn), HK, hd, hd, pos, cfg.rope_freq_base, 0)
    v = mv("blk.0.attn_v.weight", HK * hd, E, xn)
    ks.append(k.astype(np.float16)); vs.append(v.astype(np.float16))
    att = np.zeros(H * hd, np.float32)
    for h in range(H):
        hk = h // (H // HK)
        q16 = q[h * hd:(h + 1) * hd].astype(np.float16).astype(np.float64)
        s = np.array([float(np.dot(kk[hk * hd:(hk + 1) * hd].astype(np.float64), q16)) for kk in ks], np.float32) / np.float32(np.sqrt(hd))
        pr = np.exp(s - s.max()); pr = (pr / pr.sum()).astype(np.float16).astype(np.float64)
        att[h * hd:(h + 1) * hd] = sum(pr[t] * vs[t][hk * hd:(hk + 1) * hd].astype(np.float64) for t in range(len(ks)))
    x1 = x0 + mv("blk.0.attn_output.weight", E, H * hd, att)
    xn2 = oracle.rms_norm(x1, f32("blk.0.ffn_norm.weight"), cfg.norm_eps)
    g, u = mv("blk.0.ffn_gate.weight", F, E, xn2), mv("blk.0.ffn_up.weight", F, E, xn2)
    h_ = (g / (1 + np.exp(-g))) * u
    x2 = x1 + mv("blk.0.ffn_down.weight", E, F, h_.astype(np.float32))
    hid = oracle.rms_norm(x2.astype(np.float32), f32("output_norm.weight"), cfg.norm_eps)
    logits = mv("output.weight", V, E, hid)
    want_full = om.eval(tok)
    ref = {"q": q, "att": att, "h": h_, "x": x2, "logits": logits}
    print(f"pos {pos}: " + "  ".join(f"{k2} {np.max(np.abs(got[k2] - ref[k2])):.3e}/{np.max(np.abs(ref[k2])):.2f}" for k2 in ("q", "att", "h", "x", "logits")),
          f" oracle-eval vs pieces {np.max(np.abs(want_full - logits)):.2e}  gpu vs oracle-eval {np.max(np.abs(lg - want_full)):.3e}")
