"""Synthetic random-init GGUF models for tests and benchmarks.

The reference ships no weights: its pods pull them at run time through
`ollama pull <image>` (reference pkg/model/pod.go:68-83) and there is no network
here, so every model this project runs is written by this module as a GGUF v3
file with the tensor names and metadata keys llama.cpp expects
(`blk.{i}.attn_q.weight` ..., `llama.block_count` ...).

Two weight sources:
  * "gauss": f32 N(0, 1/fan_in) weights pushed through the type's reference
    quantiser (`gguf.quants.quantize`; Q4_0 / Q8_0 only).
  * "blocks": valid random block bytes with scales picked so the dequantised
    weights have the same standard deviation.  Used for K-quants (the `gguf`
    package has no Q4_K/Q6_K quantiser) and for multi-GB benchmark models,
    where it runs at memory speed.
"""
from __future__ import annotations

import dataclasses
import os
import zlib

import numpy as np

import gguf
from gguf import GGMLQuantizationType as T

QK_K = 256
BLOCK = {T.Q4_0: (32, 18), T.Q8_0: (32, 34), T.Q4_K: (256, 144), T.Q6_K: (256, 210)}


@dataclasses.dataclass
class ModelConfig:
    arch: str = "llama"          # "llama" | "phi2"
    n_layer: int = 2
    n_embd: int = 256
    n_head: int = 4
    n_head_kv: int = 4
    n_ff: int = 512
    n_vocab: int = 512
    n_ctx: int = 256
    n_rot: int | None = None     # default head_dim (llama) / given (phi2)
    norm_eps: float = 1e-5
    rope_freq_base: float = 10000.0
    rope_scaling: float = 1.0    # > 1: `rope.scaling.type = linear` with this factor (positions divided by it)
    rope_factors: bool = False   # write `rope_freqs.weight` (Llama-3.1-style per-pair frequency factors)
    ftype: str = "q4_0"          # "q4_0" | "q4_k_m" | "q8_0"
    head_type: str | None = None  # default: q6_k when n_embd % 256 == 0 else body type
    name: str = "synthetic"

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head


# The public model shapes of SURVEY.md section 8.
PRESETS = {
    "tiny-llama": ModelConfig(),
    "tiny-gqa": ModelConfig(n_head=8, n_head_kv=2, n_embd=512, n_ff=1024, ftype="q4_k_m"),
    "tiny-phi2": ModelConfig(arch="phi2", n_embd=256, n_head=4, n_head_kv=4, n_ff=1024, n_rot=32),
    "tiny-phi2-80": ModelConfig(arch="phi2", n_embd=320, n_head=4, n_head_kv=4, n_ff=1280, n_rot=32),      # phi-2's head_dim 80
    "tiny-rope31": ModelConfig(rope_freq_base=500000.0, rope_scaling=4.0, rope_factors=True),   # Llama-3.1-style rope_freqs + linear scaling
    "wide-llama-2l": ModelConfig(n_layer=2, n_embd=4096, n_head=32, n_head_kv=32, n_ff=11008, n_vocab=2048, n_ctx=512),
    "wide-llama-2l-long": ModelConfig(n_layer=2, n_embd=4096, n_head=32, n_head_kv=32, n_ff=11008, n_vocab=2048, n_ctx=4096),     # ... with room for a long context      # 7B's widths, two layers (SURVEY 8c iv)
    "wide13-llama-1l": ModelConfig(n_layer=1, n_embd=5120, n_head=40, n_head_kv=40, n_ff=13824, n_vocab=1024, n_ctx=256),     # 13B's widths: five activation units per lane, tiles smaller than a worker round
    "tiny-gqa8": ModelConfig(n_head=16, n_head_kv=2, n_embd=2048, n_ff=2048),      # 70B's head layout (8 query heads per KV head, head_dim 128), Q4_0
    "llama2-7b": ModelConfig(n_layer=32, n_embd=4096, n_head=32, n_head_kv=32, n_ff=11008, n_vocab=32000, n_ctx=4096, name="llama2-7b"),
    "mistral-7b": ModelConfig(n_layer=32, n_embd=4096, n_head=32, n_head_kv=8, n_ff=14336, n_vocab=32000, n_ctx=4096, ftype="q4_k_m", name="mistral-7b"),
    "llama2-13b": ModelConfig(n_layer=40, n_embd=5120, n_head=40, n_head_kv=40, n_ff=13824, n_vocab=32000, n_ctx=4096, name="llama2-13b"),
    "llama2-70b": ModelConfig(n_layer=80, n_embd=8192, n_head=64, n_head_kv=8, n_ff=28672, n_vocab=32000, n_ctx=4096, name="llama2-70b"),
    "phi-2": ModelConfig(arch="phi2", n_layer=32, n_embd=2560, n_head=32, n_head_kv=32, n_ff=10240, n_vocab=51200, n_ctx=2048, n_rot=32, name="phi-2"),
}


def _rng(seed: int, name: str) -> np.random.Generator:
    return np.random.default_rng([seed, zlib.crc32(name.encode())])


def _f16_bytes(a: np.ndarray) -> np.ndarray:
    return a.astype(np.float16).view(np.uint8).reshape(a.shape + (2,))


def random_blocks(qtype: T, rows: int, k: int, std: float, rng: np.random.Generator) -> np.ndarray:
    """Valid random blocks of `qtype`, shape (rows, bytes_per_row) uint8, dequantised std ~ `std`."""
    bs, ts = BLOCK[qtype]
    assert k % bs == 0, (qtype, k)
    nb = rows * (k // bs)
    out = np.empty((nb, ts), dtype=np.uint8)
    jit = rng.uniform(0.8, 1.2, size=nb).astype(np.float32)
    if qtype == T.Q4_0:
        sign = rng.integers(0, 2, size=nb).astype(np.float32) * 2 - 1
        out[:, 0:2] = _f16_bytes(sign * jit * (std / 4.61))
        out[:, 2:] = rng.integers(0, 256, size=(nb, 16), dtype=np.uint8)
    elif qtype == T.Q8_0:
        out[:, 0:2] = _f16_bytes(jit * (std / 73.6))
        out[:, 2:] = rng.integers(0, 256, size=(nb, 32), dtype=np.uint8)
    elif qtype == T.Q4_K:
        d = jit * (std / (40.0 * 4.61))
        out[:, 0:2] = _f16_bytes(d)
        out[:, 2:4] = _f16_bytes(d * 7.5)
        sc = rng.integers(16, 64, size=(nb, 8), dtype=np.int32)
        mn = np.clip(sc + rng.integers(-2, 3, size=(nb, 8)), 0, 63).astype(np.int32)
        s = np.zeros((nb, 12), dtype=np.int32)
        s[:, 0:4] = sc[:, 0:4] | ((sc[:, 4:8] >> 4) << 6)
        s[:, 4:8] = mn[:, 0:4] | ((mn[:, 4:8] >> 4) << 6)
        s[:, 8:12] = (sc[:, 4:8] & 0xF) | ((mn[:, 4:8] & 0xF) << 4)
        out[:, 4:16] = s.astype(np.uint8)
        out[:, 16:] = rng.integers(0, 256, size=(nb, 128), dtype=np.uint8)
    elif qtype == T.Q6_K:
        out[:, 0:192] = rng.integers(0, 256, size=(nb, 192), dtype=np.uint8)
        mag = rng.integers(32, 128, size=(nb, 16), dtype=np.int32)
        sgn = rng.integers(0, 2, size=(nb, 16), dtype=np.int32) * 2 - 1
        out[:, 192:208] = (mag * sgn).astype(np.int8).view(np.uint8)
        out[:, 208:210] = _f16_bytes(jit * (std / (82.0 * 18.47)))
    else:
        raise ValueError(qtype)
    return out.reshape(rows, (k // bs) * ts)


def make_weight(qtype: T, rows: int, k: int, std: float, rng: np.random.Generator, source: str) -> np.ndarray:
    if source == "gauss" and qtype in (T.Q4_0, T.Q8_0):
        w = (rng.standard_normal((rows, k), dtype=np.float32) * np.float32(std))
        return gguf.quants.quantize(w, qtype)
    return random_blocks(qtype, rows, k, std, rng)


def _more_bits(i: int, n: int) -> bool:
    # llama.cpp's Q4_K_M rule for attn_v / ffn_down [upstream use_more_bits]
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_types(cfg: ModelConfig) -> dict[str, T]:
    """ggml type of every quantised tensor for the file type."""
    body = {"q4_0": T.Q4_0, "q8_0": T.Q8_0, "q4_k_m": T.Q4_K}[cfg.ftype]
    kq_ok = cfg.n_embd % QK_K == 0
    head = {"q6_k": T.Q6_K, "q4_0": T.Q4_0, "q8_0": T.Q8_0, "q4_k": T.Q4_K, None: (T.Q6_K if kq_ok else body)}[cfg.head_type]
    types = {"token_embd": body, "output": head}
    for i in range(cfg.n_layer):
        hi = T.Q6_K if (cfg.ftype == "q4_k_m" and _more_bits(i, cfg.n_layer)) else body
        p = f"blk.{i}."
        if cfg.arch == "llama":
            types.update({p + "attn_q": body, p + "attn_k": body, p + "attn_v": hi, p + "attn_output": body,
                          p + "ffn_gate": body, p + "ffn_up": body,
                          p + "ffn_down": hi if cfg.n_ff % QK_K == 0 or hi != T.Q6_K else body})
        else:
            types.update({p + "attn_qkv": body, p + "attn_output": body, p + "ffn_up": body, p + "ffn_down": body})
    return types


def write_gguf(path: str, cfg: ModelConfig, seed: int = 0x5EED, source: str = "gauss", head_scale: float = 4.0) -> str:
    """Write a random-init model; returns `path`."""
    E, H, HK, hd, F, V = cfg.n_embd, cfg.n_head, cfg.n_head_kv, cfg.head_dim, cfg.n_ff, cfg.n_vocab
    kv = HK * hd
    n_rot = cfg.n_rot or hd
    w = gguf.GGUFWriter(path, cfg.arch)
    w.add_name(cfg.name)
    w.add_block_count(cfg.n_layer)
    w.add_context_length(cfg.n_ctx)
    w.add_embedding_length(E)
    w.add_feed_forward_length(F)
    w.add_head_count(H)
    w.add_head_count_kv(HK)
    w.add_rope_dimension_count(n_rot)
    w.add_rope_freq_base(cfg.rope_freq_base)
    if cfg.rope_scaling != 1.0:
        w.add_rope_scaling_type(gguf.RopeScalingType.LINEAR)
        w.add_rope_scaling_factor(cfg.rope_scaling)
    if cfg.arch == "llama":
        w.add_layer_norm_rms_eps(cfg.norm_eps)
    else:
        w.add_layer_norm_eps(cfg.norm_eps)
    w.add_vocab_size(V)
    w.add_file_type({"q4_0": 2, "q8_0": 7, "q4_k_m": 15}[cfg.ftype])
    if V >= 300:
        # A SentencePiece-style vocabulary that can spell any text: <unk> <s> </s>, the 256 byte-fallback pieces, then
        # filler word pieces (so that a server can be driven with text against a random-init model).
        toks = ["<unk>", "<s>", "</s>"] + [f"<0x{b:02X}>" for b in range(256)] + [f"\u2581w{i}" for i in range(V - 259)]
        w.add_tokenizer_model("llama")
        w.add_token_list(toks)
        w.add_token_scores([0.0] * 259 + [-float(i) for i in range(V - 259)])
        w.add_token_types([2, 3, 3] + [6] * 256 + [1] * (V - 259))
        w.add_bos_token_id(1); w.add_eos_token_id(2); w.add_unk_token_id(0); w.add_add_bos_token(True)
    types = tensor_types(cfg)

    def quant(name: str, rows: int, k: int, std: float):
        t = types[name]
        w.add_tensor(name + ".weight", make_weight(t, rows, k, std, _rng(seed, name), source), raw_dtype=t)

    def vec(name: str, n: int, mean: float, std: float):
        w.add_tensor(name, (mean + std * _rng(seed, name).standard_normal(n)).astype(np.float32))

    quant("token_embd", V, E, 1.0)
    for i in range(cfg.n_layer):
        p = f"blk.{i}."
        vec(p + "attn_norm.weight", E, 1.0, 0.02)
        if cfg.arch == "llama":
            quant(p + "attn_q", H * hd, E, E ** -0.5)
            quant(p + "attn_k", kv, E, E ** -0.5)
            quant(p + "attn_v", kv, E, E ** -0.5)
            quant(p + "attn_output", E, H * hd, (H * hd) ** -0.5)
            vec(p + "ffn_norm.weight", E, 1.0, 0.02)
            quant(p + "ffn_gate", F, E, E ** -0.5)
            quant(p + "ffn_up", F, E, E ** -0.5)
            quant(p + "ffn_down", E, F, F ** -0.5)
        else:
            vec(p + "attn_norm.bias", E, 0.0, 0.02)
            quant(p + "attn_qkv", H * hd + 2 * kv, E, E ** -0.5)
            vec(p + "attn_qkv.bias", H * hd + 2 * kv, 0.0, 0.02)
            quant(p + "attn_output", E, H * hd, (H * hd) ** -0.5)
            vec(p + "attn_output.bias", E, 0.0, 0.02)
            quant(p + "ffn_up", F, E, E ** -0.5)
            vec(p + "ffn_up.bias", F, 0.0, 0.02)
            quant(p + "ffn_down", E, F, F ** -0.5)
            vec(p + "ffn_down.bias", E, 0.0, 0.02)
    vec("output_norm.weight", E, 1.0, 0.02)
    if cfg.rope_factors:       # low frequencies stretched by up to 8x, high ones kept (the shape of the Llama-3.1 factors)
        w.add_tensor("rope_freqs.weight", np.linspace(1.0, 8.0, n_rot // 2).astype(np.float32))
    if cfg.arch == "phi2":
        vec("output_norm.bias", E, 0.0, 0.02)
    # head scaled so logits have std of a few units: keeps greedy margins well above tolerance
    quant("output", V, E, head_scale * E ** -0.5)
    if cfg.arch == "phi2":
        vec("output.bias", V, 0.0, 0.02)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return path


def cached_model(preset: str, directory: str | None = None, **kw) -> str:
    """Write the preset once under `directory` (default $B2_MODEL_DIR or /tmp/b2gguf-models)."""
    directory = directory or os.environ.get("B2_MODEL_DIR", "/tmp/b2gguf-models")
    os.makedirs(directory, exist_ok=True)
    cfg = PRESETS[preset]
    source = kw.get("source", "gauss")
    path = os.path.join(directory, f"{preset}-{cfg.ftype}-{source}-{kw.get('seed', 0x5EED):x}.gguf")
    if not os.path.exists(path):
        tmp = path + f".tmp{os.getpid()}"
        write_gguf(tmp, cfg, **kw)
        os.replace(tmp, path)
    return path


def prompt_tokens(n: int, n_vocab: int, seed: int = 1234) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, n_vocab, size=n, dtype=np.int32)
