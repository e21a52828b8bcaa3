"""ctypes front end of the CPU oracle (oracle/ggml_ref.c) -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
legs import this module.  Parity status: see oracle/ggml_ref.h ("parity unpinned"
against the reference, which holds no vectors for this path; block formats pinned
against the in-container `gguf` package through tests/golden/).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, Q8_K = 0, 1, 2, 8, 12, 14, 15
ARCH = {"llama": 0, "phi2": 1}


def build(force: bool = False) -> str:
    src = [os.path.join(HERE, f) for f in ("ggml_ref.c", "ggml_ref.h", "Makefile")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in src):
        subprocess.run(["make", "-C", HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return LIB


class Tensor(C.Structure):
    _fields_ = [("type", C.c_int), ("data", C.c_void_p)]


class Layer(C.Structure):
    _fields_ = [(n, Tensor) for n in ("attn_norm", "attn_norm_b", "wq", "wk", "wv", "wo", "bq", "bo", "ffn_norm",
                                      "w_gate", "w_up", "w_down", "b_up", "b_down")]


class Model(C.Structure):
    _fields_ = [("arch", C.c_int)] + [(n, C.c_int) for n in (
        "n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab", "n_ctx", "n_rot")] + [
        ("norm_eps", C.c_float), ("rope_freq_base", C.c_float),
        ("tok_embd", Tensor), ("out_norm", Tensor), ("out_norm_b", Tensor), ("output", Tensor), ("output_b", Tensor),
        ("layers", C.POINTER(Layer)), ("rope_freq_scale", C.c_float), ("rope_factors", C.c_void_p)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp, i64, fp = C.c_void_p, C.c_int64, C.POINTER(C.c_float)
        L.ref_row_size.restype = C.c_size_t
        L.ref_row_size.argtypes = [C.c_int, i64]
        L.ref_dequantize_row.argtypes = [C.c_int, vp, vp, i64]
        for n in ("q4_0", "q8_0", "q8_K"):
            getattr(L, "ref_quantize_row_" + n).argtypes = [vp, vp, i64]
        L.ref_make_q4_K.argtypes = [vp, vp, i64]
        L.ref_make_q6_K.argtypes = [vp, vp, i64]
        for n in ("q4_0_q8_0", "q8_0_q8_0", "q4_K_q8_K", "q6_K_q8_K"):
            f = getattr(L, "ref_vec_dot_" + n)
            f.restype = C.c_float
            f.argtypes = [i64, vp, vp]
        L.ref_mul_mat_vec.argtypes = [C.c_int, vp, i64, i64, vp, vp, C.c_int, C.c_int]
        L.ref_rms_norm.argtypes = [vp, vp, vp, i64, C.c_float]
        L.ref_layer_norm.argtypes = [vp, vp, vp, vp, i64, C.c_float]
        L.ref_rope.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]
        L.ref_rope_ext.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, vp]
        L.ref_softmax.argtypes = [vp, i64]
        L.ref_attention_canon.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp]
        L.ref_ctx_new.restype = vp
        L.ref_ctx_new.argtypes = [C.POINTER(Model), C.c_int, C.c_int]
        L.ref_ctx_free.argtypes = [vp]
        L.ref_ctx_set_kv_f16.argtypes = [vp, C.c_int, vp, vp, C.c_int]
        L.ref_eval.restype = C.c_int
        L.ref_eval.argtypes = [vp, C.c_int32, C.c_int, vp]
        L.ref_last_hidden.restype = fp
        L.ref_last_hidden.argtypes = [vp]
        L.ref_num_threads.restype = C.c_int
        L.ref_set_num_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a: np.ndarray) -> int:
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def row_size(t: int, k: int) -> int:
    return lib().ref_row_size(t, k)


def dequantize(t: int, blocks: np.ndarray, k: int) -> np.ndarray:
    """blocks: uint8 (rows, row_size) or flat -> float32 (rows, k)."""
    b = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, row_size(t, k))
    y = np.empty((b.shape[0], k), dtype=np.float32)
    for r in range(b.shape[0]):
        lib().ref_dequantize_row(t, _p(b[r]), _p(y[r]), k)
    return y


def quantize(t: int, x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, x.shape[-1])
    k = x.shape[1]
    fn = {Q4_0: lib().ref_quantize_row_q4_0, Q8_0: lib().ref_quantize_row_q8_0, Q8_K: lib().ref_quantize_row_q8_K,
          Q4_K: lib().ref_make_q4_K, Q6_K: lib().ref_make_q6_K}[t]
    out = np.zeros((x.shape[0], row_size(t, k)), dtype=np.uint8)
    for r in range(x.shape[0]):
        fn(_p(x[r]), _p(out[r]), k)
    return out


def mul_mat_vec(t: int, W: np.ndarray, rows: int, k: int, x: np.ndarray, faithful: int = 1, threads: int = 0) -> np.ndarray:
    W = np.ascontiguousarray(W, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty(rows, dtype=np.float32)
    lib().ref_mul_mat_vec(t, _p(W), rows, k, _p(x), _p(y), int(faithful), threads)
    return y


def rms_norm(x, g, eps):
    x = np.ascontiguousarray(x, np.float32); g = np.ascontiguousarray(g, np.float32)
    y = np.empty_like(x)
    lib().ref_rms_norm(_p(x), _p(g), _p(y), x.size, eps)
    return y


def rope(x, n_head, head_dim, n_rot, pos, freq_base, mode):
    x = np.array(x, dtype=np.float32, order="C")
    lib().ref_rope(_p(x), n_head, head_dim, n_rot, pos, freq_base, mode)
    return x


def rope_ext(x, n_head, head_dim, n_rot, pos, freq_base, mode, freq_scale=1.0, factors=None):
    x = np.ascontiguousarray(x, dtype=np.float32).copy()
    f = None if factors is None else np.ascontiguousarray(factors, dtype=np.float32)
    lib().ref_rope_ext(_p(x), n_head, head_dim, n_rot, pos, freq_base, mode, freq_scale, _p(f) if f is not None else None)
    return x


def attention_canon(q, k, v, n_head, n_head_kv, hd, scale):
    """Canonical-mode attention of one token: q f32 [n_head*hd], k / v f16 [n][n_head_kv*hd]."""
    q = np.ascontiguousarray(q, np.float32); k = np.ascontiguousarray(k, np.float16); v = np.ascontiguousarray(v, np.float16)
    out = np.empty(n_head * hd, dtype=np.float32)
    lib().ref_attention_canon(_p(q), _p(k), _p(v), k.shape[0], n_head, n_head_kv, hd, scale, _p(out))
    return out


class OracleModel:
    """A GGUF file opened with the `gguf` package's reader and evaluated by the C oracle."""

    def __init__(self, path: str, faithful: int = 1, threads: int = 0, n_ctx: int | None = None):
        """faithful: 0 ideal f32, 1 ggml generic C order, 2 canonical (f64 term accumulation; see ggml_ref.h),
        3 ggml AVX2 kernel order for Q4_0/Q8_0 matrices (a second legitimate ggml build: noise-floor measurements)."""
        import gguf
        self.reader = r = gguf.GGUFReader(path)
        self.tensors = {t.name: t for t in r.tensors}

        def kv(key, default=None):
            f = r.fields.get(key)
            if f is None:
                return default
            return f.contents()

        arch = kv("general.architecture")
        self.arch = arch
        m = Model()
        m.arch = ARCH[arch]
        m.n_layer = kv(f"{arch}.block_count")
        m.n_embd = kv(f"{arch}.embedding_length")
        m.n_head = kv(f"{arch}.attention.head_count")
        m.n_head_kv = kv(f"{arch}.attention.head_count_kv", m.n_head)
        m.head_dim = m.n_embd // m.n_head
        m.n_ff = kv(f"{arch}.feed_forward_length")
        m.n_ctx = n_ctx or kv(f"{arch}.context_length")
        m.n_rot = kv(f"{arch}.rope.dimension_count", m.head_dim)
        m.norm_eps = kv(f"{arch}.attention.layer_norm_rms_epsilon", None) or kv(f"{arch}.attention.layer_norm_epsilon", 1e-5)
        m.rope_freq_base = kv(f"{arch}.rope.freq_base", 10000.0)
        m.n_vocab = int(self.tensors["token_embd.weight"].shape[1])
        self._keep = []
        # rope.scaling.type: 0 none, 1 linear, 2 yarn (gguf.RopeScalingType); stored as a string
        st = kv(f"{arch}.rope.scaling.type", "none")
        if st not in ("none", "linear"):
            raise NotImplementedError(f"rope scaling type {st}")
        m.rope_freq_scale = 1.0 / float(kv(f"{arch}.rope.scaling.factor", 1.0)) if st == "linear" else 1.0
        rf = self.tensors.get("rope_freqs.weight")
        if rf is not None:
            a = np.ascontiguousarray(rf.data, dtype=np.float32)
            self._keep.append(a)
            m.rope_factors = a.ctypes.data

        def T(name, optional=False):
            t = self.tensors.get(name)
            if t is None:
                assert optional, name
                return Tensor(0, None)
            a = np.ascontiguousarray(t.data)
            self._keep.append(a)
            return Tensor(int(t.tensor_type), a.ctypes.data)

        m.tok_embd = T("token_embd.weight")
        m.out_norm = T("output_norm.weight")
        m.out_norm_b = T("output_norm.bias", True)
        m.output = T("output.weight", True)
        if not m.output.data:
            m.output = m.tok_embd
        m.output_b = T("output.bias", True)
        layers = (Layer * m.n_layer)()
        for i in range(m.n_layer):
            p, L = f"blk.{i}.", layers[i]
            L.attn_norm = T(p + "attn_norm.weight")
            L.attn_norm_b = T(p + "attn_norm.bias", True)
            if arch == "phi2":
                L.wq = T(p + "attn_qkv.weight")
                L.bq = T(p + "attn_qkv.bias", True)
            else:
                L.wq, L.wk, L.wv = T(p + "attn_q.weight"), T(p + "attn_k.weight"), T(p + "attn_v.weight")
                L.ffn_norm = T(p + "ffn_norm.weight")
                L.w_gate = T(p + "ffn_gate.weight")
            L.wo = T(p + "attn_output.weight")
            L.bo = T(p + "attn_output.bias", True)
            L.w_up = T(p + "ffn_up.weight")
            L.w_down = T(p + "ffn_down.weight")
            L.b_up = T(p + "ffn_up.bias", True)
            L.b_down = T(p + "ffn_down.bias", True)
        self._layers = layers
        m.layers = C.cast(layers, C.POINTER(Layer))
        self.m = m
        self.faithful = faithful
        self.ctx = lib().ref_ctx_new(C.byref(m), int(faithful), threads)
        self.pos = 0

    def close(self):
        if self.ctx:
            lib().ref_ctx_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_kv(self, il: int, k: np.ndarray, v: np.ndarray):
        """k, v: float16 arrays [n_tok][n_head_kv*head_dim] (e.g. read back from the GPU cache)."""
        k = np.ascontiguousarray(k, dtype=np.float16); v = np.ascontiguousarray(v, dtype=np.float16)
        lib().ref_ctx_set_kv_f16(self.ctx, il, _p(k), _p(v), k.shape[0])

    def eval(self, token: int, pos: int | None = None, want_logits: bool = True):
        pos = self.pos if pos is None else pos
        logits = np.empty(self.m.n_vocab, dtype=np.float32) if want_logits else None
        rc = lib().ref_eval(self.ctx, int(token), pos, _p(logits) if want_logits else None)
        if rc != 0:
            raise ValueError(f"ref_eval({token}, {pos}) failed: {rc}")
        self.pos = pos + 1
        return logits

    def hidden(self) -> np.ndarray:
        return np.ctypeslib.as_array(lib().ref_last_hidden(self.ctx), shape=(self.m.n_embd,)).copy()

    def generate(self, prompt, n_new: int):
        """Greedy decode; returns (tokens, logits of every generated step)."""
        logits = None
        for t in prompt:
            logits = self.eval(int(t))
        toks, all_logits = [], []
        for _ in range(n_new):
            nxt = int(np.argmax(logits))
            toks.append(nxt); all_logits.append(logits)
            logits = self.eval(nxt)
        return toks, all_logits
