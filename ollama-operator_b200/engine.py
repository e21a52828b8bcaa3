"""ctypes binding of libb200gguf.so (include/b2gguf.h).

Host-side mirror used by the tests, bench.py and the stand-in server; the reference's own
host language is Go (cgo binding shown in INTEGRATION.md), which this image cannot compile.
Nothing here computes: every call goes to the CUDA library, and a missing library or a
missing GPU raises instead of falling back.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200gguf.so")

TYPE_F32, TYPE_F16, TYPE_Q4_0, TYPE_Q8_0, TYPE_Q4_K, TYPE_Q6_K = 0, 1, 2, 8, 12, 14
LOAD_NO_PDL, LOAD_NO_GRAPH, LOAD_NO_MEGA, LOAD_NO_BATCH = 1, 2, 4, 8

# every symbol include/b2gguf.h declares (tests check the library exports all of them)
SYMBOLS = [
    "b2_init", "b2_shutdown", "b2_strerror", "b2_model_load", "b2_model_free", "b2_model_get_info",
    "b2_seq_new", "b2_seq_assume", "b2_seq_free", "b2_seq_len", "b2_seq_truncate", "b2_prefill", "b2_decode", "b2_decode_greedy",
    "b2_get_stats", "b2_mega_trace", "b2_profile_step", "b2_dequantize", "b2_dequantize_repacked", "b2_gemv", "b2_gemm", "b2_gemm_tc", "b2_gemv_noalloc", "b2_gemv_bench", "b2_attn_decode", "b2_attn_bench", "b2_kv_read", "b2_debug_read",
    "b2_store_resolve", "b2_store_list", "b2_tokenizer_load", "b2_tokenizer_free", "b2_tokenizer_get_info", "b2_tokenize", "b2_detokenize", "b2_token_piece",
]


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b2 error {code}: {msg}")
        self.code = code


class LoadOpts(C.Structure):
    _fields_ = [("max_batch", C.c_int32), ("max_ctx", C.c_int32), ("tp_size", C.c_int32), ("flags", C.c_int32)]


class ModelInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("arch", "n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab", "n_ctx",
                                         "n_rot", "max_batch", "max_ctx")] + [
        ("norm_eps", C.c_float), ("rope_freq_base", C.c_float),
        ("weight_bytes", C.c_uint64), ("step_weight_bytes", C.c_uint64), ("kv_bytes_per_token", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("steps", C.c_uint64), ("kernel_launches", C.c_uint64), ("gpu_ms", C.c_double), ("last_steps", C.c_uint64),
                ("last_launches", C.c_uint64)]


_lib = None


def lib() -> C.CDLL:
    """Load the CUDA library; fails loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is missing: run ./build.sh (or __graft_entry__.build()); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        vp, ip, fp = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float)
        L.b2_strerror.restype = C.c_char_p
        L.b2_init.argtypes = [C.c_int, ip, C.POINTER(vp)]
        L.b2_shutdown.argtypes = [vp]
        L.b2_shutdown.restype = None
        L.b2_model_load.argtypes = [vp, C.c_char_p, C.POINTER(LoadOpts), C.POINTER(vp)]
        L.b2_model_free.argtypes = [vp]
        L.b2_model_free.restype = None
        L.b2_model_get_info.argtypes = [vp, C.POINTER(ModelInfo)]
        L.b2_seq_new.argtypes = [vp, ip]
        L.b2_seq_free.argtypes = [vp, C.c_int]
        L.b2_seq_len.argtypes = [vp, C.c_int]
        L.b2_seq_truncate.argtypes = [vp, C.c_int, C.c_int]
        L.b2_seq_assume.argtypes = [vp, C.c_int, C.c_int]
        L.b2_prefill.argtypes = [vp, C.c_int, vp, C.c_int, vp]
        L.b2_decode.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.b2_decode_greedy.argtypes = [vp, C.c_int, C.c_int32, C.c_int, vp]
        L.b2_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.b2_mega_trace.argtypes = [vp, C.c_int, C.c_int32, vp, C.c_int, ip, ip]
        L.b2_profile_step.argtypes = [vp, C.c_int, C.c_int32, vp, vp]
        L.b2_dequantize.argtypes = [C.c_int, vp, C.c_size_t, vp]
        L.b2_dequantize_repacked.argtypes = [C.c_int, vp, C.c_int64, C.c_int64, vp]
        L.b2_gemv.argtypes = [C.c_int, vp, C.c_int64, C.c_int64, vp, vp, C.c_float, vp]
        L.b2_gemm.argtypes = [C.c_int, vp, C.c_int64, C.c_int64, vp, C.c_int, vp, C.c_float, vp, C.c_int, fp]
        L.b2_gemm_tc.argtypes = [C.c_int, vp, C.c_int64, C.c_int64, vp, C.c_int, vp, C.c_float, C.c_int, vp, C.c_int, fp]
        L.b2_gemv_noalloc.argtypes = [C.c_int, vp, C.c_int64, C.c_int64, vp, vp, C.c_float, vp]
        L.b2_gemv_bench.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int, vp, fp]
        L.b2_attn_decode.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, vp]
        L.b2_attn_bench.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, fp]
        L.b2_kv_read.argtypes = [vp, C.c_int, C.c_int, vp, vp, C.c_int]
        L.b2_debug_read.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int]
        _lib = L
    return _lib


def _check(rc: int) -> int:
    if rc < 0:
        raise B2Error(rc, lib().b2_strerror().decode(errors="replace"))
    return rc


def _ptr(a):
    return None if a is None else a.ctypes.data


# ---- hooks -------------------------------------------------------------------------------------------
def dequantize(ggml_type: int, blocks: np.ndarray, n_elem: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(n_elem, dtype=np.float32)
    _check(lib().b2_dequantize(ggml_type, _ptr(blocks), n_elem, _ptr(out)))
    return out


def dequantize_repacked(ggml_type: int, blocks: np.ndarray, rows: int, k: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty((rows, k), dtype=np.float32)
    _check(lib().b2_dequantize_repacked(ggml_type, _ptr(blocks), rows, k, _ptr(out)))
    return out


def gemv(ggml_type: int, W: np.ndarray, rows: int, k: int, x: np.ndarray, gain: np.ndarray | None = None, eps: float = 0.0) -> np.ndarray:
    W = np.ascontiguousarray(W, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float32)
    y = np.empty(rows, dtype=np.float32)
    _check(lib().b2_gemv(ggml_type, _ptr(W), rows, k, _ptr(x), _ptr(g), eps, _ptr(y)))
    return y


def gemm(ggml_type: int, W: np.ndarray, rows: int, k: int, x: np.ndarray, gain: np.ndarray | None = None, eps: float = 0.0, iters: int = 0):
    """Batched W4A8 product: x [n_cols][k] -> y [n_cols][rows]; returns (y, ms per launch)."""
    W = np.ascontiguousarray(W, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float32)
    y = np.empty((x.shape[0], rows), dtype=np.float32)
    ms = C.c_float(0)
    _check(lib().b2_gemm(ggml_type, _ptr(W), rows, k, _ptr(x), x.shape[0], _ptr(g), eps, _ptr(y), iters, C.byref(ms)))
    return y, ms.value


def gemm_tc(ggml_type: int, W: np.ndarray, rows: int, k: int, x: np.ndarray, gain: np.ndarray | None = None, eps: float = 0.0, acc32: bool = False,
            iters: int = 0):
    """The batched product on the tcgen05 tensor cores: x [n_cols <= 128][k] -> y [n_cols][rows]; returns (y, ms per launch).
    acc32: ggml's generic f32 accumulation order instead of the canonical f64 one."""
    W = np.ascontiguousarray(W, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float32)
    y = np.empty((x.shape[0], rows), dtype=np.float32)
    ms = C.c_float(0)
    _check(lib().b2_gemm_tc(ggml_type, _ptr(W), rows, k, _ptr(x), x.shape[0], _ptr(g), eps, int(acc32), _ptr(y), iters, C.byref(ms)))
    return y, ms.value


NOALLOC_SCRATCH_BYTES = 8 << 20


def gemv_noalloc(ggml_type: int, W: np.ndarray, rows: int, k: int, x: np.ndarray, gain: np.ndarray | None = None, eps: float = 0.0) -> np.ndarray:
    """b2_gemv for callers that must not allocate device memory: same result, staging in the
    library's static 8 MiB device array (larger problems raise B2Error -6; calls must not overlap)."""
    W = np.ascontiguousarray(W, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float32)
    y = np.empty(rows, dtype=np.float32)
    _check(lib().b2_gemv_noalloc(ggml_type, _ptr(W), rows, k, _ptr(x), _ptr(g), eps, _ptr(y)))
    return y


def gemv_bench(ggml_type: int, rows: int, k: int, iters: int = 50, tune=None) -> float:
    """Mean device milliseconds per launch of the GEMV kernel on resident data."""
    ms = C.c_float(0)
    t = None if tune is None else np.asarray(list(tune) + [0] * (5 - len(tune)), dtype=np.int32)
    _check(lib().b2_gemv_bench(ggml_type, rows, k, iters, _ptr(t), C.byref(ms)))
    return ms.value


def attn_decode(q: np.ndarray, k: np.ndarray, v: np.ndarray, n_head: int, n_head_kv: int, head_dim: int, scale: float,
                cluster: int = 0) -> np.ndarray:
    q = np.ascontiguousarray(q, dtype=np.float32)
    k = np.ascontiguousarray(k, dtype=np.float16)
    v = np.ascontiguousarray(v, dtype=np.float16)
    out = np.empty(n_head * head_dim, dtype=np.float32)
    _check(lib().b2_attn_decode(_ptr(q), _ptr(k), _ptr(v), k.shape[0], n_head, n_head_kv, head_dim, scale, cluster, _ptr(out)))
    return out


def attn_bench(n_ctx: int, n_head: int, n_head_kv: int, head_dim: int, cluster: int = 0, n_layers: int = 32, iters: int = 128) -> float:
    """Mean device milliseconds per launch of the decode-attention kernel."""
    ms = C.c_float(0)
    _check(lib().b2_attn_bench(n_ctx, n_head, n_head_kv, head_dim, cluster, n_layers, iters, C.byref(ms)))
    return ms.value


# ---- engine -------------------------------------------------------------------------------------------
class Engine:
    """The CUDA device(s) bound through b2_init: one index, or a list for tensor parallelism."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        devs = [device] if isinstance(device, int) else list(device)
        ids = (C.c_int * len(devs))(*devs)
        _check(lib().b2_init(len(devs), ids, C.byref(self._h)))

    def load(self, path: str, max_batch: int = 1, max_ctx: int = 0, flags: int = 0, tp_size: int = 1) -> "Model":
        return Model(self, path, max_batch, max_ctx, flags, tp_size)

    def close(self):
        if self._h:
            lib().b2_shutdown(self._h)
            self._h = C.c_void_p()


class Model:
    def __init__(self, eng: Engine, path: str, max_batch: int = 1, max_ctx: int = 0, flags: int = 0, tp_size: int = 1):
        self.eng = eng
        self._h = C.c_void_p()
        opts = LoadOpts(max_batch, max_ctx, tp_size, flags)
        _check(lib().b2_model_load(eng._h, os.fsencode(path), C.byref(opts), C.byref(self._h)))
        self.info = ModelInfo()
        _check(lib().b2_model_get_info(self._h, C.byref(self.info)))

    def close(self):
        if self._h:
            lib().b2_model_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def seq_assume(self, seq: int, n_tokens: int) -> None:
        _check(lib().b2_seq_assume(self._h, seq, n_tokens))

    def seq_new(self) -> int:
        s = C.c_int(-1)
        _check(lib().b2_seq_new(self._h, C.byref(s)))
        return s.value

    def seq_free(self, seq: int):
        _check(lib().b2_seq_free(self._h, seq))

    def seq_len(self, seq: int) -> int:
        return _check(lib().b2_seq_len(self._h, seq))

    def seq_truncate(self, seq: int, n_keep: int):
        _check(lib().b2_seq_truncate(self._h, seq, n_keep))

    def prefill(self, seq: int, toks, want_logits: bool = True):
        toks = np.ascontiguousarray(toks, dtype=np.int32)
        logits = np.empty(self.info.n_vocab, dtype=np.float32) if want_logits else None
        _check(lib().b2_prefill(self._h, seq, _ptr(toks), toks.size, _ptr(logits)))
        return logits

    def decode(self, seqs, last_toks, want_logits: bool = False):
        seqs = np.ascontiguousarray(seqs, dtype=np.int32)
        last = np.ascontiguousarray(last_toks, dtype=np.int32)
        nxt = np.empty(seqs.size, dtype=np.int32)
        logits = np.empty((seqs.size, self.info.n_vocab), dtype=np.float32) if want_logits else None
        _check(lib().b2_decode(self._h, seqs.size, _ptr(seqs), _ptr(last), _ptr(nxt), _ptr(logits)))
        return (nxt, logits) if want_logits else nxt

    def decode_one(self, seq: int, last_tok: int) -> int:
        """b2_decode for one sequence without logits, through preallocated host buffers: what a server calls per
        generated token (no numpy conversions on the way, the C ABI call is the same)."""
        buf = self.__dict__.get("_one")
        if buf is None:
            buf = self._one = ((C.c_int32 * 1)(), (C.c_int32 * 1)(), (C.c_int32 * 1)())
        buf[0][0] = seq
        buf[1][0] = last_tok
        _check(lib().b2_decode(self._h, 1, buf[0], buf[1], buf[2], None))
        return buf[2][0]

    def decode_greedy(self, seq: int, n_steps: int, first_tok: int = -1) -> np.ndarray:
        out = np.empty(n_steps, dtype=np.int32)
        _check(lib().b2_decode_greedy(self._h, seq, first_tok, n_steps, _ptr(out)))
        return out

    def kv_read(self, seq: int, layer: int, n_tok: int):
        kvd = self.info.n_head_kv * self.info.head_dim
        k = np.empty((n_tok, kvd), dtype=np.float16)
        v = np.empty((n_tok, kvd), dtype=np.float16)
        _check(lib().b2_kv_read(self._h, seq, layer, _ptr(k), _ptr(v), n_tok))
        return k, v

    PROFILE_CLASSES = ("step_begin", "gemv_qkv", "attention", "gemv_wo", "gemv_gate_up", "gemv_down", "gemv_head", "step_end", "persistent_layers")

    def profile_step(self, seq: int, tok: int) -> dict:
        """One decode step with per-kernel-class device times: {class: (ms, launches)}."""
        ms = np.zeros(9, dtype=np.float32)
        nl = np.zeros(9, dtype=np.int32)
        _check(lib().b2_profile_step(self._h, seq, tok, _ptr(ms), _ptr(nl)))
        return {c: (float(ms[i]), int(nl[i])) for i, c in enumerate(self.PROFILE_CLASSES)}

    DEBUG_BUFFERS = {"x": 0, "q": 1, "att": 2, "h": 3, "logits": 4, "att2": 5}

    def debug_read(self, seq: int, what: str) -> np.ndarray:
        i = self.info
        n = {"x": i.n_embd, "q": i.n_head * i.head_dim, "att": i.n_head * i.head_dim, "h": i.n_ff, "logits": i.n_vocab, "att2": i.n_embd}[what]
        out = np.empty(n, dtype=np.float32)
        _check(lib().b2_debug_read(self._h, seq, self.DEBUG_BUFFERS[what], _ptr(out), n))
        return out

    def mega_trace(self, seq: int, tok: int) -> np.ndarray:
        """Per-CTA, per-stage time stamps (ns) of one step of the persistent kernel: [cta][stage][16]
        (0-3 stage entered / barrier passed / prologue done / stage done; 4-11 attention phases)."""
        cap = 160 * (self.info.n_layer * 6) * 16
        out = np.zeros(cap, dtype=np.uint64)
        nc, ns = C.c_int(0), C.c_int(0)
        _check(lib().b2_mega_trace(self._h, seq, tok, _ptr(out), cap, C.byref(nc), C.byref(ns)))
        return out[: nc.value * ns.value * 16].reshape(nc.value, ns.value, 16)

    def stats(self) -> Stats:
        s = Stats()
        _check(lib().b2_get_stats(self._h, C.byref(s)))
        return s
