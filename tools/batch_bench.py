"""Throughput of the batched paths on a synthetic model: chunked prefill and multi-sequence decode.
usage: python tools/batch_bench.py [preset] [ctx]"""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200"); E = p.engine
preset = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b"
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
d = os.environ.get("B2_MODEL_DIR", "/tmp/b2_models"); os.makedirs(d, exist_ok=True)
path = p.synth.cached_model(preset, d, source="blocks")
eng = E.Engine(0)
for flags, name in ((0, "batched"), (E.LOAD_NO_BATCH, "per-token")):
    m = eng.load(path, max_batch=32, max_ctx=ctx, flags=flags)
    V = m.info.n_vocab
    s = m.seq_new()
    toks = p.synth.prompt_tokens(512, V)
    m.prefill(s, toks[:64]); m.seq_truncate(s, 0)
    t0 = time.perf_counter(); m.prefill(s, toks); t1 = time.perf_counter()
    st = m.stats()
    print(f"{name}: prefill 512 tokens: gpu {st.gpu_ms:.1f} ms ({512/st.gpu_ms*1e3:.0f} tok/s), wall {(t1-t0)*1e3:.1f} ms, launches {st.last_launches}", flush=True)
    m.seq_free(s)
    for B in (1, 2, 3, 4, 8, 16, 32):
        seqs = [m.seq_new() for _ in range(B)]
        for q in seqs: m.prefill(q, toks[:128])
        tok = [1] * B
        for _ in range(3): tok = list(m.decode(seqs, tok))
        ms = []; t0 = time.perf_counter()
        for _ in range(10):
            tok = list(m.decode(seqs, tok)); ms.append(m.stats().gpu_ms)
        wall = (time.perf_counter() - t0) / 10 * 1e3
        g = float(np.median(ms))
        print(f"{name}: decode B={B}: gpu {g:.2f} ms/step ({B/g*1e3:.0f} tok/s), wall {wall:.2f} ms ({B/wall*1e3:.0f} tok/s)", flush=True)
        for q in seqs: m.seq_free(q)
    m.close()
