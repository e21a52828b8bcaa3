"""Tensor-parallel decode throughput on a synthetic model (BASELINE.json configs 4-5).
usage: python tools/tp_bench.py <preset> <tp,tp,...> [ctx] [steps]
For every tp the same prompt is prefilled and `steps` tokens are generated greedily on the device;
the token lists of all runs must be identical (the all-reduce keeps single-GPU bits)."""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200"); E = p.engine
preset = sys.argv[1] if len(sys.argv) > 1 else "llama2-13b"
tps = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1,2").split(",")]
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 512
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 64
d = os.environ.get("B2_MODEL_DIR", "/tmp/b2_models"); os.makedirs(d, exist_ok=True)
t0 = time.time()
path = p.synth.cached_model(preset, d, source="blocks")
print(f"# {preset}: gguf ready in {time.time()-t0:.1f}s, {os.path.getsize(path)/1e9:.2f} GB", flush=True)
peak = 6550.7
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
ref = None
for tp in tps:
    eng = E.Engine(list(range(tp)))
    t0 = time.time()
    m = eng.load(path, max_batch=2 + int(os.environ.get("TP_BENCH_BATCH", "0")), max_ctx=max(ctx + steps + 72, 640), tp_size=tp)
    load_s = time.time() - t0
    info = m.info
    s = m.seq_new()
    t0 = time.time()
    if os.environ.get("TP_BENCH_ASSUME"):       # synthetic context: the (zeroed) cache rows are declared valid, only a short prompt is evaluated
        m.seq_assume(s, ctx - 16)
        m.prefill(s, p.synth.prompt_tokens(16, info.n_vocab))
    else:
        m.prefill(s, p.synth.prompt_tokens(ctx, info.n_vocab))
    pre_s = time.time() - t0
    m.decode_greedy(s, 8)
    toks = m.decode_greedy(s, steps)
    ms = m.stats().gpu_ms / steps
    last = int(toks[-1]); t0 = time.perf_counter()
    for _ in range(32):
        last = int(m.decode([s], [last])[0])
    e2e_ms = (time.perf_counter() - t0) / 32 * 1e3
    nb = int(os.environ.get("TP_BENCH_BATCH", "0"))
    batch = None
    if nb > 1:          # nb sequences decoded together (batched columns), each at the same synthetic context
        seqs = []
        for i in range(nb):
            q = m.seq_new(); m.seq_assume(q, ctx - 16); m.prefill(q, p.synth.prompt_tokens(16, info.n_vocab, seed=50 + i)); seqs.append(q)
        tk = [1] * nb
        for _ in range(3): tk = [int(x) for x in m.decode(seqs, tk)]
        g = []
        for _ in range(8):
            tk = [int(x) for x in m.decode(seqs, tk)]; g.append(m.stats().gpu_ms)
        batch = {"n_seq": nb, "ms_per_step": round(float(np.median(g)), 3), "tok_s": round(nb / float(np.median(g)) * 1e3, 1), "tokens": tk[:4]}
        t0 = time.perf_counter(); q = m.seq_new(); m.prefill(q, p.synth.prompt_tokens(512, info.n_vocab, seed=77))
        batch["prefill512_ms"] = round(m.stats().gpu_ms, 1)
    step_bytes = info.step_weight_bytes + (ctx + 8 + steps // 2) * info.kv_bytes_per_token
    frac = step_bytes / tp / (ms * 1e-3) / 1e9 / peak
    same = None if ref is None else bool(np.array_equal(ref, toks))
    if ref is None: ref = toks
    print(json.dumps({"model": preset, "tp": tp, "ctx": ctx, "steps": steps, "ms_per_step": round(ms, 4), "tok_s": round(1e3 / ms, 1),
                      "e2e_tok_s": round(1e3 / e2e_ms, 1), "per_gpu_bytes_per_step": int(step_bytes / tp), "per_gpu_hbm_frac": round(frac, 4),
                      "tokens_equal_first_run": same, "load_s": round(load_s, 1), "prefill_s": round(pre_s, 1), "batch": batch}), flush=True)
    m.close(); eng.close()
