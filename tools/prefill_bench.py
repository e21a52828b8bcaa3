"""Prefill throughput of the engine (b2_prefill, CUDA-event time of the whole call) for a few prompt lengths."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200")
model = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b"
path = p.synth.cached_model(model, source="blocks")
eng = p.engine.Engine(0)
m = eng.load(path, max_batch=1, max_ctx=2200)
for n in (512, 512, 1024, 2048, 100):
    seq = m.seq_new(); toks = p.synth.prompt_tokens(n, m.info.n_vocab)
    m.prefill(seq, toks); ms = m.stats().gpu_ms
    print(f"prefill {n:5d} tokens {ms:8.2f} ms {n / ms * 1e3:8.0f} tok/s", flush=True)
    m.seq_free(seq)
