"""One 32-sequence decode step of llama2-7B between cudaProfilerStart/Stop (for `ncu --profile-from-start off`)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
p = importlib.import_module("ollama-operator_b200"); E = p.engine
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 128
d = os.environ.get("B2_MODEL_DIR", "/tmp/b2_models"); os.makedirs(d, exist_ok=True)
path = p.synth.cached_model("llama2-7b", d, source="blocks")
eng = E.Engine(0)
m = eng.load(path, max_batch=B, max_ctx=ctx + 64)
seqs = [m.seq_new() for _ in range(B)]
for i, s in enumerate(seqs):
    m.prefill(s, p.synth.prompt_tokens(ctx, m.info.n_vocab, seed=i))
tok = [1] * B
for _ in range(3):
    tok = list(m.decode(seqs, tok))
torch.cuda.profiler.start()
tok = list(m.decode(seqs, tok))
torch.cuda.profiler.stop()
print("gpu ms", m.stats().gpu_ms, "launches", m.stats().last_launches)
