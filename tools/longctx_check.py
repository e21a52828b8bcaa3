"""Decode token by token across a range of positions and report where (and after how long) a step fails (debug aid).
usage: python tools/longctx_check.py <prompt tokens> <max_ctx> [greedy|decode]"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200")
path = p.synth.cached_model(os.environ.get("B2_CHECK_MODEL", "llama2-7b"), source="blocks")
eng = p.engine.Engine(0)
ctx = int(sys.argv[1]); mx = int(sys.argv[2])
m = eng.load(path, max_batch=1, max_ctx=mx)
seq = m.seq_new()
m.prefill(seq, p.synth.prompt_tokens(ctx, m.info.n_vocab), want_logits=False)
tok, t0 = 1, time.time()
try:
    while m.seq_len(seq) < mx - 2:
        t0 = time.time()
        tok = int(m.decode([seq], [tok])[0])
    print("ok up to", m.seq_len(seq))
except Exception as e:
    print("FAILED after len", m.seq_len(seq), "step took %.2f s" % (time.time() - t0), e)
