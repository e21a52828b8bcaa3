"""Every scheduling variant of the engine must produce the same bits: prefill logits and greedy tokens of one model
under the default path, without the tensor-core prefill, without batching, without the persistent kernel."""
import importlib, os, subprocess, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 2 and sys.argv[2] == "--child":
    p = importlib.import_module("ollama-operator_b200"); E = p.engine
    model, flags, n = sys.argv[1], int(sys.argv[3]), int(sys.argv[4])
    path = p.synth.cached_model(model, source="blocks")
    eng = E.Engine(0); m = eng.load(path, max_batch=1, max_ctx=n + 80, flags=flags)
    s = m.seq_new(); lg = m.prefill(s, p.synth.prompt_tokens(n, m.info.n_vocab))
    toks = [int(np.argmax(lg))] + [int(t) for t in m.decode_greedy(s, 40)]
    print(json.dumps({"crc": int(np.frombuffer(lg.tobytes(), np.uint32).sum() & 0xFFFFFFFF), "toks": toks}))
    sys.exit(0)
model = sys.argv[1] if len(sys.argv) > 1 else "llama2-13b"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100
ref = None
for name, flags, env in [("default", 0, {}), ("no tensor-core prefill", 0, {"B2_NO_TC": "1"}), ("no batching", 8, {}), ("no persistent kernel", 4, {}), ("neither", 12, {})]:
    out = subprocess.run([sys.executable, __file__, model, "--child", str(flags), str(n)], env={**os.environ, **env}, capture_output=True, text=True)
    try:
        r = json.loads(out.stdout.strip().splitlines()[-1])
    except Exception:
        print(name, "FAILED", out.stderr[-400:]); continue
    if ref is None: ref = r
    print(f"{name:26s} logits crc {r['crc']:10d} tokens {'same' if r['toks'] == ref['toks'] else 'DIFFER at ' + str(next(i for i in range(41) if r['toks'][i] != ref['toks'][i]))} bits {'same' if r['crc'] == ref['crc'] else 'DIFFER'}")
