"""Per-kernel-class times of one decode step of shard 0 under tensor parallelism (classes 3 and 5 include the all-reduce)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200"); E = p.engine
preset = sys.argv[1]; tps = [int(x) for x in sys.argv[2].split(",")]
d = os.environ.get("B2_MODEL_DIR", "/tmp/b2_models"); os.makedirs(d, exist_ok=True)
path = p.synth.cached_model(preset, d, source="blocks")
for tp in tps:
    eng = E.Engine(list(range(tp)))
    m = eng.load(path, max_batch=1, max_ctx=512, tp_size=tp, flags=E.LOAD_NO_MEGA)
    s = m.seq_new(); m.prefill(s, p.synth.prompt_tokens(64, m.info.n_vocab))
    m.decode_greedy(s, 8); m.decode_greedy(s, 32)
    print(preset, "tp", tp, "ms/step", round(m.stats().gpu_ms / 32, 4), "launches/step", m.stats().last_launches / 32, flush=True)
    for _ in range(2):
        pr = m.profile_step(s, 5)
    print({k: (round(v[0] * 1e3 / max(v[1], 1), 2), v[1]) for k, v in pr.items() if v[1]}, "us per launch", flush=True)
    m.close(); eng.close()
