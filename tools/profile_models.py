import importlib, os, sys
sys.path.insert(0, "/root/repo")
p = importlib.import_module("ollama-operator_b200"); E = p.engine
d = "/tmp/b2_models"; os.makedirs(d, exist_ok=True)
for preset in sys.argv[1:]:
    path = p.synth.cached_model(preset, d, source="blocks")
    eng = E.Engine(0)
    m = eng.load(path, max_batch=1, max_ctx=1024)
    s = m.seq_new(); m.prefill(s, p.synth.prompt_tokens(512, m.info.n_vocab))
    m.decode_greedy(s, 8); m.decode_greedy(s, 32)
    print(preset, "ms/step", m.stats().gpu_ms / 32, "launches/step", m.stats().last_launches / 32)
    print({k: v for k, v in m.profile_step(s, 5).items() if v[1]})
    m.close(); eng.close()
