"""Per-stage timing of the persistent decode kernel on llama2-7B (tuning aid)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
p = importlib.import_module("ollama-operator_b200")
model = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b"
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 512
path = p.synth.cached_model(model, source="blocks")
eng = p.engine.Engine(0)
m = eng.load(path, max_batch=1, max_ctx=ctx + 64)
seq = m.seq_new()
m.prefill(seq, p.synth.prompt_tokens(ctx, m.info.n_vocab), want_logits=False)
m.decode_greedy(seq, 8)
tr = m.mega_trace(seq, 5).astype(np.int64)
ncta, nst, _ = tr.shape
att = tr[:, 15 * 5 + 1, 4:10]
act = att[:, 0] > 0
print('attention phases (us, mean over active CTAs): ' + ' '.join(f'{x:.2f}' for x in (np.diff(att[act], axis=1).mean(0) / 1.9e3)), ' total', ((att[act][:, 5] - att[act][:, 0]).mean() / 1.9e3).round(2), 'active ctas', int(act.sum()))
t0 = tr[:, 0, 0].min()
names = ["qkv", "attn", "wo", "up", "down"]
print("ctas", ncta, "stages", nst, "total us", (tr[:, :, 3].max() - t0) / 1e3)
for L in (0, 1, 15, 31):
    for k in range(5):
        s = L * 5 + k
        ent, bar, pro, done = tr[:, s, 0], tr[:, s, 1], tr[:, s, 2], tr[:, s, 3]
        print(f"L{L:2d} {names[k]:5s} enter {((ent.min() - t0) / 1e3):8.2f}  barrier wait min {((bar - ent).min() / 1e3):5.2f} mean {((bar - ent).mean() / 1e3):6.2f} max {((bar - ent).max() / 1e3):6.2f}"
              f"  prologue {((pro - bar).mean() / 1e3):6.2f}  body mean {((done - pro).mean() / 1e3):6.2f} max {((done - pro).max() / 1e3):6.2f}"
              f"  stage span {((done.max() - ent.min()) / 1e3):6.2f}")
per = np.zeros(5)
for s in range(nst):
    per[s % 5] += (tr[:, s, 3].max() - tr[:, s, 0].min()) / 1e3
print("sum of stage spans per kind (us):", dict(zip(names, per.round(1))), "sum", per.sum().round(1))

for k, nm in ((0, "qkv"), (3, "up"), (4, "down")):
    g = tr[:, 15 * 5 + k, :]
    pts = np.stack([g[:, 10], g[:, 4], g[:, 5], g[:, 6], g[:, 7], g[:, 8], g[:, 9]], 1)
    pts = np.where(pts == 0, np.nan, pts)
    print(f"prologue {nm}: barrier->loads issued, ->x arrived+squares, ->stats, ->quantised, ->bar, ->xregs (us):", np.round(np.nanmean(np.diff(pts, axis=1), 0) / 1.9e3, 2), '(SM cycles / 1.9 GHz)')


for k, nm in ((0, "qkv"), (2, "wo"), (3, "up"), (4, "down")):
    g = tr[:, 15 * 5 + k, :]
    loop = (g[:, 11] - g[:, 9]) / 1.9e3
    print(f"body {nm}: set-up before the row loop {((g[:, 14] - g[:, 9]) / 1.9e3).mean():.2f} us, loop {((g[:, 11] - g[:, 14]) / 1.9e3).mean():.2f}, wait for the other warps {((g[:, 15] - g[:, 11]) / 1.9e3).mean():.2f}, epilogue {((g[:, 3] - g[:, 2]) / 1e3 - (g[:, 15] - g[:, 9]) / 1.9e3).mean():.2f}")
    tail = (g[:, 3] - g[:, 2]) / 1e3 - loop
    print(f"body {nm}: warp 0 waits for tiles {(g[:, 12] / 1.9e3).mean():.2f} us, computes {(g[:, 13] / 1.9e3).mean():.2f} us")
    print(f"body {nm}: warp 0 row loop mean {loop.mean():.2f} max {loop.max():.2f} us; rest of the body (wait for the other warps + epilogue) mean {tail.mean():.2f} max {tail.max():.2f} us")

# arrival spread: when each CTA enters / finishes a stage of layer 15, relative to the first one (us)
for k in range(5):
    s = 15 * 5 + k
    ent, done = tr[:, s, 0] / 1e3, tr[:, s, 3] / 1e3
    e, d = ent - ent.min(), done - done.min()
    late = np.argsort(-d)[:6]
    print(f"spread {names[k]:5s}: enter p50 {np.percentile(e, 50):5.2f} p90 {np.percentile(e, 90):5.2f} max {e.max():5.2f} | done p50 {np.percentile(d, 50):5.2f} p90 {np.percentile(d, 90):5.2f} max {d.max():5.2f}  latest CTAs {late.tolist()}")
for L in (7, 15, 23):
    g = tr[:, L * 5 + 3, :]
    body = (g[:, 3] - g[:, 2]) / 1e3
    wait = g[:, 12] / 1.9e3
    order = np.argsort(-body)
    print(f"up L{L}: slowest CTAs (cta body_us tile_wait_us):", [(int(c), round(float(body[c]), 2), round(float(wait[c]), 2)) for c in order[:8]], " median body", round(float(np.median(body)), 2), "median wait", round(float(np.median(wait)), 2))
