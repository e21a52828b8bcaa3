"""Times the fused GEMV kernel alone over the llama2-7B shapes and a few launch configurations
(tune = nb, g, rpg, ns, ctas_per_sm; 0 = heuristic).  Prints GB/s of algorithmic weight bytes."""
import importlib
import itertools
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
E = importlib.import_module("ollama-operator_b200").engine

SHAPES = [("qkv", 2, 12288, 4096), ("wo", 2, 4096, 4096), ("gate_up", 2, 22016, 4096), ("down", 2, 4096, 11008), ("head_q6k", 14, 32000, 4096)]
BPW = {2: 18 / 32, 8: 34 / 32, 12: 144 / 256, 14: 210 / 256}
only = sys.argv[1:] or None
for name, t, rows, k in SHAPES:
    if only and name not in only:
        continue
    nbytes = rows * k * BPW[t]
    tunes = [None]
    if t == 2:
        nbs = [4, 2, 1] if k == 4096 else [4, 2]
        tunes += [(nb, g, rpg, ns, c) for nb in nbs for g in ([4, 8] if k == 4096 else [2, 4]) for rpg in (2, 4) for ns in (3, 6) for c in (1, 2)]
    else:
        tunes += [(nb, g, rpg, ns, c) for nb in (1,) for g in (4, 8) for rpg in (2, 4) for ns in (3, 6) for c in (1, 2)]
    res = []
    for tune in tunes:
        try:
            ms = E.gemv_bench(t, rows, k, iters=100, tune=tune)
            res.append((nbytes / ms / 1e6, ms, tune))
        except Exception as ex:
            print(name, tune, "ERR", ex, flush=True)
    res.sort(key=lambda r: -r[0])
    for gbs, ms, tune in res[:6] + [r for r in res if r[2] is None]:
        print(f"{name:9s} {rows}x{k} tune={tune} {ms * 1e3:8.2f} us {gbs:8.1f} GB/s", flush=True)

if not only or "attn" in only:
    for (nh, nkv, hd) in [(32, 32, 128), (32, 8, 128), (64, 8, 128)]:
        for ctx in (128, 512, 2048):
            for cl in (1, 2, 4, 8):
                ms = E.attn_bench(ctx, nh, nkv, hd, cl)
                kvb = ctx * nkv * hd * 4
                print(f"attn heads {nh}/{nkv} hd {hd} ctx {ctx:5d} cluster {cl}: {ms * 1e3:8.2f} us  {kvb / ms / 1e6:8.1f} GB/s", flush=True)
