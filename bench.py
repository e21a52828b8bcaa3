#!/usr/bin/env python
"""bench.py -- decode throughput of the B200 GGUF engine on BASELINE.json's headline config.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one decode step (one new token for one sequence: every weight matrix streamed
once, attention over the cached context).  Workload at N=1 = BASELINE.json configs[1]:
llama2-7B Q4_0 (synthetic random-init GGUF, Q6_K head as in the shipped files), batch 1,
decode after a 512-token prefill.  With N>1 every rank is an independent replica on its own
B200 (the reference's only parallelism: `spec.replicas`, pkg/model/model.go:72) -- weak
scaling, no data-path collective; `value` is the sum over replicas.

  value   device-resident greedy loop (b2_decode_greedy): tokens, positions and KV stay in HBM
  e2e     the same steps through b2_decode with HOST token buffers (H2D + D2H every step)
  roofline dominant kernel = the persistent all-layers kernel (one launch per step); achieved =
          its algorithmic bytes (layer weights + the KV rows attention reads) / its launch
          duration, measured live with CUDA events on the engine's stream (b2_profile_step),
          against the measured burst HBM peak; `step` adds the whole-step figure
  prefill, batch32  the other half of configs[1]: the 512-token prompt through the batched
          kernels, and 32 sequences decoded together through b2_decode
  cpu_baseline  the CPU oracle (oracle/ggml_ref.c, a port of the llama.cpp CPU path) timed on
          the host cores for a bounded sample of the same workload; also used to check the
          GPU logits/tokens at full model size

`--impl reference` times that CPU path alone with all host threads (the reference's own
container cannot run here: no Go/docker/ollama).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decode_tokens_per_s"
UNIT = "tok/s"


def pkg():
    return importlib.import_module("ollama-operator_b200")


def job_times(local_ms, dist=None, device="cpu"):
    """Whole-job duration of every timed region: the MAX over ranks (the replicas run side by side,
    the job is done when the slowest is).  local_ms: this rank's milliseconds per region."""
    import torch
    t = torch.tensor(list(local_ms), dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def whole_job_rate(world, steps_per_rank, ms):
    """Units all ranks processed divided by the job's duration (weak scaling: every rank does steps_per_rank)."""
    return world * steps_per_rank / (ms / 1e3)


def oracle_mod():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    L = oracle.lib()
    # torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arm is meant to use every host
    # core this process may run on, so the count is set through the oracle binding, not taken from the env.
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    L.ref_set_num_threads(int(os.environ.get("B2_REF_THREADS", n)))
    return oracle


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.t0 = self.t1 = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def window_begin(self):
        self.t0 = time.time()

    def window_end(self):
        self.t1 = time.time()

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        # samples taken inside the timed regions (the sampler itself starts before the warm-up, nvidia-smi needs ~0.1 s to come up)
        rows = [r for t, r in self.rows if (self.t0 is None or t >= self.t0) and (self.t1 is None or t <= self.t1 + 0.02)]
        sm = [float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def model_path(args):
    p = pkg()
    t0 = time.time()
    path = p.synth.cached_model(args.model, source="blocks")
    return path, time.time() - t0


def run_reference(args):
    """CPU arm: the oracle port of the llama.cpp CPU path, all host threads, same config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    p, o = pkg(), oracle_mod()
    path, _ = model_path(args)
    cfg = p.synth.PRESETS[args.model]
    om = o.OracleModel(path, faithful=True, n_ctx=args.ctx + args.steps + args.warmup + 8)
    kvd = cfg.n_head_kv * cfg.head_dim
    rng = np.random.default_rng(99)
    # A 512-token CPU prefill would take minutes; decode speed does not depend on the cached
    # values, so the context is a synthetic f16 prefix.
    for il in range(cfg.n_layer):
        kv = (rng.standard_normal((2, args.ctx, kvd)) * 0.5).astype(np.float16)
        om.set_kv(il, kv[0], kv[1])
    tok, pos = 1, args.ctx
    for _ in range(args.warmup):
        tok = int(np.argmax(om.eval(tok, pos))); pos += 1
    steps, t0 = 0, time.perf_counter()
    budget_s = float(os.environ.get("B2_REF_BUDGET_S", "150"))
    while steps < args.steps:
        tok = int(np.argmax(om.eval(tok, pos))); pos += 1; steps += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    v = steps / dt
    cores = o.lib().ref_num_threads()
    sample = f"{steps} decode steps at ctx {args.ctx} (synthetic f16 KV prefix), {args.warmup} warm-up"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8xint4->f32",
        "data": "synthetic", "config": workload_config(args, cfg),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(args, cfg, step_weight_bytes=None):
    if step_weight_bytes is None:      # CPU arm: same formula as the engine's b2_model_info.step_weight_bytes, from the preset
        step_weight_bytes = preset_weight_bytes(cfg)
    return {"workload": f"{args.model} {cfg.ftype.upper()} (Q6_K head) batch-1 decode after a {args.ctx}-token prefill",
            "batch": 1, "context": args.ctx, "parallelism": f"replicas x{args.gpus} (one engine per GPU, no collective)",
            "l2": f"inputs larger than L2: {step_weight_bytes / 1e9:.2f} GB of weights streamed per step vs 126 MB L2"}


def preset_weight_bytes(cfg):
    """Bytes of weights one decode step streams (layer matrices + head; norms and the embedding row are <0.1 %)."""
    bpw = {"q4_0": 18 / 32, "q8_0": 34 / 32, "q4_k_m": 144 / 256}[cfg.ftype]
    e, f, kv = cfg.n_embd, cfg.n_ff, cfg.n_head_kv * cfg.head_dim
    per_layer = e * (cfg.n_head * cfg.head_dim) * 2 + 2 * e * kv + (3 if cfg.arch == "llama" else 2) * e * f
    head = cfg.n_vocab * e * (210 / 256 if e % 256 == 0 else bpw)
    return cfg.n_layer * per_layer * bpw + head


def run_gpu(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        import datetime
        # rank 0 measures the tensor-parallel engine at the end while the others wait in a barrier: allow for a 70B load
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(minutes=40))
        # a host-side group for the final rendezvous: an NCCL barrier would park a spinning kernel on every waiting rank's
        # GPU, and the tensor-parallel engine rank 0 runs at the end needs those GPUs whole (cooperative launches)
        host_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(minutes=40))
    p = pkg()
    cfg = p.synth.PRESETS[args.model]
    if rank == 0 or world == 1:
        path, gen_s = model_path(args)
    if dist:
        dist.barrier()
        path, gen_s = model_path(args)      # cached by rank 0
    E = p.engine
    eng = E.Engine(local)
    t0 = time.time()
    m = eng.load(path, max_batch=1, max_ctx=args.ctx + 2 * (args.steps + args.warmup) + 192)
    load_s = time.time() - t0
    info = m.info
    prompt = p.synth.prompt_tokens(args.ctx, info.n_vocab)
    seq = m.seq_new()
    m.prefill(seq, prompt[:64]); m.seq_truncate(seq, 0)      # first launches of the batched kernels (module load, attributes) outside the timing
    logits0 = m.prefill(seq, prompt)
    prefill_ms = m.stats().gpu_ms

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident loop: `value` ------------------------------------------------------------
    sampler = ClockSampler(local)
    sampler.start()
    wtoks = m.decode_greedy(seq, args.warmup)
    barrier()
    sampler.window_begin()
    toks = m.decode_greedy(seq, args.steps)
    st = m.stats()
    barrier()
    gpu_ms, launches = st.gpu_ms, int(st.last_launches)
    ctx_mid = args.ctx + args.warmup + args.steps // 2
    # ---- end to end through b2_decode with host buffers: `e2e` ---------------------------------------
    # ... over the SAME positions as `value`: the sequence is cut back to three steps before the timed region started (the
    # cache rows are simply written again, with the same bits), so that both numbers see the same context lengths
    last = int(toks[-1])
    if args.warmup >= 4:
        m.seq_truncate(seq, args.ctx + args.warmup - 3)
        last = int(wtoks[args.warmup - 4])
    for _ in range(3):
        last = m.decode_one(seq, last)
    barrier()
    t0 = time.perf_counter()
    e2e_gpu_ms = 0.0
    for _ in range(args.steps):
        last = m.decode_one(seq, last)       # b2_decode(1 sequence, host token in, host token out); the step's device time is read back too
        e2e_gpu_ms += m.stats().gpu_ms
    barrier()
    e2e_s = time.perf_counter() - t0
    sampler.window_end()
    clocks = sampler.stop()       # samples span both timed regions (device-resident loop and end-to-end loop)

    gpu_ms_max, e2e_ms_max = job_times([gpu_ms, e2e_s * 1e3], dist, f"cuda:{local}")
    if rank != 0:
        m.close(); eng.close()
        if dist:
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
            dist.barrier(group=host_group)      # rank 0 measures the tensor-parallel engine over all GPUs meanwhile (CPU-side wait)
            dist.destroy_process_group()
        return

    value = whole_job_rate(world, args.steps, gpu_ms_max)
    e2e = whole_job_rate(world, args.steps, e2e_ms_max)
    peak, peak_src = peaks()
    # ---- dominant kernel ----------------------------------------------------------------------------------
    # With the persistent kernel one launch per step does all layers: its duration is measured live with
    # CUDA events on the engine's stream (b2_profile_step) and compared with its algorithmic bytes (all
    # layer weights + the KV rows attention reads).  Without it the dominant kernel is the fused GEMV of
    # the largest matrix, timed alone over L2-cold operands.
    prof = {}
    for _ in range(5):
        pr = m.profile_step(seq, last)
        for k, v in pr.items():
            prof.setdefault(k, []).append(v[0])
    prof_ms = {k: float(np.median(v)) for k, v in prof.items()}
    rows_up = 2 * cfg.n_ff if cfg.arch == "llama" else cfg.n_ff
    body_type = {"q4_0": E.TYPE_Q4_0, "q8_0": E.TYPE_Q8_0, "q4_k_m": E.TYPE_Q4_K}[cfg.ftype]
    bpw = {E.TYPE_Q4_0: 18 / 32, E.TYPE_Q8_0: 34 / 32, E.TYPE_Q4_K: 144 / 256}[body_type]
    head_bytes = info.n_vocab * info.n_embd * 210 / 256 if info.n_embd % 256 == 0 else info.n_vocab * info.n_embd * bpw
    step_bytes = info.step_weight_bytes + ctx_mid * info.kv_bytes_per_token
    step_gbs = step_bytes / (gpu_ms / args.steps * 1e-3) / 1e9
    if prof_ms.get("persistent_layers", 0) > 0:
        k_ms = prof_ms["persistent_layers"]
        k_bytes = info.step_weight_bytes - head_bytes + m.seq_len(seq) * info.kv_bytes_per_token
        k_name = f"mega_kernel<{'K-quants' if cfg.ftype == 'q4_k_m' else cfg.ftype.upper()}> all {cfg.n_layer} layers of one decode step (weights + KV reads)"
    else:
        k_ms = E.gemv_bench(body_type, rows_up, cfg.n_embd, iters=200)
        k_bytes = rows_up * cfg.n_embd * bpw
        k_name = f"gemv_kernel gate|up {rows_up}x{cfg.n_embd}"
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    traffic = None            # DRAM bytes of one launch of that kernel from the committed ncu capture (same model, shorter context)
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tr["kernel"] in k_name and args.model == "llama2-7b":
            traffic = tr["dram_bytes_per_launch"] + (m.seq_len(seq) - tr["context"]) * info.kv_bytes_per_token
    except (OSError, KeyError, ValueError):
        pass
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": gpu_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int8xint4->f32", "data": "synthetic", "config": workload_config(args, cfg, info.step_weight_bytes),
        "clocks": clocks, "gpu_launches": launches,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 4,
                "gpu_ms_per_step": e2e_gpu_ms / args.steps},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "traffic_source": None if traffic is None else "dram__bytes_read+write of one ncu --set full capture of this kernel (profiles/traffic.json, taken at another context length) + the KV rows this context adds; not measured in this run",
                     "kernel": k_name, "bytes_per_launch": k_bytes, "ms_per_launch": k_ms,
                     "peak_source": peak_src,
                     "step": {"bytes": step_bytes, "achieved": step_gbs, "frac": step_gbs / peak}},
        "profile_step_ms": {k: round(v, 4) for k, v in prof_ms.items() if v > 0},
        "prefill": {"tokens": args.ctx, "ms": prefill_ms, "tok_s": args.ctx / (prefill_ms * 1e-3),
                    "note": "prompt chunks of up to 512 columns per weight pass through the tcgen05 kernel (kind::i8 MMA per 32-weight block, TMEM accumulators, f64 term accumulation: bit-identical to token-by-token)"},
        "setup_s": {"write_gguf": round(gen_s, 1), "load": round(load_s, 1)},
    }
    if world == 1 and not args.no_cpu:
        out["cpu_baseline"], out["parity"] = cpu_baseline(args, p, m, seq, path, cfg)
    m.close()
    if world == 1 and not args.no_batch:
        out["batch32"] = batch_decode(args, p, eng, path, info)
    eng.close()
    if not args.no_tp and os.environ.get("B2_BENCH_TP", "1") != "0":
        # every other rank has released its engine (they returned above after the MAX of the timings); rank 0 now drives
        # all of the job's GPUs itself -- the tensor-parallel engine is one process with one shard per GPU
        try:
            out["tp"] = tp_section(args, p, world)
        except Exception as e:          # the replica headline above stands on its own
            out["tp"] = {"error": f"{type(e).__name__}: {e}"}
    print(json.dumps(out), flush=True)          # the line is out before any teardown that a faulted context could stall
    if dist:
        try:
            dist.barrier(group=host_group)
            dist.destroy_process_group()
        except Exception:
            pass


def tp_run(p, path, tp, ctx, steps, n_batch, peak):
    """One tensor-parallel engine (tp GPUs inside this process): greedy decode at a synthetic `ctx`-token context."""
    E = p.engine
    eng = E.Engine(list(range(tp)))
    m = eng.load(path, max_batch=1 + n_batch, max_ctx=ctx + steps + 80, tp_size=tp)
    info = m.info
    s = m.seq_new()
    m.seq_assume(s, ctx - 16)            # the cache rows are declared valid (zeros): reaching the context costs no prefill
    m.prefill(s, p.synth.prompt_tokens(16, info.n_vocab))
    m.decode_greedy(s, 4)
    toks = m.decode_greedy(s, steps)
    ms = m.stats().gpu_ms / steps
    step_bytes = info.step_weight_bytes + (ctx + 4 + steps // 2) * info.kv_bytes_per_token
    out = {"tp": tp, "tok_s": 1e3 / ms, "ms_per_step": ms, "per_gpu_hbm_frac": step_bytes / tp / (ms * 1e-3) / 1e9 / peak}
    if n_batch > 1:
        seqs = []
        for i in range(n_batch):
            q = m.seq_new(); m.seq_assume(q, ctx - 16); m.prefill(q, p.synth.prompt_tokens(16, info.n_vocab, seed=50 + i)); seqs.append(q)
        tk = [1] * n_batch
        for _ in range(2):
            tk = [int(x) for x in m.decode(seqs, tk)]
        g = []
        for _ in range(6):
            tk = [int(x) for x in m.decode(seqs, tk)]; g.append(m.stats().gpu_ms)
        out["batch"] = {"n_seq": n_batch, "ms_per_step": float(np.median(g)), "tok_s": n_batch / float(np.median(g)) * 1e3, "tokens": tk}
    m.close(); eng.close()
    return out, [int(t) for t in toks]


def tp_section(args, p, n):
    """BASELINE.json configs 4-5 inside the driver's record: llama2-70B Q4_0 split over the job's n GPUs (one process,
    one engine shard per GPU, exchange over NVLink peer memory), batch 1 and 16 sequences at a 2048-token context,
    against the same model on one GPU; plus parity of the split models with their single-GPU bits and the oracle."""
    peak, _ = peaks()
    steps = int(os.environ.get("B2_TP_STEPS", "32"))
    t0 = time.time()
    path70 = p.synth.cached_model("llama2-70b", source="blocks")
    gen_s = time.time() - t0
    one, toks1 = tp_run(p, path70, 1, 2048, steps, 16 if n == 1 else 0, peak)
    out = {"model": "llama2-70b Q4_0, decode at a 2048-token context (synthetic cache rows)", "steps": steps, "write_gguf_s": round(gen_s, 1), "tp1": one}
    if n > 1:
        run, toks = tp_run(p, path70, n, 2048, steps, 16, peak)
        run["speedup_vs_tp1"] = run["tok_s"] / one["tok_s"]
        run["tokens_equal_tp1"] = toks == toks1
        out["tpN"] = run
        out["tp_parity"] = tp_parity(p, n)
    return out


def tp_parity(p, n):
    """The split model must give the single-GPU bits: logits of a prefill + 24 teacher-forced steps on a small GQA model
    (8 query heads per KV head, as 70B) against tp = 1 and against the canonical oracle; 13B greedy tokens against tp = 1."""
    E, o = p.engine, oracle_mod()
    res = {}
    path = p.synth.cached_model("tiny-gqa8")
    toks = p.synth.prompt_tokens(9, p.synth.PRESETS["tiny-gqa8"].n_vocab)
    tp = 2 if n >= 2 else 1           # tiny-gqa8 has two KV heads: at most two shards
    logs = {}
    for t in (1, tp):
        eng = E.Engine(list(range(t))); m = eng.load(path, max_batch=1, max_ctx=128, tp_size=t)
        s = m.seq_new(); lg = [m.prefill(s, toks)]
        tok = int(np.argmax(lg[0]))
        for _ in range(24):
            nxt, l = m.decode([s], [tok], want_logits=True); lg.append(l[0]); tok = int(np.argmax(l[0]))
        logs[t] = np.stack(lg); m.close(); eng.close()
    om = o.OracleModel(path, faithful=2)
    for t_ in toks:
        w = om.eval(int(t_))
    ref = [w]; tok = int(np.argmax(w))
    for _ in range(24):
        w = om.eval(tok); ref.append(w); tok = int(np.argmax(w))
    ref = np.stack(ref)
    res["tiny-gqa8"] = {"tp": tp, "logits_bits_equal_tp1": bool(np.array_equal(logs[tp].view(np.uint32), logs[1].view(np.uint32))),
                        "tokens_equal_tp1": bool(np.array_equal(logs[tp].argmax(1), logs[1].argmax(1))),
                        "oracle_max_abs": float(np.max(np.abs(logs[tp] - ref)))}
    path13 = p.synth.cached_model("llama2-13b", source="blocks")
    t13 = {}
    for t in (1, n):
        eng = E.Engine(list(range(t))); m = eng.load(path13, max_batch=1, max_ctx=256, tp_size=t)
        s = m.seq_new(); m.prefill(s, p.synth.prompt_tokens(32, m.info.n_vocab)); t13[t] = [int(x) for x in m.decode_greedy(s, 32)]
        m.close(); eng.close()
    res["llama2-13b"] = {"tp": n, "tokens_equal_tp1": t13[n] == t13[1], "steps": 32}
    return res


def batch_decode(args, p, eng, path, info, n_seq=32, steps=16):
    """BASELINE.json configs[1], second half: 32 sequences (different prompts) decoded together
    through b2_decode with host token buffers, after a prefill of args.ctx tokens each."""
    m = eng.load(path, max_batch=n_seq, max_ctx=args.ctx + steps + 8)
    seqs, tok = [], []
    pre_ms = 0.0
    for i in range(n_seq):
        s = m.seq_new()
        lg = m.prefill(s, p.synth.prompt_tokens(args.ctx, info.n_vocab, seed=4000 + i))
        pre_ms += m.stats().gpu_ms
        seqs.append(s); tok.append(int(np.argmax(lg)))
    for _ in range(3):
        tok = [int(t) for t in m.decode(seqs, tok)]
    t0 = time.perf_counter()
    gpu_ms = 0.0
    for _ in range(steps):
        tok = [int(t) for t in m.decode(seqs, tok)]
        gpu_ms += m.stats().gpu_ms
    wall = time.perf_counter() - t0
    ctx_mid = args.ctx + 3 + steps // 2
    step_bytes = info.step_weight_bytes + n_seq * ctx_mid * info.kv_bytes_per_token
    m.close()
    peak, _ = peaks()
    return {"n_seq": n_seq, "steps": steps, "tok_s": n_seq * steps / wall, "ms_per_step": wall / steps * 1e3, "gpu_ms_per_step": gpu_ms / steps,
            "prefill_tok_s": n_seq * args.ctx / (pre_ms * 1e-3), "step_bytes": step_bytes,
            "hbm_frac": step_bytes / (gpu_ms / steps * 1e-3) / 1e9 / peak}


def cpu_baseline(args, p, m, seq, path, cfg):
    """Bounded CPU sample: continue the SAME sequence on the host from the GPU's KV cache.
    Timing uses the oracle's ggml-generic path (AVX2 integer dots, all host threads); the
    parity check uses its order-free canonical mode, which the engine must match bit for bit
    up to the 2e-2 logit tolerance of BASELINE.json's north star."""
    o = oracle_mod()
    n = m.seq_len(seq)
    n_time = int(os.environ.get("B2_CPU_STEPS", "12"))
    n_par = int(os.environ.get("B2_PARITY_STEPS", "64"))
    kv = [m.kv_read(seq, il, n) for il in range(cfg.n_layer)]
    # ---- parity at full model size: every side evaluates the same token at the same position (teacher-forced with
    #      the canonical oracle's arg-max), each with its own KV cache, for n_par steps:
    #        canonical  (mode 2, order-free)      <- the engine must match this bit for bit
    #        ggml order (mode 1, generic C path)  <- the comparator BASELINE.json's north star names (2e-2, same tokens)
    #        AVX2 order (mode 3)                  <- a second ggml build: its distance from mode 1 is the noise floor
    oms = {}
    for name, mode in (("canonical", 2), ("ggml", 1), ("avx2", 3)):
        om = o.OracleModel(path, faithful=mode, n_ctx=n + n_par + 2)
        for il, (k, v) in enumerate(kv):
            om.set_kv(il, k, v)
        oms[name] = om
    tok, pos = 7, n
    err = {"canonical": 0.0, "ggml": 0.0, "noise": 0.0}
    same = {"canonical": 0, "ggml": 0, "noise": 0}
    first_div = {"ggml": None, "noise": None}
    for step in range(n_par):
        g_next, g_logits = m.decode([seq], [tok], want_logits=True)
        lg = {name: om.eval(tok, pos) for name, om in oms.items()}
        for name in ("canonical", "ggml"):
            err[name] = max(err[name], float(np.max(np.abs(g_logits[0] - lg[name]))))
            same[name] += int(int(g_next[0]) == int(np.argmax(lg[name])))
        err["noise"] = max(err["noise"], float(np.max(np.abs(lg["avx2"] - lg["ggml"]))))
        same["noise"] += int(int(np.argmax(lg["avx2"])) == int(np.argmax(lg["ggml"])))
        for name, a_, b_ in (("ggml", g_logits[0], lg["ggml"]), ("noise", lg["avx2"], lg["ggml"])):
            if first_div[name] is None and int(np.argmax(a_)) != int(np.argmax(b_)):
                top = np.sort(b_)[-2:]
                first_div[name] = {"step": step, "top2_margin": float(top[1] - top[0])}
        tok, pos = int(np.argmax(lg["canonical"])), pos + 1
    for om in oms.values():
        om.close()
    parity = {"oracle": "canonical (order-free) mode, teacher-forced", "steps": n_par, "logits_max_abs_err": err["canonical"], "tolerance": 2e-2,
              "greedy_tokens_equal": f"{same['canonical']}/{n_par}",
              "parity_ggml_order": {"oracle": "ggml generic C order (mode 1), same teacher-forced tokens", "steps": n_par,
                                    "logits_max_abs_err": err["ggml"], "greedy_tokens_equal": f"{same['ggml']}/{n_par}",
                                    "first_divergence": first_div["ggml"]},
              "ggml_build_noise": {"what": "oracle in ggml's AVX2 kernel order (mode 3) against the generic C order (mode 1): two legitimate ggml builds",
                                   "steps": n_par, "logits_max_abs_diff": err["noise"], "greedy_tokens_equal": f"{same['noise']}/{n_par}",
                                   "first_divergence": first_div["noise"]}}
    # ---- timing: ggml-generic path ----------------------------------------------------------------
    om = o.OracleModel(path, faithful=1, n_ctx=n + n_time + 4)
    for il, (k, v) in enumerate(kv):
        om.set_kv(il, k, v)
    tok = int(np.argmax(om.eval(7, n)))              # warm-up (also faults the weights in)
    t0 = time.perf_counter()
    for i in range(n_time):
        tok = int(np.argmax(om.eval(tok, n + 1 + i)))
    dt = time.perf_counter() - t0
    base = {"value": n_time / dt, "unit": UNIT, "cores": o.lib().ref_num_threads(), "kind": "port",
            "sample": f"{n_time} decode steps at ctx {n}+ continuing the GPU's sequence (KV cache copied from HBM), ggml-generic path"}
    om.close()
    return base, parity


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama2-7b")
    ap.add_argument("--ctx", type=int, default=512)
    ap.add_argument("--no-batch", action="store_true", help="skip the 32-sequence decode measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-tp", action="store_true", help="skip the tensor-parallel section (llama2-70B over the job's GPUs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
