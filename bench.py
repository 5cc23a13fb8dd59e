#!/usr/bin/env python
"""Benchmark of the tutoring-inference hot path (BASELINE.json).

Headline workload = configs[2]: GPT-2 small (124M), bf16 weights + KV cache, 64 requests per GPU, 512-token prompts,
256 new tokens each, greedy, requests batch-sharded over N GPUs (one replica per GPU, no collective).
One "step" = one whole generate() call for the batch (prefill + 255 decode forwards).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--no-extras]

Prints ONE JSON line (rank 0).  `value` = generated tokens/s with the prompts already resident in HBM;
`e2e` = the same through the reference-facing host API (CPU LongTensor in, CPU LongTensor out, pinned).
`legs` carries the other figures of BASELINE.json's metric as measurements of their own - configs[0] (bs-1 fp32),
configs[1] (BERT gate), configs[3] (GPT-2 large), configs[4] (corpus search; sharded over the N GPUs with the NCCL
exchange when N > 1) - each timed with CUDA events on the stream its kernels run on, with its roofline fraction, a
clock sample, and the reference's own CPU call timed beside it.
`--impl reference` times the reference's own CPU path (transformers fp32 on the host cores, via oracle/) on a
bounded sample of the same workload; both arms use the same estimator (cpu_cfg2) and print the same `config`.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, T_PROMPT, NEW = 64, 512, 256
METRIC = "gpt2_small_bs64_greedy_decode_tokens_per_s"
WORKLOAD = ("configs[2]: GPT-2 small bf16, 64 requests/GPU x (512-token prompt -> 256 new tokens), greedy, "
            "random-init weights, synthetic ids")


def log(msg):
    """progress on stderr (the JSON line is the only thing on stdout)"""
    if os.environ.get("TUTOR_BENCH_VERBOSE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} +{time.perf_counter() - T0:.1f}s] {msg}", file=sys.stderr, flush=True)


T0 = time.perf_counter()


def config(world, new=NEW):
    """The same dict in both arms (the reference arm computes in fp32 on the host; see `dtype`)."""
    return {"workload": WORKLOAD, "requests_per_gpu": B_PER_GPU, "prompt_len": T_PROMPT, "new_tokens": new,
            "parallelism": f"dp{world} (independent replicas, no collective)"}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        burst = float(p.get("bf16_tflops", 1590.0))
        return {"hbm": float(p["hbm_gbs"]), "tc_burst": burst, "tc_sustained": float(p.get("bf16_tflops_sustained", burst)),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm": 6650.0, "tc_burst": 1590.0, "tc_sustained": 1590.0, "source": "fallback (B200_PROFILING.md)"}


def measured_traffic(kernel):
    """DRAM bytes per launch of `kernel` from this round's `ncu --set full` capture (profiles/r2_ncu_traffic.json,
    written by tools/ncu_traffic.py from the .ncu-rep); None if there is no capture for it."""
    path = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
    try:
        rec = json.load(open(path)).get(kernel)
        return None if rec is None else {"bytes": rec["dram_read"] + rec["dram_write"], "source": "profiles/r2_ncu_traffic.json"}
    except (OSError, ValueError, KeyError):
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                out, _ = self.proc.communicate(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
                out = ""
            self.lines = [l for l in out.splitlines() if l.strip()]

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------- CPU reference side
def host_threads():
    """All host cores for the one rank that times the CPU path (torchrun exports OMP_NUM_THREADS=1 for every rank, which
    would leave the reference on one core)."""
    import torch
    n = os.cpu_count() or 1
    torch.set_num_threads(n)
    return torch.get_num_threads()


class _StepStamps:
    """A no-op logits processor that timestamps every call: generate() calls it once per forward (prefill included), so
    one stock `model.generate` call yields the prefill time and every decode-step time without subtracting two runs."""

    def __init__(self):
        self.t = []

    def __call__(self, input_ids, scores):
        self.t.append(time.perf_counter())
        return scores


CPU_ROWS, CPU_DECODE_STEPS = 32, 32


def cpu_cfg2(model=None, rows=CPU_ROWS, n_dec=CPU_DECODE_STEPS):
    """The reference's CPU path (transformers fp32, oracle.generate == REF/tutoring_server.py:21-29 kwargs) on a bounded
    sample of configs[2]: `rows` of the 64 requests, the full 512-token prefill and `n_dec` decode steps in ONE stock
    generate() call, every forward timestamped.  Tokens/s for the 256-token workload = rows*256 / (t_prefill + 255 *
    median decode step); the decode step is sampled at cache lengths 512..512+n_dec (on the CPU the KV reads are ~2 % of a
    step's work, so cache growth to 767 moves it by less than the run-to-run spread)."""
    import torch
    from transformers import LogitsProcessorList

    from oracle import reference_path as rp
    torch.set_grad_enabled(False)
    if model is None:
        model = rp.build_gpt2("small", seed=0)
    ids = rp.gpt2_prompt_ids(B_PER_GPU, T_PROMPT, seed=1)[:rows]
    stamps = _StepStamps()
    t0 = time.perf_counter()
    rp.generate(model, ids, max_new_tokens=1 + n_dec, repetition_penalty=1.0, logits_processor=LogitsProcessorList([stamps]))
    t_end = time.perf_counter()
    t_prefill = stamps.t[0] - t0
    steps = [b - a for a, b in zip(stamps.t[:-1], stamps.t[1:])]
    t_step = statistics.median(steps)
    total = t_prefill + (NEW - 1) * t_step
    return {"value": rows * NEW / total, "sample_ms": (t_end - t0) * 1e3, "prefill_ms": t_prefill * 1e3,
            "decode_step_ms": t_step * 1e3, "extrapolated_ms_per_step": total * (B_PER_GPU / rows) * 1e3,
            "sample": (f"{rows} of 64 rows: 512-token prefill + {n_dec} decode steps in one stock model.generate call "
                       f"(transformers {__import__('transformers').__version__} fp32, sdpa), each forward timestamped; "
                       f"extrapolated to 256 new tokens; torch threads {torch.get_num_threads()} of {os.cpu_count()} cpus")}


def cpu_cfg0(model):
    """configs[0] on the CPU, whole workload: bs 1, 32-token prompt -> 64 new tokens, penalty 1.2 (median of 3)."""
    from oracle import reference_path as rp
    ids = rp.gpt2_prompt_ids(1, 32, seed=1)
    rp.generate(model, ids, max_new_tokens=8, repetition_penalty=1.2)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        rp.generate(model, ids, max_new_tokens=64, repetition_penalty=1.2)
        ts.append(time.perf_counter() - t0)
    return 64 / statistics.median(ts)


def cpu_cfg1(pairs=4):
    """configs[1] as the reference runs it (REF/lms_server.py:97-104, 1262-1266): per request two bs-1 BERT forwards
    (128-token query, 512-token assignment) + cosine, on `pairs` of the 32 pairs; queries/s."""
    from oracle import reference_path as rp
    t0 = time.perf_counter()
    model = rp.build_bert(seed=0)
    load_s = time.perf_counter() - t0  # outside the timed region; the reference pays a model load per request (lms_server.py:1258-1260)
    q, a = rp.bert_ids(32, 128, seed=2), rp.bert_ids(32, 512, seed=5)
    rp.cosine_similarity(rp.embeddings_from_ids(model, q[:1]), rp.embeddings_from_ids(model, a[:1]))
    t0 = time.perf_counter()
    for i in range(pairs):
        rp.cosine_similarity(rp.embeddings_from_ids(model, q[i:i + 1]), rp.embeddings_from_ids(model, a[i:i + 1]))
    return pairs / (time.perf_counter() - t0), load_s


def cpu_cfg4(rows=100_000):
    """configs[4] on the CPU: torch cosine + topk of 32 queries over a `rows`-row sample; seconds per 1 M rows."""
    import torch

    from oracle import reference_path as rp
    g = torch.Generator().manual_seed(3)
    corpus, q = torch.randn(rows, 768, generator=g), torch.randn(32, 768, generator=g)
    rp.corpus_topk(q, corpus[:1000], 10)
    t0 = time.perf_counter()
    rp.corpus_topk(q, corpus, 10)
    return (time.perf_counter() - t0) * (1_000_000 / rows)


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = host_threads()
    t0 = time.perf_counter()
    from oracle import reference_path as rp
    model = rp.build_gpt2("small", seed=0)
    for _ in range(1 if args.warmup else 0):
        cpu_cfg2(model, rows=2, n_dec=2)  # warms the allocator / kernel selection; the samples below are the timed ones
    samples = [cpu_cfg2(model) for _ in range(max(1, min(args.steps, 3)))]
    s = sorted(samples, key=lambda r: r["value"])[len(samples) // 2]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": s["value"], "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": s["sample_ms"],
        "extrapolated_ms_per_step": s["extrapolated_ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config(args.gpus),
        "method": {"timing": "host perf_counter inside one model.generate call per step; CPU only; ms_per_step is the "
                             "sample that was timed, extrapolated_ms_per_step the whole 64 x (512 -> +256) batch",
                   "samples_tokens_per_s": [r["value"] for r in samples], "prefill_ms": s["prefill_ms"],
                   "decode_step_ms": s["decode_step_ms"]},
        "cpu_baseline": {"value": s["value"], "unit": "tokens/s", "cores": threads, "kind": "reference", "sample": s["sample"]},
        "e2e": {"value": s["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0}))


# ------------------------------------------------------------------------------------------------------ GPU legs
def event_timer(torch, stream):
    """ms between two CUDA events recorded on `stream` around fn() x n (synchronised on both sides)."""
    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    return timed


def attn_kernel_roofline(lib, check, torch, pk):
    """Dominant kernel of the step (86 % of its bytes): the KV-streaming decode attention AS THE STEP LAUNCHES IT (c_attn
    split-K partials summed + bias, K/V row appended, attention over the cache), timed alone with CUDA events on the
    launching stream at the workload's mid-point cache length, rotating over 12 layers of cache (1.5 GB > L2)."""
    B, H, L, ks = B_PER_GPU, 12, 12, 2  # ks: the K splits of c_attn at this batch size (decode_plan: 18 x 4 tiles x 2)
    S = T_PROMPT + NEW // 2  # tokens including the one being decoded: 639 cached rows + the new one
    Smax, d = 768, H * 64
    g = torch.Generator(device="cuda").manual_seed(0)
    k = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
    v = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
    part = torch.randn(ks, B, 3 * d, device="cuda", generator=g) * 0.3
    bias = torch.randn(3 * d, device="cuda", generator=g) * 0.1
    out = torch.empty(B, d, device="cuda", dtype=torch.bfloat16)
    side = torch.cuda.Stream()

    def launch(l, st):
        check(lib.tutor_attn_decode_fused(part.data_ptr(), ks, bias.data_ptr(), k[l].data_ptr(), v[l].data_ptr(),
                                          out.data_ptr(), B, H, Smax, S, st))
    for l in range(L):
        launch(l, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    # the 12 launches are captured into one CUDA graph so that the host (Python + ctypes, ~10 us per call) is not
    # what is being timed; events bracket 10 replays on the launching stream
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            for l in range(L):
                launch(l, side.cuda_stream)
        graph.replay()
        side.synchronize()
        reps = 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(reps):
            graph.replay()
        e1.record(side)
        side.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * L)
    # K+V rows read, partials + bias read, new K/V row and the output written
    bytes_per_launch = B * H * (S - 1) * 64 * 2 * 2 + ks * B * 3 * d * 4 + 3 * d * 4 + B * 2 * d * 2 + B * d * 2
    achieved = bytes_per_launch / (us * 1e-6) / 1e9
    tr = measured_traffic("attn_decode_tma_kernel<1,64,2,1>")
    return {"kernel": "attn_decode_tma_kernel<fused> (cp.async.bulk + mbarrier ring; the instance the decode step runs)",
            "bound": "hbm", "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"],
            "traffic": tr["bytes"] if tr else None, "traffic_source": tr["source"] if tr else None,
            "us_per_launch": us, "bytes_per_launch": bytes_per_launch, "peak_source": pk["source"],
            "shape": f"B={B} H={H} len={S} hd=64 bf16, ks={ks}"}


def leg_cfg0(pkg, rp, torch, hf_small, pk, local, with_cpu):
    eng = pkg.B200GPT2LMHeadModel.from_hf(hf_small, dtype="fp32", device=local)
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))
    timed = event_timer(torch, stream)
    ids = rp.gpt2_prompt_ids(1, 32, seed=1)
    gen = lambda: eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    pre = lambda: eng.generate(ids, max_new_tokens=1, repetition_penalty=1.2)
    for _ in range(3):
        gen(); pre()
    with ClockSampler(local) as clocks:
        ms = timed(gen, 10)
        ms_pre = timed(pre, 10)
    ms_fwd = (ms - ms_pre) / 63
    gb = sum(eng.decode_bytes(1, 32 + i) for i in range(1, 64)) / 63 / 1e9
    res = {"workload": "configs[0]: GPT-2 small fp32, bs 1, 32-token prompt -> 64 new tokens, penalty 1.2, e2e from host ids",
           "tokens_per_s": 64 / (ms * 1e-3), "ms_per_generate": ms, "prefill_ms": ms_pre, "decode_forward_us": ms_fwd * 1e3,
           "timing": "CUDA events on the engine stream around 10 generate() calls",
           "roofline": {"bound": "hbm", "achieved": gb / (ms_fwd * 1e-3), "peak": pk["hbm"], "unit": "GB/s",
                        "frac": gb / (ms_fwd * 1e-3) / pk["hbm"], "algorithmic_gb_per_forward": gb, "what": "decode forward"},
           "clocks": clocks.summary()}
    if with_cpu:
        v = cpu_cfg0(hf_small)
        res["cpu_baseline"] = {"value": v, "unit": "tokens/s", "kind": "reference",
                               "sample": "the whole workload through oracle.generate (model.generate as REF/tutoring_server.py:21-29), median of 3"}
    return res


def leg_cfg1(pkg, rp, torch, pk, local, with_cpu):
    bert = pkg.B200BertModel.from_hf(rp.build_bert(seed=0), dtype="bf16", device=local)
    stream = torch.cuda.ExternalStream(bert.lib.tutor_bert_stream(bert._h), device=torch.device("cuda", local))
    timed = event_timer(torch, stream)
    q, a = rp.bert_ids(32, 128, seed=2), rp.bert_ids(32, 512, seed=5)
    qd, ad = q.cuda(), a.cuda()
    gate_dev = lambda: pkg.cosine_similarity_batch(bert.embed(qd), bert.embed(ad))
    gate_host = lambda: pkg.cosine_similarity_batch(bert.embed(q), bert.embed(a))
    for _ in range(3):
        gate_dev(); gate_host()
    with ClockSampler(local) as clocks:
        ms = timed(gate_dev, 20)
        ms_host = timed(gate_host, 20)
    tflop = 3.81
    res = {"workload": "configs[1]: BERT-base bf16, 32 pairs (128-token query, 512-token assignment) -> mean-pool -> cosine",
           "queries_per_s": 32 / (ms * 1e-3), "ms_per_batch": ms, "e2e_queries_per_s": 32 / (ms_host * 1e-3),
           "e2e_ms_per_batch": ms_host, "timing": "CUDA events on the engine stream around 20 batches (ids in HBM / e2e from host ids)",
           "roofline": {"bound": "tensor", "achieved": tflop / (ms * 1e-3), "peak": pk["tc_sustained"], "unit": "TFLOP/s",
                        "frac": tflop / (ms * 1e-3) / pk["tc_sustained"], "frac_of_burst_peak": tflop / (ms * 1e-3) / pk["tc_burst"],
                        "algorithmic_tflop": tflop},
           "clocks": clocks.summary()}
    if with_cpu:
        v, load_s = cpu_cfg1()
        res["cpu_baseline"] = {"value": v, "unit": "queries/s", "kind": "reference",
                               "sample": "4 of the 32 pairs, each as the reference does it: two bs-1 BERT forwards + cosine (REF/lms_server.py:97-104)",
                               "model_construction_s": load_s,
                               "note": "the reference also loads BERT from disk on every request (REF/lms_server.py:1258-1260); "
                                       "model_construction_s (random-init BertModel, no checkpoint offline) is outside the timed region"}
    return res


def leg_cfg4(pkg, rp, torch, pk, local, world, rank, with_cpu):
    """configs[4]: 1 M x 768 fp32 corpus, 32 queries, top-10.  One GPU scans everything; N GPUs scan 1M/N rows each and
    exchange their candidates with one ncclAllGather (the path's only collective), the whole search as one CUDA graph."""
    import torch.distributed as dist
    N, Q, k = 1_000_000, 32, 10
    lo, hi = pkg.shard_bounds(N, world, rank)
    g = torch.Generator().manual_seed(3)
    rows = []
    for start in range(0, N, 100_000):  # every rank draws the same stream and keeps its slice
        blk = torch.randn(100_000, 768, generator=g)
        a, b = max(lo, start), min(hi, start + 100_000)
        if a < b:
            rows.append(blk[a - start:b - start].cuda())
    queries = torch.randn(Q, 768, generator=g)
    shard = pkg.CorpusShard(torch.cat(rows), row_lo=lo)
    del rows
    qd = queries.cuda()
    log(f"cfg4: shard rows [{lo}, {hi}) resident")
    if world > 1:
        pkg.init_exchange()
        log("cfg4: exchange communicator up")
    shard.search(qd, k)
    torch.cuda.synchronize()
    log("cfg4: first search done")
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        shard.search(qd, k)
        side.synchronize()
        with torch.cuda.graph(graph, stream=side):
            gs, gi = shard.search(qd, k)
        graph.replay()
        side.synchronize()
        log("cfg4: graph captured and replayed once")
        if world > 1:
            dist.barrier()
        with ClockSampler(local) as clocks:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(side)
            for _ in range(50):
                graph.replay()
            e1.record(side)
            side.synchronize()
    us = torch.tensor([e0.elapsed_time(e1) * 1e3 / 50], device="cuda")
    if world > 1:
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
    us = us.item()
    log(f"cfg4: timed, {us:.1f} us per search")
    gs, gi = gs.clone(), gi.clone()
    del graph  # NCCL keeps a communicator alive while a captured graph still refers to it: ncclCommDestroy would wait
    if world > 1:  # every rank tears its communicator down together, before rank 0 goes off to check against the oracle
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        pkg.close_exchange()
        log("cfg4: exchange communicator closed")
    res = None
    if rank == 0:
        host_threads()
        bytes_per_gpu = (hi - lo) * 768 * 4
        res = {"workload": f"configs[4]: 1M x 768 fp32 corpus sharded over {world} GPU(s), 32 queries, top-10"
                           + ("; per-GPU scan + top-k, one ncclAllGather of packed candidates, merge; one CUDA graph" if world > 1 else ""),
               "us_per_search": us, "queries_per_s": Q / (us * 1e-6), "timing": "CUDA events around 50 graph replays, max over ranks",
               "roofline": {"bound": "hbm", "achieved": bytes_per_gpu / (us * 1e-6) / 1e9, "peak": pk["hbm"], "unit": "GB/s per GPU",
                            "frac": bytes_per_gpu / (us * 1e-6) / 1e9 / pk["hbm"], "algorithmic_bytes_per_gpu": bytes_per_gpu},
               "clocks": clocks.summary()}
        g = torch.Generator().manual_seed(3)
        full = torch.cat([torch.randn(100_000, 768, generator=g) for _ in range(N // 100_000)])
        t0 = time.perf_counter()
        rs, ri = rp.corpus_topk(queries, full, k)
        t_cpu = time.perf_counter() - t0
        res["indices_identical_to_oracle"] = bool(torch.equal(gi.cpu().long(), ri))
        res["max_score_error"] = float((gs.cpu() - rs).abs().max())
        if with_cpu:
            res["cpu_baseline"] = {"value": Q / t_cpu, "unit": "queries/s", "kind": "reference",
                                   "sample": "the whole 1 M-row search: torch normalise + matmul + topk on the host (oracle.corpus_topk)"}
    return res


def leg_tokenize(pkg):
    """The host step either side of the GPU work: prompt templating + GPT-2 BPE encode,
    decode of the 150-token reply, BERT WordPiece of a query and of an assignment text - us per request through the C
    library, with transformers' own tokenizers (the reference's) timed beside them.  Vocabularies: tests/golden/vocab."""
    import random
    from transformers import BertTokenizer, GPT2Tokenizer
    d = os.path.join(ROOT, "tests", "golden", "vocab")
    v, m, bv = os.path.join(d, "gpt2_vocab.json"), os.path.join(d, "gpt2_merges.txt"), os.path.join(d, "bert_vocab.txt")
    mine_g, hf_g = pkg.B200GPT2Tokenizer(v, m), GPT2Tokenizer(vocab=v, merges=m)
    mine_b = pkg.B200BertTokenizer(bv)
    hf_b = BertTokenizer(vocab={l.rstrip("\n"): i for i, l in enumerate(open(bv, encoding="utf-8"))})
    rng = random.Random(3)
    words = open(os.path.join(ROOT, "SURVEY.md"), encoding="utf-8").read().split()
    text = lambda n: " ".join(rng.choice(words) for _ in range(n))
    prompts = ["Answer the following question as a helpful tutor: " + text(20) for _ in range(200)]
    replies = [mine_g.encode(text(110))[:150] for _ in range(200)]
    docs = [text(600) for _ in range(50)]

    def clock(fn, items):
        fn(items[0])
        t0 = time.perf_counter()
        for x in items:
            fn(x)
        return (time.perf_counter() - t0) / len(items) * 1e6
    kw = dict(return_tensors="pt", padding=True, truncation=True, max_length=512)
    return {"workload": "prompt template + GPT-2 BPE encode; decode of a 150-token reply; BERT WordPiece of a 20-word query and "
                        "a 600-word assignment text (cut at 512 ids); synthetic vocabularies (tests/golden/vocab)",
            "us_per_request": {"gpt2_encode": clock(lambda p: mine_g.encode(p, return_tensors="pt"), prompts),
                               "gpt2_decode": clock(lambda r: mine_g.decode(r, skip_special_tokens=True), replies),
                               "bert_query": clock(lambda p: mine_b(p, **kw), prompts), "bert_assignment": clock(lambda t: mine_b(t, **kw), docs)},
            "cpu_baseline": {"kind": "reference", "unit": "us/request", "sample": "transformers GPT2Tokenizer / BertTokenizer on the same inputs",
                             "gpt2_encode": clock(lambda p: hf_g.encode(p, return_tensors="pt"), prompts),
                             "gpt2_decode": clock(lambda r: hf_g.decode(r, skip_special_tokens=True), replies),
                             "bert_query": clock(lambda p: hf_b(p, **kw), prompts), "bert_assignment": clock(lambda t: hf_b(t, **kw), docs)}}


def leg_cfg3(pkg, rp, torch, pk, local):
    """configs[3]: GPT-2 large bf16, B=16 - (ii) prefill-only over the full 1024-token context (one forward, first token
    only; 24.74 TFLOP) and (i) 896-token prompt -> +128 tokens end to end (1024 total: wpe has 1024 rows)."""
    large = pkg.B200GPT2LMHeadModel.from_hf(rp.build_gpt2("large", seed=0), dtype="bf16", device=local)
    stream = torch.cuda.ExternalStream(large.stream, device=torch.device("cuda", local))
    timed = event_timer(torch, stream)
    ids_full, ids_896 = rp.gpt2_prompt_ids(16, 1024, seed=1).cuda(), rp.gpt2_prompt_ids(16, 896, seed=1)
    pre = lambda: large.generate(ids_full, max_new_tokens=1, do_sample=False)
    pre896 = lambda: large.generate(ids_896, max_new_tokens=1, do_sample=False)
    gen = lambda: large.generate(ids_896, max_new_tokens=128, do_sample=False)
    for _ in range(2):
        pre(); pre896()
    gen()
    with ClockSampler(local) as clocks:
        ms_pre = timed(pre, 3)
        ms_pre896 = timed(pre896, 3)
        ms_gen = timed(gen, 3)
    ms_fwd = (ms_gen - ms_pre896) / 127
    gb = sum(large.decode_bytes(16, 896 + i) for i in range(1, 128)) / 127 / 1e9
    return {"workload": "configs[3]: GPT-2 large bf16, B=16; prefill of the full 1024-token context, and 896 -> +128 tokens e2e",
            "prefill_1024_ms": ms_pre, "e2e_896_plus_128_ms": ms_gen, "e2e_new_tokens_per_s": 16 * 128 / (ms_gen * 1e-3),
            "decode_forward_us": ms_fwd * 1e3, "timing": "CUDA events on the engine stream around 3 generate() calls each",
            "roofline": {"bound": "tensor", "achieved": 24.74 / (ms_pre * 1e-3), "peak": pk["tc_sustained"], "unit": "TFLOP/s",
                         "frac": 24.74 / (ms_pre * 1e-3) / pk["tc_sustained"], "frac_of_burst_peak": 24.74 / (ms_pre * 1e-3) / pk["tc_burst"],
                         "what": "prefill, 24.74 TFLOP"},
            "roofline_decode": {"bound": "hbm", "achieved": gb / (ms_fwd * 1e-3), "peak": pk["hbm"], "unit": "GB/s",
                                "frac": gb / (ms_fwd * 1e-3) / pk["hbm"], "algorithmic_gb_per_forward": gb},
            "clocks": clocks.summary()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg0 / cfg1 / cfg3 / cfg4 legs and the CPU baselines")
    ap.add_argument("--new-tokens", type=int, default=NEW, help=argparse.SUPPRESS)  # profiling runs only
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    import lms_b200
    from oracle import reference_path as rp  # random-init weights + synthetic ids only (not on the timed path)

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    new = args.new_tokens
    pk = peaks()
    lib, check = lms_b200._lib.load(), lms_b200._lib.check

    hf = rp.build_gpt2("small", seed=0)
    eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16", device=local)
    ids_host = rp.gpt2_prompt_ids(B_PER_GPU, T_PROMPT, seed=1 + rank).pin_memory()
    ids_dev = ids_host.cuda()
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    run_dev = lambda: eng.generate(ids_dev, max_new_tokens=new, do_sample=False, repetition_penalty=1.0)
    run_host = lambda: eng.generate(ids_host, max_new_tokens=new, do_sample=False, repetition_penalty=1.0)
    warmup = max(3, args.warmup)
    for _ in range(warmup):
        out = run_dev()
    assert tuple(out.shape) == (B_PER_GPU, T_PROMPT + new)
    n0 = lib.tutor_launch_count()
    with ClockSampler(local) as clocks:
        ms = timed(run_dev, args.steps)
    launches = lib.tutor_launch_count() - n0
    ms_step = ms / args.steps
    tokens = world * B_PER_GPU * new
    value = tokens / (ms_step * 1e-3)
    run_host()
    ms_e2e = timed(run_host, args.steps) / args.steps
    e2e = tokens / (ms_e2e * 1e-3)

    # prefill alone (max_new_tokens=1) so the decode-only rate and the whole-step roofline can be stated
    pre = lambda: eng.generate(ids_dev, max_new_tokens=1, do_sample=False)
    pre()
    ms_prefill = timed(pre, 3) / 3
    ms_decode = max(1e-6, ms_step - ms_prefill) / max(1, new - 1)
    step_bytes = sum(eng.decode_bytes(B_PER_GPU, T_PROMPT + i) for i in range(1, new)) / max(1, new - 1)
    step_gbs = step_bytes / (ms_decode * 1e-3) / 1e9

    log(f"headline timed: {value:.0f} tok/s")
    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "config": config(world, new),
            "method": {"l2": "inputs larger than L2 (weights 247 MB + KV cache 1.2-1.8 GB per step vs 126 MB L2)",
                       "timing": "CUDA events on the engine stream, barrier+synchronize both sides, max over ranks"},
            "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": world * B_PER_GPU * T_PROMPT * 4,
                    "d2h_bytes_per_step": world * (B_PER_GPU * (T_PROMPT + new) * 4 + 8), "ms_per_step": ms_e2e,
                    "api": "B200GPT2LMHeadModel.generate(cpu LongTensor) -> cpu LongTensor (tutor_gpt2_generate)"},
            "gpu_launches": int(launches) * world,
            "clocks": clocks.summary(),
            "decode": {"ms_per_forward": ms_decode, "tokens_per_s_per_gpu": B_PER_GPU / (ms_decode * 1e-3),
                       "prefill_ms": ms_prefill, "algorithmic_gb_per_forward": step_bytes / 1e9, "achieved_gbs": step_gbs,
                       "frac_of_hbm_peak": step_gbs / pk["hbm"]},
            # the metric's "% HBM roofline": the WHOLE decode forward (weights + KV read + KV write) over its measured time
            "roofline_step": {"what": "configs[2] decode forward, averaged over cache lengths 512..766", "bound": "hbm",
                              "achieved": step_gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": step_gbs / pk["hbm"],
                              "frac_of_nominal_8tbs": step_gbs / 8000.0,
                              "algorithmic_gb": step_bytes / 1e9, "ms": ms_decode, "peak_source": pk["source"]},
        }
        try:
            line["roofline"] = attn_kernel_roofline(lib, check, torch, pk)
        except Exception as e:  # keep the headline line even if the side measurement fails
            line["roofline"] = {"error": repr(e)}
    del eng
    torch.cuda.empty_cache()
    legs = {}
    if not args.no_extras:
        if world > 1:  # the only collective on the path: sharded corpus search inside the NCCL group
            try:
                r = leg_cfg4(lms_b200, rp, torch, pk, local, world, rank, with_cpu=False)
                if rank == 0:
                    legs["cfg4_sharded"] = r
            except Exception as e:
                legs["cfg4_sharded"] = {"error": repr(e)}
        elif rank == 0:
            threads = host_threads()
            try:
                s = cpu_cfg2(hf, rows=2, n_dec=2) and cpu_cfg2(hf)  # one warm-up, one timed sample
                line["cpu_baseline"] = {"value": s["value"], "unit": "tokens/s", "cores": threads, "kind": "reference",
                                        "sample": s["sample"], "extrapolated_ms_per_step": s["extrapolated_ms_per_step"]}
            except Exception as e:
                line["cpu_baseline"] = {"error": repr(e)}
            for name, fn in (("cfg0", lambda: leg_cfg0(lms_b200, rp, torch, hf, pk, local, True)),
                             ("cfg1", lambda: leg_cfg1(lms_b200, rp, torch, pk, local, True)),
                             ("cfg4", lambda: leg_cfg4(lms_b200, rp, torch, pk, local, 1, 0, True)),
                             ("cfg3", lambda: leg_cfg3(lms_b200, rp, torch, pk, local)),
                             ("tokenize", lambda: leg_tokenize(lms_b200))):
                try:
                    legs[name] = fn()
                except Exception as e:
                    legs[name] = {"error": repr(e)}
                torch.cuda.empty_cache()
    if rank == 0:
        if legs:
            line["legs"] = legs
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
