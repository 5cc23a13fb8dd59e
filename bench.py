#!/usr/bin/env python
"""Benchmark of the tutoring-inference hot path (BASELINE.json).

Headline workload = configs[2]: GPT-2 small (124M), bf16 weights + KV cache, 64 requests per GPU, 512-token prompts,
256 new tokens each, greedy, requests batch-sharded over N GPUs (one replica per GPU, no collective).
One "step" = one whole generate() call for the batch (prefill + 255 decode forwards).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--no-extras]

Prints ONE JSON line (rank 0).  `value` = generated tokens/s with the prompts already resident in HBM;
`e2e` = the same through the reference-facing host API (CPU LongTensor in, CPU LongTensor out, pinned).
`--impl reference` times the reference's own CPU path (transformers fp32 on the host cores, via oracle/) on a
bounded sample of the same workload.
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


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return float(p["hbm_gbs"]), float(p.get("bf16_tflops", 1590.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1590.0, "fallback (B200_PROFILING.md)"


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


def cpu_sample(steps=1, rows=8, new=8):
    """The reference's CPU path (transformers fp32, oracle.generate == REF/tutoring_server.py:21-29 kwargs) on a bounded
    sample: `rows` of the 64 requests, the full 512-token prefill and `new` decode steps, extrapolated to the
    256-token workload as B*NEW / (B/rows * (t_prefill + (NEW-1) * t_step))."""
    import torch

    from oracle import reference_path as rp
    torch.set_grad_enabled(False)
    model = rp.build_gpt2("small", seed=0)
    ids = rp.gpt2_prompt_ids(B_PER_GPU, T_PROMPT, seed=1)[:rows]
    vals = []
    for _ in range(steps):
        t0 = time.perf_counter()
        rp.generate(model, ids, max_new_tokens=1, repetition_penalty=1.0)
        t_prefill = time.perf_counter() - t0
        t0 = time.perf_counter()
        rp.generate(model, ids, max_new_tokens=new, repetition_penalty=1.0)
        t_step = max(1e-9, (time.perf_counter() - t0 - t_prefill) / (new - 1))
        vals.append(rows * NEW / (t_prefill + (NEW - 1) * t_step))
    sample = (f"{rows} of 64 rows: 512-token prefill + {new} decode steps of oracle.generate (transformers "
              f"{__import__('transformers').__version__} fp32, sdpa), extrapolated to 256 new tokens; "
              f"torch threads {torch.get_num_threads()} of {os.cpu_count()} cpus")
    return statistics.median(vals), torch.get_num_threads(), sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    for _ in range(args.warmup and 1):
        cpu_sample(1, rows=2, new=2)
    v, cores, sample = cpu_sample(max(1, min(args.steps, 3)))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": B_PER_GPU * NEW / v * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "timing": "host perf_counter; CPU only"},
        "cpu_baseline": {"value": v, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0}))


def attn_kernel_roofline(lib, check, torch, hbm_peak, peak_src):
    """Dominant kernel of the step (86 % of its bytes): KV-streaming decode attention, timed alone with CUDA events
    on the launching stream at the workload's mid-point cache length, rotating over 12 layers of cache (1.5 GB > L2)."""
    B, H, S, L = B_PER_GPU, 12, T_PROMPT + NEW // 2, 12
    Smax = 768
    g = torch.Generator(device="cuda").manual_seed(0)
    k = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
    v = torch.randn(L, B, H, Smax, 64, device="cuda", generator=g).bfloat16()
    q = torch.randn(B, H * 64, device="cuda", generator=g).bfloat16()
    out = torch.empty_like(q)
    st = torch.cuda.current_stream().cuda_stream

    def launch(l):
        check(lib.tutor_attn_decode(1, q.data_ptr(), k[l].data_ptr(), v[l].data_ptr(), out.data_ptr(), B, H, Smax, S, st))
    for l in range(L):
        launch(l)
    torch.cuda.synchronize()
    # the 12 launches are captured into one CUDA graph so that the host (Python + ctypes, ~10 us per call) is not
    # what is being timed; events bracket 10 replays on the launching stream
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        st = side.cuda_stream
        with torch.cuda.graph(graph, stream=side):
            for l in range(L):
                launch(l)
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
    bytes_per_launch = B * H * S * 64 * 2 * 2 + 2 * B * H * 64 * 2  # K+V rows read, q read, out written
    achieved = bytes_per_launch / (us * 1e-6) / 1e9
    return {"kernel": "attn_decode_tma_kernel<bf16> (cp.async.bulk + mbarrier ring)", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
            "frac": achieved / hbm_peak,
            # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this kernel at
            # this shape (profiles/r1d_ncu_attn_decode_tma_full.txt): 125.95 MB read + 5.5 MB written
            "traffic": 131.5e6, "us_per_launch": us, "bytes_per_launch": bytes_per_launch,
            "peak_source": peak_src, "shape": f"B={B} H={H} S={S} hd=64 bf16"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg0 / cfg1 side measurements and the CPU baseline")
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
    hbm_peak, tc_peak, peak_src = peaks()
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
    for _ in range(max(3, args.warmup)):
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

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "requests_per_gpu": B_PER_GPU, "prompt_len": T_PROMPT, "new_tokens": new,
                   "parallelism": f"dp{world} (independent replicas, no collective)",
                   "l2": "inputs larger than L2 (weights 247 MB + KV cache 1.2-1.8 GB per step vs 126 MB L2)",
                   "timing": "CUDA events on the engine stream, barrier+synchronize both sides, max over ranks"},
        "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": world * B_PER_GPU * T_PROMPT * 4,
                "d2h_bytes_per_step": world * (B_PER_GPU * (T_PROMPT + new) * 4 + 8), "ms_per_step": ms_e2e,
                "api": "B200GPT2LMHeadModel.generate(cpu LongTensor) -> cpu LongTensor (tutor_gpt2_generate)"},
        "gpu_launches": int(launches) * world,
        "clocks": clocks.summary(),
        "decode": {"ms_per_forward": ms_decode, "tokens_per_s_per_gpu": B_PER_GPU / (ms_decode * 1e-3),
                   "prefill_ms": ms_prefill, "algorithmic_gb_per_forward": step_bytes / 1e9, "achieved_gbs": step_gbs,
                   "frac_of_hbm_peak": step_gbs / hbm_peak},
    }
    try:
        line["roofline"] = attn_kernel_roofline(lib, check, torch, hbm_peak, peak_src)
    except Exception as e:  # keep the headline line even if the side measurement fails
        line["roofline"] = {"error": repr(e)}
    if world == 1 and not args.no_extras:
        try:
            v, cores, sample = cpu_sample(1)
            line["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": cores, "kind": "reference", "sample": sample}
        except Exception as e:
            line["cpu_baseline"] = {"error": repr(e)}
        try:
            line["extras"] = extras(lms_b200, rp, torch, hf, tc_peak, hbm_peak)
        except Exception as e:
            line["extras"] = {"error": repr(e)}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def extras(pkg, rp, torch, hf_small, tc_peak, hbm_peak):
    """The two other figures BASELINE.json's metric names, at 1 GPU: configs[0] (bs=1 fp32 decode) and configs[1]
    (BERT relevance queries/s)."""
    res = {}
    eng = pkg.B200GPT2LMHeadModel.from_hf(hf_small, dtype="fp32")
    ids = rp.gpt2_prompt_ids(1, 32, seed=1)
    for _ in range(3):
        eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    t0 = time.perf_counter()
    n = 5
    for _ in range(n):
        eng.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    dt = (time.perf_counter() - t0) / n
    gb = sum(eng.decode_bytes(1, 32 + i) for i in range(1, 64)) / 1e9
    res["cfg0_gpt2_small_fp32_bs1"] = {"tokens_per_s": 64 / dt, "ms_per_generate": dt * 1e3, "timing": "host wall, e2e",
                                        "algorithmic_gb": gb, "frac_of_hbm_peak": gb / dt / hbm_peak}
    del eng
    bert_hf = rp.build_bert(seed=0)
    bert = pkg.B200BertModel.from_hf(bert_hf, dtype="bf16")
    q, a = rp.bert_ids(32, 128, seed=2), rp.bert_ids(32, 512, seed=5)
    qd, ad = q.cuda(), a.cuda()
    def gate_dev():
        return pkg.cosine_similarity_batch(bert.embed(qd), bert.embed(ad))
    def gate_host():
        return pkg.cosine_similarity_batch(bert.embed(q), bert.embed(a))
    for _ in range(3):
        gate_dev(); gate_host()
    torch.cuda.synchronize()
    n = 10
    t0 = time.perf_counter()
    for _ in range(n):
        gate_dev()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n
    t0 = time.perf_counter()
    for _ in range(n):
        gate_host()
    dt_host = (time.perf_counter() - t0) / n
    res["cfg1_bert_base_bf16_32pairs"] = {"queries_per_s": 32 / dt, "ms_per_batch": dt * 1e3,
                                           "e2e_queries_per_s": 32 / dt_host, "e2e_ms_per_batch": dt_host * 1e3,
                                           "timing": "host wall around synchronous calls; ids resident in HBM / e2e from host ids",
                                           "tflops": 3.81 / dt, "frac_of_bf16_peak": 3.81 / dt / tc_peak}
    del bert
    # configs[4]: 1M x 768 fp32 corpus on one GPU, 32 queries, top-10 (the 8-GPU form shards the rows)
    g = torch.Generator(device="cuda").manual_seed(3)
    corpus = torch.randn(1_000_000, 768, device="cuda", generator=g)
    shard = pkg.CorpusShard(corpus)
    qs = torch.randn(32, 768, device="cuda", generator=g)
    for _ in range(3):
        shard.topk(qs, 10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        shard.topk(qs, 10)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    gbs = 1_000_000 * 768 * 4 / (us * 1e-6) / 1e9
    res["cfg4_corpus_1M_x768_fp32_32q_top10"] = {"us_per_search": us, "queries_per_s": 32 / (us * 1e-6), "achieved_gbs": gbs,
                                                  "frac_of_hbm_peak": gbs / hbm_peak, "timing": "CUDA events"}
    del corpus, shard
    # configs[3]: GPT-2 large bf16, B=16 - (ii) prefill-only over the full 1024-token context (one forward, first token
    # only; 24.74 TFLOP) and (i) 896-token prompt -> +128 tokens end to end (1024 total: wpe has 1024 rows)
    torch.cuda.empty_cache()
    large = pkg.B200GPT2LMHeadModel.from_hf(rp.build_gpt2("large", seed=0), dtype="bf16")
    ids_full, ids_896 = rp.gpt2_prompt_ids(16, 1024, seed=1).cuda(), rp.gpt2_prompt_ids(16, 896, seed=1)
    for _ in range(2):
        large.generate(ids_full, max_new_tokens=1, do_sample=False)
    torch.cuda.synchronize()
    n = 3
    t0 = time.perf_counter()
    for _ in range(n):
        large.generate(ids_full, max_new_tokens=1, do_sample=False)
    dt_pre = (time.perf_counter() - t0) / n
    large.generate(ids_896, max_new_tokens=128, do_sample=False)
    t0 = time.perf_counter()
    for _ in range(n):
        large.generate(ids_896, max_new_tokens=128, do_sample=False)
    dt_gen = (time.perf_counter() - t0) / n
    res["cfg3_gpt2_large_bf16_bs16"] = {"prefill_1024_ms": dt_pre * 1e3, "prefill_tflops": 24.74 / dt_pre,
                                        "prefill_frac_of_bf16_peak": 24.74 / dt_pre / tc_peak,
                                        "e2e_896_plus_128_ms": dt_gen * 1e3, "e2e_new_tokens_per_s": 16 * 128 / dt_gen,
                                        "timing": "host wall around synchronous generate() calls (ids in HBM for the "
                                                  "prefill figure, host ids for the end-to-end one)"}
    del large
    return res


if __name__ == "__main__":
    main()
