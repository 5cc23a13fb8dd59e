"""-m gpu: GPT-2 generation parity against the committed goldens (made by running the UNMODIFIED reference
modules, oracle/make_golden.py) and against the oracle run live on the box's host cores.  Gates (BASELINE.json
north_star): fp32 mode -> token ids bit-identical, logits within 1e-3 relative; bf16 mode -> agreement rate stated."""
import numpy as np
import pytest
import torch

from oracle import reference_path as rp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small_hf():
    return rp.build_gpt2("small", seed=0)


@pytest.fixture(scope="module")
def small_f32(pkg, small_hf):
    return pkg.B200GPT2LMHeadModel.from_hf(small_hf, dtype="fp32")


@pytest.fixture(scope="module")
def small_bf16(pkg, small_hf):
    return pkg.B200GPT2LMHeadModel.from_hf(small_hf, dtype="bf16")


@pytest.fixture(scope="module")
def tiny_hf():
    return rp.build_gpt2("tiny", seed=3)


@pytest.mark.parametrize("pen,key", [(1.0, "p10"), (1.2, "p12")])
def test_cfg0_tokens_bit_identical(small_f32, golden, pen, key):
    """configs[0]: GPT-2 small fp32, B=1, 32-token prompt -> 64 new tokens."""
    g = golden("gpt2_cfg0")
    prompt = torch.from_numpy(g["prompt"])
    out = small_f32.generate(prompt, max_new_tokens=64, do_sample=False, repetition_penalty=pen)
    assert out.dtype == torch.int64 and tuple(out.shape) == (1, 96)
    ref = g[f"ids_{key}"]
    mism = np.nonzero(out.numpy() != ref)[1]
    assert mism.size == 0, f"first mismatch at position {mism[:1]}; min oracle top-1/top-2 gap {g[f'gap_{key}'].min():.2e}"


def test_cfg0_logits_within_1e3(small_f32, golden):
    g = golden("gpt2_cfg0")
    prompt = torch.from_numpy(g["prompt"])
    small_f32.prefill(prompt, max_new_tokens=64, repetition_penalty=1.2)
    worst = 0.0
    for advance, row in ((0, 0), (31, 1), (32, 2)):  # golden rows hold the logits of forwards 0, 31 and 63
        if advance:
            small_f32.step(advance)
        logits = small_f32.last_logits()[0].cpu().numpy()
        ref = g["logits_p12"][row]
        worst = max(worst, float(np.abs(logits - ref).max() / np.abs(ref).max()))
    ids = small_f32.read_ids()
    assert np.array_equal(ids.numpy(), g["ids_p12"][:, :ids.shape[1]])
    assert worst < 1e-3, worst
    print(f"cfg0 max relative logit error {worst:.3e}")


def test_reference_call_as_written(small_f32, golden):
    """generate_llm_response exactly as REF/tutoring_server.py:15-31 with the engine injected as `model`."""
    g = golden("gpt2_as_written")
    tok = rp.FakeGPT2Tokenizer(prompt_len=32, seed=1)
    for i in range(3):
        text = rp.generate_llm_response(str(g[f"query_{i}"]), small_f32, tok)
        assert text == str(g[f"text_{i}"])


def test_tiny_batch_eos_padding_and_max_length(pkg, tiny_hf, golden):
    g = golden("gpt2_tiny")
    eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="fp32")
    prompt = torch.from_numpy(g["prompt"])
    eos = int(g["eos"])
    for pen, key in ((1.0, "p10"), (1.2, "p12")):
        out = eng.generate(prompt, max_new_tokens=24, repetition_penalty=pen, eos_token_id=eos, pad_token_id=eos)
        assert np.array_equal(out.numpy(), g[f"ids_{key}"]), key
    out = eng.generate(prompt, max_length=20, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    assert np.array_equal(out.numpy(), g["ids_maxlen20"])
    one = eng.generate(prompt[:1], max_new_tokens=24, eos_token_id=int(g["eos_row0"]), pad_token_id=0)
    assert np.array_equal(one.numpy(), g["ids_row0_eos_early"])  # stops early: every row finished
    # device-resident ids in, device ids out
    dev = eng.generate(prompt.cuda(), max_new_tokens=24, repetition_penalty=1.2, eos_token_id=eos, pad_token_id=eos)
    assert dev.is_cuda and np.array_equal(dev.cpu().numpy(), g["ids_p12"])


def test_tiny_live_oracle_many_shapes(pkg, tiny_hf):
    """ragged shapes against the oracle run live (HF on the host cores): B in {1,3,9}, prompt lengths 1..70."""
    eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="fp32")
    for B, T, new in ((1, 1, 5), (3, 7, 20), (9, 70, 33), (2, 64, 64)):
        ids = rp.gpt2_prompt_ids(B, T, seed=B * 100 + T, vocab=1000)
        ref = rp.generate(tiny_hf, ids, max_new_tokens=new, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
        out = eng.generate(ids, max_new_tokens=new, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
        assert torch.equal(out, ref), (B, T, new)


def test_errors_match_reference(small_f32):
    with pytest.raises(ValueError, match="max_length"):
        small_f32.generate(torch.zeros((1, 150), dtype=torch.int64), max_length=150)
    with pytest.raises(NotImplementedError):
        small_f32.generate(torch.zeros((1, 4), dtype=torch.int64), max_length=8, do_sample=True)


def test_concurrent_callers(small_f32, golden):
    """The reference serves 10 gRPC handler threads on one model (REF/tutoring_server.py:40)."""
    from concurrent.futures import ThreadPoolExecutor
    g = golden("gpt2_cfg0")
    prompt = torch.from_numpy(g["prompt"])
    with ThreadPoolExecutor(4) as ex:
        outs = list(ex.map(lambda _: small_f32.generate(prompt, max_new_tokens=16, repetition_penalty=1.2), range(8)))
    for o in outs:
        assert np.array_equal(o.numpy(), g["ids_p12"][:, :48])


def test_bf16_mode_agreement_cfg0(small_bf16, golden):
    """bf16 throughput mode vs the fp32 oracle: agreement is reported, and the prefix before the first
    divergence must be long (stated tolerance: >= 8 identical leading tokens on configs[0])."""
    g = golden("gpt2_cfg0")
    out = small_bf16.generate(torch.from_numpy(g["prompt"]), max_new_tokens=64, repetition_penalty=1.0)
    same = out.numpy()[0, 32:] == g["ids_p10"][0, 32:]
    first_div = int(np.argmin(same)) if not same.all() else 64
    print(f"bf16 vs fp32 oracle on cfg0: {same.mean():.3f} token agreement, first divergence at step {first_div}")
    assert first_div >= 8
    small_bf16.prefill(torch.from_numpy(g["prompt"]), max_new_tokens=64)
    logits = small_bf16.last_logits()[0].float().cpu().numpy()
    err = float(np.abs(logits - g["logits_p10"][0]).max() / np.abs(g["logits_p10"][0]).max())
    print(f"bf16 step-0 relative logit error {err:.3e}")
    assert err < 3e-2


def test_cfg2_full_size(small_bf16, golden):
    """configs[2] at full size: B=64, 512-token prompts, +256 tokens, bf16.  Checked against the fp32 oracle's
    ids (committed golden): first generated token agreement and overall agreement rate; plus size-independent
    properties (prompt preserved, ids in range, determinism across two runs)."""
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)
    out = small_bf16.generate(ids, max_new_tokens=256, repetition_penalty=1.0)
    assert tuple(out.shape) == (64, 768)
    assert torch.equal(out[:, :512], ids)
    assert int(out.min()) >= 0 and int(out.max()) < 50257
    again = small_bf16.generate(ids, max_new_tokens=256, repetition_penalty=1.0)
    assert torch.equal(out, again), "bf16 generation must be run-to-run deterministic"
    g = golden("gpt2_cfg2")
    same = out[:, 512:].numpy() == g["ids_p10"]
    first = same[:, 0].mean()
    prefix = np.where(same.all(axis=1), 256, np.argmin(same, axis=1))
    print(f"cfg2 bf16 vs fp32 oracle: first-token agreement {first:.3f}, overall {same.mean():.3f}, "
          f"median identical prefix {np.median(prefix):.0f} tokens")
    # stated bf16 tolerance for configs[2]: a greedy sequence stays on the oracle's path until the
    # first near-tie flips, after which the two contexts differ for good - hence a prefix length, not a per-token rate
    assert first >= 0.9
    assert same.mean() >= CFG2_MIN_AGREEMENT and np.median(prefix) >= CFG2_MIN_MEDIAN_PREFIX
    # ... and next to transformers' OWN bf16 run of the same model (model.bfloat16() on the host, rows 0..15, 128 tokens;
    # SURVEY section 8d): how far a bf16 run strays from the fp32 oracle is a property of bf16, not of this engine
    hb = golden("gpt2_cfg2_hf_bf16")["ids_p10"]
    R, S = hb.shape
    ours, hf_same = same[:R, :S], hb == g["ids_p10"][:R, :S]
    pre = lambda m: np.where(m.all(axis=1), m.shape[1], np.argmin(m, axis=1))
    print(f"rows 0..{R - 1}, {S} tokens: engine bf16 vs fp32 oracle {ours.mean():.3f} (median prefix {np.median(pre(ours)):.0f}), "
          f"transformers bf16 vs fp32 oracle {hf_same.mean():.3f} (median prefix {np.median(pre(hf_same)):.0f}), "
          f"engine bf16 vs transformers bf16 {(out[:R, 512:512 + S].numpy() == hb).mean():.3f}")
    assert ours.mean() >= hf_same.mean() - 0.15 and np.median(pre(ours)) >= 0.5 * np.median(pre(hf_same))


CFG2_MIN_AGREEMENT, CFG2_MIN_MEDIAN_PREFIX = 0.40, 16  # measured on B200: 0.601 overall, median prefix 46
CFG2_DECODE_LOGIT_TOL = 2e-2                            # max |logit - oracle| / max |oracle| per row; measured 6.4e-3 .. 7.3e-3


def test_cfg2_bf16_decode_logits_teacher_forced(small_bf16, golden):
    """The benchmarked path itself - the bf16 decode forward at B=64 over a 512..766-row cache - against the fp32 oracle:
    the engine prefills the ORACLE's ids up to decode step s, runs ONE decode forward, and its logits are compared with
    the oracle's teacher-forced logits for that position (tests/golden/gpt2_cfg2_logits.npz, made by
    oracle/make_golden.py from the unmodified transformers path).  Stated tolerance: 2e-2 of the row's largest logit, and >= 0.85 of the
    rows must pick the oracle's next token."""
    g, gl = golden("gpt2_cfg2"), golden("gpt2_cfg2_logits")
    seq = torch.cat([rp.gpt2_prompt_ids(64, 512, seed=1), torch.from_numpy(g["ids_p10"].astype(np.int64))], dim=1)
    checked = 0
    for i, (b, s) in enumerate(zip(gl["rows"], gl["steps"])):
        b, s = int(b), int(s)
        small_bf16.prefill(seq[:, :512 + s - 1], max_new_tokens=2)  # selects token 512+s-1 (from prefill logits)
        fed = small_bf16.read_ids()[:, -1]
        small_bf16.step(1)                                          # the decode forward under test: cache length 512+s-1
        logits = small_bf16.last_logits().float().cpu().numpy()
        ok_rows = (fed == seq[:, 512 + s - 1]).numpy()              # rows whose decode input is the oracle's token
        nxt = (logits.argmax(axis=1) == seq[:, 512 + s].numpy())[ok_rows].mean()
        print(f"decode step {s}: {ok_rows.mean():.3f} of the rows were fed the oracle's token; teacher-forced next-token "
              f"agreement of those rows {nxt:.3f}")
        assert ok_rows.mean() >= 0.9 and nxt >= 0.85
        if ok_rows[b]:
            ref = gl[f"logits_{i}"]
            err = float(np.abs(logits[b] - ref).max() / np.abs(ref).max())
            print(f"  row {b}, cache length {512 + s - 1}: relative logit error {err:.3e}, "
                  f"arg-max {'same' if logits[b].argmax() == ref.argmax() else 'differs'}")
            assert err < CFG2_DECODE_LOGIT_TOL
            checked += 1
    assert checked >= 3


def test_cfg2_fp32_parity_subset(small_f32, golden):
    """fp32 mode on the first 8 rows of configs[2]: ids identical to the oracle's wherever the oracle's own
    top-1/top-2 gap exceeds fp32 reordering noise (1e-4)."""
    g = golden("gpt2_cfg2")
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)[:8]
    out = small_f32.generate(ids, max_new_tokens=64, repetition_penalty=1.2)
    ref, gap = g["ids_p12"][:8, :64], g["gap_p12"][:64, :8].T
    mine = out[:, 512:].numpy()
    excused = 0
    for b in range(8):
        mism = np.nonzero(mine[b] != ref[b])[0]
        if mism.size:
            assert gap[b, mism[0]] < 1e-4, (b, mism[0], gap[b, mism[0]])
            excused += 1
    print(f"cfg2 fp32 subset: {8 - excused} of 8 rows identical over 64 tokens, {excused} diverge at an oracle near-tie (< 1e-4)")
    assert excused <= 2


def test_fused_decode_path_matches_unfused(pkg, tiny_hf, monkeypatch):
    """bf16 decode runs split-K GEMMs with fused reductions; TUTOR_NO_FUSED=1 selects the plain kernel chain.
    Same weights, same prompt: logits must agree to bf16 noise and the greedy ids must be the same."""
    ids = rp.gpt2_prompt_ids(5, 9, seed=11, vocab=1000)
    fused = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
    monkeypatch.setenv("TUTOR_NO_FUSED", "1")
    plain = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
    monkeypatch.delenv("TUTOR_NO_FUSED")
    monkeypatch.setenv("TUTOR_NO_CLUSTER", "1")  # the 8-launch-per-layer chain with separate reduce kernels
    splitk = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
    monkeypatch.delenv("TUTOR_NO_CLUSTER")
    splitk.prefill(ids, max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    splitk.step(12)
    ref_logits = splitk.last_logits().float().cpu()
    logits = []
    for eng in (fused, plain):
        eng.prefill(ids, max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
        eng.step(12)
        logits.append(eng.last_logits().float().cpu())
    err = ((logits[0] - logits[1]).abs().max() / logits[1].abs().max()).item()
    print(f"fused vs unfused bf16 decode: relative logit difference {err:.3e}")
    assert err < 2e-2
    err_c = ((logits[0] - ref_logits).abs().max() / ref_logits.abs().max()).item()
    print(f"cluster (LN-operand / DSMEM reduce) vs split-K + reduce kernels: relative logit difference {err_c:.3e}")
    assert err_c < 2e-2
    a = fused.generate(ids, max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    b = plain.generate(ids, max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    assert (a == b).float().mean().item() > 0.9
    # key-range splits of the decode attention (last-arriving CTA merges): same numbers as one CTA per (b,h)
    long_ids = rp.gpt2_prompt_ids(3, 150, seed=12, vocab=1000)
    outs = []
    for splits in ("1", "4"):
        monkeypatch.setenv("TUTOR_ATTN_SPLITS", splits)
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
        eng.prefill(long_ids, max_new_tokens=40, eos_token_id=999, pad_token_id=999)
        eng.step(20)
        outs.append((eng.last_logits().float().cpu(), eng.read_ids()))
    monkeypatch.delenv("TUTOR_ATTN_SPLITS")
    err = ((outs[0][0] - outs[1][0]).abs().max() / outs[0][0].abs().max()).item()
    print(f"attention splits 1 vs 4: relative logit difference {err:.3e}")
    assert err < 5e-3 and (outs[0][1] == outs[1][1]).float().mean().item() > 0.95


def test_cfg3_gpt2_large_prefill_heavy(pkg, golden):
    """configs[3]: GPT-2 large (774M), B=16, 896-token prompts + 128 new tokens = 1024 positions (the most GPT-2 has),
    bf16 tcgen05 prefill.  Against the fp32 oracle's ids (committed golden): step-0 logits of row 0 within the stated
    bf16 tolerance (5e-2 relative), first generated token agreement >= 0.8; prompt preserved; deterministic."""
    g = golden("gpt2_cfg3")
    hf = rp.build_gpt2("large", seed=0)
    eng = pkg.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
    del hf
    ids = rp.gpt2_prompt_ids(16, 896, seed=1)
    out = eng.generate(ids, max_new_tokens=128, repetition_penalty=1.0)
    assert tuple(out.shape) == (16, 1024) and torch.equal(out[:, :896], ids)
    same = out[:, 896:].numpy() == g["ids_p10"]
    print(f"cfg3 bf16 vs fp32 oracle: first-token agreement {same[:, 0].mean():.3f}, overall {same.mean():.3f}")
    assert same[:, 0].mean() >= 0.8 and same.mean() >= 0.5  # measured 0.938 / 0.794
    assert torch.equal(out, eng.generate(ids, max_new_tokens=128, repetition_penalty=1.0))
    eng.prefill(ids, max_new_tokens=128)
    logits = eng.last_logits()[0].float().cpu().numpy()
    err = float(np.abs(logits - g["logits_step0_row0"]).max() / np.abs(g["logits_step0_row0"]).max())
    print(f"cfg3 bf16 step-0 relative logit error {err:.3e}")
    assert err < 5e-2
    # the decode path, teacher-forced: prefill the oracle's ids up to decode step s, one decode forward, and the arg-max
    # of every row that was fed the oracle's token must be the oracle's next token (stated: >= 0.8 of the rows)
    seq = torch.cat([ids, torch.from_numpy(g["ids_p10"].astype(np.int64))], dim=1)
    for s in (1, 64, 127):
        eng.prefill(seq[:, :896 + s - 1], max_new_tokens=2)
        fed = eng.read_ids()[:, -1]
        eng.step(1)
        nxt = eng.read_ids()[:, -1]
        ok_rows = fed == seq[:, 896 + s - 1]
        agree = (nxt == seq[:, 896 + s])[ok_rows].float().mean().item()
        print(f"cfg3 decode step {s}: {ok_rows.float().mean():.3f} rows fed the oracle's token, next-token agreement {agree:.3f}")
        assert ok_rows.float().mean() >= 0.8 and agree >= 0.8
    with pytest.raises(Exception, match="n_positions"):
        eng.generate(rp.gpt2_prompt_ids(1, 1000, seed=2), max_new_tokens=128)  # would index past wpe


def test_left_padded_batch_matches_hf(pkg, tiny_hf, golden):
    """Prompts of different lengths in one batch (attention_mask = zeros then ones): ids identical to transformers'
    generate with the same mask, and each row equal to its own unpadded batch-1 generation."""
    g = golden("gpt2_padded")
    ids, mask = torch.from_numpy(g["ids"]), torch.from_numpy(g["mask"])
    for dtype in ("fp32",):
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype=dtype)
        for pen, key in ((1.0, "p10"), (1.2, "p12")):
            out = eng.generate(ids, attention_mask=mask, max_new_tokens=20, repetition_penalty=pen, eos_token_id=999,
                               pad_token_id=999)
            assert np.array_equal(out.numpy(), g[f"batch_{key}"]), key
            for b, p in enumerate(g["npad"].tolist()):
                one = eng.generate(ids[b:b + 1, p:], max_new_tokens=20, repetition_penalty=pen, eos_token_id=999,
                                   pad_token_id=999)
                assert np.array_equal(one.numpy(), g[f"row{b}_{key}"])
    b16 = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")  # fused decode path with padding
    out = b16.generate(ids, attention_mask=mask, max_new_tokens=20, repetition_penalty=1.0, eos_token_id=999, pad_token_id=999)
    agree = (out.numpy() == g["batch_p10"]).mean()
    print(f"left-padded bf16 vs fp32 oracle token agreement {agree:.3f}")
    assert agree > 0.8
    with pytest.raises(NotImplementedError):
        eng.generate(ids, attention_mask=torch.flip(mask, dims=[1]), max_new_tokens=4)  # right padding


def test_launch_modes_are_bit_identical(pkg, small_hf, monkeypatch):
    """CUDA graph replay, eager launches, programmatic dependent launch on/off: the same kernels in the same order,
    so the ids must be identical - a guard against ordering hazards of the PDL chain (the length counter, the
    split-K partial buffers and the KV rows are all produced and consumed by neighbouring kernels)."""
    ids = rp.gpt2_prompt_ids(16, 200, seed=31)
    outs = {}
    for name, env in (("default", {}), ("no_pdl", {"TUTOR_NO_PDL": "1"}), ("no_graph", {"TUTOR_NO_GRAPH": "1"}),
                      ("no_pdl_no_graph", {"TUTOR_NO_PDL": "1", "TUTOR_NO_GRAPH": "1"})):
        for k in ("TUTOR_NO_PDL", "TUTOR_NO_GRAPH"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        eng = pkg.B200GPT2LMHeadModel.from_hf(small_hf, dtype="bf16")
        runs = [eng.generate(ids, max_new_tokens=48, repetition_penalty=1.2) for _ in range(3)]
        assert all(torch.equal(r, runs[0]) for r in runs), name
        outs[name] = runs[0]
        del eng
    for name, o in outs.items():
        assert torch.equal(o, outs["no_pdl_no_graph"]), name


@pytest.mark.parametrize("B,T", [(1, 1), (1, 37), (3, 5), (17, 65), (33, 9), (100, 20), (128, 3)])
def test_bf16_ragged_shapes_against_oracle_logits(pkg, tiny_hf, B, T):
    """bf16 throughput mode over batch sizes that select every tile width of the swap-AB GEMM (16/32/64/128) and
    prompt lengths that leave ragged attention / GEMM tiles: prefill logits within the stated bf16 tolerance
    (3e-2 relative) of the fp32 oracle, decode runs, output deterministic."""
    eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
    ids = rp.gpt2_prompt_ids(B, T, seed=B * 1000 + T, vocab=1000)
    with torch.no_grad():
        ref = tiny_hf(ids).logits[:, -1].float()
    eng.prefill(ids, max_new_tokens=12, eos_token_id=999, pad_token_id=999)
    got = eng.last_logits().float().cpu()
    err = ((got - ref).abs().max() / ref.abs().max()).item()
    assert err < 3e-2, err
    a = eng.generate(ids, max_new_tokens=12, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    b = eng.generate(ids, max_new_tokens=12, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    assert tuple(a.shape) == (B, T + 12) and torch.equal(a, b) and torch.equal(a[:, :T], ids)
    assert int(a.min()) >= 0 and int(a.max()) < 1000


@pytest.mark.parametrize("B,T,new", [(3, 65, 20), (16, 70, 30), (4, 199, 40), (2, 225, 30)])
def test_prefill_rows_rounded_up_to_32_change_nothing(pkg, tiny_hf, monkeypatch, B, T, new):
    """bf16 prompts whose length is not a multiple of 32 are prefilled as if right-padded to the next one (the c_attn
    GEMM keeps its TMA cache scatter): prefill logits and every generated id equal those of the unpadded prefill, also
    when the cache rows the pad positions wrote are reached and overwritten by the decode steps."""
    ids = rp.gpt2_prompt_ids(B, T, seed=B + T, vocab=1000)
    res = {}
    for name, env in (("padded", None), ("plain", "1")):
        monkeypatch.delenv("TUTOR_NO_PREFILL_PAD", raising=False)
        if env:
            monkeypatch.setenv("TUTOR_NO_PREFILL_PAD", env)
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
        eng.prefill(ids, max_new_tokens=new, eos_token_id=999, pad_token_id=999)
        logits = eng.last_logits().float().cpu()
        out = eng.generate(ids, max_new_tokens=new, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
        res[name] = (logits, out)
        del eng
    # same arithmetic per row in both runs except the c_attn epilogue flavour (same accumulation order): tight tolerance
    assert (res["padded"][0] - res["plain"][0]).abs().max().item() <= 1e-3 * res["plain"][0].abs().max().item()
    assert torch.equal(res["padded"][1], res["plain"][1])


@pytest.mark.parametrize("B,T", [(1, 9), (5, 40), (16, 70), (33, 20), (64, 130), (100, 20)])
def test_persistent_decode_kernel_matches_launch_chain(pkg, tiny_hf, monkeypatch, B, T):
    """The opt-in persistent decode kernel (TUTOR_MEGA=1: LayerNorm folded into the weights, TMA operands, split-K
    closed through L2 counters, queue-fed attention streams) against the default launch chain: same prompt, one decode
    step (so both see the same context), logits within the bf16 tolerance; generation is run-to-run deterministic."""
    ids = rp.gpt2_prompt_ids(B, T, seed=21 + B, vocab=1000)
    logits, outs = [], []
    for mode in ("0", "1"):
        monkeypatch.setenv("TUTOR_MEGA", mode)
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
        eng.prefill(ids, max_new_tokens=12, eos_token_id=999, pad_token_id=999)
        eng.step(1)
        logits.append(eng.last_logits().float().cpu())
        outs.append([eng.generate(ids, max_new_tokens=12, eos_token_id=999, pad_token_id=999) for _ in range(2)])
    monkeypatch.delenv("TUTOR_MEGA")
    err = ((logits[0] - logits[1]).abs().max() / logits[0].abs().max()).item()
    print(f"persistent kernel vs launch chain, B={B}: relative logit difference {err:.3e}")
    assert err < 2e-2
    with torch.no_grad():
        ref = tiny_hf(outs[1][0][:, :T + 1]).logits[:, -1].float()  # fp32 oracle at the context of that step
    assert ((logits[1] - ref).abs().max() / ref.abs().max()).item() < 5e-2 or not torch.equal(outs[0][0][:, :T + 1], outs[1][0][:, :T + 1])
    assert torch.equal(outs[1][0], outs[1][1]), "the persistent kernel must be run-to-run deterministic"


@pytest.mark.parametrize("B,T,new,pad", [(1, 9, 24, 0), (2, 33, 24, 0), (1, 200, 40, 0), (2, 200, 30, 57), (2, 130, 20, 129)])
def test_cooperative_fp32_decode_matches_launch_chain_and_oracle(pkg, tiny_hf, monkeypatch, B, T, new, pad):
    """The cooperative fp32 decode kernel (one or two sequences, csrc/decode_f32_persistent.cu) against the launch chain
    (TUTOR_F32_PERSISTENT=0) and the oracle: identical ids, including contexts long enough for several key ranges per
    head (partial softmax states merged by the last CTA to arrive) and a left-padded second row."""
    ids = rp.gpt2_prompt_ids(B, T, seed=41 + T, vocab=1000)
    mask = None
    if pad:
        mask = torch.ones_like(ids)
        mask[B - 1, :pad] = 0
        ids[B - 1, :pad] = 999
    kw = dict(max_new_tokens=new, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    outs = []
    for flag in ("0", "1"):
        monkeypatch.setenv("TUTOR_F32_PERSISTENT", flag)
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="fp32")
        outs.append([eng.generate(ids, attention_mask=mask, **kw) for _ in range(2)])
    monkeypatch.delenv("TUTOR_F32_PERSISTENT")
    assert torch.equal(outs[1][0], outs[1][1]), "run-to-run deterministic"
    assert torch.equal(outs[0][0], outs[1][0]), "cooperative kernel and launch chain disagree"
    ref = rp.generate(tiny_hf, ids, attention_mask=mask, **kw) if pad else rp.generate(tiny_hf, ids, **kw)
    assert torch.equal(outs[1][0], ref)


def test_cooperative_fp32_decode_on_gpt2_small_two_sequences(pkg, small_hf, monkeypatch):
    """d = 768, two sequences, a context that needs several key ranges per head: ids identical to the launch chain."""
    ids = rp.gpt2_prompt_ids(2, 260, seed=5)
    outs = []
    for flag in ("0", "1"):
        monkeypatch.setenv("TUTOR_F32_PERSISTENT", flag)
        eng = pkg.B200GPT2LMHeadModel.from_hf(small_hf, dtype="fp32")
        outs.append(eng.generate(ids, max_new_tokens=24, repetition_penalty=1.2))
    monkeypatch.delenv("TUTOR_F32_PERSISTENT")
    assert torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("B,T", [(3, 17), (40, 60), (100, 20)])
def test_folded_layernorm_decode_chain_matches_default(pkg, tiny_hf, monkeypatch, B, T):
    """TUTOR_FOLD=1 (LayerNorm folded into the c_attn / c_fc weights, bf16 copy of x as the TMA operand, rstd and mean
    applied where the split-K sum is closed) against the default chain: one decode step from the same prompt, logits
    within the bf16 tolerance; deterministic."""
    ids = rp.gpt2_prompt_ids(B, T, seed=61 + B, vocab=1000)
    logits, outs = [], []
    for mode in ("0", "1"):
        monkeypatch.setenv("TUTOR_FOLD", mode)
        eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
        eng.prefill(ids, max_new_tokens=12, eos_token_id=999, pad_token_id=999)
        eng.step(1)
        logits.append(eng.last_logits().float().cpu())
        outs.append([eng.generate(ids, max_new_tokens=12, eos_token_id=999, pad_token_id=999) for _ in range(2)])
    monkeypatch.delenv("TUTOR_FOLD")
    err = ((logits[0] - logits[1]).abs().max() / logits[0].abs().max()).item()
    print(f"folded LayerNorm vs operand-load LayerNorm, B={B}: relative logit difference {err:.3e}")
    assert err < 2e-2
    assert torch.equal(outs[1][0], outs[1][1])


def test_long_prompt_then_short_prompt_on_one_bf16_engine(pkg, tiny_hf):
    """The captured decode graph bakes in how many prompt rows the attention may stream ahead of its wait.  A long prompt
    followed by a short one on the SAME engine must not replay the long prompt's bound (advisor finding, round 1): the
    short generation equals a fresh engine's."""
    long_ids = rp.gpt2_prompt_ids(3, 140, seed=31, vocab=1000)
    short_ids = rp.gpt2_prompt_ids(3, 20, seed=32, vocab=1000)
    kw = dict(max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16")
    eng.generate(long_ids, **kw)
    again = eng.generate(short_ids, **kw)
    fresh = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16").generate(short_ids, **kw)
    assert torch.equal(again, fresh)
    assert torch.equal(eng.generate(long_ids, **kw), pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="bf16").generate(long_ids, **kw))


def test_allocation_failure_leaves_the_engine_usable(pkg, tiny_hf):
    """A request whose KV cache cannot be allocated fails with an error, and the next (small) request on the same engine
    still runs correctly - capacity bookkeeping is reset before anything is freed (engine_gpt2.cu ensure_capacity)."""
    eng = pkg.B200GPT2LMHeadModel.from_hf(tiny_hf, dtype="fp32")
    small = rp.gpt2_prompt_ids(2, 9, seed=33, vocab=1000)
    kw = dict(max_new_tokens=10, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    before = eng.generate(small, **kw)
    huge = torch.zeros((1_500_000, 4), dtype=torch.int64)  # 2 layers x 2 x 1.5 M rows x 2 heads x 256 x 64 fp32 = 393 GB of cache
    with pytest.raises(pkg._lib.TutorError):
        eng.generate(huge, max_new_tokens=250, eos_token_id=999, pad_token_id=999)
    after = eng.generate(small, **kw)
    assert torch.equal(before, after)
    assert torch.equal(after, rp.generate(tiny_hf, small, **kw))
