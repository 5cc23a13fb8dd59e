"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Generates tests/golden/*.npz by running the UNMODIFIED reference modules
(/root/reference/GUI_RAFT_LLM_SourceCode/{tutoring_server,lms_server}.py), imported
under the three shims SURVEY.md §8c describes, and checks that oracle/reference_path.py
returns the same values.  Runs only in the build container (the reference is not on
the GPU box); the fixtures it writes are committed.

    python -m oracle.make_golden [--only NAME ...]
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile
import time
import types

import numpy as np
import torch

from oracle import reference_path as rp

REF_DIR = "/root/reference/GUI_RAFT_LLM_SourceCode"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class _PresetBertTokenizer:
    """Tokenizer shim: text is a key into preset id rows (truncation honoured)."""

    def __init__(self):
        self.table = {}

    def __call__(self, text, return_tensors="pt", padding=True, truncation=True, max_length=512):
        ids = self.table[text][:, :max_length]
        return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids)}


def import_reference():
    """Import the reference's two hot-path modules without network, PyPDF2 or vocab files."""
    import transformers

    sys.modules.setdefault("PyPDF2", types.ModuleType("PyPDF2"))  # REF/lms_server.py:9
    gpt2_model = rp.build_gpt2("small", seed=0)
    tok = rp.FakeGPT2Tokenizer(prompt_len=32, seed=1)
    orig = (transformers.GPT2LMHeadModel.from_pretrained, transformers.GPT2Tokenizer.from_pretrained)
    transformers.GPT2LMHeadModel.from_pretrained = classmethod(lambda cls, name, *a, **k: gpt2_model)
    transformers.GPT2Tokenizer.from_pretrained = classmethod(lambda cls, name, *a, **k: tok)
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp()
    os.chdir(tmp)  # lms_server creates ./uploads on import (REF/lms_server.py:30-32)
    sys.path.insert(0, REF_DIR)
    try:
        import lms_server  # noqa
        import tutoring_server  # noqa
    finally:
        os.chdir(cwd)
        sys.path.remove(REF_DIR)
        transformers.GPT2LMHeadModel.from_pretrained, transformers.GPT2Tokenizer.from_pretrained = orig
    return tutoring_server, lms_server


def gaps(scores):
    """top-1/top-2 gap per step and row for a [steps,B,V] score tensor."""
    t2 = torch.topk(scores, 2, dim=-1).values
    return (t2[..., 0] - t2[..., 1]).numpy()


def make_cfg0(tut, lms):
    """configs[0]: GPT-2 small fp32, B=1, 32-token prompt -> 64 new tokens."""
    model = tut.model
    ids = rp.gpt2_prompt_ids(1, 32, seed=1)
    out = {"prompt": ids.numpy()}
    for name, pen in (("p10", 1.0), ("p12", 1.2)):
        with torch.no_grad():
            ref = model.generate(ids, max_new_tokens=64, num_return_sequences=1, temperature=0.7, top_k=50,
                                 top_p=0.9, repetition_penalty=pen)
        seq, logits, scores = rp.generate_with_logits(model, ids, max_new_tokens=64, repetition_penalty=pen)
        assert torch.equal(ref, seq), "oracle.generate != reference-style model.generate"
        out[f"ids_{name}"] = seq.numpy()
        out[f"gap_{name}"] = gaps(scores)
        out[f"logits_{name}"] = logits[[0, 31, 63], 0].numpy()  # raw fp32 logits at three steps
        out[f"scores_{name}"] = scores[[0, 31, 63], 0].numpy()
    known = [23612, 26724, 7723, 7723, 28384, 32473, 12040, 35616, 32638, 45260, 309, 38118, 43456, 191, 34581, 7226]
    assert out["ids_p12"][0, 32:48].tolist() == known, "SURVEY §8c known answer not reproduced"
    np.savez_compressed(os.path.join(OUT, "gpt2_cfg0.npz"), **out)
    print("cfg0 min gaps", out["gap_p10"].min(), out["gap_p12"].min())


def make_as_written(tut, lms):
    """The reference call exactly as written: generate_llm_response(query) -> str (max_length=150)."""
    queries = ["What is Raft consensus?", "Explain leader election", ""]
    rows = {}
    for i, q in enumerate(queries):
        text = tut.generate_llm_response(q)
        mine = rp.generate_llm_response(q, tut.model, tut.tokenizer)
        assert text == mine, "oracle.generate_llm_response != reference"
        prompt = tut.tokenizer.encode(rp.PROMPT_TEMPLATE.format(query=q), return_tensors="pt")
        rows[f"prompt_{i}"] = prompt.numpy()
        rows[f"text_{i}"] = np.array(text)
        rows[f"query_{i}"] = np.array(q)
    np.savez_compressed(os.path.join(OUT, "gpt2_as_written.npz"), **rows)


def make_tiny(tut, lms):
    """Scaled-down GPT-2 (2 layers, d=128): batch>1, EOS/padding of finished rows, max_length semantics."""
    model = rp.build_gpt2("tiny", seed=3)
    ids = rp.gpt2_prompt_ids(4, 8, seed=4, vocab=1000)
    out = {"prompt": ids.numpy()}
    free = rp.generate(model, ids, max_new_tokens=24, repetition_penalty=1.2, eos_token_id=999, pad_token_id=999)
    # choose as EOS a token that row 1 emits at step 5 so that rows finish at different times
    eos = int(free[1, 8 + 5])
    out["eos"] = np.array(eos)
    for name, pen in (("p10", 1.0), ("p12", 1.2)):
        out[f"ids_{name}"] = rp.generate(model, ids, max_new_tokens=24, repetition_penalty=pen,
                                         eos_token_id=eos, pad_token_id=eos).numpy()
    out["ids_maxlen20"] = rp.generate(model, ids, max_length=20, repetition_penalty=1.2,
                                      eos_token_id=999, pad_token_id=999).numpy()
    free10 = rp.generate(model, ids[:1], max_new_tokens=24, repetition_penalty=1.0, eos_token_id=999, pad_token_id=999)
    eos0 = int(free10[0, 8 + 3])
    one = rp.generate(model, ids[:1], max_new_tokens=24, repetition_penalty=1.0, eos_token_id=eos0, pad_token_id=0)
    assert one.shape[1] < 32, "early stop not exercised"
    out["ids_row0_eos_early"] = one.numpy()  # all rows finished -> output shorter than max
    out["eos_row0"] = np.array(eos0)
    # numpy restatement cross-check
    from oracle import restate_numpy as rn
    net = rn.GPT2Numpy(rn.sd_numpy(model), n_head=2)
    mine = rn.greedy_generate(net, ids.numpy(), max_new_tokens=24, repetition_penalty_=1.2, eos_token_id=eos)
    assert np.array_equal(mine, out["ids_p12"]), "numpy restatement != HF on tiny config"
    np.savez_compressed(os.path.join(OUT, "gpt2_tiny.npz"), **out)
    print("tiny eos", eos, "shapes", {k: v.shape for k, v in out.items()})


def make_bert(tut, lms):
    """configs[1]: 32 pairs, 128-token query vs 512-token assignment, through the reference's own functions."""
    model = rp.build_bert(seed=0)
    tok = _PresetBertTokenizer()
    q_ids, a_ids = rp.bert_ids(32, 128, seed=2), rp.bert_ids(32, 520, seed=5)  # 520 > 512: truncation exercised
    a_ids[:, 511] = 102  # what a real tokenizer leaves after truncation: [SEP] at the cut
    qe, ae, cos = [], [], []
    t0 = time.time()
    for i in range(32):
        tok.table[f"q{i}"], tok.table[f"a{i}"] = q_ids[i:i + 1], a_ids[i:i + 1]
        e1 = lms.compute_embeddings(f"q{i}", model, tok)
        e2 = lms.compute_embeddings(f"a{i}", model, tok)
        c = lms.cosine_similarity(e1, e2)
        assert torch.equal(e1, rp.compute_embeddings(f"q{i}", model, tok))
        assert c == rp.cosine_similarity(e1, e2)
        qe.append(e1[0].numpy()); ae.append(e2[0].numpy()); cos.append(c)
    print("bert bs-1 loop", time.time() - t0, "s; cos range", min(cos), max(cos))
    batched = rp.embeddings_from_ids(model, q_ids)
    print("batched vs bs1 max diff", (batched - torch.tensor(np.stack(qe))).abs().max().item())
    np.savez_compressed(os.path.join(OUT, "bert_cfg1.npz"), q_ids=q_ids.numpy().astype(np.int32),
                        a_ids=a_ids[:, :512].numpy().astype(np.int32), q_emb=np.stack(qe), a_emb=np.stack(ae),
                        cos=np.array(cos, dtype=np.float64))


def make_bert_tiny(tut, lms):
    """Small BERT (2 layers, d=128, 4 heads) for fast unit tests; hidden states kept."""
    cfgkw = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                 vocab_size=2000, max_position_embeddings=128)
    model = rp.build_bert(seed=6, **cfgkw)
    ids = rp.bert_ids(3, 40, seed=7, vocab=2000)
    with torch.no_grad():
        h = model(input_ids=ids).last_hidden_state
    from oracle import restate_numpy as rn
    mine = rn.bert_forward(rn.sd_numpy(model), ids.numpy(), n_head=2)
    err = np.abs(mine - h.numpy()).max()
    assert err < 2e-5, err
    np.savez_compressed(os.path.join(OUT, "bert_tiny.npz"), ids=ids.numpy().astype(np.int32), hidden=h.numpy(),
                        emb=h.mean(dim=1).numpy())
    print("bert tiny numpy-vs-HF max err", err)


def make_cfg2(tut, lms):
    """configs[2] parity target: fp32 oracle ids for B=64, 512-token prompts, +256 tokens."""
    model = tut.model
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)
    out = {}
    for name, pen in (("p10", 1.0), ("p12", 1.2)):
        t0 = time.time()
        seq, logits, scores = rp.generate_with_logits(model, ids, max_new_tokens=256, repetition_penalty=pen)
        out[f"ids_{name}"] = seq[:, 512:].numpy().astype(np.int32)
        out[f"gap_{name}"] = gaps(scores).astype(np.float32)  # [256,64]
        print("cfg2", name, time.time() - t0, "s; min gap", out[f"gap_{name}"].min())
    np.savez_compressed(os.path.join(OUT, "gpt2_cfg2.npz"), **out)
    make_cfg2_logits(tut, lms)


CFG2_TF = ((0, 1), (21, 64), (42, 128), (63, 255))  # (row, decode step): cache lengths 512, 575, 639, 766


def make_cfg2_logits(tut, lms, seq=None):
    """Teacher-forced fp32 logits of configs[2] for the bf16 decode gate: for (row b, decode step s) the oracle's logits
    at the forward that consumes token 512+s-1 (cache length 512+s-1) - what one engine decode step must reproduce after
    a prefill of the oracle's own ids[:, :512+s-1]."""
    model = tut.model
    if seq is None:  # from the committed ids (the generation itself takes many CPU-minutes)
        g = np.load(os.path.join(OUT, "gpt2_cfg2.npz"))
        seq = torch.cat([rp.gpt2_prompt_ids(64, 512, seed=1), torch.from_numpy(g["ids_p10"].astype(np.int64))], dim=1)
    out = {"rows": np.array([b for b, _ in CFG2_TF]), "steps": np.array([s for _, s in CFG2_TF])}
    for i, (b, s) in enumerate(CFG2_TF):
        with torch.no_grad():
            lg = model(seq[b:b + 1, :512 + s]).logits[0, -1].float()
        assert int(lg.argmax()) == int(seq[b, 512 + s]), "teacher-forced logits do not reproduce the oracle's token"
        out[f"logits_{i}"] = lg.numpy()
    np.savez_compressed(os.path.join(OUT, "gpt2_cfg2_logits.npz"), **out)
    print("cfg2 teacher-forced logits for", CFG2_TF)


def make_cfg2_hf_bf16(tut, lms):
    """What transformers' OWN bf16 run does on configs[2]: the same model cast with .bfloat16(), same
    stock generate call on the host, rows 0..15, 128 new tokens.  The GPU test compares the engine's bf16 agreement with the
    fp32 oracle to this run's agreement with the fp32 oracle over the same rows and steps."""
    import copy
    model = copy.deepcopy(tut.model).bfloat16()
    ids = rp.gpt2_prompt_ids(64, 512, seed=1)[:16]
    t0 = time.time()
    seq = rp.generate(model, ids, max_new_tokens=128, repetition_penalty=1.0)
    print("cfg2 HF bf16 (16 rows, 128 tokens)", time.time() - t0, "s")
    np.savez_compressed(os.path.join(OUT, "gpt2_cfg2_hf_bf16.npz"), ids_p10=seq[:, 512:].numpy().astype(np.int32))


def make_cfg3(tut, lms):
    """configs[3] parity target: GPT-2 large, B=16, 896-token prompts, +128 tokens (1024 total)."""
    model = rp.build_gpt2("large", seed=0)
    ids = rp.gpt2_prompt_ids(16, 896, seed=1)
    t0 = time.time()
    seq, logits, scores = rp.generate_with_logits(model, ids, max_new_tokens=128, repetition_penalty=1.0)
    print("cfg3", time.time() - t0, "s")
    np.savez_compressed(os.path.join(OUT, "gpt2_cfg3.npz"), ids_p10=seq[:, 896:].numpy().astype(np.int32),
                        gap_p10=gaps(scores).astype(np.float32),
                        logits_step0_row0=logits[0, 0].numpy())


def make_padded(tut, lms):
    """Left-padded batch (prompts of different lengths, as the micro-batching queue builds them): HF generate with
    attention_mask, position_ids = cumsum(mask)-1; plus each row generated alone, unpadded."""
    model = rp.build_gpt2("tiny", seed=3)
    T, npad = 12, [0, 3, 7, 1]
    ids = rp.gpt2_prompt_ids(4, T, seed=21, vocab=1000)
    mask = torch.ones_like(ids)
    for b, p in enumerate(npad):
        ids[b, :p] = 999
        mask[b, :p] = 0
    out = {"ids": ids.numpy(), "mask": mask.numpy(), "npad": np.array(npad)}
    for name, pen in (("p10", 1.0), ("p12", 1.2)):
        out[f"batch_{name}"] = rp.generate(model, ids, attention_mask=mask, max_new_tokens=20, repetition_penalty=pen,
                                           eos_token_id=999, pad_token_id=999).numpy()
        for b, p in enumerate(npad):
            out[f"row{b}_{name}"] = rp.generate(model, ids[b:b + 1, p:], max_new_tokens=20, repetition_penalty=pen,
                                                eos_token_id=999, pad_token_id=999).numpy()
    same = [np.array_equal(out["batch_p10"][b, npad[b]:], out[f"row{b}_p10"][0]) for b in range(4)]
    print("padded batch rows equal their unpadded bs-1 generation (HF itself):", same)
    np.savez_compressed(os.path.join(OUT, "gpt2_padded.npz"), **out)


MAKERS = dict(padded=make_padded, cfg0=make_cfg0, as_written=make_as_written, tiny=make_tiny, bert=make_bert, bert_tiny=make_bert_tiny,
              cfg2=make_cfg2, cfg2_logits=make_cfg2_logits, cfg2_hf_bf16=make_cfg2_hf_bf16, cfg3=make_cfg3)
# every fixture under tests/golden/, in the order they are cheapest to make (cfg2 and cfg3 take many CPU-minutes each;
# cfg2 writes gpt2_cfg2_logits.npz too, `cfg2_logits` alone rebuilds it from the committed ids)
ALL = ["cfg0", "as_written", "tiny", "padded", "bert", "bert_tiny", "cfg2", "cfg2_hf_bf16", "cfg3"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=ALL, help="fixtures to (re)generate; default: all of them")
    args = ap.parse_args()
    os.makedirs(OUT, exist_ok=True)
    torch.set_grad_enabled(False)
    tut, lms = import_reference()
    print("reference imported:", tut.__file__, lms.__file__,
          "| torch", torch.__version__, "threads", torch.get_num_threads())
    for name in args.only:
        t0 = time.time()
        MAKERS[name](tut, lms)
        print(f"[{name}] done in {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
