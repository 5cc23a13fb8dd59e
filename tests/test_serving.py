"""The Tutoring RPC boundary.  CPU part: the wire format built here is byte-identical to the
reference's generated messages (checked against /root/reference when it is mounted) and the servicer keeps the
reference's contract.  GPU part: the whole path - gRPC on localhost, 10 handler threads, engine behind it -
returns exactly the text the oracle's generate_llm_response produces."""
import os
import sys

import numpy as np
import pytest

from oracle import reference_path as rp

REF = "/root/reference/GUI_RAFT_LLM_SourceCode"


class EchoModel:
    """Stands in for the engine on CPU: appends three fixed ids, like a generate() that stops after 3 tokens."""
    def generate(self, inputs, **kw):
        import torch
        assert kw == dict(max_length=150, num_return_sequences=1, temperature=0.7, top_k=50, top_p=0.9,
                          repetition_penalty=1.2)  # the reference's kwargs, REF/tutoring_server.py:23-28
        return torch.cat([inputs, torch.tensor([[7, 8, 9]])], dim=1)


def test_wire_format_matches_reference_stubs(pkg):
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    req = serving.QueryRequest(token="tok", query="What is Raft?")
    resp = serving.QueryResponse(success=True, response="answer")
    assert serving.QueryRequest.FromString(req.SerializeToString()).query == "What is Raft?"
    if not os.path.isdir(REF):
        pytest.skip("reference not mounted: cross-check against lms_pb2 skipped")
    sys.path.insert(0, REF)
    try:
        import lms_pb2
    finally:
        sys.path.remove(REF)
    assert lms_pb2.QueryRequest(token="tok", query="What is Raft?").SerializeToString() == req.SerializeToString()
    assert lms_pb2.QueryResponse(success=True, response="answer").SerializeToString() == resp.SerializeToString()
    assert lms_pb2.QueryResponse.FromString(resp.SerializeToString()).response == "answer"


def test_service_contract_over_localhost_grpc(pkg):
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    tok = rp.FakeGPT2Tokenizer(prompt_len=32, seed=1)
    server = serving.serve(EchoModel(), tok, address="127.0.0.1:0")
    try:
        r = serving.call(f"127.0.0.1:{server.bound_port}", "t", "Explain leader election")
        assert r.success is True
        prompt = tok.encode(rp.PROMPT_TEMPLATE.format(query="Explain leader election"))
        assert r.response == tok.decode(prompt + [7, 8, 9], skip_special_tokens=True)  # prompt text is included
        if os.path.isdir(REF):  # the reference's own stub talks to this server
            import grpc
            sys.path.insert(0, REF)
            try:
                import lms_pb2, lms_pb2_grpc
            finally:
                sys.path.remove(REF)
            with grpc.insecure_channel(f"127.0.0.1:{server.bound_port}") as ch:
                r2 = lms_pb2_grpc.TutoringStub(ch).GetLLMAnswer(lms_pb2.QueryRequest(token="t", query="Explain leader election"))
            assert r2.success and r2.response == r.response
    finally:
        server.stop(0)


@pytest.mark.gpu
def test_grpc_drop_in_with_engine(pkg, golden):
    from concurrent.futures import ThreadPoolExecutor
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    g = golden("gpt2_as_written")
    eng = pkg.B200GPT2LMHeadModel.from_hf(rp.build_gpt2("small", seed=0), dtype="fp32")
    tok = rp.FakeGPT2Tokenizer(prompt_len=32, seed=1)
    server = serving.serve(eng, tok, address="127.0.0.1:0")
    try:
        addr = f"127.0.0.1:{server.bound_port}"
        queries = [str(g[f"query_{i}"]) for i in range(3)]  # read here: an NpzFile must not be read from several threads
        texts = [str(g[f"text_{i}"]) for i in range(3)]
        with ThreadPoolExecutor(6) as ex:  # concurrent clients, like the LMS leader's handler threads
            rs = list(ex.map(lambda i: serving.call(addr, "t", queries[i % 3]), range(6)))
        for i, r in enumerate(rs):
            assert r.success and r.response == texts[i % 3]  # identical to the unmodified reference's text
        # serve() put the micro-batching queue in front of the engine: the six concurrent requests did not run one by one
        assert isinstance(server.model, serving.BatchingModel) and max(server.model.batches) > 1
    finally:
        server.stop(0)


class RowModel:
    """CPU stand-in with the engine's generate contract for left-padded batches: row-wise deterministic 'generation'
    (next token = (sum of the row's real tokens + step) % 50), EOS = 7."""
    def __init__(self):
        from types import SimpleNamespace
        self.generation_config = SimpleNamespace(max_length=20, eos_token_id=7, pad_token_id=None)
        self.calls = []

    def _resolve(self, T, max_length, max_new_tokens, eos, pad):
        import lms_b200
        return lms_b200.B200GPT2LMHeadModel._resolve(self, T, max_length, max_new_tokens, eos, pad)

    def generate(self, ids, attention_mask=None, max_new_tokens=None, max_length=None, repetition_penalty=1.0,
                 eos_token_id=None, pad_token_id=None, **kw):
        import torch
        B, T = ids.shape
        mask = attention_mask if attention_mask is not None else torch.ones_like(ids)
        new = max_new_tokens if max_new_tokens is not None else max_length - T
        self.calls.append(B)
        out = torch.full((B, T + new), 7, dtype=torch.int64)
        out[:, :T] = ids
        for b in range(B):
            base = int((ids[b] * mask[b]).sum())
            done = False
            for s in range(new):
                tok = 7 if done else (base + s) % 50
                out[b, T + s] = tok
                done = done or tok == 7
        return out


def test_batching_queue_returns_batch1_results(pkg):
    """10 concurrent batch-1 callers with different prompt lengths and budgets: coalesced into few batches, each
    caller gets exactly what its own batch-1 call returns (own length, cut after its EOS)."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    eng = RowModel()
    bm = serving.BatchingModel(eng, max_batch=16, max_wait_ms=150)
    prompts = [torch.randint(1, 40, (1, n), generator=torch.Generator().manual_seed(n)) for n in (3, 9, 5, 12, 7, 4, 11, 6, 8, 10)]
    def one(p):
        return bm.generate(p, max_length=30, num_return_sequences=1, temperature=0.7, top_k=50, top_p=0.9,
                           repetition_penalty=1.2)
    try:
        with ThreadPoolExecutor(10) as ex:
            outs = list(ex.map(one, prompts))
    finally:
        bm.close()
    assert sum(bm.batches) == 10 and len(bm.batches) <= 3  # coalesced
    solo = RowModel()
    for p, o in zip(prompts, outs):
        ref = solo.generate(p, max_length=30)[0]
        hit = (ref[p.shape[1]:] == 7).nonzero()
        ref = ref[:p.shape[1] + int(hit[0]) + 1] if hit.numel() else ref
        assert o.shape[0] == 1 and torch.equal(o[0], ref)
    with pytest.raises(ValueError, match="max_length"):
        serving.BatchingModel(RowModel()).generate(torch.ones((1, 30), dtype=torch.int64), max_length=30)


def test_relevance_gate_caches_assignment_embedding(pkg):
    import torch
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    calls = []
    class Tok:
        def __call__(self, text, **kw):
            ids = rp.bert_ids(1, 6 + len(text) % 5, seed=len(text), vocab=2000)
            return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids)}
    hf = rp.build_bert(seed=6, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                       vocab_size=2000, max_position_embeddings=128)
    class Counting:
        def __call__(self, **inputs):
            calls.append(inputs["input_ids"].shape[1])
            return hf(**inputs)
    gate = serving.RelevanceGate(Counting(), Tok(), threshold=0.6)
    gate._bert = type("B", (), dict(compute_embeddings=staticmethod(rp.compute_embeddings),
                                    cosine_similarity=staticmethod(rp.cosine_similarity),
                                    NOT_RELATED_MESSAGE=rp.NOT_RELATED_MESSAGE))
    text = "the assignment text " * 3
    r1 = gate.check("first question", text)
    r2 = gate.check("second question?", text)
    assert len(calls) == 3 and gate.hits == 1 and gate.misses == 1  # assignment embedded once, two query forwards
    ref = rp.cosine_similarity(rp.compute_embeddings("first question", hf, Tok()), rp.compute_embeddings(text, hf, Tok()))
    assert abs(r1[1] - ref) < 1e-6 and r1[0] == rp.relevance_decision(ref)
    assert gate.answer("q", text, lambda q: "tutor says hi") in ("tutor says hi", rp.NOT_RELATED_MESSAGE)
    gate.threshold = 2.0  # nothing is relevant: the canned reply of REF/lms_server.py:1270
    assert gate.answer("q", text, lambda q: "tutor says hi") == rp.NOT_RELATED_MESSAGE


@pytest.mark.gpu
def test_batching_queue_with_engine_equals_batch1(pkg):
    """Engine behind the queue: every caller's ids equal its own batch-1 engine call (fp32 mode)."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")
    eng = pkg.B200GPT2LMHeadModel.from_hf(rp.build_gpt2("tiny", seed=3), dtype="fp32")
    eng.generation_config.eos_token_id = 999
    prompts = [rp.gpt2_prompt_ids(1, n, seed=40 + n, vocab=1000) for n in (5, 17, 9, 30, 12, 8)]
    solo = [eng.generate(p, max_length=48, repetition_penalty=1.2) for p in prompts]
    bm = serving.BatchingModel(eng, max_batch=8, max_wait_ms=300)
    try:
        with ThreadPoolExecutor(6) as ex:
            outs = list(ex.map(lambda p: bm.generate(p, max_length=48, repetition_penalty=1.2), prompts))
    finally:
        bm.close()
    assert sum(bm.batches) == 6 and len(bm.batches) <= 2
    for o, s in zip(outs, solo):
        assert torch.equal(o, s)


@pytest.mark.gpu
def test_assignment_index_chunks_embeds_and_finds(pkg):
    """whole-assignment ingestion.  Chunk 0 equals the reference's (truncated) embedding; a query
    that repeats a later chunk finds that chunk with cosine ~1."""
    import zlib

    import torch

    serving = __import__("importlib").import_module("distributed-lms-raft-llm_b200.serving")

    class Tok:  # deterministic "word pieces": one id per word, framed by [CLS]=101 / [SEP]=102
        def __call__(self, text, truncation=True, max_length=512, **kw):
            body = [1000 + zlib.crc32(w.encode()) % 900 for w in text.split()]
            if truncation:
                body = body[:max_length - 2]
            ids = torch.tensor([[101] + body + [102]], dtype=torch.int64)
            return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids)}

    hf = rp.build_bert(seed=6, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                       vocab_size=2000, max_position_embeddings=64)
    bert = pkg.B200BertModel.from_hf(hf, dtype="fp32")
    index = serving.AssignmentIndex(bert, Tok(), chunk_tokens=64)
    words = [f"w{i}" for i in range(62 * 3 + 20)]  # 3 full chunks and a 20-word tail
    text = " ".join(words)
    assert index.ingest("hw1", text) == 4 and len(index) == 4
    assert index.ingest("hw2", " ".join(f"x{i}" for i in range(30))) == 1
    # chunk 0 is the reference's own embedding of the (truncated) text
    ref0 = rp.compute_embeddings(text, hf, lambda t, **kw: Tok()(t, truncation=True, max_length=64))
    got0 = index._corpus().rows[0].cpu()
    assert (got0 - ref0[0]).abs().max().item() < 1e-4
    # a query made of chunk 2's words scores ~1 against ("hw1", 2)
    hits = index.search(" ".join(words[124:186]), k=3)
    assert hits[0][:2] == ("hw1", 2) and abs(hits[0][2] - 1.0) < 1e-4
    assert hits[0][2] > hits[1][2]
    ok, best = index.is_relevant(" ".join(f"x{i}" for i in range(30)), threshold=0.99)
    assert ok and best[:2] == ("hw2", 0)


@pytest.mark.gpu
def test_relevance_gate_over_b200_bert(pkg):
    """SURVEY section 8f rank 2 on the GPU: RelevanceGate over B200BertModel (load once, assignment embedding cached)
    gives the reference's similarity (REF/lms_server.py:1262-1266, via the oracle on the same weights, <= 1e-4), the same
    relevant / not-relevant decision at the reference's 0.6 and across a threshold sweep, one BERT forward per request
    after the first, and the canned reply of REF/lms_server.py:1270 for an unrelated query."""
    import torch
    from importlib import import_module
    serving = import_module("distributed-lms-raft-llm_b200.serving")

    class Tok:  # deterministic ids per text (no vocabulary offline), the reference's tokenizer call shape
        def __call__(self, text, **kw):
            n = min(kw.get("max_length", 512), 8 + (sum(text.encode()) % 90))
            ids = rp.bert_ids(1, n, seed=sum(text.encode()), vocab=2000)
            return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids)}

    hf = rp.build_bert(seed=6, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                       vocab_size=2000, max_position_embeddings=128)
    bert = pkg.B200BertModel.from_hf(hf, dtype="fp32")
    launches = pkg._lib.load().tutor_launch_count
    gate = serving.RelevanceGate(bert, Tok(), threshold=0.6)
    text = "write a report about distributed consensus and log replication " * 2
    queries = ["how does raft elect a leader", "what is log replication?", "best pizza topping", "x"]
    sims = []
    for i, qtext in enumerate(queries):
        n0 = launches()
        relevant, sim = gate.check(qtext, text)
        n1 = launches()
        ref = rp.cosine_similarity(rp.compute_embeddings(qtext, hf, Tok()), rp.compute_embeddings(text, hf, Tok()))
        assert abs(sim - ref) < 1e-4, (qtext, sim, ref)
        assert relevant == rp.relevance_decision(ref)
        sims.append((sim, ref))
        if i == 1:
            per_forward = n1 - n0
        if i >= 2:
            assert n1 - n0 == per_forward, "a cached assignment costs exactly one BERT forward per request"
    assert gate.misses == 1 and gate.hits == len(queries) - 1
    for sim, ref in sims:  # decisions identical wherever the threshold falls
        for thr in (ref - 1e-3, ref + 1e-3, 0.0, 0.6, 0.99):
            gate.threshold = thr
            assert (not (sim < thr)) == (not (ref < thr))
    gate.threshold = 2.0
    assert gate.answer(queries[0], text, lambda q: "tutor says hi") == rp.NOT_RELATED_MESSAGE
    gate.threshold = -2.0
    assert gate.answer(queries[0], text, lambda q: "tutor says hi") == "tutor says hi"
