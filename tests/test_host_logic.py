"""CPU: host-side logic of the Python boundary (argument resolution, sharding) — no kernels."""
import pytest
import torch

from types import SimpleNamespace


def bare_gpt2(pkg):
    m = pkg.B200GPT2LMHeadModel.__new__(pkg.B200GPT2LMHeadModel)
    m._h = None
    m.generation_config = SimpleNamespace(max_length=20, eos_token_id=50256, pad_token_id=None)
    return m


def test_generate_length_resolution(pkg):
    m = bare_gpt2(pkg)
    assert m._resolve(32, 150, None, None, None) == (150, 50256, 50256)  # the reference call: pad := eos
    assert m._resolve(32, None, 64, None, None) == (96, 50256, 50256)
    assert m._resolve(5, None, None, None, None)[0] == 20                # transformers' default max_length
    assert m._resolve(5, 10, None, 7, 3) == (10, 7, 3)
    with pytest.raises(ValueError, match="Input length of input_ids is 150"):
        m._resolve(150, 150, None, None, None)                             # REF prompt >= 150 tokens


def test_generate_rejects_what_the_reference_never_does(pkg):
    m = bare_gpt2(pkg)
    ids = torch.zeros((1, 4), dtype=torch.int64)
    with pytest.raises(NotImplementedError):
        m.generate(ids, max_length=8, do_sample=True)
    with pytest.raises(ValueError):
        m.generate(ids, max_length=8, num_return_sequences=2)
    with pytest.raises(NotImplementedError):
        m.generate(ids, max_length=8, attention_mask=torch.tensor([[1, 1, 1, 0]]))  # right padding (left is supported)


@pytest.mark.parametrize("n,ws", [(1_000_000, 8), (10, 3), (7, 8), (0, 2)])
def test_shard_bounds_partition(pkg, n, ws):
    spans = [pkg.shard_bounds(n, ws, r) for r in range(ws)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    sizes = [hi - lo for lo, hi in spans]
    assert max(sizes) - min(sizes) <= 1


def test_relevance_decision_boundary(pkg):
    assert pkg.is_relevant(0.6) and pkg.is_relevant(0.61) and not pkg.is_relevant(0.5999)
    assert pkg.SIMILARITY_THRESHOLD == 0.6


def test_assignment_index_chunking_frames_every_window():
    """serving.AssignmentIndex.chunk_ids (no GPU): [CLS] + <= chunk_tokens-2 word pieces + [SEP] windows that cover the
    whole text in order; chunk 0 equals the reference's truncated tokenisation (REF/lms_server.py:98)."""
    import importlib

    import torch
    serving = importlib.import_module("distributed-lms-raft-llm_b200.serving")

    class Tok:
        def __call__(self, text, truncation=True, max_length=512, **kw):
            body = [1000 + int(w) for w in text.split()]
            if truncation:
                body = body[:max_length - 2]
            return {"input_ids": torch.tensor([[101] + body + [102]], dtype=torch.int64)}

    index = serving.AssignmentIndex(model=None, tokenizer=Tok(), chunk_tokens=10)
    text = " ".join(str(i) for i in range(21))  # 21 word pieces -> windows of 8, 8, 5
    chunks = index.chunk_ids(text)
    assert [c.numel() for c in chunks] == [10, 10, 7]
    assert all(int(c[0]) == 101 and int(c[-1]) == 102 for c in chunks)
    assert torch.cat([c[1:-1] for c in chunks]).tolist() == [1000 + i for i in range(21)]
    assert torch.equal(chunks[0], Tok()(text, truncation=True, max_length=10)["input_ids"][0])
    assert [c.numel() for c in index.chunk_ids("7")] == [3]
