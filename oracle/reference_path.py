"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference's tutoring-inference path, restated call site by call site on top
of the same third-party code the reference itself runs (``transformers 5.5.0``,
``torch 2.11.0``, CPU, fp32).  ``/root/reference`` is not needed at run time, so
this file travels to the GPU box; ``oracle/make_golden.py`` checks — in the build
container, where the reference is mounted — that every function here returns
exactly what the unmodified reference module returns on the same inputs.

REF = /root/reference/GUI_RAFT_LLM_SourceCode
"""
from __future__ import annotations

import warnings

import torch
import torch.nn.functional as F

# Reference generate kwargs, REF/tutoring_server.py:21-29 (the sampling knobs are
# inert because do_sample is never set; repetition_penalty IS active).
REFERENCE_GENERATE_KWARGS = dict(
    max_length=150,
    num_return_sequences=1,
    temperature=0.7,
    top_k=50,
    top_p=0.9,
    repetition_penalty=1.2,
)
# REF/tutoring_server.py:18
PROMPT_TEMPLATE = (
    "You are an intelligent assistant. Answer the following question in detail:\n"
    "Question: {query}\nAnswer:"
)
SIMILARITY_THRESHOLD = 0.6  # REF/lms_server.py:1267
NOT_RELATED_MESSAGE = (  # REF/lms_server.py:1270
    "Your query does not relate to your assignment. Please ask a question related to your assignment."
)

GPT2_ARCH = {
    "small": dict(n_embd=768, n_layer=12, n_head=12),
    "large": dict(n_embd=1280, n_layer=36, n_head=20),
    # scaled-down shapes for fast CPU/GPU unit tests (not a BASELINE config)
    "tiny": dict(n_embd=128, n_layer=2, n_head=2, vocab_size=1000, n_positions=256),
}


def build_gpt2(arch: str = "small", seed: int = 0, **overrides):
    """Random-init GPT2LMHeadModel, fp32, eval (SURVEY.md §8c golden recipe)."""
    from transformers import GPT2Config, GPT2LMHeadModel

    cfg = dict(GPT2_ARCH[arch])
    cfg.update(overrides)
    torch.manual_seed(seed)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = GPT2LMHeadModel(GPT2Config(**cfg)).eval()
    return model


def build_bert(seed: int = 0, **overrides):
    """Random-init BertModel (bert-base-uncased architecture), fp32, eval."""
    from transformers import BertConfig, BertModel

    torch.manual_seed(seed)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = BertModel(BertConfig(**overrides)).eval()
    return model


def gpt2_prompt_ids(batch: int, length: int, seed: int = 1, vocab: int = 50257):
    """Synthetic prompts; excludes the EOS id (vocab-1) like SURVEY.md §8c."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab - 1, (batch, length), generator=g)


def bert_ids(batch: int, length: int, seed: int = 2, vocab: int = 30522):
    """Synthetic WordPiece ids: [CLS] ... [SEP], all-ones mask, zero type ids."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1000, vocab, (batch, length), generator=g)
    ids[:, 0] = 101
    ids[:, -1] = 102
    return ids


class FakeGPT2Tokenizer:
    """Stands in for GPT2Tokenizer (vocab files are absent offline).

    ``encode`` maps the text to a deterministic synthetic id row; ``decode`` renders
    ids as space-separated integers so the round trip is checkable.
    """

    def __init__(self, prompt_len: int = 32, seed: int = 1, vocab: int = 50257):
        self.prompt_len, self.seed, self.vocab = prompt_len, seed, vocab
        self.eos_token_id = vocab - 1

    def encode(self, text, return_tensors=None):
        h = (sum(text.encode()) * 2654435761 + self.seed) % (2**31)
        ids = gpt2_prompt_ids(1, self.prompt_len, seed=h, vocab=self.vocab)
        return ids if return_tensors == "pt" else ids[0].tolist()

    def decode(self, ids, skip_special_tokens=False):
        ids = [int(i) for i in ids]
        if skip_special_tokens:
            ids = [i for i in ids if i != self.eos_token_id]
        return " ".join(str(i) for i in ids)


class FakeBertTokenizer:
    """Stands in for BertTokenizer: text -> deterministic ids, truncation at max_length."""

    def __init__(self, vocab: int = 30522):
        self.vocab = vocab

    def __call__(self, text, return_tensors="pt", padding=True, truncation=True, max_length=512):
        n_words = max(1, len(text.split()))
        length = min(max_length, n_words + 2) if truncation else n_words + 2
        h = (sum(text.encode()) * 2654435761) % (2**31)
        ids = bert_ids(1, length, seed=h, vocab=self.vocab)
        return {
            "input_ids": ids,
            "token_type_ids": torch.zeros_like(ids),
            "attention_mask": torch.ones_like(ids),
        }


# --------------------------------------------------------------------------- GPT-2
def generate(model, input_ids, **kwargs):
    """``model.generate`` exactly as REF/tutoring_server.py:21-29 calls it.

    With no kwargs this is the reference call.  Bench configs override
    ``max_new_tokens`` / ``repetition_penalty`` / ``do_sample=False``.
    """
    kw = dict(REFERENCE_GENERATE_KWARGS)
    if "max_new_tokens" in kwargs:
        kw.pop("max_length")
    kw.update(kwargs)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return model.generate(input_ids, **kw)


def generate_with_logits(model, input_ids, **kwargs):
    """Same call, also returning per-step raw fp32 logits [steps, B, V] and
    post-processor scores (HF/generation/utils.py:2762-2773)."""
    out = generate(model, input_ids, return_dict_in_generate=True, output_logits=True,
                   output_scores=True, **kwargs)
    return out.sequences, torch.stack(out.logits), torch.stack(out.scores)


def generate_llm_response(query, model, tokenizer):
    """REF/tutoring_server.py:15-31 with the module globals passed in."""
    prompt = PROMPT_TEMPLATE.format(query=query)
    inputs = tokenizer.encode(prompt, return_tensors="pt")
    outputs = generate(model, inputs)
    return tokenizer.decode(outputs[0], skip_special_tokens=True)


# ---------------------------------------------------------------------------- BERT
def embeddings_from_ids(model, input_ids):
    """REF/lms_server.py:99-101 on already-tokenised input: unmasked mean over T."""
    inputs = {
        "input_ids": input_ids,
        "token_type_ids": torch.zeros_like(input_ids),
        "attention_mask": torch.ones_like(input_ids),
    }
    with torch.no_grad():
        outputs = model(**inputs)
    return outputs.last_hidden_state.mean(dim=1)


def compute_embeddings(text, model, tokenizer):
    """REF/lms_server.py:97-101."""
    inputs = tokenizer(text, return_tensors="pt", padding=True, truncation=True, max_length=512)
    with torch.no_grad():
        outputs = model(**inputs)
    return outputs.last_hidden_state.mean(dim=1)


def cosine_similarity(emb1, emb2):
    """REF/lms_server.py:103-104."""
    return F.cosine_similarity(emb1, emb2).item()


def relevance_decision(similarity: float, threshold: float = SIMILARITY_THRESHOLD) -> bool:
    """REF/lms_server.py:1269: True = forward the query to the tutoring server."""
    return not (similarity < threshold)


def corpus_topk(queries, corpus, k):
    """cfg4 (an extension of REF/lms_server.py:1266 to a corpus): cosine of every
    query against every corpus row, top-k by score (ties -> lower index)."""
    qn = queries / queries.norm(dim=1, keepdim=True).clamp_min(1e-8)
    cn = corpus / corpus.norm(dim=1, keepdim=True).clamp_min(1e-8)
    scores = qn @ cn.T
    vals, idx = torch.topk(scores, k, dim=1)
    return vals, idx
