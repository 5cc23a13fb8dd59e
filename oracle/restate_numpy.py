"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

A from-the-papers numpy restatement of the arithmetic the reference's two call
sites execute inside ``transformers`` — an independent second opinion on the
oracle in ``reference_path.py`` and the per-kernel checker for unit tests.

Every function cites the third-party line it restates:
HF = site-packages/transformers (5.5.0).  Weights come in as the HF ``state_dict``
(numpy fp32), so the weight-name contract of the engine's loader is exercised too.
"""
from __future__ import annotations

import math

import numpy as np


def layer_norm(x, w, b, eps):
    """nn.LayerNorm: biased variance over the last axis (HF/models/gpt2/modeling_gpt2.py:252)."""
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)
    return (x - mu) / np.sqrt(var + eps) * w + b


def gelu_new(x):
    """NewGELUActivation, HF/activations.py:59-66."""
    return 0.5 * x * (1.0 + np.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x**3)))


def gelu_erf(x):
    """F.gelu exact (erf) form, HF/activations.py:70-89."""
    erf = np.vectorize(math.erf, otypes=[np.float32])
    return (0.5 * x * (1.0 + erf(x / math.sqrt(2.0)))).astype(np.float32)


def softmax(x):
    x = x - x.max(-1, keepdims=True)
    e = np.exp(x)
    return e / e.sum(-1, keepdims=True)


def sd_numpy(model):
    return {k: v.detach().cpu().float().numpy() for k, v in model.state_dict().items()}


# --------------------------------------------------------------------------- GPT-2
class GPT2Numpy:
    """GPT2LMHeadModel.forward with a preallocated KV cache (HF/models/gpt2/modeling_gpt2.py:522-706)."""

    def __init__(self, sd, n_head, eps=1e-5):
        self.sd, self.H, self.eps = sd, n_head, eps
        self.L = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("transformer.h."))
        self.d = sd["transformer.wte.weight"].shape[1]
        self.k = [None] * self.L
        self.v = [None] * self.L

    def reset(self):
        self.k = [None] * self.L
        self.v = [None] * self.L

    def forward(self, ids, past_len):
        """ids [B,T] -> last-position logits [B,V] (logits_to_keep=1, HF/generation/utils.py:2490)."""
        sd, H = self.sd, self.H
        B, T = ids.shape
        hd = self.d // H
        pos = np.arange(past_len, past_len + T)
        x = sd["transformer.wte.weight"][ids] + sd["transformer.wpe.weight"][pos][None]  # :577-585
        for l in range(self.L):
            p = f"transformer.h.{l}."
            a = layer_norm(x, sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], self.eps)  # :273
            # Conv1D: x @ W[in,out] + b  (HF/pytorch_utils.py:119-123)
            qkv = a @ sd[p + "attn.c_attn.weight"] + sd[p + "attn.c_attn.bias"]
            q, k, v = np.split(qkv, 3, axis=-1)  # :185
            sh = lambda t: t.reshape(B, T, H, hd).transpose(0, 2, 1, 3)
            q, k, v = sh(q), sh(k), sh(v)
            if self.k[l] is not None:  # DynamicLayer.update, HF/cache_utils.py:119-120
                k = np.concatenate([self.k[l], k], axis=2)
                v = np.concatenate([self.v[l], v], axis=2)
            self.k[l], self.v[l] = k, v
            s = (q @ k.transpose(0, 1, 3, 2)) * (hd**-0.5)  # sdpa, scale = head_dim^-0.5
            S = k.shape[2]
            causal = np.arange(S)[None, :] <= (np.arange(T)[:, None] + (S - T))
            s = np.where(causal[None, None], s, -np.inf)
            o = softmax(s) @ v
            o = o.transpose(0, 2, 1, 3).reshape(B, T, self.d)
            x = x + (o @ sd[p + "attn.c_proj.weight"] + sd[p + "attn.c_proj.bias"])  # :282
            m = layer_norm(x, sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], self.eps)  # :304
            h = gelu_new(m @ sd[p + "mlp.c_fc.weight"] + sd[p + "mlp.c_fc.bias"])  # :239-240
            x = x + (h @ sd[p + "mlp.c_proj.weight"] + sd[p + "mlp.c_proj.bias"])  # :307
        x = layer_norm(x[:, -1], sd["transformer.ln_f.weight"], sd["transformer.ln_f.bias"], self.eps)
        return x @ sd["transformer.wte.weight"].T  # tied lm_head, :646-651


def repetition_penalty(scores, ids_so_far, penalty):
    """RepetitionPenaltyLogitsProcessor, HF/generation/logits_process.py:404-410."""
    out = scores.copy()
    for b in range(scores.shape[0]):
        seen = np.unique(ids_so_far[b])
        s = scores[b, seen]
        out[b, seen] = np.where(s < 0, s * penalty, s / penalty)
    return out


def greedy_generate(net: GPT2Numpy, ids, max_new_tokens=None, max_length=None, repetition_penalty_=1.0,
                    eos_token_id=50256, pad_token_id=None):
    """GenerationMixin._sample greedy branch (HF/generation/utils.py:2736-2805)."""
    ids = np.asarray(ids, dtype=np.int64)
    B, T = ids.shape
    total = T + max_new_tokens if max_new_tokens is not None else max_length
    pad = eos_token_id if pad_token_id is None else pad_token_id
    net.reset()
    unfinished = np.ones(B, dtype=bool)
    logits = net.forward(ids, 0)
    while True:
        scores = repetition_penalty(logits.astype(np.float32), ids, repetition_penalty_) \
            if repetition_penalty_ != 1.0 else logits
        nxt = scores.argmax(-1)  # first maximal index, like torch.argmax
        nxt = np.where(unfinished, nxt, pad)
        ids = np.concatenate([ids, nxt[:, None]], axis=1)
        unfinished &= nxt != eos_token_id
        if ids.shape[1] >= total or not unfinished.any():
            return ids
        logits = net.forward(ids[:, -1:], ids.shape[1] - 1)


# ---------------------------------------------------------------------------- BERT
def bert_forward(sd, ids, n_head=12, eps=1e-12):
    """BertModel.forward -> last_hidden_state (HF/models/bert/modeling_bert.py:628-691);
    all-ones mask, zero token types, pooler skipped (its output is unused by the reference)."""
    B, T = ids.shape
    e = "embeddings."
    x = sd[e + "word_embeddings.weight"][ids] + sd[e + "token_type_embeddings.weight"][0] \
        + sd[e + "position_embeddings.weight"][:T][None]  # :102-109
    x = layer_norm(x, sd[e + "LayerNorm.weight"], sd[e + "LayerNorm.bias"], eps)
    d = x.shape[-1]
    hd = d // n_head
    L = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("encoder.layer."))
    for l in range(L):
        p = f"encoder.layer.{l}."
        lin = lambda t, n: t @ sd[p + n + ".weight"].T + sd[p + n + ".bias"]  # nn.Linear: W[out,in]
        sh = lambda t: t.reshape(B, T, n_head, hd).transpose(0, 2, 1, 3)
        q, k, v = sh(lin(x, "attention.self.query")), sh(lin(x, "attention.self.key")), sh(lin(x, "attention.self.value"))
        o = softmax((q @ k.transpose(0, 1, 3, 2)) * hd**-0.5) @ v  # :192-205
        o = o.transpose(0, 2, 1, 3).reshape(B, T, d)
        x = layer_norm(lin(o, "attention.output.dense") + x,
                       sd[p + "attention.output.LayerNorm.weight"], sd[p + "attention.output.LayerNorm.bias"], eps)
        h = gelu_erf(lin(x, "intermediate.dense").astype(np.float32))  # :339-342
        x = layer_norm(lin(h, "output.dense") + x,
                       sd[p + "output.LayerNorm.weight"], sd[p + "output.LayerNorm.bias"], eps)  # :352-356
    return x


def mean_pool(last_hidden_state):
    """REF/lms_server.py:101: unmasked mean over every position."""
    return last_hidden_state.mean(axis=1)


def cosine(a, b, eps=1e-8):
    """F.cosine_similarity(dim=1, eps=1e-8), REF/lms_server.py:104."""
    na = np.maximum(np.sqrt((a * a).sum(-1)), eps)
    nb = np.maximum(np.sqrt((b * b).sum(-1)), eps)
    return (a * b).sum(-1) / (na * nb)
