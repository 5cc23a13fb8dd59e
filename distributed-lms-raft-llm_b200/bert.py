"""Drop-ins for the reference's BERT relevance gate (REF/lms_server.py:97-104, 1258-1270).

* ``B200BertModel`` is callable like the ``BertModel`` the reference builds at REF/lms_server.py:1259:
  ``model(**inputs).last_hidden_state`` has shape [B, T, 768].
* ``compute_embeddings(text, model, tokenizer)`` and ``cosine_similarity(emb1, emb2)`` keep the reference's
  signatures, return types ([1,768] fp32 CPU tensor; Python float) and error behaviour (none caught).
  With a ``B200BertModel`` the mean over tokens is taken on the GPU and only the [B,768] embedding crosses
  PCIe; the cosine runs in the fused GPU kernel as well.
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import torch

from . import _lib

SIMILARITY_THRESHOLD = 0.6  # REF/lms_server.py:1267
NOT_RELATED_MESSAGE = ("Your query does not relate to your assignment. Please ask a question related to your "
                       "assignment.")  # REF/lms_server.py:1270


class B200BertModel:
    def __init__(self, config, dtype: str = "fp32", device: int = 0):
        """``config``: a transformers BertConfig (bert-base-uncased architecture; head_dim must be 64)."""
        self.lib = _lib.load()
        self.config = config
        self.dtype = dtype
        self.device_index = device
        cfg = _lib.BertConfig(config.num_hidden_layers, config.num_attention_heads, config.hidden_size,
                              config.intermediate_size, config.vocab_size, config.max_position_embeddings,
                              config.layer_norm_eps, _lib.DTYPES[dtype])
        if getattr(config, "hidden_act", "gelu") != "gelu":
            raise NotImplementedError("only the exact-erf GELU of bert-base-uncased is built")
        self._h = C.c_void_p()
        _lib.check(self.lib.tutor_bert_create(C.byref(cfg), device, C.byref(self._h)))

    @classmethod
    def from_hf(cls, hf_model, dtype: str = "fp32", device: int = 0):
        self = cls(hf_model.config, dtype=dtype, device=device)
        self.load_state_dict(hf_model.state_dict())
        return self

    def load_state_dict(self, state_dict):
        _lib.load_state_dict(self._h, self.lib.tutor_bert_load_tensor, state_dict)
        _lib.check(self.lib.tutor_bert_check_loaded(self._h))
        return self

    def eval(self):
        return self

    def to(self, *a, **k):
        return self

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self.lib.tutor_bert_destroy(h)
            except Exception:
                pass

    @staticmethod
    def _check_inputs(input_ids, token_type_ids, attention_mask):
        if input_ids is None or input_ids.dim() != 2:
            raise ValueError("input_ids must be a [batch, seq] tensor")
        if attention_mask is not None and not bool(torch.all(attention_mask == 1)):
            raise NotImplementedError("padded batches are not supported: the reference embeds one text per call")
        if token_type_ids is not None and bool(torch.any(token_type_ids != 0)):
            raise NotImplementedError("non-zero token_type_ids are not on the reference path")

    def embed(self, input_ids, return_hidden: bool = False):
        """[B,T] ids -> mean-pooled embedding [B, hidden] (and optionally last_hidden_state), CPU fp32."""
        if input_ids.is_cuda:
            ids = input_ids.to(torch.int32).contiguous()
            emb = torch.empty((ids.shape[0], self.config.hidden_size), dtype=torch.float32, device=ids.device)
            torch.cuda.current_stream(ids.device).synchronize()
            _lib.check(self.lib.tutor_bert_embed_dev(self._h, ids.data_ptr(), ids.shape[0], ids.shape[1], emb.data_ptr()))
            return emb
        ids = input_ids.to(torch.int64).contiguous()
        B, T = ids.shape
        emb = torch.empty((B, self.config.hidden_size), dtype=torch.float32)
        hidden = torch.empty((B, T, self.config.hidden_size), dtype=torch.float32) if return_hidden else None
        _lib.check(self.lib.tutor_bert_embed(self._h, ids.data_ptr(), B, T, emb.data_ptr(),
                                             hidden.data_ptr() if return_hidden else None))
        return (emb, hidden) if return_hidden else emb

    @torch.no_grad()
    def __call__(self, input_ids=None, token_type_ids=None, attention_mask=None, **kwargs):
        """BertModel.forward look-alike (HF/models/bert/modeling_bert.py:628-691)."""
        self._check_inputs(input_ids, token_type_ids, attention_mask)
        emb, hidden = self.embed(input_ids.cpu(), return_hidden=True)
        return SimpleNamespace(last_hidden_state=hidden, pooler_output=None, mean_embedding=emb)


def compute_embeddings(text, model, tokenizer):
    """REF/lms_server.py:97-101, same signature and return ([1,768] fp32)."""
    inputs = tokenizer(text, return_tensors="pt", padding=True, truncation=True, max_length=512)
    if isinstance(model, B200BertModel):
        model._check_inputs(inputs["input_ids"], inputs.get("token_type_ids"), inputs.get("attention_mask"))
        return model.embed(inputs["input_ids"])
    with torch.no_grad():
        outputs = model(**inputs)
    return outputs.last_hidden_state.mean(dim=1)


def cosine_similarity(emb1, emb2):
    """REF/lms_server.py:103-104: F.cosine_similarity(dim=1, eps=1e-8).item(), on the GPU."""
    lib = _lib.load()
    if emb1.is_cuda:
        a, b = emb1.float().contiguous(), emb2.float().contiguous()
        out = torch.empty((a.shape[0],), dtype=torch.float32, device=a.device)
        torch.cuda.current_stream(a.device).synchronize()
        _lib.check(lib.tutor_cosine_pairs(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.shape[0], a.shape[1], None))
        torch.cuda.synchronize(a.device)
        return out.item()
    a, b = emb1.float().contiguous(), emb2.float().contiguous()
    out = torch.empty((a.shape[0],), dtype=torch.float32)
    _lib.check(lib.tutor_cosine_pairs_host(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.shape[0], a.shape[1]))
    return out.item()  # like the reference: raises unless exactly one pair


def cosine_similarity_batch(emb1, emb2):
    """Vector form of the same comparison for B pairs (bench / batched gate)."""
    lib = _lib.load()
    a, b = emb1.float().contiguous(), emb2.float().contiguous()
    out = torch.empty((a.shape[0],), dtype=torch.float32, device=a.device)
    if a.is_cuda:
        torch.cuda.current_stream(a.device).synchronize()
        _lib.check(lib.tutor_cosine_pairs(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.shape[0], a.shape[1], None))
        torch.cuda.synchronize(a.device)
    else:
        _lib.check(lib.tutor_cosine_pairs_host(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.shape[0], a.shape[1]))
    return out


def is_relevant(similarity: float, threshold: float = SIMILARITY_THRESHOLD) -> bool:
    """REF/lms_server.py:1269: the query is forwarded unless ``similarity < threshold``."""
    return not (similarity < threshold)
