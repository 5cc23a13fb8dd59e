"""Drop-in for the module-global ``model`` of the reference's tutoring server.

The reference calls ``model.generate(inputs, max_length=150, num_return_sequences=1, temperature=0.7,
top_k=50, top_p=0.9, repetition_penalty=1.2)`` (REF/tutoring_server.py:21-29) and decodes ``outputs[0]``.
``B200GPT2LMHeadModel.generate`` keeps that signature and those semantics — greedy (the sampling knobs are
inert without ``do_sample=True``, HF/generation/utils.py:1198-1223), RepetitionPenaltyLogitsProcessor,
MaxLengthCriteria counting the prompt, EosTokenCriteria with pad:=eos — and runs the whole loop on the GPU
through ``tutor_gpt2_generate`` (include/tutor_b200.h).
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import torch

from . import _lib


class B200GPT2LMHeadModel:
    def __init__(self, config, dtype: str = "fp32", device: int = 0):
        """``config``: a transformers GPT2Config (or anything with the same attributes)."""
        self.lib = _lib.load()
        self.config = config
        self.dtype = dtype
        self.device_index = device
        self.n_positions = int(config.n_positions)
        self.generation_config = SimpleNamespace(max_length=20, eos_token_id=getattr(config, "eos_token_id", 50256),
                                                 pad_token_id=getattr(config, "pad_token_id", None))
        cfg = _lib.Gpt2Config(config.n_layer, config.n_head, config.n_embd, config.vocab_size, config.n_positions,
                              config.layer_norm_epsilon, _lib.DTYPES[dtype])
        self._h = C.c_void_p()
        _lib.check(self.lib.tutor_gpt2_create(C.byref(cfg), device, C.byref(self._h)))

    @classmethod
    def from_hf(cls, hf_model, dtype: str = "fp32", device: int = 0):
        """Build from a transformers GPT2LMHeadModel (the object ``from_pretrained`` returns)."""
        self = cls(hf_model.config, dtype=dtype, device=device)
        self.load_state_dict(hf_model.state_dict())
        return self

    def load_state_dict(self, state_dict):
        _lib.load_state_dict(self._h, self.lib.tutor_gpt2_load_tensor, state_dict)
        _lib.check(self.lib.tutor_gpt2_check_loaded(self._h))
        return self

    # nn.Module look-alikes the reference's surroundings may touch
    def eval(self):
        return self

    def to(self, *a, **k):
        return self

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self.lib.tutor_gpt2_destroy(h)
            except Exception:
                pass

    # ------------------------------------------------------------------ generate
    def _resolve(self, T, max_length, max_new_tokens, eos_token_id, pad_token_id):
        if max_new_tokens is not None:
            total = T + int(max_new_tokens)  # HF/generation/utils.py:1540-1552
        else:
            total = int(max_length) if max_length is not None else self.generation_config.max_length
        if T >= total:  # the ValueError transformers raises (HF/generation/utils.py:1575-1581)
            raise ValueError(
                f"Input length of input_ids is {T}, but `max_length` is set to {total}. This can lead to unexpected "
                "behavior. You should consider increasing `max_length` or, better yet, setting `max_new_tokens`.")
        eos = self.generation_config.eos_token_id if eos_token_id is None else eos_token_id
        if isinstance(eos, (list, tuple)):
            if len(eos) != 1:
                raise NotImplementedError("a single eos_token_id is supported")
            eos = eos[0]
        pad = pad_token_id if pad_token_id is not None else self.generation_config.pad_token_id
        if pad is None:
            pad = eos if eos is not None else 0  # "Setting pad_token_id to eos_token_id" (utils.py:1718-1726)
        return total, (-1 if eos is None else int(eos)), int(pad)

    @torch.no_grad()
    def generate(self, inputs=None, max_length=None, max_new_tokens=None, num_return_sequences=1, temperature=None,
                 top_k=None, top_p=None, repetition_penalty=1.0, do_sample=False, attention_mask=None,
                 pad_token_id=None, eos_token_id=None, input_ids=None, **kwargs):
        if inputs is None:
            inputs = input_ids
        if inputs is None or inputs.dim() != 2 or inputs.dtype not in (torch.int64, torch.int32):
            raise ValueError("generate expects a [batch, seq] integer tensor of token ids")
        if do_sample:
            raise NotImplementedError("the reference path is greedy (do_sample is never set); sampling is not built")
        if num_return_sequences != 1:
            raise ValueError("Greedy methods without beam search do not support `num_return_sequences` different than 1")
        if kwargs.get("num_beams", 1) != 1:
            raise NotImplementedError("beam search is not on the reference path")
        B, T = inputs.shape
        npad = None
        if attention_mask is not None and not bool(torch.all(attention_mask == 1)):
            # left padding only (what decoder-only generation requires; transformers warns about right padding)
            mask = attention_mask.to("cpu", torch.int64)
            npad = (mask == 0).sum(dim=1)
            expect = (torch.arange(T)[None, :] >= npad[:, None]).to(torch.int64)
            if tuple(mask.shape) != (B, T) or not torch.equal(mask, expect) or bool(torch.any(npad >= T)):
                raise NotImplementedError("only left-padded batches are supported (attention_mask = zeros then ones)")
            npad = npad.to(torch.int32).contiguous()
        total, eos, pad = self._resolve(T, max_length, max_new_tokens, eos_token_id, pad_token_id)
        n = C.c_int(0)
        if inputs.is_cuda and npad is None:
            ids = inputs.to(torch.int32).contiguous()
            out = torch.empty((B, total), dtype=torch.int32, device=inputs.device)
            torch.cuda.current_stream(inputs.device).synchronize()
            _lib.check(self.lib.tutor_gpt2_generate_dev(self._h, ids.data_ptr(), B, T, total, float(repetition_penalty),
                                                        eos, pad, out.data_ptr(), C.byref(n)))
            return out[:, :n.value].to(torch.int64)
        ids = inputs.to("cpu", torch.int64).contiguous()
        out = torch.empty((B * total,), dtype=torch.int64)
        _lib.check(self.lib.tutor_gpt2_generate_padded(self._h, ids.data_ptr(), None if npad is None else npad.data_ptr(),
                                                       B, T, total, float(repetition_penalty), eos, pad, out.data_ptr(),
                                                       C.byref(n)))
        return out[:B * n.value].view(B, n.value)

    # ------------------------------------------------------------------ stepwise (tests, profiling)
    def prefill(self, inputs, max_length=None, max_new_tokens=None, repetition_penalty=1.0, eos_token_id=None,
                pad_token_id=None):
        ids = inputs.to("cpu", torch.int64).contiguous()
        B, T = ids.shape
        total, eos, pad = self._resolve(T, max_length, max_new_tokens, eos_token_id, pad_token_id)
        _lib.check(self.lib.tutor_gpt2_prefill(self._h, ids.data_ptr(), B, T, total, float(repetition_penalty), eos, pad))
        self._cur = (B, total)

    def step(self, n: int = 1):
        _lib.check(self.lib.tutor_gpt2_step(self._h, n))

    def read_ids(self):
        B, total = self._cur
        out = torch.empty((B * total,), dtype=torch.int64)
        n, stop = C.c_int(0), C.c_int(0)
        _lib.check(self.lib.tutor_gpt2_read_ids(self._h, out.data_ptr(), B * total, C.byref(n), C.byref(stop)))
        return out[:B * n.value].view(B, n.value)

    def last_logits(self):
        """fp32 logits [B, V] of the most recent forward (materialised only on request)."""
        B, _ = self._cur
        out = torch.empty((B, self.config.vocab_size), dtype=torch.float32, device=f"cuda:{self.device_index}")
        torch.cuda.synchronize(out.device)
        _lib.check(self.lib.tutor_gpt2_logits(self._h, out.data_ptr()))
        return out

    def decode_bytes(self, B: int, S: int) -> float:
        return float(self.lib.tutor_gpt2_decode_bytes(self._h, B, S))

    @property
    def stream(self):
        return self.lib.tutor_gpt2_stream(self._h)
