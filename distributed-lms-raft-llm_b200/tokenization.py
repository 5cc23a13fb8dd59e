"""Host tokenizers with the call shapes the reference uses, backed by the C library (csrc/tokenizer.cpp).

REF/tutoring_server.py:20,30   ``tokenizer.encode(prompt, return_tensors="pt")`` / ``tokenizer.decode(ids, skip_special_tokens=True)``
REF/lms_server.py:98           ``tokenizer(text, return_tensors="pt", padding=True, truncation=True, max_length=512)``

Drop-ins for ``GPT2Tokenizer.from_pretrained(...)`` / ``BertTokenizer.from_pretrained(...)`` built from the same files
(vocab.json + merges.txt, vocab.txt); ids are identical to transformers' (tests/test_tokenizer.py)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def _ids_list(ids):
    if isinstance(ids, torch.Tensor):
        ids = ids.reshape(-1).tolist()
    return [int(i) for i in ids]


class B200GPT2Tokenizer:
    def __init__(self, vocab_file: str, merges_file: str):
        self.lib = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self.lib.tutor_bpe_create(vocab_file.encode(), merges_file.encode(), C.byref(self._h)))
        self.eos_token_id = self.bos_token_id = self.unk_token_id = self.lib.tutor_bpe_eos_id(self._h)
        self.eos_token = self.bos_token = self.unk_token = "<|endoftext|>"
        self.pad_token_id = None

    def __len__(self):
        return self.lib.tutor_bpe_vocab_size(self._h)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self.lib.tutor_bpe_destroy(h)

    def encode(self, text: str, return_tensors=None, **kwargs):
        raw = text.encode("utf-8")
        cap = len(raw) + 8  # byte-level BPE never yields more ids than bytes
        buf = (C.c_int32 * cap)()
        n = C.c_int(0)
        _lib.check(self.lib.tutor_bpe_encode(self._h, raw, len(raw), buf, cap, C.byref(n)))
        ids = list(buf[:n.value])
        return torch.tensor([ids], dtype=torch.int64) if return_tensors == "pt" else ids

    def decode(self, token_ids, skip_special_tokens: bool = False, **kwargs) -> str:
        ids = _ids_list(token_ids)
        arr = (C.c_int32 * max(1, len(ids)))(*ids)
        cap = 16 * len(ids) + 16
        while True:
            out = C.create_string_buffer(cap)
            n = C.c_int(0)
            rc = self.lib.tutor_bpe_decode(self._h, arr, len(ids), 1 if skip_special_tokens else 0, out, cap, C.byref(n))
            if rc != 0 and n.value > cap:  # a vocabulary with very long tokens: retry with what it asked for
                cap = n.value
                continue
            _lib.check(rc)
            return out.raw[:n.value].decode("utf-8", errors="replace")


class B200BertTokenizer:
    def __init__(self, vocab_file: str):
        self.lib = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self.lib.tutor_wordpiece_create(vocab_file.encode(), C.byref(self._h)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self.lib.tutor_wordpiece_destroy(h)

    def encode_ids(self, text: str, truncation: bool = True, max_length: int = 512):
        raw = text.encode("utf-8")
        cap = 3 * len(raw) + 8  # never more ids than bytes + 2
        buf = (C.c_int32 * cap)()
        n = C.c_int(0)
        _lib.check(self.lib.tutor_wordpiece_encode(self._h, raw, len(raw), max_length if truncation else 0, buf, cap,
                                                   C.byref(n)))
        return list(buf[:n.value])

    def __call__(self, text: str, return_tensors="pt", padding=True, truncation=True, max_length=512, **kwargs):
        """One text -> {"input_ids", "token_type_ids", "attention_mask"} of shape [1, T] (padding a single text is a no-op)."""
        ids = self.encode_ids(text, truncation=truncation, max_length=max_length)
        if return_tensors == "pt":
            t = torch.tensor([ids], dtype=torch.int64)
            return {"input_ids": t, "token_type_ids": torch.zeros_like(t), "attention_mask": torch.ones_like(t)}
        return {"input_ids": ids, "token_type_ids": [0] * len(ids), "attention_mask": [1] * len(ids)}
