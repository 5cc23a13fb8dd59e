"""ctypes binding of libtutor_b200.so (include/tutor_b200.h).  There is no fallback: if the library
cannot be built or loaded, importing the engine classes fails loudly."""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

c_void_p, c_int, c_float, c_char_p, c_size_t = C.c_void_p, C.c_int, C.c_float, C.c_char_p, C.c_size_t
P = C.POINTER


class Gpt2Config(C.Structure):
    _fields_ = [("n_layer", C.c_int32), ("n_head", C.c_int32), ("n_embd", C.c_int32), ("vocab_size", C.c_int32),
                ("n_positions", C.c_int32), ("layer_norm_epsilon", C.c_float), ("dtype", C.c_int32)]


class BertConfig(C.Structure):
    _fields_ = [("num_hidden_layers", C.c_int32), ("num_attention_heads", C.c_int32), ("hidden_size", C.c_int32),
                ("intermediate_size", C.c_int32), ("vocab_size", C.c_int32), ("max_position_embeddings", C.c_int32),
                ("layer_norm_eps", C.c_float), ("dtype", C.c_int32)]


DT_F32, DT_BF16 = 0, 1
DTYPES = {"fp32": DT_F32, "float32": DT_F32, "f32": DT_F32, "bf16": DT_BF16, "bfloat16": DT_BF16}

# name -> (restype, argtypes); mirrors include/tutor_b200.h one to one
SIGNATURES = {
    "tutor_last_error": (c_char_p, []),
    "tutor_version": (c_int, []),
    "tutor_launch_count": (C.c_uint64, []),
    "tutor_gpt2_create": (c_int, [P(Gpt2Config), c_int, P(c_void_p)]),
    "tutor_gpt2_destroy": (None, [c_void_p]),
    "tutor_gpt2_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, P(C.c_int64), c_int]),
    "tutor_gpt2_check_loaded": (c_int, [c_void_p]),
    "tutor_gpt2_generate": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_int, c_void_p, P(c_int)]),
    "tutor_gpt2_generate_padded": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_int,
                                           c_void_p, P(c_int)]),
    "tutor_gpt2_generate_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_int, c_void_p,
                                        P(c_int)]),
    "tutor_gpt2_prefill": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_int]),
    "tutor_gpt2_step": (c_int, [c_void_p, c_int]),
    "tutor_gpt2_read_ids": (c_int, [c_void_p, c_void_p, c_int, P(c_int), P(c_int)]),
    "tutor_gpt2_logits": (c_int, [c_void_p, c_void_p]),
    "tutor_gpt2_decode_bytes": (C.c_double, [c_void_p, c_int, c_int]),
    "tutor_gpt2_stream": (c_void_p, [c_void_p]),
    "tutor_gpt2_decode_trace": (c_int, [c_void_p, c_int, c_void_p, c_int]),
    "tutor_bert_create": (c_int, [P(BertConfig), c_int, P(c_void_p)]),
    "tutor_bert_destroy": (None, [c_void_p]),
    "tutor_bert_load_tensor": (c_int, [c_void_p, c_char_p, c_void_p, P(C.c_int64), c_int]),
    "tutor_bert_check_loaded": (c_int, [c_void_p]),
    "tutor_bert_embed": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "tutor_bert_embed_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_bert_stream": (c_void_p, [c_void_p]),
    "tutor_cosine_pairs": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_cosine_pairs_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "tutor_row_inv_norm": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_cosine_topk_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "tutor_cosine_topk": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
                                  c_void_p, c_void_p]),
    "tutor_topk_merge": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "tutor_nccl_load": (c_int, [c_char_p]),
    "tutor_nccl_unique_id": (c_int, [c_void_p]),
    "tutor_nccl_comm_create": (c_int, [c_void_p, c_int, c_int, P(c_void_p)]),
    "tutor_nccl_comm_destroy": (None, [c_void_p]),
    "tutor_topk_merge_allgather_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "tutor_topk_merge_allgather": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                           c_void_p]),
    "tutor_layernorm": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_linear_bf16_residual": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                           c_void_p]),
    "tutor_layernorm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_linear": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                             c_int, c_void_p]),
    "tutor_attn_prefill": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                   c_void_p]),
    "tutor_attn_decode": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "tutor_attn_decode_fused": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                        c_void_p]),
    "tutor_lmhead_scratch_bytes": (c_size_t, [c_int]),
    "tutor_lmhead_argmax": (c_int, [c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_float, c_void_p,
                                    c_void_p, c_void_p]),
    "tutor_meanpool": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "tutor_row_tile_stats": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tutor_decode_ln_linear_partial": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int, c_int,
                                               c_int, c_int, c_void_p, c_void_p]),
    "tutor_decode_ln_linear_act": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int,
                                           c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "tutor_decode_linear_residual": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                             c_void_p, c_void_p]),
    "tutor_bpe_create": (c_int, [c_char_p, c_char_p, P(c_void_p)]),
    "tutor_bpe_destroy": (None, [c_void_p]),
    "tutor_bpe_eos_id": (c_int, [c_void_p]),
    "tutor_bpe_vocab_size": (c_int, [c_void_p]),
    "tutor_bpe_encode": (c_int, [c_void_p, c_char_p, C.c_int64, c_void_p, c_int, P(c_int)]),
    "tutor_bpe_decode": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, P(c_int)]),
    "tutor_wordpiece_create": (c_int, [c_char_p, P(c_void_p)]),
    "tutor_wordpiece_destroy": (None, [c_void_p]),
    "tutor_wordpiece_encode": (c_int, [c_void_p, c_char_p, C.c_int64, c_int, c_void_p, c_int, P(c_int)]),
    "tutor_lmhead_argmax_static": (c_int, [c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_float, c_void_p,
                                           c_void_p]),
}


def lib_path() -> str:
    return _build.LIB


def load():
    """Build (when nvcc is present and the sources are newer) and load the shared library."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if os.path.exists(_build.NVCC):
            _build.build()
        if not os.path.exists(_build.LIB):
            raise RuntimeError(f"{_build.LIB} is missing and nvcc is not available to build it; "
                               "the tutoring engine has no CPU fallback")
        lib = C.CDLL(_build.LIB)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        _lib = lib
        return lib


class TutorError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"tutor_b200 error {code}: {message}")
        self.code, self.message = code, message


def check(rc: int):
    if rc != 0:
        raise TutorError(rc, load().tutor_last_error().decode("utf-8", "replace"))


def load_state_dict(handle, load_fn, state_dict):
    """Feed every entry of a transformers state_dict to tutor_*_load_tensor (host fp32, contiguous)."""
    import torch

    for name, t in state_dict.items():
        if not torch.is_floating_point(t):
            continue  # position_ids / token_type_ids / attn.bias (bool) buffers are not weights
        a = t.detach().to("cpu", torch.float32).contiguous()
        shape = (C.c_int64 * a.dim())(*a.shape)
        check(load_fn(handle, name.encode(), a.data_ptr(), shape, a.dim()))
