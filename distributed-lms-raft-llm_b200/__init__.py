"""B200-native engine for the tutoring-inference hot path of naggender2/distributed-lms-raft-llm:
GPT-2 greedy generation (REF/tutoring_server.py:15-31) and the BERT mean-pooled-embedding cosine relevance
gate (REF/lms_server.py:97-104, 1237-1274).  Hand-written sm_100a CUDA behind a C ABI (include/tutor_b200.h);
this package is the Python host side that mirrors the reference's call signatures.  No CPU fallback.
"""
from . import _lib
from .bert import (NOT_RELATED_MESSAGE, SIMILARITY_THRESHOLD, B200BertModel, compute_embeddings, cosine_similarity,
                   cosine_similarity_batch, is_relevant)
from .gpt2 import B200GPT2LMHeadModel
from .tokenization import B200BertTokenizer, B200GPT2Tokenizer
from .relevance import CorpusShard, close_exchange, exchange_candidates, init_exchange, shard_bounds

__all__ = ["B200GPT2LMHeadModel", "B200BertModel", "B200GPT2Tokenizer", "B200BertTokenizer", "compute_embeddings", "cosine_similarity", "cosine_similarity_batch",
           "is_relevant", "CorpusShard", "exchange_candidates", "init_exchange", "close_exchange", "shard_bounds", "SIMILARITY_THRESHOLD",
           "NOT_RELATED_MESSAGE", "_lib"]
