"""The tutoring server's gRPC surface with the B200 engine behind it.

The reference's `TutoringService.GetLLMAnswer` (REF/tutoring_server.py:33-37) unwraps `request.query`, calls
`generate_llm_response` (REF/tutoring_server.py:15-31: prompt template -> tokenizer.encode -> model.generate ->
tokenizer.decode) and always answers `success=True`.  This module keeps that behaviour and the wire format of
`lms.Tutoring/GetLLMAnswer` (REF/lms.proto:95-103, 123-125):

* `generate_llm_response(query, model, tokenizer)` - the reference function with its two module globals passed in;
* `install(tutoring_module, engine)` - the one-line injection a maintainer does on the unmodified module;
* `TutoringService` / `serve()` - a servicer registered under the same full method name with messages built from a
  descriptor declared here (field numbers and types from lms.proto), so an unmodified `lms_pb2_grpc.TutoringStub`
  can call it.  Nothing is copied from the generated stubs; with the reference on the path, its own
  `add_TutoringServicer_to_server` works just as well.
"""
from __future__ import annotations

from concurrent import futures

PROMPT_TEMPLATE = ("You are an intelligent assistant. Answer the following question in detail:\n"
                   "Question: {query}\nAnswer:")  # REF/tutoring_server.py:18
GENERATE_KWARGS = dict(max_length=150, num_return_sequences=1, temperature=0.7, top_k=50, top_p=0.9,
                       repetition_penalty=1.2)  # REF/tutoring_server.py:23-28
FULL_METHOD = "/lms.Tutoring/GetLLMAnswer"


def generate_llm_response(query, model, tokenizer):
    """REF/tutoring_server.py:15-31 with `model` / `tokenizer` as arguments instead of module globals."""
    prompt = PROMPT_TEMPLATE.format(query=query)
    inputs = tokenizer.encode(prompt, return_tensors="pt")
    outputs = model.generate(inputs, **GENERATE_KWARGS)
    return tokenizer.decode(outputs[0], skip_special_tokens=True)


def install(tutoring_module, engine):
    """Swap the reference module's global `model` (REF/tutoring_server.py:11) for the engine; everything else in the
    module - the prompt, the kwargs, the servicer - keeps running unmodified."""
    tutoring_module.model = engine
    return tutoring_module


def _messages():
    """QueryRequest{string token=1; string query=2} / QueryResponse{bool success=1; string response=2}."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    pool = descriptor_pool.DescriptorPool()
    fd = descriptor_pb2.FileDescriptorProto(name="tutor_b200_lms_subset.proto", package="lms", syntax="proto3")
    F = descriptor_pb2.FieldDescriptorProto
    req = fd.message_type.add(name="QueryRequest")
    req.field.add(name="token", number=1, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
    req.field.add(name="query", number=2, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
    resp = fd.message_type.add(name="QueryResponse")
    resp.field.add(name="success", number=1, type=F.TYPE_BOOL, label=F.LABEL_OPTIONAL)
    resp.field.add(name="response", number=2, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
    pool.Add(fd)
    get = message_factory.GetMessageClass
    return get(pool.FindMessageTypeByName("lms.QueryRequest")), get(pool.FindMessageTypeByName("lms.QueryResponse"))


QueryRequest, QueryResponse = _messages()


class TutoringService:
    """Same contract as REF/tutoring_server.py:33-37: unwrap the query, generate, always success=True; exceptions
    propagate (gRPC reports StatusCode.UNKNOWN, as with the reference)."""

    def __init__(self, model, tokenizer):
        self.model, self.tokenizer = model, tokenizer

    def GetLLMAnswer(self, request, context):
        response = generate_llm_response(request.query, self.model, self.tokenizer)
        return QueryResponse(success=True, response=response)


def add_to_server(servicer, server):
    import grpc

    handler = grpc.unary_unary_rpc_method_handler(servicer.GetLLMAnswer, request_deserializer=QueryRequest.FromString,
                                                  response_serializer=QueryResponse.SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler("lms.Tutoring", {"GetLLMAnswer": handler}),))


def serve(model, tokenizer, address="[::]:50054", max_workers=10, batching=True, max_batch=64, max_wait_ms=4.0):
    """REF/tutoring_server.py:39-49 (same port, same 10-thread pool); returns the started server.

    The engine serialises its callers (one generation at a time per engine instance), so with the reference's ten handler
    threads each request would wait for the ones before it at ~20 ms apiece.  By default the model is therefore wrapped in
    a BatchingModel: concurrent requests run as ONE left-padded batch and each caller gets exactly its batch-1 result.
    `batching=False` passes the model through untouched (e.g. when the caller already wrapped it)."""
    import grpc

    if batching and not isinstance(model, BatchingModel) and hasattr(model, "_resolve"):
        model = BatchingModel(model, max_batch=max_batch, max_wait_ms=max_wait_ms)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    add_to_server(TutoringService(model, tokenizer), server)
    port = server.add_insecure_port(address)
    server.start()
    server.bound_port = port
    server.model = model
    return server


def call(address, token, query, timeout=60.0):
    """What `tutoring_stub.GetLLMAnswer(lms_pb2.QueryRequest(token=..., query=...))` (REF/lms_server.py:1273) sends."""
    import grpc

    with grpc.insecure_channel(address) as channel:
        rpc = channel.unary_unary(FULL_METHOD, request_serializer=QueryRequest.SerializeToString,
                                  response_deserializer=QueryResponse.FromString)
        return rpc(QueryRequest(token=token, query=query), timeout=timeout)


# ------------------------------------------------------------------------------------------------------------
# the product's real traffic is batch-1 requests on 10 gRPC handler threads
# (REF/tutoring_server.py:40).  BatchingModel coalesces concurrent `generate` calls into one left-padded batch so the
# batched decode kernels see a batch, and hands every caller exactly what its own batch-1 call would have returned.
import threading
import time
from collections import OrderedDict


class BatchingModel:
    """Drop-in for the tutoring server's `model`: `generate(inputs[1,T], **kwargs)` from many threads.

    Requests that arrive within `max_wait_ms` of each other (and share repetition_penalty / eos / pad) run as one
    batch: rows are left-padded to the longest prompt, the engine masks the pad slots and numbers positions from the
    first real token (tutor_gpt2_generate_padded), and each caller gets `[1, its own length]` back - its prompt, its
    own `max_length` budget, cut after its first EOS.  Pad slots are filled with the row's own first token, so the
    repetition penalty sees exactly the ids a batch-1 call would see.
    """

    def __init__(self, engine, max_batch: int = 64, max_wait_ms: float = 4.0):
        self.engine, self.max_batch, self.max_wait = engine, max_batch, max_wait_ms / 1e3
        self._cv = threading.Condition()
        self._queue = []
        self._stop = False
        self.batches = []  # sizes of the batches actually run (observability / tests)
        self._worker = threading.Thread(target=self._run, daemon=True)
        self._worker.start()

    def close(self):
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        self._worker.join(timeout=5)

    def generate(self, inputs, max_length=None, max_new_tokens=None, repetition_penalty=1.0, eos_token_id=None,
                 pad_token_id=None, **kwargs):
        if inputs.dim() != 2 or inputs.shape[0] != 1:
            return self.engine.generate(inputs, max_length=max_length, max_new_tokens=max_new_tokens,
                                        repetition_penalty=repetition_penalty, eos_token_id=eos_token_id,
                                        pad_token_id=pad_token_id, **kwargs)
        T = inputs.shape[1]
        total, eos, pad = self.engine._resolve(T, max_length, max_new_tokens, eos_token_id, pad_token_id)  # may raise
        req = dict(ids=inputs[0].to("cpu"), new=total - T, key=(float(repetition_penalty), eos, pad),
                   done=threading.Event(), out=None, err=None)
        with self._cv:
            self._queue.append(req)
            self._cv.notify_all()
        req["done"].wait()
        if req["err"] is not None:
            raise req["err"]
        return req["out"]

    def _run(self):
        import torch
        while True:
            with self._cv:
                while not self._queue and not self._stop:
                    self._cv.wait()
                if self._stop and not self._queue:
                    return
                deadline = time.monotonic() + self.max_wait
                while len(self._queue) < self.max_batch and time.monotonic() < deadline and not self._stop:
                    self._cv.wait(timeout=max(0.0, deadline - time.monotonic()))
                key = self._queue[0]["key"]
                # a padded batch runs max(T) + max(new) positions for every row: admit a request only while that
                # stays inside the model's position table, so rows that are legal on their own never fail together
                limit = getattr(self.engine, "n_positions", None)
                batch, bt, bn = [], 0, 0
                for r in self._queue:
                    if r["key"] != key or len(batch) >= self.max_batch:
                        continue
                    t, n = max(bt, r["ids"].numel()), max(bn, r["new"])
                    if batch and limit is not None and t + n > limit + 1:
                        continue
                    batch.append(r)
                    bt, bn = t, n
                self._queue = [r for r in self._queue if all(r is not b for b in batch)]
            try:
                pen, eos, pad = key
                T = max(r["ids"].numel() for r in batch)
                new = max(r["new"] for r in batch)
                ids = torch.empty((len(batch), T), dtype=torch.int64)
                mask = torch.ones((len(batch), T), dtype=torch.int64)
                for b, r in enumerate(batch):
                    p = T - r["ids"].numel()
                    ids[b, :p] = r["ids"][0]  # pad with the row's own first token: invisible to the repetition penalty
                    ids[b, p:] = r["ids"]
                    mask[b, :p] = 0
                out = self.engine.generate(ids, attention_mask=mask, max_new_tokens=new, repetition_penalty=pen,
                                           eos_token_id=None if eos < 0 else eos, pad_token_id=pad)
                self.batches.append(len(batch))
                for b, r in enumerate(batch):
                    p, t = T - r["ids"].numel(), r["ids"].numel()
                    row = out[b, p:min(out.shape[1], T + r["new"])]
                    gen = row[t:]
                    hit = (gen == eos).nonzero()
                    if eos >= 0 and hit.numel():
                        row = row[:t + int(hit[0]) + 1]  # a batch-1 call stops right after its first EOS
                    r["out"] = row.clone()[None]
            except Exception as e:  # every waiting caller sees the failure, like a batch-1 call would
                for r in batch:
                    r["err"] = e
            for r in batch:
                r["done"].set()


# ------------------------------------------------------------------------------------------------------------
# the reference reloads BERT from disk and re-embeds the (unchanged) assignment text on every
# request (REF/lms_server.py:1258-1263).  RelevanceGate loads once and caches assignment embeddings.
class RelevanceGate:
    """`check(query, assignment_text)` = REF/lms_server.py:1262-1269 with one BERT forward per request instead of a
    model load plus two forwards.  Same functions underneath (compute_embeddings, cosine_similarity), same 0.6."""

    def __init__(self, model, tokenizer, threshold: float = 0.6, cache_size: int = 1024):
        from . import bert as _bert
        self._bert = _bert
        self.model, self.tokenizer, self.threshold = model, tokenizer, threshold
        self._cache, self._cache_size, self._lock = OrderedDict(), cache_size, threading.Lock()
        self.hits = self.misses = 0

    def assignment_embedding(self, text):
        with self._lock:
            if text in self._cache:
                self._cache.move_to_end(text)
                self.hits += 1
                return self._cache[text]
        emb = self._bert.compute_embeddings(text, self.model, self.tokenizer)
        with self._lock:
            self.misses += 1
            self._cache[text] = emb
            while len(self._cache) > self._cache_size:
                self._cache.popitem(last=False)
        return emb

    def check(self, query, assignment_text):
        q = self._bert.compute_embeddings(query, self.model, self.tokenizer)
        similarity = self._bert.cosine_similarity(q, self.assignment_embedding(assignment_text))
        return (not (similarity < self.threshold)), similarity

    def answer(self, query, assignment_text, ask_tutor):
        """The tail of LMSService.GetLLMAnswer (REF/lms_server.py:1266-1274): canned reply or forward to the tutor."""
        relevant, _ = self.check(query, assignment_text)
        return ask_tutor(query) if relevant else self._bert.NOT_RELATED_MESSAGE


# ------------------------------------------------------------------------------------------------------------
# assignment ingestion -> chunk -> embed -> corpus.  The reference extracts the text of an
# uploaded assignment (REF/lms_server.py:21-27, 911-927) but only ever looks at the first 512 tokens of the student's
# first assignment (:1254, :98).  AssignmentIndex embeds the whole text in 512-token chunks at upload time and answers
# relevance queries against all of them with the corpus scan (configs[4]).
class AssignmentIndex:
    """Chunked assignment embeddings in HBM.

    A chunk is what `compute_embeddings` feeds BERT for a text that fills its 512-token window: [CLS] + up to
    `chunk_tokens - 2` word pieces + [SEP], mean-pooled over every position (REF/lms_server.py:97-101).  Chunk 0 of a
    text is therefore exactly the embedding the reference computes for it.  Equal-length chunks are embedded as one
    batch; rows live in a `CorpusShard` that is rebuilt lazily after an upload."""

    def __init__(self, model, tokenizer, chunk_tokens: int = 512, embed_batch: int = 32):
        if chunk_tokens < 3:
            raise ValueError("chunk_tokens must leave room for [CLS], one word piece and [SEP]")
        self.model, self.tokenizer = model, tokenizer
        self.chunk_tokens, self.embed_batch = chunk_tokens, embed_batch
        self.meta = []       # row -> (doc_id, chunk number)
        self._rows = []      # list of [n, hidden] CUDA tensors, in row order
        self._shard = None
        self._lock = threading.Lock()

    def chunk_ids(self, text):
        """Token ids of `text` without truncation, cut into [CLS] body [SEP] windows of at most chunk_tokens ids."""
        import torch
        enc = self.tokenizer(text, return_tensors="pt", padding=False, truncation=False)
        ids = enc["input_ids"][0].to(torch.int64)
        if ids.numel() < 3:
            raise ValueError("the tokenizer must frame the text with [CLS] ... [SEP]")
        cls, sep, body = ids[:1], ids[-1:], ids[1:-1]
        step = self.chunk_tokens - 2
        return [torch.cat([cls, body[i:i + step], sep]) for i in range(0, max(1, body.numel()), step)]

    def ingest(self, doc_id, text) -> int:
        """Embed every chunk of `text`; returns the number of chunks added."""
        import torch
        chunks = self.chunk_ids(text)
        embs = [None] * len(chunks)
        by_len = {}
        for i, c in enumerate(chunks):
            by_len.setdefault(c.numel(), []).append(i)
        for _, idxs in by_len.items():  # equal-length chunks share a batch (no padding: the mean is unmasked)
            for j in range(0, len(idxs), self.embed_batch):
                part = idxs[j:j + self.embed_batch]
                e = self.model.embed(torch.stack([chunks[i] for i in part]).cuda())
                for row, i in enumerate(part):
                    embs[i] = e[row]
        rows = torch.stack(embs).float().cuda()
        with self._lock:
            self.meta.extend((doc_id, n) for n in range(len(chunks)))
            self._rows.append(rows)
            self._shard = None
        return len(chunks)

    def __len__(self):
        return len(self.meta)

    def _corpus(self):
        import torch
        from .relevance import CorpusShard
        with self._lock:
            if self._shard is None:
                if not self._rows:
                    raise RuntimeError("AssignmentIndex is empty")
                self._rows = [torch.cat(self._rows)]
                self._shard = CorpusShard(self._rows[0])
            return self._shard

    def search(self, query, k: int = 5):
        """[(doc_id, chunk number, cosine)] of the k chunks most similar to `query`, best first."""
        from . import bert as _bert
        q = _bert.compute_embeddings(query, self.model, self.tokenizer)
        scores, idx = self._corpus().topk(q.cuda(), min(k, len(self.meta)))
        return [(*self.meta[int(i)], float(s)) for s, i in zip(scores[0].tolist(), idx[0].tolist()) if i >= 0]

    def is_relevant(self, query, threshold: float = 0.6):
        """The gate of REF/lms_server.py:1266-1269 against the best-matching chunk instead of the first 512 tokens."""
        best = self.search(query, 1)
        return bool(best) and not (best[0][2] < threshold), (best[0] if best else None)
