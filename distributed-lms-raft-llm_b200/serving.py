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


def serve(model, tokenizer, address="[::]:50054", max_workers=10):
    """REF/tutoring_server.py:39-49 (same port, same 10-thread pool); returns the started server."""
    import grpc

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    add_to_server(TutoringService(model, tokenizer), server)
    port = server.add_insecure_port(address)
    server.start()
    server.bound_port = port
    return server


def call(address, token, query, timeout=60.0):
    """What `tutoring_stub.GetLLMAnswer(lms_pb2.QueryRequest(token=..., query=...))` (REF/lms_server.py:1273) sends."""
    import grpc

    with grpc.insecure_channel(address) as channel:
        rpc = channel.unary_unary(FULL_METHOD, request_serializer=QueryRequest.SerializeToString,
                                  response_deserializer=QueryResponse.FromString)
        return rpc(QueryRequest(token=token, query=query), timeout=timeout)
