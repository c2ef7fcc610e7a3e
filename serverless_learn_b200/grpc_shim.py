"""gRPC front for the role cores, on the reference's own wire surface.

The image has no protoc / gRPC C++ toolchain, so the C++ shims of INTEGRATION.md cannot be
compiled here; this module stands in for them with grpcio: the message and service descriptors of
src/protos/serverless_learn.proto (package `serverless_learn`, 3 services, 6 RPCs, 11 messages —
proto:5-87) are built at run time, and every handler is one call into the C ABI
(libsl_roles.so / libsl_gpu.so), exactly as the one-line C++ shims would be.  Payload bytes
(`Update.delta` packed doubles, `Chunk.data`) go from the protobuf buffers to the ABI untouched.

Test / demonstration code: nothing here computes.
"""
from __future__ import annotations

import ctypes as C
from concurrent import futures

import grpc
import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from . import roles

_T = descriptor_pb2.FieldDescriptorProto


def _build_messages():
    fd = descriptor_pb2.FileDescriptorProto(name="serverless_learn_rt.proto", package="serverless_learn",
                                            syntax="proto3")

    def msg(name, *fields):
        m = fd.message_type.add(name=name)
        for fname, num, ftype, rep in fields:
            m.field.add(name=fname, number=num, type=ftype,
                        label=_T.LABEL_REPEATED if rep else _T.LABEL_OPTIONAL)

    msg("WorkerBirthInfo", ("addr", 1, _T.TYPE_STRING, False))            # proto:17-19
    msg("RegisterBirthAck", ("ok", 1, _T.TYPE_BOOL, False))               # proto:22-24
    msg("Push", ("recipient_addr", 1, _T.TYPE_STRING, False),             # proto:36-39
        ("file_num", 2, _T.TYPE_UINT32, False))
    msg("PushOutcome", ("ok", 1, _T.TYPE_BOOL, False))                    # proto:42-44
    msg("Chunk", ("data", 1, _T.TYPE_BYTES, False))                       # proto:59-61
    msg("ReceiveFileAck", ("ok", 1, _T.TYPE_BOOL, False))                 # proto:64-66
    msg("PeerList", ("peer_addrs", 1, _T.TYPE_STRING, True))              # proto:69-71
    msg("FlowFeedback")                                                   # proto:73-75
    msg("LoadFeedback")                                                   # proto:77-79
    msg("Update", ("delta", 1, _T.TYPE_DOUBLE, True))                     # proto:81-83
    msg("Empty")                                                          # proto:85-87
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {m.name: message_factory.GetMessageClass(pool.FindMessageTypeByName("serverless_learn." + m.name))
            for m in fd.message_type}


M = _build_messages()


def _method(service, name):
    return f"/serverless_learn.{service}/{name}"          # proto:5,8,27,47


def _unary(fn, req, resp):
    return grpc.unary_unary_rpc_method_handler(fn, request_deserializer=M[req].FromString,
                                               response_serializer=M[resp].SerializeToString)


def _delta_array(update) -> np.ndarray:
    return np.asarray(update.delta, dtype=np.float64)


# --------------------------------------------------------------------------------- servers
class WorkerService:
    """WorkerImpl (src/worker.cc:41-104) over a roles.Worker."""

    def __init__(self, worker: roles.Worker):
        self.w = worker

    def ReceiveFile(self, request_iterator, context):                      # src/worker.cc:49-61
        L, p = self.w.L, self.w.p
        ok = bool(L.slr_worker_receive_begin(p))
        for chunk in request_iterator:                                     # while (reader->Read(&chunk))
            data = chunk.data
            ok = bool(L.slr_worker_receive_chunk(p, data, len(data))) and ok
        ok = bool(L.slr_worker_receive_end(p)) and ok
        return M["ReceiveFileAck"](ok=ok)                                  # always Status::OK

    def CheckUp(self, request, context):                                   # src/worker.cc:64-77
        self.w.check_up(list(request.peer_addrs))
        return M["FlowFeedback"]()

    def ExchangeUpdates(self, request, context):                           # src/worker.cc:81-100
        reply = self.w.exchange_updates(_delta_array(request))
        return M["Update"](delta=reply if reply is not None else [])

    def handler(self):
        return grpc.method_handlers_generic_handler("serverless_learn.Worker", {
            "ReceiveFile": grpc.stream_unary_rpc_method_handler(
                self.ReceiveFile, request_deserializer=M["Chunk"].FromString,
                response_serializer=M["ReceiveFileAck"].SerializeToString),
            "CheckUp": _unary(self.CheckUp, "PeerList", "FlowFeedback"),
            "ExchangeUpdates": _unary(self.ExchangeUpdates, "Update", "Update"),
        })


class MasterService:
    """MasterImpl (src/master.cc:70-118) over a roles.Master."""

    def __init__(self, master: roles.Master):
        self.m = master

    def RegisterBirth(self, request, context):                             # src/master.cc:79-91
        return M["RegisterBirthAck"](ok=self.m.register_birth(request.addr))

    def ExchangeUpdates(self, request, context):                           # src/master.cc:95-114
        reply = self.m.exchange_updates(_delta_array(request))
        return M["Update"](delta=reply if reply is not None else [])

    def handler(self):
        return grpc.method_handlers_generic_handler("serverless_learn.Master", {
            "RegisterBirth": _unary(self.RegisterBirth, "WorkerBirthInfo", "RegisterBirthAck"),
            "ExchangeUpdates": _unary(self.ExchangeUpdates, "Update", "Update"),
        })


class FileServerService:
    """FileServerImpl (src/file_server.cc:95-134) over a roles.FileServer: DoPush opens a channel to
    the recipient and client-streams file 0 in CHUNK_SIZE pieces (src/file_server.cc:60-87)."""

    def __init__(self, fs: roles.FileServer):
        self.fs = fs

    def _chunks(self, file_num=0):
        k = 0
        while True:
            ch = self.fs.chunk_of(file_num, k)
            if ch.size == 0:
                return
            yield M["Chunk"](data=ch.tobytes())
            k += 1

    def DoPush(self, request, context):                                    # src/file_server.cc:103-119
        if self.fs.file_length_of(request.file_num) < 0:                   # reference: exit(1), :107-110
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, "received push with unknown file_number")
        with grpc.insecure_channel(request.recipient_addr, options=_CHANNEL_OPTS) as ch:
            call = ch.stream_unary(_method("Worker", "ReceiveFile"),
                                   request_serializer=M["Chunk"].SerializeToString,
                                   response_deserializer=M["ReceiveFileAck"].FromString)
            try:
                ack = call(self._chunks(request.file_num))
                ok = bool(ack.ok)
            except grpc.RpcError:
                ok = False
        return M["PushOutcome"](ok=ok)

    def CheckUp(self, request, context):                                   # src/file_server.cc:122-130
        return M["LoadFeedback"]()

    def handler(self):
        return grpc.method_handlers_generic_handler("serverless_learn.FileServer", {
            "DoPush": _unary(self.DoPush, "Push", "PushOutcome"),
            "CheckUp": _unary(self.CheckUp, "Empty", "LoadFeedback"),
        })


_CHANNEL_OPTS = [("grpc.max_receive_message_length", 16 << 20), ("grpc.max_send_message_length", 16 << 20)]


def serve(service, addr: str) -> grpc.Server:
    """ServerBuilder + insecure creds (src/master.cc:204-215 and twins)."""
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=_CHANNEL_OPTS)
    server.add_generic_rpc_handlers((service.handler(),))
    port = server.add_insecure_port(addr)
    if port == 0:
        raise RuntimeError(f"cannot listen on {addr}")
    server.start()
    return server


# ---------------------------------------------------------------------------------- clients
class Stub:
    """The reference's FileServerStub / WorkerStub / MasterStub calls, one channel per call as the
    reference does (src/worker.cc:210, src/master.cc:258,284)."""

    @staticmethod
    def _call(addr, service, method, req, resp_type):
        with grpc.insecure_channel(addr, options=_CHANNEL_OPTS) as ch:
            fn = ch.unary_unary(_method(service, method), request_serializer=type(req).SerializeToString,
                                response_deserializer=M[resp_type].FromString)
            return fn(req, timeout=120)

    @classmethod
    def register_birth(cls, master_addr, my_addr) -> bool:                 # src/worker.cc:117-129
        return cls._call(master_addr, "Master", "RegisterBirth", M["WorkerBirthInfo"](addr=my_addr),
                         "RegisterBirthAck").ok

    @classmethod
    def exchange_updates(cls, addr, service, delta) -> np.ndarray:         # src/worker.cc:145-172
        back = cls._call(addr, service, "ExchangeUpdates", M["Update"](delta=np.asarray(delta, dtype=np.float64)),
                         "Update")
        return _delta_array(back)

    @classmethod
    def do_push(cls, fs_addr, recipient_addr, file_num=0) -> bool:         # src/master.cc:131-144
        return cls._call(fs_addr, "FileServer", "DoPush",
                         M["Push"](recipient_addr=recipient_addr, file_num=file_num), "PushOutcome").ok

    @classmethod
    def check_up_worker(cls, addr, peers) -> None:                         # src/master.cc:177-196
        cls._call(addr, "Worker", "CheckUp", M["PeerList"](peer_addrs=list(peers)), "FlowFeedback")

    @classmethod
    def check_up_file_server(cls, addr) -> None:                           # src/master.cc:148-160
        cls._call(addr, "FileServer", "CheckUp", M["Empty"](), "LoadFeedback")
