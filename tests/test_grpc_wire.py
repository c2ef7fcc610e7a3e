"""The reference's wire surface (src/protos/serverless_learn.proto) served by the role cores
through grpcio (serverless_learn_b200/grpc_shim.py).  CPU part: message bytes and the file
server's 100-chunk client stream over a real localhost channel.  GPU part: a whole job —
RegisterBirth, CheckUp, DoPush -> ReceiveFile into HBM, ExchangeUpdates — over the wire."""
import socket

import grpc
import numpy as np
import pytest


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_update_message_bytes_match_the_oracle(orc):
    from serverless_learn_b200 import grpc_shim as g
    d = np.array([1.0, -2.5, 0.0])
    assert g.M["Update"](delta=d).SerializeToString() == orc.update_encode(d)       # B.5
    assert g.M["Update"]().SerializeToString() == b""
    big = np.random.default_rng(0).standard_normal(1024)
    raw = g.M["Update"](delta=big).SerializeToString()
    assert len(raw) == 8195 and np.array_equal(orc.update_decode(raw), big)
    assert g._method("Master", "ExchangeUpdates") == "/serverless_learn.Master/ExchangeUpdates"
    assert set(g.M) == {"WorkerBirthInfo", "RegisterBirthAck", "Push", "PushOutcome", "Chunk",
                        "ReceiveFileAck", "PeerList", "FlowFeedback", "LoadFeedback", "Update", "Empty"}


def test_file_server_streams_file0_over_grpc(orc, file0):
    """DoPush -> WorkerStub::ReceiveFile (src/file_server.cc:60-87,103-119) against a sink worker:
    100 Chunk messages of 1,000,000 bytes, byte-exact."""
    from serverless_learn_b200 import grpc_shim as g, roles
    got = {"n": 0, "lens": [], "hash": 0xcbf29ce484222325, "sum": 0}

    def receive_file(request_iterator, context):
        for chunk in request_iterator:
            b = np.frombuffer(chunk.data, dtype=np.uint8)
            got["n"] += 1
            got["lens"].append(b.size)
            got["sum"] += int(b.sum(dtype=np.uint64))
            got.setdefault("parts", []).append(b)
        return g.M["ReceiveFileAck"](ok=True)

    class Sink:
        def handler(self):
            return grpc.method_handlers_generic_handler("serverless_learn.Worker", {
                "ReceiveFile": grpc.stream_unary_rpc_method_handler(
                    receive_file, request_deserializer=g.M["Chunk"].FromString,
                    response_serializer=g.M["ReceiveFileAck"].SerializeToString)})

    waddr, faddr = f"127.0.0.1:{_free_port()}", f"127.0.0.1:{_free_port()}"
    fs = roles.FileServer()
    s1, s2 = g.serve(Sink(), waddr), g.serve(g.FileServerService(fs), faddr)
    try:
        assert g.Stub.do_push(faddr, waddr, 0) is True
        assert got["n"] == 100 and set(got["lens"]) == {1_000_000}
        assert got["sum"] == 12_749_417_848
        assert np.array_equal(np.concatenate(got["parts"]), file0[0])
        g.Stub.check_up_file_server(faddr)
        with pytest.raises(grpc.RpcError) as e:                 # reference: exit(1) on file_num != 0
            g.Stub.do_push(faddr, waddr, 1)
        assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
        assert g.Stub.do_push(faddr, f"127.0.0.1:{_free_port()}", 0) is False    # nobody listening
    finally:
        s1.stop(0)
        s2.stop(0)
        fs.close()


@pytest.mark.gpu
def test_job_over_the_wire(orc, file0):
    from serverless_learn_b200 import grpc_shim as g, roles
    import ctypes as C
    F = 1024
    maddr, faddr = f"127.0.0.1:{_free_port()}", f"127.0.0.1:{_free_port()}"
    waddrs = [f"127.0.0.1:{_free_port()}" for _ in range(2)]
    master, fs = roles.Master(F), roles.FileServer()
    workers = [roles.Worker(a, F, 4096) for a in waddrs]
    servers = [g.serve(g.MasterService(master), maddr), g.serve(g.FileServerService(fs), faddr)]
    servers += [g.serve(g.WorkerService(w), a) for w, a in zip(workers, waddrs)]
    try:
        for a in waddrs:                                            # src/worker.cc:249-252
            assert g.Stub.register_birth(maddr, a)
        assert master.workers() == waddrs
        for a in master.workers():                                  # src/master.cc:240-266
            g.Stub.check_up_worker(a, master.workers())
        assert workers[1].peers() == waddrs
        for a in master.workers():                                  # src/master.cc:231-235
            assert g.Stub.do_push(faddr, a, 0)
        L = roles._lib.load()
        for w in workers:
            s, ws = C.c_ulonglong(), C.c_ulonglong()
            assert L.sl_shard_checksum(w.handle, C.byref(s), C.byref(ws)) == 0
            assert s.value == 12_749_417_848 and ws.value == orc.weighted_sum(file0[0])
        # worker 0 ticks, gossips with worker 1 and with the master over the wire
        om = [np.zeros(F), np.zeros(F)]
        oo = [np.zeros(F), np.zeros(F)]
        mm, mo = np.zeros(F), np.zeros(F)
        L2 = roles._lib.load()
        for _ in range(3):
            workers[0].train_tick()
            orc.train_tick(om[0])
        dp = C.POINTER(C.c_double)
        delta = np.empty(F)
        assert L2.sl_delta_snapshot(workers[0].handle, delta.ctypes.data_as(dp), F) == 0
        assert np.array_equal(delta, orc.delta(om[0], oo[0]))
        back = g.Stub.exchange_updates(waddrs[1], "Worker", delta)  # src/worker.cc:145-172
        assert np.array_equal(back, orc.merge_callee(om[1], oo[1], delta))
        assert L2.sl_apply_reply(workers[0].handle, back.ctypes.data_as(dp), F) == 0
        orc.apply_reply(om[0], oo[0], back)
        delta = orc.delta(om[1], oo[1])                             # worker 1 -> master (proto:13)
        back = g.Stub.exchange_updates(maddr, "Master", delta)
        assert np.array_equal(back, orc.merge_callee(mm, mo, delta))
        assert np.array_equal(workers[0].model(), om[0]) and np.array_equal(workers[1].model(), om[1])
        assert np.array_equal(master.model(), mm)
        # length mismatch: handler still answers OK with an empty Update (ok=false convention)
        assert g.Stub.exchange_updates(maddr, "Master", np.zeros(7)).size == 0
    finally:
        for s in servers:
            s.stop(0)
        for w in workers:
            w.close()
        master.close()
        fs.close()
