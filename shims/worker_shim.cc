// worker_shim.cc — the reference's `worker` program (src/worker.cc) with every handler and loop body
// replaced by one call into the role cores (include/sl_roles.h) over the sm_100a library (include/sl_gpu.h).
// The proto, the RPC surface and the three threads of main() stay as they are (src/worker.cc:233-258).
//
// Needs gRPC and the protoc output of src/protos/serverless_learn.proto, which this repository's image does
// not have: the file is kept compilable (`g++ -fsyntax-only` against shims/standin/, tests/test_shims_syntax.py)
// and is built where gRPC exists with the reference's own Makefile flags plus
//   -I$(B200)/include -L$(B200)/serverless_learn_b200 -lsl_roles -lsl_gpu.
#if __has_include(<grpcpp/grpcpp.h>)
#include <grpcpp/grpcpp.h>

#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "serverless_learn.grpc.pb.h"
#include "sl_roles.h"

using grpc::ClientContext;
using grpc::ServerBuilder;
using grpc::ServerContext;
using grpc::ServerReader;
using grpc::Status;
using serverless_learn::Chunk;
using serverless_learn::FlowFeedback;
using serverless_learn::Master;
using serverless_learn::PeerList;
using serverless_learn::ReceiveFileAck;
using serverless_learn::RegisterBirthAck;
using serverless_learn::Update;       // repair R1: the reference's using-list omits it (src/worker.cc:26-33)
using serverless_learn::Worker;
using serverless_learn::WorkerBirthInfo;

#ifndef MASTER_ADDR
#define MASTER_ADDR "localhost:50052"   // src/serverless_learn.h:5
#endif
#ifndef GOSSIP_INTERVAL
#define GOSSIP_INTERVAL 5000            // src/serverless_learn.h:10
#endif
#ifndef SL_MODEL_LENGTH
#define SL_MODEL_LENGTH 1024            // repair R6: the reference's vectors can never grow from empty
#endif
#ifndef SL_BATCH
#define SL_BATCH 4096
#endif
#ifndef SL_STEPS_PER_CALL
#define SL_STEPS_PER_CALL 64
#endif

static slr_worker* g_worker = nullptr;  // takes the place of model_state / old_state / LEARN_RATE (repair R4)

// ---- optional: the gossip round as ONE kernel between worker processes of a box (SL_GOSSIP_P2P=1) -------------
// The proto stays as it is.  Control messages are Updates no model can produce: their first element is a NaN
// with a payload of our own.  {kSigAttach, lambda, 16 doubles = the 128 handle bytes} -> {kSigAttached, peer};
// {kSigRing, peer} -> {kSigDone} once the callee's kernel has finished the round in both replicas.  A peer that
// does not know them answers with a wrong-length Update and is gossiped with the reference's way from then on.
static const uint64_t kSigAttach = 0x7FF8534C50325001ull, kSigAttached = 0x7FF8534C50325002ull,
                      kSigRing = 0x7FF8534C50325003ull, kSigDone = 0x7FF8534C50325004ull;
static const double kGossipLambda = 0.5;  // LEARN_RATE, src/master.cc:60
static inline uint64_t bits_of(double v) { uint64_t b; std::memcpy(&b, &v, 8); return b; }
static inline double sig(uint64_t b) { double v; std::memcpy(&v, &b, 8); return v; }
static std::mutex g_attached_mu;
static std::map<std::string, int> g_attached;   // callee: handle bytes -> peer; initiator: address -> peer at that address (-1 = plain)

class WorkerImpl final : public Worker::Service {
  public:
    // src/worker.cc:49-61: every Chunk goes to the GPU instead of being dropped
    Status ReceiveFile(ServerContext*, ServerReader<Chunk>* reader, ReceiveFileAck* ack) override {
        Chunk chunk;
        bool ok = slr_worker_receive_begin(g_worker) == 1;
        while (reader->Read(&chunk)) ok = ok && slr_worker_receive_chunk(g_worker, chunk.data().data(), chunk.data().size()) == 1;
        ack->set_ok(slr_worker_receive_end(g_worker) == 1 && ok);
        return Status::OK;
    }
    // src/worker.cc:64-77
    Status CheckUp(ServerContext*, const PeerList* peer_list, FlowFeedback*) override {
        std::vector<const char*> addrs;
        for (int i = 0; i < peer_list->peer_addrs_size(); ++i) addrs.push_back(peer_list->peer_addrs(i).c_str());
        slr_worker_check_up(g_worker, addrs.data(), (int)addrs.size());
        return Status::OK;
    }
    // src/worker.cc:81-100: model += lambda*delta; reply = model - old; old = model — one kernel
    Status ExchangeUpdates(ServerContext*, const Update* update, Update* return_update) override {
        sl_handle* h = slr_worker_handle(g_worker);
        if (update->delta_size() == 18 && bits_of(update->delta(0)) == kSigAttach) {
            const std::string key(reinterpret_cast<const char*>(update->delta().data() + 2), 128);
            std::lock_guard<std::mutex> lk(g_attached_mu);
            auto it = g_attached.find(key);
            int peer = it == g_attached.end() ? -1 : it->second;
            if (peer < 0 && sl_state_attach(h, key.data(), SL_MODEL_LENGTH, update->delta(1), &peer) == 0) g_attached[key] = peer;
            if (peer >= 0) {
                return_update->add_delta(sig(kSigAttached));
                return_update->add_delta((double)peer);
            }
            return Status::OK;   // empty reply: the initiator keeps to host buffers
        }
        if (update->delta_size() == 2 && bits_of(update->delta(0)) == kSigRing) {
            if (sl_gossip_p2p_remote(h, (int)update->delta(1)) == 0) return_update->add_delta(sig(kSigDone));
            else std::cerr << sl_last_error(h) << std::endl;
            return Status::OK;
        }
        return_update->mutable_delta()->Resize(update->delta_size(), 0.0);
        if (slr_worker_exchange_updates(g_worker, update->delta().data(), update->delta_size(),
                                        return_update->mutable_delta()->mutable_data(), return_update->delta_size()) != 1) {
            std::cerr << sl_last_error(slr_worker_handle(g_worker)) << std::endl;
            return_update->clear_delta();
        }
        return Status::OK;  // the reference never returns a failing Status (src/worker.cc:60,76,99)
    }
};

// src/worker.cc:181-192
static void run_service(std::string addr) {
    WorkerImpl service;
    ServerBuilder builder;
    builder.AddListeningPort(addr, grpc::InsecureServerCredentials());
    builder.RegisterService(&service);
    std::unique_ptr<grpc::Server> server(builder.BuildAndStart());
    server->Wait();
}

// One round with the worker process at `addr` as a single kernel on ITS GPU (include/sl_gpu.h, sl_state_*).
// false = that peer cannot (other box, older program): the caller falls back to :205-215 through host buffers.
static bool gossip_p2p_round(sl_handle* h, const std::string& addr) {
    auto stub = Worker::NewStub(grpc::CreateChannel(addr, grpc::InsecureChannelCredentials()));
    int peer;
    {
        std::lock_guard<std::mutex> lk(g_attached_mu);
        auto it = g_attached.find(addr);
        if (it == g_attached.end()) {
            Update hello, answer;
            unsigned char hs[128];
            peer = -1;
            if (sl_state_export(h, hs) == 0) {
                hello.mutable_delta()->Resize(18, 0.0);
                hello.set_delta(0, sig(kSigAttach));
                hello.set_delta(1, kGossipLambda);
                std::memcpy(hello.mutable_delta()->mutable_data() + 2, hs, 128);
                ClientContext context;
                if (stub->ExchangeUpdates(&context, hello, &answer).ok() && answer.delta_size() == 2 &&
                    bits_of(answer.delta(0)) == kSigAttached)
                    peer = (int)answer.delta(1);
            }
            g_attached[addr] = peer;
        } else {
            peer = it->second;
        }
    }
    if (peer < 0) return false;
    Update ring, done;
    ring.add_delta(sig(kSigRing));
    ring.add_delta((double)peer);
    if (sl_state_lend(h) != 0) return false;     // my training thread waits from here (:221-231 beside :194-219)
    ClientContext context;
    const bool ok = stub->ExchangeUpdates(&context, ring, &done).ok() && done.delta_size() == 1 &&
                    bits_of(done.delta(0)) == kSigDone;
    sl_state_reclaim(h);
    return ok;
}

// src/worker.cc:194-219 with the stub of :145-172: partner = rand() % N over the CheckUp peer list
static void periodically_gossip() {
    const int D = SL_MODEL_LENGTH;
    sl_handle* h = slr_worker_handle(g_worker);
    const char* p2p_env = std::getenv("SL_GOSSIP_P2P");
    const bool p2p = p2p_env && p2p_env[0] == '1';
    while (true) {
        const int n = slr_worker_peer_count(g_worker);
        if (n > 0) {                                                          // repair R7: no peer, no round
            const char* lucky_addr = slr_worker_peer_addr(g_worker, rand() % n);  // :200-201
            if (p2p && gossip_p2p_round(h, lucky_addr)) {
                std::this_thread::sleep_for(std::chrono::milliseconds(GOSSIP_INTERVAL));
                continue;
            }
            Update update, back;
            update.mutable_delta()->Resize(D, 0.0);
            sl_delta_snapshot(h, update.mutable_delta()->mutable_data(), D);   // :205-208
            auto stub = Worker::NewStub(grpc::CreateChannel(lucky_addr, grpc::InsecureChannelCredentials()));  // :210
            ClientContext context;
            Status status = stub->ExchangeUpdates(&context, update, &back);    // :153
            if (status.ok() && back.delta_size() == D) sl_apply_reply(h, back.delta().data(), D);  // :161-162, :215
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(GOSSIP_INTERVAL));  // :217
    }
}

// src/worker.cc:221-231: the minibatch steps that take the slot of `model_state[i] += 1`
static void simulate_training() {
    std::vector<double> loss(SL_STEPS_PER_CALL);
    while (true) {
        if (slr_worker_train_steps(g_worker, SL_STEPS_PER_CALL, loss.data()) != 1)   // no shard yet, or an error
            std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
}

int main(int argc, char** argv) {
    if (argc != 2) {
        std::cerr << "usage: worker ADDR" << std::endl;  // src/worker.cc:235-238
        return 1;
    }
    std::string addr(argv[1]);
    if (slr_worker_create(/*device*/0, addr.c_str(), SL_MODEL_LENGTH, SL_MODEL_LENGTH, SL_BATCH, /*eta*/0.1,
                          /*lambda = LEARN_RATE, src/master.cc:60*/0.5, &g_worker) != 0) {
        std::cerr << sl_last_error(nullptr) << std::endl;
        return 1;
    }
    std::thread service_thread(run_service, addr);               // :242
    std::thread gossip_thread(periodically_gossip);              // :244
    std::thread simulated_training_thread(simulate_training);    // :246
    {   // :249-253
        auto master = Master::NewStub(grpc::CreateChannel(MASTER_ADDR, grpc::InsecureChannelCredentials()));
        WorkerBirthInfo birth;
        RegisterBirthAck ack;
        ClientContext context;
        birth.set_addr(addr);
        master->RegisterBirth(&context, birth, &ack);
    }
    service_thread.join();
    return 0;
}
#endif  // __has_include(<grpcpp/grpcpp.h>)
