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
#include <iostream>
#include <memory>
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

// src/worker.cc:194-219 with the stub of :145-172: partner = rand() % N over the CheckUp peer list
static void periodically_gossip() {
    const int D = SL_MODEL_LENGTH;
    sl_handle* h = slr_worker_handle(g_worker);
    while (true) {
        const int n = slr_worker_peer_count(g_worker);
        if (n > 0) {                                                          // repair R7: no peer, no round
            const char* lucky_addr = slr_worker_peer_addr(g_worker, rand() % n);  // :200-201
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
