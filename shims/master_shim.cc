// master_shim.cc — the reference's `master` program (src/master.cc): handlers as one-line calls into the role
// cores (include/sl_roles.h).  See worker_shim.cc for how this file is built and checked.
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
using grpc::Status;
using serverless_learn::Chunk;
using serverless_learn::Empty;
using serverless_learn::FileServer;
using serverless_learn::FlowFeedback;
using serverless_learn::LoadFeedback;
using serverless_learn::Master;
using serverless_learn::PeerList;
using serverless_learn::Push;
using serverless_learn::PushOutcome;
using serverless_learn::RegisterBirthAck;
using serverless_learn::Update;       // repair R1 (src/master.cc:30-40 omits it)
using serverless_learn::Worker;
using serverless_learn::WorkerBirthInfo;

#ifndef MASTER_ADDR
#define MASTER_ADDR "localhost:50052"        // src/serverless_learn.h:5
#endif
#ifndef FILE_SERVER_ADDR
#define FILE_SERVER_ADDR "localhost:50053"   // src/serverless_learn.h:8
#endif
#ifndef PUSH_INTERVAL
#define PUSH_INTERVAL 5000                   // src/master.cc:43
#endif
#ifndef SL_MODEL_LENGTH
#define SL_MODEL_LENGTH 1024
#endif

static slr_master* g_master = nullptr;  // takes the place of model_state / old_state / LEARN_RATE (src/master.cc:58-60)

class MasterImpl final : public Master::Service {
  public:
    // src/master.cc:79-91
    Status RegisterBirth(ServerContext*, const WorkerBirthInfo* birth, RegisterBirthAck* ack) override {
        ack->set_ok(slr_master_register_birth(g_master, birth->addr().c_str()) == 1);
        return Status::OK;
    }
    // src/master.cc:95-114
    Status ExchangeUpdates(ServerContext*, const Update* update, Update* return_update) override {
        return_update->mutable_delta()->Resize(update->delta_size(), 0.0);
        if (slr_master_exchange_updates(g_master, update->delta().data(), update->delta_size(),
                                        return_update->mutable_delta()->mutable_data(), return_update->delta_size()) != 1) {
            std::cerr << sl_last_error(slr_master_handle(g_master)) << std::endl;
            return_update->clear_delta();
        }
        return Status::OK;
    }
};

// src/master.cc:204-215
static void run_service() {
    MasterImpl service;
    ServerBuilder builder;
    builder.AddListeningPort(MASTER_ADDR, grpc::InsecureServerCredentials());
    builder.RegisterService(&service);
    std::unique_ptr<grpc::Server> server(builder.BuildAndStart());
    server->Wait();
}

// src/master.cc:220-237: file 0 to every registered worker, in registration order, through the file server
static void periodically_request_pushes() {
    auto fs = FileServer::NewStub(grpc::CreateChannel(FILE_SERVER_ADDR, grpc::InsecureChannelCredentials()));
    while (true) {
        const int n = slr_master_worker_count(g_master);
        for (int i = 0; i < n; ++i) {
            Push push;
            PushOutcome outcome;
            ClientContext context;
            push.set_recipient_addr(slr_master_worker_addr(g_master, i));
            push.set_file_num(0);                                   // :233 (a sharded job sends i + 1)
            fs->DoPush(&context, push, &outcome);
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(PUSH_INTERVAL));
    }
}

// src/master.cc:240-266: every worker receives the full peer list
static void periodically_do_checkups() {
    while (true) {
        const int n = slr_master_worker_count(g_master);
        PeerList peers;
        for (int i = 0; i < n; ++i) peers.add_peer_addrs(slr_master_worker_addr(g_master, i));
        for (int i = 0; i < n; ++i) {
            auto w = Worker::NewStub(grpc::CreateChannel(slr_master_worker_addr(g_master, i), grpc::InsecureChannelCredentials()));
            FlowFeedback fb;
            ClientContext context;
            w->CheckUp(&context, peers, &fb);                       // failures are ignored, as in :190-195
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(PUSH_INTERVAL));
    }
}

int main() {
    if (slr_master_create(/*device*/0, SL_MODEL_LENGTH, /*B*/4096, /*eta*/0.1, /*lambda*/0.5, &g_master) != 0) {
        std::cerr << sl_last_error(nullptr) << std::endl;
        return 1;
    }
    std::thread service_thread(run_service);                          // src/master.cc:296
    std::thread push_thread(periodically_request_pushes);             // :298
    std::thread checkup_thread(periodically_do_checkups);             // :300
    service_thread.join();                                            // :302
    return 0;
}
#endif  // __has_include(<grpcpp/grpcpp.h>)
