// file_server_shim.cc — the reference's `file_server` program (src/file_server.cc) over FileServerCore
// (include/sl_roles.h): file 0 restated byte for byte, further files (training shards) by Push.file_num.
// See worker_shim.cc for how this file is built and checked.
#if __has_include(<grpcpp/grpcpp.h>)
#include <grpcpp/grpcpp.h>

#include <iostream>
#include <memory>
#include <string>

#include "serverless_learn.grpc.pb.h"
#include "sl_roles.h"

using grpc::ClientContext;
using grpc::ClientWriter;
using grpc::ServerBuilder;
using grpc::ServerContext;
using grpc::Status;
using serverless_learn::Chunk;
using serverless_learn::Empty;
using serverless_learn::FileServer;
using serverless_learn::LoadFeedback;
using serverless_learn::Push;
using serverless_learn::PushOutcome;
using serverless_learn::ReceiveFileAck;
using serverless_learn::Worker;

#ifndef FILE_SERVER_ADDR
#define FILE_SERVER_ADDR "localhost:50053"   // src/serverless_learn.h:8
#endif

static slr_file_server* g_fs = nullptr;

class FileServerImpl final : public FileServer::Service {
  public:
    // src/file_server.cc:103-119 with the client-streaming loop of WorkerStub::ReceiveFile (:60-87)
    Status DoPush(ServerContext*, const Push* push, PushOutcome* outcome) override {
        if (slr_file_server_file_length_of(g_fs, push->file_num()) < 0) {
            std::cerr << "unknown file number" << std::endl;      // the reference exits here (:107-110)
            outcome->set_ok(false);
            return Status::OK;
        }
        auto worker = Worker::NewStub(grpc::CreateChannel(push->recipient_addr(), grpc::InsecureChannelCredentials()));
        ClientContext context;
        ReceiveFileAck ack;
        std::unique_ptr<ClientWriter<Chunk>> writer(worker->ReceiveFile(&context, &ack));
        bool ok = true;
        for (long long k = 0; ok; ++k) {
            const void* data = nullptr;
            const long long len = slr_file_server_chunk_of(g_fs, push->file_num(), k, &data);
            if (len <= 0) break;                                    // past the end
            Chunk chunk;
            chunk.set_data(data, (size_t)len);                      // :70-71: 1,000,000-byte chunks
            ok = writer->Write(chunk);
        }
        writer->WritesDone();
        Status status = writer->Finish();
        outcome->set_ok(ok && status.ok() && ack.ok());
        return Status::OK;
    }
    // src/file_server.cc:122-130
    Status CheckUp(ServerContext*, const Empty*, LoadFeedback*) override {
        slr_file_server_check_up(g_fs);
        return Status::OK;
    }
};

int main() {
    if (slr_file_server_create(&g_fs) != 0) {        // builds file 0 exactly as src/file_server.cc:152-156 does
        std::cerr << "file server: out of memory" << std::endl;
        return 1;
    }
    FileServerImpl service;                          // src/file_server.cc:137-148
    ServerBuilder builder;
    builder.AddListeningPort(FILE_SERVER_ADDR, grpc::InsecureServerCredentials());
    builder.RegisterService(&service);
    std::unique_ptr<grpc::Server> server(builder.BuildAndStart());
    server->Wait();
    return 0;
}
#endif  // __has_include(<grpcpp/grpcpp.h>)
