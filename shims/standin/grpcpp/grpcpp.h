// API-shaped stand-in for <grpcpp/grpcpp.h>: ONLY for `g++ -fsyntax-only` of shims/*.cc in images without gRPC
// (tests/test_shims_syntax.py).  Declarations only, shaped after the calls the reference makes
// (src/master.cc:14-25, src/worker.cc:14-25, src/file_server.cc:14-27); nothing here can be linked.
#pragma once
#include <memory>
#include <string>
namespace grpc {
class Status {
  public:
    static const Status& OK;
    bool ok() const;
};
class ServerContext {};
class ClientContext {};
class ChannelCredentials {};
class ServerCredentials {};
class Channel {};
class Service {};
class Server {
  public:
    void Wait();
};
std::shared_ptr<ChannelCredentials> InsecureChannelCredentials();
std::shared_ptr<ServerCredentials> InsecureServerCredentials();
std::shared_ptr<Channel> CreateChannel(const std::string& target, const std::shared_ptr<ChannelCredentials>& creds);
class ServerBuilder {
  public:
    ServerBuilder& AddListeningPort(const std::string& addr, std::shared_ptr<ServerCredentials> creds);
    ServerBuilder& RegisterService(Service* service);
    std::unique_ptr<Server> BuildAndStart();
};
template <class R>
class ServerReader {
  public:
    bool Read(R* msg);
};
template <class W>
class ClientWriter {
  public:
    bool Write(const W& msg);
    bool WritesDone();
    Status Finish();
};
}  // namespace grpc
