// API-shaped stand-in for the protoc / grpc_cpp_plugin output of src/protos/serverless_learn.proto (the
// messages and services of proto:8-87), for `g++ -fsyntax-only` only — see grpcpp/grpcpp.h beside it.
#pragma once
#include <grpcpp/grpcpp.h>

#include <memory>
#include <string>
namespace google { namespace protobuf {
template <class T>
class RepeatedField {
  public:
    int size() const;
    const T* data() const;
    T* mutable_data();
    void Resize(int n, const T& v);
    void Clear();
};
}}  // namespace google::protobuf
namespace serverless_learn {
#define SL_STANDIN_OK_MSG(NAME) \
    class NAME {                \
      public:                   \
        bool ok() const;        \
        void set_ok(bool v);    \
    };
SL_STANDIN_OK_MSG(RegisterBirthAck)
SL_STANDIN_OK_MSG(PushOutcome)
SL_STANDIN_OK_MSG(ReceiveFileAck)
class Empty {};
class FlowFeedback {};
class LoadFeedback {};
class WorkerBirthInfo {
  public:
    const std::string& addr() const;
    void set_addr(const std::string& v);
};
class Push {
  public:
    const std::string& recipient_addr() const;
    void set_recipient_addr(const std::string& v);
    unsigned file_num() const;
    void set_file_num(unsigned v);
};
class Chunk {
  public:
    const std::string& data() const;
    void set_data(const void* p, size_t n);
};
class PeerList {
  public:
    int peer_addrs_size() const;
    const std::string& peer_addrs(int i) const;
    void add_peer_addrs(const std::string& v);
};
class Update {
  public:
    int delta_size() const;
    double delta(int i) const;
    void set_delta(int i, double v);
    void add_delta(double v);
    const google::protobuf::RepeatedField<double>& delta() const;
    google::protobuf::RepeatedField<double>* mutable_delta();
    void clear_delta();
};
class Master {
  public:
    class Stub {
      public:
        grpc::Status RegisterBirth(grpc::ClientContext*, const WorkerBirthInfo&, RegisterBirthAck*);
        grpc::Status ExchangeUpdates(grpc::ClientContext*, const Update&, Update*);
    };
    static std::unique_ptr<Stub> NewStub(const std::shared_ptr<grpc::Channel>&);
    class Service : public grpc::Service {
      public:
        virtual ~Service();
        virtual grpc::Status RegisterBirth(grpc::ServerContext*, const WorkerBirthInfo*, RegisterBirthAck*);
        virtual grpc::Status ExchangeUpdates(grpc::ServerContext*, const Update*, Update*);
    };
};
class FileServer {
  public:
    class Stub {
      public:
        grpc::Status DoPush(grpc::ClientContext*, const Push&, PushOutcome*);
        grpc::Status CheckUp(grpc::ClientContext*, const Empty&, LoadFeedback*);
    };
    static std::unique_ptr<Stub> NewStub(const std::shared_ptr<grpc::Channel>&);
    class Service : public grpc::Service {
      public:
        virtual ~Service();
        virtual grpc::Status DoPush(grpc::ServerContext*, const Push*, PushOutcome*);
        virtual grpc::Status CheckUp(grpc::ServerContext*, const Empty*, LoadFeedback*);
    };
};
class Worker {
  public:
    class Stub {
      public:
        std::unique_ptr<grpc::ClientWriter<Chunk>> ReceiveFile(grpc::ClientContext*, ReceiveFileAck*);
        grpc::Status CheckUp(grpc::ClientContext*, const PeerList&, FlowFeedback*);
        grpc::Status ExchangeUpdates(grpc::ClientContext*, const Update&, Update*);
    };
    static std::unique_ptr<Stub> NewStub(const std::shared_ptr<grpc::Channel>&);
    class Service : public grpc::Service {
      public:
        virtual ~Service();
        virtual grpc::Status ReceiveFile(grpc::ServerContext*, grpc::ServerReader<Chunk>*, ReceiveFileAck*);
        virtual grpc::Status CheckUp(grpc::ServerContext*, const PeerList*, FlowFeedback*);
        virtual grpc::Status ExchangeUpdates(grpc::ServerContext*, const Update*, Update*);
    };
};
}  // namespace serverless_learn
