// roles.h — host-side role cores: the reference's three programs minus gRPC, over the C ABI
// of sl_gpu.h.  Same names, argument meaning and error behaviour as the reference's handlers
// and loops (file:line cited per member), so a gRPC shim is one line per handler.
#pragma once
#include <stdint.h>

#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/sl_gpu.h"

namespace slr {

class WorkerCore;

// addr -> worker; stands in for grpc::CreateChannel(addr) (src/worker.cc:210,
// src/master.cc:258,284, src/file_server.cc:111-113)
class Directory {
  public:
    void add(const std::string& addr, WorkerCore* w) {
        std::lock_guard<std::mutex> lk(mu_);
        map_[addr] = w;
    }
    WorkerCore* find(const std::string& addr) {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = map_.find(addr);
        return it == map_.end() ? nullptr : it->second;
    }
    void remove(const std::string& addr) {  // the worker's server went away: calls to it fail from now on
        std::lock_guard<std::mutex> lk(mu_);
        map_.erase(addr);
    }

  private:
    std::mutex mu_;
    std::map<std::string, WorkerCore*> map_;
};

// src/worker.cc
class WorkerCore {
  public:
    WorkerCore(int device, const std::string& addr, int D, int F, int B, double eta, double lambda);
    ~WorkerCore();
    bool ok() const { return h_ != nullptr; }
    sl_handle* handle() { return h_; }
    const std::string& addr() const { return addr_; }

    // WorkerImpl::ReceiveFile, src/worker.cc:49-61 — split at the reader->Read loop
    bool ReceiveFileBegin();
    bool ReceiveFileChunk(const void* data, size_t n);
    bool ReceiveFileEnd();  // ack->ok
    // WorkerImpl::CheckUp, src/worker.cc:64-77: replace the peer list
    void CheckUp(const std::vector<std::string>& peer_addrs);
    std::vector<std::string> peers();
    // WorkerImpl::ExchangeUpdates, src/worker.cc:81-100
    bool ExchangeUpdates(const double* delta, int n, double* reply, int n_out);
    // one iteration of periodically_gossip, src/worker.cc:194-219 (no sleep)
    int GossipOnce(Directory& dir);
    // one iteration of simulate_training, src/worker.cc:221-231 (no sleep)
    bool TrainTick();
    bool TrainSteps(int n_steps, double* loss_out);

  private:
    sl_handle* h_ = nullptr;
    std::string addr_;
    int D_;
    std::vector<std::string> peer_list_;  // src/worker.cc:35
    std::mutex peer_list_mutex_;          // src/worker.cc:37
};

// src/file_server.cc
class FileServerCore {
  public:
    static constexpr long long kDummyFileLength = 100LL * 1000 * 1000;  // :40
    static constexpr long long kChunkSize = 1LL * 1000 * 1000;          // :46
    FileServerCore();  // fills file 0, :152-156
    const std::string& file() const { return file_; }
    long long chunk(long long k, const void** data) const;  // :68-76
    // Beyond the reference: further files, addressed by Push.file_num
    // (proto:36-39).  File k > 0 is a training shard in the stream layout sl_shard_* expects
    // (all of X row-major fp32, then the labels): rows [row0, row0+n_rows) of the synthetic
    // matrix of SURVEY.md §8(d).  Unknown file numbers keep the reference's behaviour.
    // with_header: the file starts with the 64-byte sl_shard_header (sl_gpu.h) describing it
    bool AddSyntheticShard(uint32_t file_num, uint64_t row0, uint64_t n_rows, uint32_t F, uint64_t seed,
                           bool with_header = false);
    const std::string* find_file(uint32_t file_num) const;
    long long chunk_of(uint32_t file_num, long long k, const void** data) const;
    // FileServerImpl::DoPush :103-119 -> WorkerStub::ReceiveFile :60-87.  -1: file_num != 0
    int DoPush(Directory& dir, const std::string& recipient_addr, uint32_t file_num);
    bool CheckUp() const { return true; }  // :122-130

  private:
    std::string file_;  // :43
    std::map<uint32_t, std::string> more_files_;
};

// src/master.cc
class MasterCore {
  public:
    MasterCore(int device, int D, int B, double eta, double lambda);
    ~MasterCore();
    bool ok() const { return h_ != nullptr; }
    sl_handle* handle() { return h_; }
    bool RegisterBirth(const std::string& addr);  // :79-91, append-only, no de-duplication
    std::vector<std::string> workers();
    bool ExchangeUpdates(const double* delta, int n, double* reply, int n_out);  // :95-114
    int RequestPushes(FileServerCore& fs, Directory& dir);  // one round of :220-237
    // same round with a shard assignment instead of "file 0 to everyone" (§8f n2): the i-th
    // registered worker gets file i+1
    int RequestPushesSharded(FileServerCore& fs, Directory& dir);
    int DoCheckups(FileServerCore& fs, Directory& dir);     // one round of :240-266

    // ---- beyond the reference: elastic membership, off unless enabled ----
    // The reference never removes a worker (:84-85 only appends), pushes file 0 to everybody every
    // 5 s (:231-235) and gives a late joiner no model.  With `EnableElastic(k)`:
    //  * a worker that fails k check-ups in a row is dropped from the list (and from the peer lists
    //    the next check-up sends); its shard number returns to the free pool;
    //  * `AdmitPending` gives every not-yet-admitted worker the lowest free shard (file_num >= 1,
    //    pushed once through the file server) and the master's current model — over the existing
    //    wire surface: one Worker.ExchangeUpdates with delta = model/lambda, reply ignored, which
    //    leaves the joiner with model = old = the master's model (exact when lambda is a power of 2);
    //  * `GossipOnce` is the master-side initiator the reference has as dead code
    //    (periodically_send_updates, :268-293), so that the master's model follows the workers'.
    void EnableElastic(int max_missed) { elastic_ = true; max_missed_ = max_missed < 1 ? 1 : max_missed; }
    int AdmitPending(FileServerCore& fs, Directory& dir);  // returns the number admitted, <0: error
    int GossipOnce(Directory& dir);                         // partner index, -1 none/unreachable, -2 error
    int ShardOf(const std::string& addr);                   // 0 = none

  private:
    struct Member { uint32_t shard = 0; bool admitted = false; int missed = 0; };
    sl_handle* h_ = nullptr;
    int D_;
    std::vector<std::string> workers_;  // :63 (WorkerInfo holds only the address, :49-56)
    std::mutex workers_mutex_;          // :66
    double lambda_ = 0.5;
    bool elastic_ = false;
    int max_missed_ = 3;
    std::map<std::string, Member> members_;  // guarded by workers_mutex_
};

}  // namespace slr
