// roles.cc — role cores (see roles.h) and their C ABI (include/sl_roles.h).
#include "roles.h"

#include <stdlib.h>
#include <string.h>

#include <new>

#include "../../include/sl_roles.h"
#include "wire.h"

namespace slr {

// ---------------------------------------------------------------------------------- worker
WorkerCore::WorkerCore(int device, const std::string& addr, int D, int F, int B, double eta,
                       double lambda)
    : addr_(addr), D_(D) {
    if (sl_create(device, D, F, B, eta, lambda, &h_) != 0) h_ = nullptr;
}
WorkerCore::~WorkerCore() { sl_destroy(h_); }

bool WorkerCore::ReceiveFileBegin() { return sl_shard_begin(h_, 0) == 0; }
bool WorkerCore::ReceiveFileChunk(const void* data, size_t n) {
    return sl_shard_append(h_, data, n) == 0;  // the reference's loop body is empty (:54-56)
}
bool WorkerCore::ReceiveFileEnd() { return sl_shard_end(h_, nullptr) == 0; }

void WorkerCore::CheckUp(const std::vector<std::string>& peer_addrs) {
    std::lock_guard<std::mutex> lk(peer_list_mutex_);  // :68
    peer_list_.clear();                                // :69
    for (const std::string& peer : peer_addrs) peer_list_.push_back(peer);  // :70-72
}
std::vector<std::string> WorkerCore::peers() {
    std::lock_guard<std::mutex> lk(peer_list_mutex_);
    return peer_list_;
}

bool WorkerCore::ExchangeUpdates(const double* delta, int n, double* reply, int n_out) {
    return sl_merge(h_, delta, n, reply, n_out) == 0;  // :85-96 in one kernel
}

int WorkerCore::GossipOnce(Directory& dir) {
    std::string lucky_worker_addr;
    int lucky_worker_index;
    {
        std::lock_guard<std::mutex> lk(peer_list_mutex_);  // :199
        if (peer_list_.empty()) return -1;  // R7: the reference computes rand() % 0 here
        lucky_worker_index = rand() % (int)peer_list_.size();  // :200
        lucky_worker_addr = peer_list_[lucky_worker_index];    // :201
    }
    WorkerCore* peer = dir.find(lucky_worker_addr);                // :210-211
    if (!peer) return -1;                                          // R7: unreachable partner
    if (peer != this) {
        // both replicas live in this process: the whole round (:205-215 here, :85-96 on the peer)
        // runs as one kernel over peer-mapped memory, bit-exact with the staged path below
        return sl_gossip_p2p(h_, peer->h_) == 0 ? lucky_worker_index : -2;
    }
    std::vector<double> update((size_t)D_), update_back((size_t)D_);
    if (sl_delta_snapshot(h_, update.data(), D_) != 0) return -2;  // :205-208
    // WorkerStub::ExchangeUpdates, :145-172.  The reference applies the reply before it looks
    // at the status (:155-166); here a failed call skips the apply (R7).
    if (!peer->ExchangeUpdates(update.data(), D_, update_back.data(), D_)) return -1;
    if (sl_apply_reply(h_, update_back.data(), D_) != 0) return -2;  // :161-162 and :215
    return lucky_worker_index;
}

bool WorkerCore::TrainTick() { return sl_train_tick(h_) == 0; }  // :225-227
bool WorkerCore::TrainSteps(int n_steps, double* loss_out) {
    return sl_train_steps(h_, n_steps, loss_out) == 0;
}

// ------------------------------------------------------------------------------- file server
FileServerCore::FileServerCore() {
    // std::independent_bits_engine<std::default_random_engine, CHAR_BIT, unsigned char> restated
    // for libstdc++ (:152-153): minstd_rand0 (x <- 16807 x mod 2^31-1, seed 1), one draw per byte,
    // u = x - 1, draws with u >= floor((2^31-2)/256)*256 are rejected, byte = u & 0xFF.
    file_.resize((size_t)kDummyFileLength);
    uint64_t x = 1;
    for (long long i = 0; i < kDummyFileLength;) {  // :154
        x = (16807u * x) % 2147483647u;
        const uint64_t u = x - 1;
        if (u >= 2147483392u) continue;
        file_[(size_t)i++] = (char)(u & 0xFFu);     // :155
    }
}

long long FileServerCore::chunk(long long k, const void** data) const {
    const long long pos = k * kChunkSize;  // pos advances by substr(pos, CHUNK_SIZE).length(), :70,75
    if (k < 0 || pos >= (long long)file_.size()) return 0;
    if (data) *data = file_.data() + pos;
    const long long left = (long long)file_.size() - pos;
    return left < kChunkSize ? left : kChunkSize;
}

namespace {
inline uint64_t splitmix64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
}  // namespace

bool FileServerCore::AddSyntheticShard(uint32_t file_num, uint64_t row0, uint64_t n_rows, uint32_t F,
                                       uint64_t seed, bool with_header) {
    if (file_num == 0 || n_rows == 0 || F == 0) return false;  // file 0 stays the reference's dummy
    const size_t hb = with_header ? sizeof(sl_shard_header) : 0;
    const size_t payload = (size_t)n_rows * ((size_t)F + 1) * sizeof(float);
    std::string bytes;
    bytes.resize(hb + payload);
    float* X = reinterpret_cast<float*>(&bytes[hb]);
    float* y = X + (size_t)n_rows * F;
    for (uint64_t r = 0; r < n_rows; ++r) {
        float* xr = X + (size_t)r * F;
        for (uint32_t c = 0; c < F; ++c) {
            const uint64_t u = splitmix64(seed ^ ((row0 + r) * (uint64_t)F + c));
            xr[c] = (float)(u >> 40) * (1.0f / 16777216.0f) * 2.0f - 1.0f;
        }
        y[r] = xr[0] > 0.0f ? 1.0f : 0.0f;
    }
    if (with_header) {
        sl_shard_header hd;
        memset(&hd, 0, sizeof(hd));
        memcpy(hd.magic, SL_SHARD_MAGIC, 8);
        hd.version = 1;
        hd.F = F;
        hd.n_rows = n_rows;
        hd.row0 = row0;
        hd.flags = SL_SHARD_HAS_SUMS;
        const unsigned char* p = reinterpret_cast<const unsigned char*>(&bytes[hb]);
        unsigned long long bs = 0, ws = 0;
        for (size_t i = 0; i < payload; ++i) {
            bs += p[i];
            ws += (unsigned long long)(i + 1) * p[i];
        }
        hd.byte_sum = bs;
        hd.weighted = ws;
        memcpy(&bytes[0], &hd, sizeof(hd));
    }
    more_files_[file_num] = std::move(bytes);
    return true;
}

const std::string* FileServerCore::find_file(uint32_t file_num) const {
    if (file_num == 0) return &file_;
    auto it = more_files_.find(file_num);
    return it == more_files_.end() ? nullptr : &it->second;
}

long long FileServerCore::chunk_of(uint32_t file_num, long long k, const void** data) const {
    const std::string* f = find_file(file_num);
    if (!f) return -1;
    const long long pos = k * kChunkSize;
    if (k < 0 || pos >= (long long)f->size()) return 0;
    if (data) *data = f->data() + pos;
    const long long left = (long long)f->size() - pos;
    return left < kChunkSize ? left : kChunkSize;
}

int FileServerCore::DoPush(Directory& dir, const std::string& recipient_addr, uint32_t file_num) {
    if (!find_file(file_num)) return -1;  // :107-110: the reference prints and calls exit(1)
    WorkerCore* worker = dir.find(recipient_addr);  // :111-113
    if (!worker) return 0;
    bool ok = worker->ReceiveFileBegin();           // WorkerStub::ReceiveFile, :60-87
    for (long long k = 0; ok; ++k) {
        const void* data = nullptr;
        const long long n = chunk_of(file_num, k, &data);
        if (n == 0) break;                          // :69 while (pos < file.length())
        ok = worker->ReceiveFileChunk(data, (size_t)n);  // :72 writer->Write(chunk)
    }
    ok = worker->ReceiveFileEnd() && ok;            // :77-79 status.ok() && ack.ok()
    return ok ? 1 : 0;
}

// ------------------------------------------------------------------------------------ master
MasterCore::MasterCore(int device, int D, int B, double eta, double lambda) : D_(D), lambda_(lambda) {
    if (sl_create(device, D, D, B, eta, lambda, &h_) != 0) h_ = nullptr;
}
MasterCore::~MasterCore() { sl_destroy(h_); }

bool MasterCore::RegisterBirth(const std::string& addr) {
    std::lock_guard<std::mutex> lk(workers_mutex_);  // :83
    workers_.push_back(addr);                        // :84-85
    return true;                                     // :88 ack->set_ok(true)
}
std::vector<std::string> MasterCore::workers() {
    std::lock_guard<std::mutex> lk(workers_mutex_);
    return workers_;
}
bool MasterCore::ExchangeUpdates(const double* delta, int n, double* reply, int n_out) {
    return sl_merge(h_, delta, n, reply, n_out) == 0;  // :99-110
}
int MasterCore::RequestPushes(FileServerCore& fs, Directory& dir) {
    const std::vector<std::string> worker_addrs = workers();  // :225-230, copied under the lock
    int ok = 0;
    for (const std::string& addr : worker_addrs)              // :231-234, sequential, file 0
        if (fs.DoPush(dir, addr, 0) == 1) ++ok;
    return ok;
}
int MasterCore::RequestPushesSharded(FileServerCore& fs, Directory& dir) {
    const std::vector<std::string> worker_addrs = workers();
    int ok = 0;
    uint32_t file_num = 1;
    for (const std::string& addr : worker_addrs)
        if (fs.DoPush(dir, addr, file_num++) == 1) ++ok;
    return ok;
}
int MasterCore::DoCheckups(FileServerCore& fs, Directory& dir) {
    fs.CheckUp();                                             // :247
    const std::vector<std::string> worker_addrs = workers();  // :249-254
    int reached = 0;
    std::vector<std::string> lost;
    for (const std::string& addr : worker_addrs) {            // :256-262
        WorkerCore* w = dir.find(addr);
        if (!w) {                                             // "checkup failed", ignored (:190-195)
            if (elastic_) {
                std::lock_guard<std::mutex> lk(workers_mutex_);
                if (++members_[addr].missed >= max_missed_) lost.push_back(addr);
            }
            continue;
        }
        if (elastic_) {
            std::lock_guard<std::mutex> lk(workers_mutex_);
            members_[addr].missed = 0;
        }
        w->CheckUp(worker_addrs);  // WorkerStub::CheckUp sends the full list, self included (:184-188)
        ++reached;
    }
    if (!lost.empty()) {  // n4: drop workers that stayed silent; survivors learn it at the next round
        std::lock_guard<std::mutex> lk(workers_mutex_);
        for (const std::string& addr : lost) {
            for (size_t i = 0; i < workers_.size();)
                if (workers_[i] == addr) workers_.erase(workers_.begin() + (long)i); else ++i;
            members_.erase(addr);
        }
    }
    return reached;
}

int MasterCore::ShardOf(const std::string& addr) {
    std::lock_guard<std::mutex> lk(workers_mutex_);
    auto it = members_.find(addr);
    return it == members_.end() ? 0 : (int)it->second.shard;
}

int MasterCore::AdmitPending(FileServerCore& fs, Directory& dir) {
    if (!elastic_) return -1;
    const double lambda = lambda_;
    if (lambda == 0.0) return -2;
    int admitted = 0;
    for (const std::string& addr : workers()) {
        uint32_t shard = 0;
        {
            std::lock_guard<std::mutex> lk(workers_mutex_);
            Member& m = members_[addr];
            if (m.admitted) continue;
            if (m.shard == 0) {  // lowest shard number nobody holds
                for (uint32_t cand = 1; m.shard == 0; ++cand) {
                    bool taken = false;
                    for (const auto& kv : members_) taken = taken || kv.second.shard == cand;
                    if (!taken) m.shard = cand;
                }
            }
            shard = m.shard;
        }
        WorkerCore* w = dir.find(addr);
        if (!w) continue;                                  // not reachable yet: stays pending
        if (fs.DoPush(dir, addr, shard) != 1) continue;    // unknown file / failed push: stays pending
        // the current model, through the worker's ExchangeUpdates handler
        std::vector<double> seed((size_t)D_), ignored((size_t)D_);
        if (sl_get_model(h_, seed.data(), D_) != 0) return -2;
        for (double& v : seed) v /= lambda;
        if (!w->ExchangeUpdates(seed.data(), D_, ignored.data(), D_)) continue;
        {
            std::lock_guard<std::mutex> lk(workers_mutex_);
            members_[addr].admitted = true;
        }
        ++admitted;
    }
    return admitted;
}

int MasterCore::GossipOnce(Directory& dir) {
    // periodically_send_updates, src/master.cc:268-293 (never started by the reference's main)
    const std::vector<std::string> worker_addrs = workers();
    if (worker_addrs.empty()) return -1;
    const int lucky_worker_index = rand() % (int)worker_addrs.size();  // :274
    WorkerCore* w = dir.find(worker_addrs[(size_t)lucky_worker_index]);
    if (!w) return -1;
    std::vector<double> update((size_t)D_), update_back((size_t)D_);
    if (sl_delta_snapshot(h_, update.data(), D_) != 0) return -2;      // :279-282
    if (!w->ExchangeUpdates(update.data(), D_, update_back.data(), D_)) return -1;
    if (sl_apply_reply(h_, update_back.data(), D_) != 0) return -2;    // :289 old = model
    return lucky_worker_index;
}

}  // namespace slr

// =====================================================================================
// C ABI
// =====================================================================================
struct slr_worker { slr::WorkerCore core; std::vector<std::string> peer_cache;
    slr_worker(int dev, const char* addr, int D, int F, int B, double eta, double lam)
        : core(dev, addr, D, F, B, eta, lam) {} };
struct slr_master { slr::MasterCore core; std::vector<std::string> addr_cache;
    slr_master(int dev, int D, int B, double eta, double lam) : core(dev, D, B, eta, lam) {} };
struct slr_file_server { slr::FileServerCore core; };
struct slr_directory { slr::Directory dir; };

extern "C" {

long slr_update_encode(const double* delta, int n, void* out, long cap) {
    return slw::update_encode(delta, n, out, cap);
}
int slr_update_decode(const void* buf, long len, double* out, int cap) {
    return slw::update_decode(buf, len, out, cap);
}

int slr_directory_create(slr_directory** out) {
    *out = new (std::nothrow) slr_directory();
    return *out ? 0 : SL_ENOMEM;
}
void slr_directory_destroy(slr_directory* d) { delete d; }
int slr_directory_add(slr_directory* d, const char* addr, slr_worker* w) {
    if (!d || !addr || !w) return SL_EINVAL;
    d->dir.add(addr, &w->core);
    return 0;
}

int slr_file_server_create(slr_file_server** out) {
    *out = new (std::nothrow) slr_file_server();
    return *out ? 0 : SL_ENOMEM;
}
void slr_file_server_destroy(slr_file_server* fs) { delete fs; }
long long slr_file_server_file_length(slr_file_server* fs) { return (long long)fs->core.file().size(); }
long long slr_file_server_chunk(slr_file_server* fs, long long k, const void** data) {
    return fs->core.chunk(k, data);
}
int slr_file_server_do_push(slr_file_server* fs, slr_directory* d, const char* recipient_addr,
                            unsigned file_num) {
    return fs->core.DoPush(d->dir, recipient_addr, file_num);
}
int slr_file_server_add_synthetic_shard(slr_file_server* fs, unsigned file_num, unsigned long long row0,
                                        unsigned long long n_rows, unsigned F, unsigned long long seed) {
    return fs->core.AddSyntheticShard(file_num, row0, n_rows, F, seed) ? 1 : 0;
}
int slr_file_server_add_described_shard(slr_file_server* fs, unsigned file_num, unsigned long long row0,
                                        unsigned long long n_rows, unsigned F, unsigned long long seed) {
    return fs && fs->core.AddSyntheticShard(file_num, row0, n_rows, F, seed, true) ? 1 : 0;
}
long long slr_file_server_file_length_of(slr_file_server* fs, unsigned file_num) {
    const std::string* f = fs->core.find_file(file_num);
    return f ? (long long)f->size() : -1;
}
long long slr_file_server_chunk_of(slr_file_server* fs, unsigned file_num, long long k, const void** data) {
    return fs->core.chunk_of(file_num, k, data);
}
int slr_master_request_pushes_sharded(slr_master* m, slr_file_server* fs, slr_directory* d) {
    return m->core.RequestPushesSharded(fs->core, d->dir);
}
int slr_file_server_check_up(slr_file_server* fs) { return fs->core.CheckUp() ? 1 : 0; }

int slr_master_create(int device, int D, int B, double eta, double lambda, slr_master** out) {
    *out = nullptr;
    slr_master* m = new (std::nothrow) slr_master(device, D, B, eta, lambda);
    if (!m) return SL_ENOMEM;
    if (!m->core.ok()) { delete m; return SL_ECUDA; }
    *out = m;
    return 0;
}
void slr_master_destroy(slr_master* m) { delete m; }
sl_handle* slr_master_handle(slr_master* m) { return m->core.handle(); }
int slr_master_register_birth(slr_master* m, const char* addr) { return m->core.RegisterBirth(addr) ? 1 : 0; }
int slr_master_worker_count(slr_master* m) { return (int)m->core.workers().size(); }
const char* slr_master_worker_addr(slr_master* m, int i) {
    m->addr_cache = m->core.workers();
    return (i >= 0 && i < (int)m->addr_cache.size()) ? m->addr_cache[(size_t)i].c_str() : nullptr;
}
int slr_master_exchange_updates(slr_master* m, const double* delta, int n, double* reply, int n_out) {
    return m->core.ExchangeUpdates(delta, n, reply, n_out) ? 1 : 0;
}
int slr_master_exchange_updates_wire(slr_master* m, const void* in, long in_len, void* out,
                                     long out_cap, long* out_len) {
    // what the gRPC shim does: parse Update, run the handler, serialise the reply
    std::vector<double> delta((size_t)(in_len / 8 + 1)), reply;
    const int n = slw::update_decode(in, in_len, delta.data(), (int)delta.size());
    if (n < 0) return 0;
    reply.resize((size_t)n > 0 ? (size_t)n : 1);
    if (!m->core.ExchangeUpdates(delta.data(), n, reply.data(), n)) return 0;
    const long w = slw::update_encode(reply.data(), n, out, out_cap);
    if (w < 0) return 0;
    if (out_len) *out_len = w;
    return 1;
}
int slr_master_request_pushes(slr_master* m, slr_file_server* fs, slr_directory* d) {
    return m->core.RequestPushes(fs->core, d->dir);
}
int slr_directory_remove(slr_directory* d, const char* addr) {
    if (!d || !addr) return 0;
    d->dir.remove(addr);
    return 1;
}
int slr_master_enable_elastic(slr_master* m, int max_missed) {
    if (!m) return 0;
    m->core.EnableElastic(max_missed);
    return 1;
}
int slr_master_admit_pending(slr_master* m, slr_file_server* fs, slr_directory* d) {
    return m->core.AdmitPending(fs->core, d->dir);
}
int slr_master_gossip_once(slr_master* m, slr_directory* d) { return m->core.GossipOnce(d->dir); }
int slr_master_shard_of(slr_master* m, const char* addr) { return addr ? m->core.ShardOf(addr) : 0; }
int slr_master_do_checkups(slr_master* m, slr_file_server* fs, slr_directory* d) {
    return m->core.DoCheckups(fs->core, d->dir);
}

int slr_worker_create(int device, const char* addr, int D, int F, int B, double eta, double lambda,
                      slr_worker** out) {
    *out = nullptr;
    slr_worker* w = new (std::nothrow) slr_worker(device, addr ? addr : "", D, F, B, eta, lambda);
    if (!w) return SL_ENOMEM;
    if (!w->core.ok()) { delete w; return SL_ECUDA; }
    *out = w;
    return 0;
}
void slr_worker_destroy(slr_worker* w) { delete w; }
sl_handle* slr_worker_handle(slr_worker* w) { return w->core.handle(); }
const char* slr_worker_addr(slr_worker* w) { return w->core.addr().c_str(); }
int slr_worker_receive_begin(slr_worker* w) { return w->core.ReceiveFileBegin() ? 1 : 0; }
int slr_worker_receive_chunk(slr_worker* w, const void* data, size_t n) {
    return w->core.ReceiveFileChunk(data, n) ? 1 : 0;
}
int slr_worker_receive_end(slr_worker* w) { return w->core.ReceiveFileEnd() ? 1 : 0; }
int slr_worker_check_up(slr_worker* w, const char* const* peer_addrs, int n) {
    std::vector<std::string> v;
    for (int i = 0; i < n; ++i) v.push_back(peer_addrs[i]);
    w->core.CheckUp(v);
    return 1;
}
int slr_worker_peer_count(slr_worker* w) { return (int)w->core.peers().size(); }
const char* slr_worker_peer_addr(slr_worker* w, int i) {
    w->peer_cache = w->core.peers();
    return (i >= 0 && i < (int)w->peer_cache.size()) ? w->peer_cache[(size_t)i].c_str() : nullptr;
}
int slr_worker_exchange_updates(slr_worker* w, const double* delta, int n, double* reply, int n_out) {
    return w->core.ExchangeUpdates(delta, n, reply, n_out) ? 1 : 0;
}
int slr_worker_gossip_once(slr_worker* w, slr_directory* d) { return w->core.GossipOnce(d->dir); }
int slr_worker_train_tick(slr_worker* w) { return w->core.TrainTick() ? 1 : 0; }
int slr_worker_train_steps(slr_worker* w, int n_steps, double* loss_out) {
    return w->core.TrainSteps(n_steps, loss_out) ? 1 : 0;
}

int slr_workers_train_steps_group(slr_worker* const* ws, int W, int n_steps, double* loss_out) {
    if (!ws || W < 1 || W > 8) return 0;
    sl_handle* hs[8];
    for (int k = 0; k < W; ++k) {
        if (!ws[k]) return 0;
        hs[k] = ws[k]->core.handle();
    }
    return sl_train_steps_group(hs, W, n_steps, loss_out) == 0 ? 1 : 0;
}
int slr_workers_train_steps_multi(slr_worker* const* ws, int W, int n_steps, double* loss_out) {
    if (!ws || W < 1 || W > 8) return 0;
    sl_handle* hs[8];
    for (int k = 0; k < W; ++k) {
        if (!ws[k]) return 0;
        hs[k] = ws[k]->core.handle();
    }
    return sl_train_steps_multi(hs, W, n_steps, loss_out) == 0 ? 1 : 0;
}
}  // extern "C"
