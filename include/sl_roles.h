/*
 * sl_roles.h — C ABI over the host-side role cores (WorkerCore, MasterCore, FileServerCore):
 * the reference's three programs with the gRPC types replaced by plain buffers, on top of
 * the GPU ABI of sl_gpu.h.  Each entry cites the handler / loop it mirrors (file:line in the
 * reference).  "addr" strings play the part of the reference's "host:port" addresses; an
 * in-process directory (slr_directory_*) stands in for grpc::CreateChannel(addr).
 * Return convention: handlers that answer with an `ok` bool in the reference return 1/0.
 */
#ifndef SL_ROLES_H
#define SL_ROLES_H

#include <stddef.h>

#include "sl_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct slr_worker slr_worker;
typedef struct slr_master slr_master;
typedef struct slr_file_server slr_file_server;
typedef struct slr_directory slr_directory;

/* ---- Update wire format (src/protos/serverless_learn.proto:81-83, proto3 packed) ------- */
long slr_update_encode(const double* delta, int n, void* out, long cap); /* bytes, -1 = cap */
int slr_update_decode(const void* buf, long len, double* out, int cap);  /* count, -1 = bad */

/* ---- directory: addr -> worker (stand-in for CreateChannel, src/worker.cc:210) ---------- */
int slr_directory_create(slr_directory** out);
void slr_directory_destroy(slr_directory* d);
int slr_directory_add(slr_directory* d, const char* addr, slr_worker* w);

/* ---- file server (src/file_server.cc) ---------------------------------------------------- */
/* builds file 0 exactly as main() does (:152-156): 100,000,000 bytes of
 * independent_bits_engine<default_random_engine, 8, unsigned char>, restated */
int slr_file_server_create(slr_file_server** out);
void slr_file_server_destroy(slr_file_server* fs);
long long slr_file_server_file_length(slr_file_server* fs);
/* chunk k of the stream WorkerStub::ReceiveFile writes (:60-87); returns its length (0 past
 * the end) and a pointer valid for the file server's lifetime */
long long slr_file_server_chunk(slr_file_server* fs, long long k, const void** data);
/* FileServerImpl::DoPush (:103-119): stream the file to the recipient.  Returns the
 * PushOutcome ok flag; file_num != 0 returns -1 (the reference calls exit(1), :107-110) */
int slr_file_server_do_push(slr_file_server* fs, slr_directory* d, const char* recipient_addr,
                            unsigned file_num);
int slr_file_server_check_up(slr_file_server* fs); /* :122-130, no-op, returns 1 */
/* Beyond the reference (it only knows file 0): further files addressed by Push.file_num
 * (proto:36-39).  File k > 0 = a training shard in the stream layout sl_shard_* expects, rows
 * [row0, row0+n_rows) of the synthetic matrix; unknown numbers keep the reference's error. */
int slr_file_server_add_synthetic_shard(slr_file_server* fs, unsigned file_num, unsigned long long row0,
                                        unsigned long long n_rows, unsigned F, unsigned long long seed);
/* Same file, preceded by the 64-byte sl_shard_header of sl_gpu.h (F, n_rows, row0, payload sums):
 * the receiving worker's sl_shard_end then checks what it got. */
int slr_file_server_add_described_shard(slr_file_server* fs, unsigned file_num, unsigned long long row0,
                                        unsigned long long n_rows, unsigned F, unsigned long long seed);
long long slr_file_server_file_length_of(slr_file_server* fs, unsigned file_num); /* -1 unknown */
long long slr_file_server_chunk_of(slr_file_server* fs, unsigned file_num, long long k,
                                   const void** data);

/* ---- master (src/master.cc) ------------------------------------------------------------- */
int slr_master_create(int device, int D, int B, double eta, double lambda, slr_master** out);
void slr_master_destroy(slr_master* m);
sl_handle* slr_master_handle(slr_master* m);
int slr_master_register_birth(slr_master* m, const char* addr);       /* :79-91 */
int slr_master_worker_count(slr_master* m);
const char* slr_master_worker_addr(slr_master* m, int i);
int slr_master_exchange_updates(slr_master* m, const double* delta, int n, double* reply,
                                int n_out);                            /* :95-114 */
/* the same handler fed with serialized Update messages (what a gRPC shim would see) */
int slr_master_exchange_updates_wire(slr_master* m, const void* in, long in_len, void* out,
                                     long out_cap, long* out_len);
/* one round of periodically_request_pushes (:220-237): file 0 to every registered worker, in
 * registration order; returns the number of successful pushes */
int slr_master_request_pushes(slr_master* m, slr_file_server* fs, slr_directory* d);
/* the same round with a shard assignment: the i-th registered worker gets file i+1 */
int slr_master_request_pushes_sharded(slr_master* m, slr_file_server* fs, slr_directory* d);
/* one round of periodically_do_checkups (:240-266): every worker receives the full peer list */
int slr_master_do_checkups(slr_master* m, slr_file_server* fs, slr_directory* d);
/* Beyond the reference — elastic membership; off unless enabled, the calls
 * above keep the reference's behaviour (append-only list src/master.cc:84-85, failed check-ups
 * ignored src/master.cc:190-195, no model for a late joiner).
 *  enable_elastic : a worker that fails `max_missed` check-up rounds in a row is dropped and its
 *                   shard number freed.
 *  admit_pending  : every worker not admitted yet gets the lowest free shard (file_num >= 1,
 *                   pushed once, src/file_server.cc:103-119 path) and the master's current model
 *                   through Worker.ExchangeUpdates (src/worker.cc:81-100) with delta = model/lambda,
 *                   reply ignored.  Returns the number admitted (< 0: not enabled / error).
 *  gossip_once    : the master-side initiator the reference holds as dead code
 *                   (periodically_send_updates, src/master.cc:268-293), one round.
 *  shard_of       : shard number the master assigned to addr (0: none).
 *  slr_directory_remove : the worker's server is gone; calls to addr fail from now on. */
int slr_master_enable_elastic(slr_master* m, int max_missed);
int slr_master_admit_pending(slr_master* m, slr_file_server* fs, slr_directory* d);
int slr_master_gossip_once(slr_master* m, slr_directory* d);
int slr_master_shard_of(slr_master* m, const char* addr);
int slr_directory_remove(slr_directory* d, const char* addr);

/* ---- worker (src/worker.cc) --------------------------------------------------------------- */
int slr_worker_create(int device, const char* addr, int D, int F, int B, double eta, double lambda,
                      slr_worker** out);
void slr_worker_destroy(slr_worker* w);
sl_handle* slr_worker_handle(slr_worker* w);
const char* slr_worker_addr(slr_worker* w);
/* ReceiveFile (:49-61) as begin / one call per Chunk / end; end returns the ack's ok */
int slr_worker_receive_begin(slr_worker* w);
int slr_worker_receive_chunk(slr_worker* w, const void* data, size_t n);
int slr_worker_receive_end(slr_worker* w);
int slr_worker_check_up(slr_worker* w, const char* const* peer_addrs, int n); /* :64-77 */
int slr_worker_peer_count(slr_worker* w);
const char* slr_worker_peer_addr(slr_worker* w, int i);
int slr_worker_exchange_updates(slr_worker* w, const double* delta, int n, double* reply,
                                int n_out);                                    /* :81-100 */
/* one iteration of periodically_gossip (:194-219) without the sleep: partner = rand() % N over
 * the peer list; returns the partner index, -1 when there is no peer or the partner cannot be
 * reached (repair R7: the round is skipped), -2 on a GPU error */
/* W workers of this process that share one GPU take n_steps synchronous steps as one job
 * (sl_train_steps_group: BASELINE config 2, "4 workers ... same 4-shard job on 1 B200").  1 = ok. */
int slr_workers_train_steps_group(slr_worker* const* ws, int W, int n_steps, double* loss_out);
/* The same W workers each take n_steps LOCAL steps side by side in one kernel (sl_train_steps_multi):
 * the bodies of W training threads (src/worker.cc:221-231) in one call.  loss_out: W*n_steps, worker-major. */
int slr_workers_train_steps_multi(slr_worker* const* ws, int W, int n_steps, double* loss_out);
int slr_worker_gossip_once(slr_worker* w, slr_directory* d);
/* one iteration of simulate_training (:221-231) without the sleep: the reference's +1 ... */
int slr_worker_train_tick(slr_worker* w);
/* ... and what takes its slot in this build */
int slr_worker_train_steps(slr_worker* w, int n_steps, double* loss_out);

#ifdef __cplusplus
}
#endif
#endif /* SL_ROLES_H */
