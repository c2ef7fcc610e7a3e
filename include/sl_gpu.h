/*
 * sl_gpu.h — C ABI of the B200 (sm_100a) hot path of serverless_learn.
 *
 * The reference (sheaconlon/serverless_learn) has no plugin / FFI layer: its
 * arithmetic sits inline in gRPC handlers and thread loops over file-scope
 * globals.  Each entry point below replaces one of those inline loops; the
 * comment on it cites the loop (file:line relative to the reference's root).
 * The outer surface — roles, RPCs and messages of
 * src/protos/serverless_learn.proto — is unchanged; INTEGRATION.md shows the
 * edits a maintainer makes in master.cc / worker.cc / file_server.cc.
 *
 * Conventions
 *   - plain C, opaque handle, plain pointers and sizes; no C++ or torch types.
 *   - every function returns 0 on success and a negative SL_E* code on failure;
 *     nothing throws across the boundary.  sl_last_error() gives the text.  The
 *     reference's handlers always return Status::OK and report failure in the
 *     ack's `ok` bool (src/master.cc:90,113; src/worker.cc:60,76,99): a shim maps
 *     rc != 0 to ok=false.
 *   - all `double*` / `void*` arguments are HOST pointers owned by the caller and
 *     used only for the duration of the call (they match
 *     RepeatedField<double>::data() / mutable_data() and Chunk::data()).
 *   - one handle = one model replica (one worker, or the master) on one GPU.
 *     Calls on one handle may come from any thread (the reference touches its
 *     state from >=3 threads with no lock, src/worker.cc:242-246); they are
 *     serialised on a per-handle mutex and one CUDA stream.
 *   - there is no CPU fallback: without a usable CUDA device sl_create fails.
 */
#ifndef SL_GPU_H
#define SL_GPU_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sl_handle sl_handle;

enum {
    SL_OK = 0,
    SL_EINVAL = -1,   /* bad argument (length mismatch, shape not supported, ...) */
    SL_ECUDA = -2,    /* a CUDA runtime call failed */
    SL_ESTATE = -3,   /* call out of order (train before a shard is sealed, ...) */
    SL_ENOMEM = -4,   /* host or device allocation failed */
    SL_ETIMEOUT = -5, /* an in-kernel wait gave up (exchange watchdog) */
    SL_ENCCL = -6     /* NCCL missing or a NCCL call failed */
};

/* ---- lifetime ----------------------------------------------------------------
 * Replaces the file-scope state `model_state`, `old_state`, `LEARN_RATE`
 * (src/master.cc:58-60; used but never declared in src/worker.cc).  D is the model
 * length — the reference's vectors start empty and can never grow,
 * so D is fixed here (repair R6).  F, B, eta configure the minibatch step; the
 * model is bias-free logistic regression so D must equal F.  lambda is
 * LEARN_RATE (0.5 in the reference).  Both vectors start at zero. */
int sl_create(int device, int D, int F, int B, double eta, double lambda, sl_handle** out);
int sl_destroy(sl_handle* h);
/* Text of the last failure on this handle (h == NULL: last sl_create failure). */
const char* sl_last_error(const sl_handle* h);

/* Tuning / introspection knobs (DESIGN.md lists the keys). */
int sl_set_option(sl_handle* h, const char* key, long long value);
long long sl_get_info(sl_handle* h, const char* key);

/* Diagnostics: after sl_set_option(h, "debug_timeline", s) the next persistent launch
 * records, for steps s..s+3 of that launch and every CTA, 8 (clock64, globaltimer) pairs
 * at fixed points of the step; this reads them back ([4][grid][16] u64).  No
 * counterpart in the reference. */
int sl_debug_read(sl_handle* h, unsigned long long* out, size_t n_words);

/* ---- shard ingest --------------------------------------------------------------
 * Replaces the body of Worker::ReceiveFile (src/worker.cc:49-61), whose loop reads
 * each Chunk and drops it.  Call sl_shard_begin once per stream, sl_shard_append
 * once per Chunk (any length, any alignment; rows may straddle calls) and
 * sl_shard_end when reader->Read returns false.  Bytes go through a pinned ring to
 * a resident HBM buffer; a page-locked source is DMA'd directly and asynchronously, and —
 * the one exception to "buffers are used only during the call" — must stay valid until
 * sl_shard_end returns.  The previously sealed shard stays trainable until sl_shard_end
 * swaps the new one in, so ingest overlaps training.  The byte stream is
 * kept verbatim: n_rows = floor(len / (4*(F+1))); the first n_rows*F fp32 are X
 * (row-major), the next n_rows fp32 are the labels.  size_hint = expected length
 * in bytes (0 = unknown, the buffer then grows by doubling).
 *
 * Self-describing streams (the reference's files carry no format at all,
 * src/file_server.cc:152-156): a stream that starts with the 64-byte sl_shard_header below is
 * checked at sl_shard_end — F must match the handle, 64 + n_rows*4*(F+1) must be the stream
 * length, and with SL_SHARD_HAS_SUMS the payload's byte sum and weighted sum (the
 * sl_shard_checksum quantities, over the bytes after the header) must match — and the payload
 * after the header is what gets trained on.  A mismatch fails sl_shard_end (SL_EINVAL) and keeps
 * the previously sealed shard.  Streams without the magic are taken verbatim as before.
 * A stream that does not hold one complete row never replaces a trainable shard: sl_shard_end
 * returns SL_OK with *n_rows = 0 and the previously sealed shard stays the one that is trained on
 * (on a handle that holds no rows yet, the bytes are kept as an untrainable shard of zero rows). */
#define SL_SHARD_MAGIC "SLSHARD\1"
#define SL_SHARD_HAS_SUMS 1u
typedef struct sl_shard_header {   /* little-endian, 64 bytes */
    char magic[8];                 /* SL_SHARD_MAGIC */
    unsigned int version;          /* 1 */
    unsigned int F;                /* features per row */
    unsigned long long n_rows;
    unsigned long long row0;       /* first row of the global data set held by this shard */
    unsigned int flags;            /* SL_SHARD_HAS_SUMS */
    unsigned int reserved0;
    unsigned long long byte_sum;   /* sum of payload bytes mod 2^64 */
    unsigned long long weighted;   /* sum of (i+1)*byte[i] mod 2^64, i from 0 at the payload start */
    unsigned long long reserved1;
} sl_shard_header;
int sl_shard_begin(sl_handle* h, size_t size_hint);
int sl_shard_append(sl_handle* h, const void* bytes, size_t n);
int sl_shard_end(sl_handle* h, long long* n_rows /* nullable */);
/* Device-side check of what arrived: byte_sum = sum b[i], weighted = sum (i+1)*b[i]
 * (both mod 2^64) over the sealed shard's bytes.  Used for the file-0 parity gate
 * (src/file_server.cc:152-156). */
int sl_shard_checksum(sl_handle* h, unsigned long long* byte_sum, unsigned long long* weighted);
/* Copy n bytes at offset of the sealed shard back to the host (tests). */
int sl_shard_read(sl_handle* h, size_t offset, void* out, size_t n);
long long sl_shard_bytes(sl_handle* h);
/* Fill the shard on the device with rows [row0, row0+n_rows) of the synthetic
 * matrix of SURVEY.md §8(d) (bit-identical to the oracle's generator) and seal it. */
int sl_shard_synth(sl_handle* h, long long n_rows, unsigned long long row0,
                   unsigned long long seed);

/* ---- training ------------------------------------------------------------------
 * sl_train_tick replaces one iteration of simulate_training (src/worker.cc:225-227:
 * model[i] += 1) and exists for parity with the reference as it stands.
 * sl_train_steps is what takes that slot in this build: n_steps minibatch steps of
 * logistic regression over the resident shard, rows (t*B + j) mod n_rows, t
 * continuing from the previous call (reset by sl_shard_end).  loss_out (nullable)
 * receives n_steps mean losses.  Blocks until the steps are done.
 * Range: gradient sums travel as 64-bit fixed point (units of 2^-32 with 7 count bits in the cluster
 * kernel, 2^-36 in the older kernels), so for every column |sum of (sigmoid(z)-y)*x| is bounded: in the
 * cluster kernel each cluster's share of a minibatch (B / clusters rows, 33 clusters on a B200) must stay
 * below 2^24 / (clusters * ranks) — 5.1e5 on one GPU, 6.4e4 with 8 ranks, i.e. features up to several
 * hundred in magnitude at B = 4096; in the older kernels the whole minibatch's sum over all ranks below
 * 2^27.  Any fp32 data of ordinary scale; labels may be anywhere in [0,1].
 * A partial sum outside the range (or not finite) fails the call with SL_EINVAL instead of wrapping:
 * the model is then not usable and should be restored (sl_set_state). */
int sl_train_tick(sl_handle* h);
int sl_train_steps(sl_handle* h, int n_steps, double* loss_out);
/* Device time (CUDA events on the handle's stream) of the kernels of the last
 * sl_train_steps / sl_train_steps_sync call, in milliseconds. */
float sl_last_train_ms(sl_handle* h);

/* ---- update exchange -------------------------------------------------------------
 * sl_merge: callee side, the loop of MasterImpl::ExchangeUpdates
 * (src/master.cc:99-110) == WorkerImpl::ExchangeUpdates (src/worker.cc:85-96):
 *   model += lambda*delta_in; reply = model - old; old = model     — one kernel.
 * n_in and n_out must equal D (the reference zero-extends or reads out of bounds).
 * sl_delta_snapshot: initiator's delta build (src/worker.cc:205-208).
 * sl_apply_reply: initiator's apply (src/worker.cc:161-162) followed by
 *   old = model (src/worker.cc:215).
 * All three are bit-exact fp64 (rounded multiply, then rounded add). */
int sl_merge(sl_handle* h, const double* delta_in, int n_in, double* reply_out, int n_out);
int sl_delta_snapshot(sl_handle* h, double* out, int n);
int sl_apply_reply(sl_handle* h, const double* back, int n);
int sl_get_model(sl_handle* h, double* out, int n);
int sl_get_old(sl_handle* h, double* out, int n);
/* Overwrite the replica (tests, checkpoint restore).  old may be NULL = keep. */
int sl_set_state(sl_handle* h, const double* model, const double* old, int n);

/* ---- same-box data parallel (one handle per GPU, one process per GPU) ---------------
 * Build-defined extension (the reference only has asynchronous pairwise gossip):
 * every step each rank takes its local gradient and all ranks apply the mean.
 * sl_comm_unique_id fills 128 bytes on one rank; broadcast them by any means and
 * call sl_comm_init_rank on every rank.  NCCL is resolved with dlopen at that
 * point.  sl_allreduce_avg: model <- mean over ranks of model (one ncclAllReduce).
 * sl_train_steps_sync: n_steps of { local gradient; all-reduce; update }. */
/* Reference-faithful pairwise exchange between two replicas of ONE process (one per GPU, or two
 * on the same GPU): the whole round of periodically_gossip (src/worker.cc:194-219) and of the
 * peer's ExchangeUpdates handler (src/worker.cc:81-100) in one kernel on the callee's GPU, which
 * reads and writes the initiator's state through peer-mapped memory (NVLink) — no host staging,
 * no wire.  Bit-exact with sl_delta_snapshot -> sl_merge -> sl_apply_reply. */
int sl_gossip_p2p(sl_handle* initiator, sl_handle* callee);
/* The same one-kernel round between worker PROCESSES on one box — the reference's deployment, one worker per
 * process (src/worker.cc:233-258), where today the round is three blocking calls through host buffers.
 *   once:   initiator  sl_state_export(h, handles128)        two 64-byte IPC handles (model, old)
 *           callee     sl_state_attach(h, handles128, n, initiator's lambda, &peer)   maps them (<= 16 peers)
 *   round:  initiator  sl_state_lend(h)      waits for its own work, then keeps every other call on h out
 *           initiator -> callee: any doorbell (an ExchangeUpdates request with an empty delta will do)
 *           callee     sl_gossip_p2p_remote(h, peer)   src/worker.cc:194-219 + :81-100 in one kernel on the
 *                      callee's GPU, the initiator's state read and written through the mapping (NVLink, or
 *                      the same GPU's memory); returns when the kernel has finished
 *           callee -> initiator: the (empty) reply
 *           initiator  sl_state_reclaim(h)   same thread as sl_state_lend
 * Bit-exact with sl_delta_snapshot -> sl_merge -> sl_apply_reply and with sl_gossip_p2p.  The handles are those
 * of cudaMalloc allocations: both processes need the same box and CUDA IPC (not available inside one process —
 * use sl_gossip_p2p there).  sl_state_detach unmaps a peer that has left (the initiator must outlive the
 * mapping's last use, as with any IPC memory). */
int sl_state_export(sl_handle* h, void* handles128);
int sl_state_attach(sl_handle* h, const void* handles128, int n, double peer_lambda, int* peer_out);
int sl_state_detach(sl_handle* h, int peer);
int sl_state_lend(sl_handle* h);
int sl_state_reclaim(sl_handle* h);
int sl_gossip_p2p_remote(sl_handle* callee, int peer);

/* Fused path (preferred): sl_peer_export yields a 64-byte IPC handle of this replica's inbox;
 * gather the handles of all ranks (rank order) and pass them to sl_peer_attach on every rank
 * (1 <= nranks <= 8, one process per GPU, same box).  sl_train_steps_sync then runs ALL steps
 * in one persistent kernel per GPU: the kernels exchange 32-element slices of the exact integer
 * gradient sum through peer-mapped memory over NVLink, no host-launched collective per step.
 * Every rank must call sl_train_steps_sync with the same n_steps at about the same time (the
 * kernels wait for one another).  loss_out: with NCCL also initialised (sl_comm_init_rank) the
 * returned losses are the mean over ranks; with sl_peer_attach alone they are THIS rank's own
 * minibatch losses (the models are identical on all ranks either way).  Option "sync_path" = 1
 * forces the NCCL path; a shape the fused kernels cannot run needs NCCL too (SL_ESTATE otherwise). */
int sl_peer_export(sl_handle* h, void* handle64);
int sl_peer_attach(sl_handle* h, int nranks, int rank, const void* handles);
/* NVSwitch variant for the cluster kernel (narrow rows, F <= 1024): instead of (or besides) the IPC inbox
 * above, the caller hands in a SYMMETRIC allocation it made with its own plumbing (cuMemCreate / cuMemMap on
 * every rank, e.g. torch.distributed._symmetric_memory): bufs[p] is rank p's buffer as mapped in this process,
 * `multicast` the multicast address of the same allocation (cuMulticastCreate / cuMulticastBindMem) or NULL.
 * bytes >= 128 * F, zero-filled, 8-byte aligned; the memory must outlive the handle.  With a multicast
 * address a GPU's per-step total reaches all ranks with ONE multimem.st per column through the switch; without
 * it, with one plain store per peer.  Only stores and loads ever touch these addresses — no TMA copy reads
 * them.  No reference counterpart. */
int sl_peer_attach_symmetric(sl_handle* h, int nranks, int rank, const void* const* bufs, void* multicast,
                             size_t bytes);
/* A worker joins or leaves the synchronous group (the membership changes of src/master.cc:79-91 and :177-196
 * applied to this build-defined mode): EVERY member calls sl_peer_detach between two training calls, the callers
 * meet (any barrier), then the new set attaches again with its new nranks / ranks — sl_peer_export yields the same
 * handle as before.  A joiner first takes the group's state: sl_set_state with any member's sl_get_model (the
 * models are bit-identical on all members) and, to continue the same minibatch schedule, option "step".  Detach
 * wipes this replica's inbox; a symmetric buffer belongs to the caller, who zero-fills it before attaching it
 * again.  Replicas that trained alone before they attach need no preparation. */
int sl_peer_detach(sl_handle* h);
int sl_comm_unique_id(void* id128);
int sl_comm_init_rank(sl_handle* h, int nranks, int rank, const void* id128);
int sl_allreduce_avg(sl_handle* h);
int sl_train_steps_sync(sl_handle* h, int n_steps, double* loss_out);
/* The same synchronous rule for W worker replicas that SHARE one GPU (BASELINE config 2: "4 workers
 * ... same 4-shard job on 1 B200"): every step each worker takes the gradient of its own minibatch
 * over its own shard, the W gradients are summed and all replicas apply the mean — exactly what
 * sl_train_steps_sync computes across W GPUs, here as ONE persistent kernel over the W resident
 * shards (no copies: the kernel reads each handle's shard in place; the step's W*B rows are spread
 * over all SMs).  Requirements: 1 <= W <= 8, same device, F, B and eta, sealed shards with the
 * same number of rows, F % 4 == 0 (persistent kernel only).  The model of handles[0] is the
 * starting point; afterwards all W handles hold the new model and their step counters advance by
 * n_steps.  loss_out (nullable): n_steps losses, each the mean over the W minibatches.
 * No reference counterpart (the reference's workers are separate CPU processes). */
int sl_train_steps_group(sl_handle* const* handles, int W, int n_steps, double* loss_out);
/* W worker replicas that share one GPU each take n_steps LOCAL steps — the reference's own structure
 * (every worker process runs its own training thread, src/worker.cc:221-231,246, and only gossip
 * couples them): independent models, nothing summed across workers.  Rows the cluster kernel takes (F a
 * multiple of 16 up to 1024) run as W concurrent launches of it, one per replica on the replica's own stream
 * with 1/W of the clusters each — every worker ends bit-identical to sl_train_steps on its own with option
 * "clusters" set to that share; while one replica's step is in its exchange the others keep the row stream
 * going.  Otherwise (or with option "path" = 1 on the first handle) ONE persistent kernel: for W <= 4 and
 * F <= 1024 the replicas interleaved on every SM (every worker bit-identical to sl_train_steps on the older
 * kernel at F = 1024, within fp32 rounding of it for narrower rows), else the SMs cut into W equal partitions
 * (W <= 8, F <= 2048; bit-identical to sl_train_steps with "grid" = SMs / W).  Same requirements as
 * sl_train_steps_group.  loss_out (nullable): W * n_steps doubles, worker-major.  sl_last_train_ms of the first
 * handle: from the first launch to the last replica's end.  No reference counterpart. */
int sl_train_steps_multi(sl_handle* const* handles, int W, int n_steps, double* loss_out);

/* ---- allocation-free step ------------------------------------------------------------
 * No counterpart in the reference (which never touches a GPU): one minibatch step
 * (step index t, rows (t*B + j) mod n_rows, same three kernels as sl_train_steps' simple
 * path) for callers that must not allocate device memory.  Two kinds of caller in this
 * system are in that position: a worker that shares its B200 with other worker replicas
 * (BASELINE config 2 runs four on one GPU) — cudaMalloc/cudaFree synchronise the whole
 * device and would stall the neighbours' streams — and a worker whose resident shard was
 * sized to fill HBM, for which any further allocation fails.  So this entry point takes no
 * handle and calls no cudaMalloc/cudaFree or other allocating routine: its scratch is a
 * fixed array inside the library image (about 2.06 MiB per GPU that loads the library, held
 * until the process exits).
 *   - d_X [n_rows*F] fp32, d_y [n_rows] fp32, d_model [F] fp64 (updated in place) and
 *     d_loss (nullable, 1 fp64) are DEVICE pointers on `device`, owned by the caller.
 *     When F is a multiple of 4 the rows are read 16 bytes at a time: d_X must then be
 *     16-byte aligned (SL_EINVAL otherwise — a view into the middle of a larger buffer may not be).
 *   - limits: B <= 8192, F <= 4096; larger shapes return SL_EINVAL (see sl_last_error(NULL)).
 *   - asynchronous: the three kernels are enqueued on `stream` (a cudaStream_t, NULL = the
 *     default stream); results are ready when the stream reaches that point.
 *   - the first call on a device looks up the scratch's device address
 *     (cudaGetSymbolAddress); make that first call before starting any stream capture.
 *   - every call on a device uses the same scratch, which keeps whatever the last call left
 *     there (nothing in it needs clearing: each call overwrites all it reads).  Two calls
 *     whose kernels overlap on the GPU — different streams, or two threads — race on it;
 *     serialise them. */
int sl_lr_step_noalloc(int device, const float* d_X, const float* d_y, long long n_rows, int F,
                       int B, long long t, double eta, double* d_model, double* d_loss,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SL_GPU_H */
