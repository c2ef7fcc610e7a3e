/*
 * sl_oracle.h — CPU oracle for the serverless_learn data-parallel hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and there only as the checker or the timed CPU
 * baseline.  The product (serverless_learn_b200/) never links or calls it.
 *
 * What is restated from the reference (file:line relative to /root/reference):
 *   - file 0 byte stream            src/file_server.cc:40,46,152-156
 *   - chunking                      src/file_server.cc:60-87
 *   - merge callee                  src/master.cc:95-114 == src/worker.cc:81-100
 *   - delta build (initiator)       src/worker.cc:205-208 == src/master.cc:279-282
 *   - apply reply (initiator)       src/worker.cc:155-164, old=model at :215
 *   - placeholder training tick     src/worker.cc:221-231
 *   - partner choice rand()%N       src/worker.cc:199-202
 *   - Update wire format            src/protos/serverless_learn.proto:81-83
 * Repairs needed for the reference's code to be runnable at all (SURVEY.md
 * Appendix A): R1 import Update; R2 `updates`->`update`; R3 `->` on objects;
 * R4 declare model_state/old_state/LEARN_RATE in the worker; R5 drop the pasted
 * `return_update` line in the stub; R6 fixed dimension D and an equal-length
 * check instead of out-of-bounds reads; R7 skip a round with no peer; R8 one
 * deterministic single-threaded schedule instead of unsynchronised threads;
 * R9 somebody actually calls Master::ExchangeUpdates.
 *
 * What is NOT from the reference: the logistic-regression minibatch step.  The
 * reference has no model, loss or gradient (its "training" adds 1 to every
 * element, src/worker.cc:225-227).  orc_lr_* is a BUILD-DEFINED oracle
 * (SURVEY.md Appendix D); parity for it is unpinned by the reference.
 *
 * Pins: file 0, merge, delta, apply and the +1 tick are checked against the
 * reference's own lines compiled from /root/reference (oracle/build_ref.sh ->
 * oracle/_ref/) and against the known-answer vectors of SURVEY.md Appendix B.
 */
#ifndef SL_ORACLE_H
#define SL_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- file 0 (src/file_server.cc:152-156) ------------------------------- */
#define ORC_FILE0_LENGTH 100000000L /* src/file_server.cc:40 */
#define ORC_CHUNK_SIZE 1000000L     /* src/file_server.cc:46 */

/* Write the first n bytes of file 0 into out.  rejections (nullable, capacity
 * max_rej) receives the output positions just before which a draw was rejected.
 * Returns the number of rejections seen. */
int orc_file0(uint8_t* out, size_t n, int64_t* rejections, int max_rej);
uint64_t orc_fnv1a64(const uint8_t* p, size_t n);
uint64_t orc_bytesum(const uint8_t* p, size_t n);
/* position-weighted checksum: sum_i (i+1+base) * p[i]  mod 2^64 */
uint64_t orc_weighted_sum(const uint8_t* p, size_t n, uint64_t base);
/* number of chunks and the length of chunk k for a file of len bytes
 * (src/file_server.cc:68-76: substr(pos, CHUNK_SIZE) until pos == len) */
int64_t orc_chunk_count(int64_t len, int64_t chunk);
int64_t orc_chunk_len(int64_t len, int64_t chunk, int64_t k);

/* ---- exchange arithmetic (fp64) ---------------------------------------- */
/* callee merge, src/master.cc:105-110: model += lambda*delta_in;
 * reply = model - old; old = model.  n_in must equal n (R6) else returns -1. */
int orc_merge_callee(double* model, double* old, int n, const double* delta_in, int n_in,
                     double lambda, double* reply);
/* initiator delta, src/worker.cc:205-208 */
void orc_delta(const double* model, const double* old, int n, double* out);
/* initiator apply, src/worker.cc:161-162 then old = model (:215) */
int orc_apply_reply(double* model, double* old, int n, const double* back, int n_back,
                    double lambda);
/* placeholder tick, src/worker.cc:225-227 */
void orc_train_tick(double* model, int n);

/* ---- Update wire format (proto:81-83; proto3 packed doubles) ------------ */
/* returns bytes written (0 for n==0), or -1 if cap is too small */
long orc_update_encode(const double* delta, int n, uint8_t* out, long cap);
/* returns number of doubles decoded, or -1 on malformed input */
int orc_update_decode(const uint8_t* buf, long len, double* out, int cap);

/* ---- synthetic data -------------------------------------- */
#define ORC_SEED 0x5EEDULL
uint64_t orc_splitmix64(uint64_t x);
float orc_synth_x(uint64_t seed, uint64_t row, uint64_t col, uint64_t F);
/* fill rows [row0, row0+nrows) of the global synthetic matrix */
void orc_synth_fill(float* X, float* y, uint64_t row0, uint64_t nrows, uint64_t F,
                    uint64_t seed);

/* ---- build-defined logistic-regression step ------- */
/* one minibatch step t over a shard of n_rows rows: rows (t*B + j) mod n_rows.
 * fp32 inputs, fp64 accumulators and weights.  Returns the mean loss.  scratch
 * must hold B doubles (residuals). */
double orc_lr_step(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t,
                   double* w, double eta, double* scratch);
/* gradient only (no update): g[F] = (1/B) sum_j (p_j - y_j) x_j ; returns loss */
double orc_lr_grad(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t,
                   const double* w, double* g, double* scratch);
/* n_steps steps starting at step t0; loss_out nullable [n_steps] */
void orc_lr_steps(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t0,
                  int n_steps, double* w, double eta, double* loss_out);
/* synchronous data-parallel extension (label: build-defined): W workers, worker k
 * owns Xs[k] (n_rows[k] rows); every step all workers take a local step from the
 * same w and the new w is the mean of the W results.  loss_out[n_steps] = mean of
 * worker losses.  threads>1 runs one pthread per worker. */
void orc_lr_steps_sync(const float* const* Xs, const float* const* ys, const int64_t* n_rows,
                       int W, int F, int B, int64_t t0, int n_steps, double* w, double eta,
                       double* loss_out, int threads);
/* throughput helper for the timed CPU baseline: W independent workers (the
 * reference's structure: one training thread per worker, src/worker.cc:246), each
 * running n_steps local steps on its own shard and weights ws[k].  Returns
 * seconds of wall time for the arithmetic only. */
double orc_lr_bench_workers(const float* const* Xs, const float* const* ys,
                            const int64_t* n_rows, int W, int F, int B, int n_steps,
                            double* const* ws, double eta);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
