/*
 * sl_oracle_lr.c — BUILD-DEFINED oracle for the minibatch step.
 *
 * TEST INFRASTRUCTURE ONLY (see sl_oracle.h).  The reference has no model, loss or
 * gradient: its training thread is `model[i] = model[i] + 1` every 2 s
 * (src/worker.cc:221-231).  This file defines what takes that slot:
 * binary logistic regression without bias, D = F, on fp32 row-major data,
 *     z_j = sum_f X[r_j,f] w_f          r_j = (t*B + j) mod n_rows
 *     p_j = 1 / (1 + exp(-z_j))
 *     loss = (1/B) sum_j softplus(z_j) - y_j z_j
 *     g_f  = (1/B) sum_j (p_j - y_j) X[r_j,f]
 *     w   <- w - eta g
 * with every accumulator and the weights in fp64.  Parity of the GPU path against
 * this file is "unpinned by the reference": there is nothing in the reference to
 * compare it with.  It doubles as the timed CPU baseline (one thread per worker,
 * the reference's structure, src/worker.cc:246), so the loops are written to
 * vectorise under -O3 -march=native without -ffast-math: the dot product keeps 8
 * strided partial sums that are combined pairwise at the end.
 */
#include "sl_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <pthread.h>
#include <unistd.h>

static inline double dot_row(const float* x, const double* w, int F) {
    double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int f = 0;
    for (; f + 8 <= F; f += 8)
        for (int k = 0; k < 8; k++) a[k] += (double)x[f + k] * w[f + k];
    for (int k = 0; f < F; f++, k++) a[k] += (double)x[f] * w[f];
    return ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
}

static inline double softplus(double z) {
    return (z > 0 ? z : 0) + log1p(exp(-fabs(z)));
}

double orc_lr_grad(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t,
                   const double* w, double* g, double* scratch) {
    (void)scratch;
    double loss = 0.0;
    memset(g, 0, sizeof(double) * (size_t)F);
    int64_t r = (int64_t)(((__int128)t * B) % n_rows);
    for (int j = 0; j < B; j++) {
        const float* x = X + (size_t)r * (size_t)F;
        double z = dot_row(x, w, F);
        double p = 1.0 / (1.0 + exp(-z));
        double yy = (double)y[r];
        loss += softplus(z) - yy * z;
        double res = p - yy;
        for (int f = 0; f < F; f++) g[f] += res * (double)x[f];
        if (++r == n_rows) r = 0;
    }
    double inv = 1.0 / (double)B;
    for (int f = 0; f < F; f++) g[f] *= inv;
    return loss * inv;
}

double orc_lr_step(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t,
                   double* w, double eta, double* scratch) {
    double* g = (double*)malloc(sizeof(double) * (size_t)F);
    double loss = orc_lr_grad(X, y, n_rows, F, B, t, w, g, scratch);
    for (int f = 0; f < F; f++) w[f] -= eta * g[f];
    free(g);
    return loss;
}

void orc_lr_steps(const float* X, const float* y, int64_t n_rows, int F, int B, int64_t t0,
                  int n_steps, double* w, double eta, double* loss_out) {
    double* g = (double*)malloc(sizeof(double) * (size_t)F);
    for (int s = 0; s < n_steps; s++) {
        double loss = orc_lr_grad(X, y, n_rows, F, B, t0 + s, w, g, NULL);
        for (int f = 0; f < F; f++) w[f] -= eta * g[f];
        if (loss_out) loss_out[s] = loss;
    }
    free(g);
}

/* one thread per worker, fork/join per step (only used by tests and the sync baseline) */
struct grad_job {
    const float* X; const float* y; int64_t n_rows; int F, B; int64_t t;
    const double* w; double* g; double* loss;
};
static void* grad_main(void* p) {
    struct grad_job* j = (struct grad_job*)p;
    *j->loss = orc_lr_grad(j->X, j->y, j->n_rows, j->F, j->B, j->t, j->w, j->g, NULL);
    return NULL;
}
static void grad_all_workers(const float* const* Xs, const float* const* ys, const int64_t* n_rows,
                             int W, int F, int B, int64_t t, const double* w, double* g, double* l,
                             int threads) {
    if (threads <= 1 || W <= 1) {
        for (int k = 0; k < W; k++)
            l[k] = orc_lr_grad(Xs[k], ys[k], n_rows[k], F, B, t, w, g + (size_t)k * F, NULL);
        return;
    }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)W);
    struct grad_job* jobs = (struct grad_job*)malloc(sizeof(struct grad_job) * (size_t)W);
    for (int k = 0; k < W; k++) {
        struct grad_job j = {Xs[k], ys[k], n_rows[k], F, B, t, w, g + (size_t)k * F, &l[k]};
        jobs[k] = j;
        if (k > 0) pthread_create(&th[k], NULL, grad_main, &jobs[k]);
    }
    grad_main(&jobs[0]);
    for (int k = 1; k < W; k++) pthread_join(th[k], NULL);
    free(th);
    free(jobs);
}

struct bench_job {
    const float* X; const float* y; int64_t n_rows; int F, B, n_steps; double* w; double eta;
};
static void* bench_main(void* p) {
    struct bench_job* j = (struct bench_job*)p;
    orc_lr_steps(j->X, j->y, j->n_rows, j->F, j->B, 0, j->n_steps, j->w, j->eta, NULL);
    return NULL;
}

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

void orc_lr_steps_sync(const float* const* Xs, const float* const* ys, const int64_t* n_rows,
                       int W, int F, int B, int64_t t0, int n_steps, double* w, double eta,
                       double* loss_out, int threads) {
    double* g = (double*)malloc(sizeof(double) * (size_t)F * (size_t)W);
    double* l = (double*)malloc(sizeof(double) * (size_t)W);
    for (int s = 0; s < n_steps; s++) {
        grad_all_workers(Xs, ys, n_rows, W, F, B, t0 + s, w, g, l, threads);
        /* mean over workers of (w - eta g_k) == w - eta * mean_k g_k, summed in rank order */
        for (int f = 0; f < F; f++) {
            double acc = 0.0;
            for (int k = 0; k < W; k++) acc += g[(size_t)k * F + f];
            w[f] -= eta * (acc / (double)W);
        }
        if (loss_out) {
            double acc = 0.0;
            for (int k = 0; k < W; k++) acc += l[k];
            loss_out[s] = acc / (double)W;
        }
    }
    free(g);
    free(l);
}


double orc_lr_bench_workers(const float* const* Xs, const float* const* ys,
                            const int64_t* n_rows, int W, int F, int B, int n_steps,
                            double* const* ws, double eta) {
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)W);
    struct bench_job* jobs = (struct bench_job*)malloc(sizeof(struct bench_job) * (size_t)W);
    double t0 = now_s();
    for (int k = 0; k < W; k++) {
        struct bench_job j = {Xs[k], ys[k], n_rows[k], F, B, n_steps, ws[k], eta};
        jobs[k] = j;
        if (k > 0) pthread_create(&th[k], NULL, bench_main, &jobs[k]);
    }
    bench_main(&jobs[0]);
    for (int k = 1; k < W; k++) pthread_join(th[k], NULL);
    double dt = now_s() - t0;
    free(th);
    free(jobs);
    return dt;
}

int orc_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
