/*
 * sl_oracle.c — exact (byte / fp64) half of the CPU oracle.  See sl_oracle.h for
 * the scope note: test infrastructure only, never on the product path.
 *
 * Compile with -ffp-contract=off: the reference is built by plain `g++ -std=c++11`
 * with no -O and no -march (src/Makefile:6-8), so `model += LEARN_RATE * delta`
 * is a rounded multiply followed by a rounded add, never a fused multiply-add.
 */
#include "sl_oracle.h"

#include <string.h>

/* ---------------------------------------------------------------------------
 * file 0 — src/file_server.cc:152-156
 *   std::independent_bits_engine<std::default_random_engine, CHAR_BIT, unsigned char>
 * libstdc++: default_random_engine = minstd_rand0 = LCG(a=16807, c=0, m=2^31-1),
 * seed 1, first output is the state after one multiply.  independent_bits_engine
 * with w=8 over a generator of range R = max-min+1 = 2^31-2 takes ONE draw per
 * output (n=1, w0=8): u = x - min = x - 1; y0 = floor(R / 2^8) * 2^8 = 2147483392;
 * draws with u >= y0 are rejected; result = u & 0xFF.
 * ------------------------------------------------------------------------- */
int orc_file0(uint8_t* out, size_t n, int64_t* rejections, int max_rej) {
    const uint64_t a = 16807u, m = 2147483647u;
    const uint64_t y0 = 2147483392u;
    uint64_t x = 1;
    int nrej = 0;
    for (size_t i = 0; i < n;) {
        x = (a * x) % m;
        uint64_t u = x - 1;
        if (u >= y0) {
            if (rejections && nrej < max_rej) rejections[nrej] = (int64_t)i;
            nrej++;
            continue;
        }
        out[i++] = (uint8_t)(u & 0xFFu);
    }
    return nrej;
}

uint64_t orc_fnv1a64(const uint8_t* p, size_t n) {
    uint64_t h = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 0x100000001b3ULL;
    }
    return h;
}

uint64_t orc_bytesum(const uint8_t* p, size_t n) {
    uint64_t s = 0;
    for (size_t i = 0; i < n; i++) s += p[i];
    return s;
}

uint64_t orc_weighted_sum(const uint8_t* p, size_t n, uint64_t base) {
    uint64_t s = 0;
    for (size_t i = 0; i < n; i++) s += (uint64_t)(base + i + 1) * (uint64_t)p[i];
    return s;
}

/* src/file_server.cc:68-76 — pos advances by substr(pos, CHUNK_SIZE).length() */
int64_t orc_chunk_count(int64_t len, int64_t chunk) {
    if (len <= 0 || chunk <= 0) return 0;
    return (len + chunk - 1) / chunk;
}

int64_t orc_chunk_len(int64_t len, int64_t chunk, int64_t k) {
    int64_t pos = k * chunk;
    if (k < 0 || pos >= len) return 0;
    return (len - pos < chunk) ? (len - pos) : chunk;
}

/* ---------------------------------------------------------------------------
 * exchange arithmetic
 * ------------------------------------------------------------------------- */

/* src/master.cc:99-110 (== src/worker.cc:85-96), repairs R2 and R6. */
int orc_merge_callee(double* model, double* old, int n, const double* delta_in, int n_in,
                     double lambda, double* reply) {
    if (n_in != n) return -1; /* R6: the reference zero-extends or reads out of bounds */
    for (int i = 0; i < n; i++) {
        model[i] += lambda * delta_in[i];  /* :106 */
        reply[i] = model[i] - old[i];      /* :107 */
    }
    memcpy(old, model, sizeof(double) * (size_t)n); /* :110 old_state = model_state */
    return 0;
}

/* src/worker.cc:205-208 */
void orc_delta(const double* model, const double* old, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = model[i] - old[i];
}

/* src/worker.cc:161-162 (R3, R5) and old_state = model_state at :215 */
int orc_apply_reply(double* model, double* old, int n, const double* back, int n_back,
                    double lambda) {
    if (n_back != n) return -1;
    for (int i = 0; i < n; i++) model[i] += lambda * back[i];
    memcpy(old, model, sizeof(double) * (size_t)n);
    return 0;
}

/* src/worker.cc:225-227 */
void orc_train_tick(double* model, int n) {
    for (int i = 0; i < n; i++) model[i] = model[i] + 1;
}

/* ---------------------------------------------------------------------------
 * Update wire format — proto3 `repeated double delta = 1;` is packed:
 * tag 0x0A (field 1, wire type 2), varint byte length 8*n, then n LE doubles.
 * An empty repeated field serialises to nothing.
 * ------------------------------------------------------------------------- */
long orc_update_encode(const double* delta, int n, uint8_t* out, long cap) {
    if (n == 0) return 0;
    uint64_t len = 8ull * (uint64_t)n;
    uint8_t hdr[12];
    int h = 0;
    hdr[h++] = 0x0A;
    uint64_t v = len;
    while (v >= 0x80) {
        hdr[h++] = (uint8_t)(v | 0x80);
        v >>= 7;
    }
    hdr[h++] = (uint8_t)v;
    if ((long)(h + len) > cap) return -1;
    memcpy(out, hdr, (size_t)h);
    memcpy(out + h, delta, len); /* x86-64 is little-endian */
    return (long)(h + len);
}

int orc_update_decode(const uint8_t* buf, long len, double* out, int cap) {
    long pos = 0;
    int n = 0;
    while (pos < len) {
        uint8_t tag = buf[pos++];
        if (tag == 0x0A) { /* packed run */
            uint64_t l = 0;
            int shift = 0;
            for (;;) {
                if (pos >= len || shift > 63) return -1;
                uint8_t b = buf[pos++];
                l |= (uint64_t)(b & 0x7F) << shift;
                shift += 7;
                if (!(b & 0x80)) break;
            }
            if (l % 8 != 0 || pos + (long)l > len) return -1;
            int k = (int)(l / 8);
            if (n + k > cap) return -1;
            memcpy(out + n, buf + pos, l);
            n += k;
            pos += (long)l;
        } else if (tag == 0x09) { /* unpacked element, also legal proto3 input */
            if (pos + 8 > len || n + 1 > cap) return -1;
            memcpy(out + n, buf + pos, 8);
            n += 1;
            pos += 8;
        } else {
            return -1;
        }
    }
    return n;
}

/* ---------------------------------------------------------------------------
 * synthetic data — SURVEY.md §8(d): counter-based, identical on CPU and GPU
 * ------------------------------------------------------------------------- */
uint64_t orc_splitmix64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

float orc_synth_x(uint64_t seed, uint64_t row, uint64_t col, uint64_t F) {
    uint64_t u = orc_splitmix64(seed ^ (row * F + col));
    /* top 24 bits -> [0,1) exactly representable, then [-1,1) */
    float f = (float)(u >> 40) * (1.0f / 16777216.0f);
    return f * 2.0f - 1.0f;
}

void orc_synth_fill(float* X, float* y, uint64_t row0, uint64_t nrows, uint64_t F,
                    uint64_t seed) {
    for (uint64_t r = 0; r < nrows; r++) {
        float* xr = X + r * F;
        for (uint64_t c = 0; c < F; c++) xr[c] = orc_synth_x(seed, row0 + r, c, F);
        if (y) y[r] = xr[0] > 0.0f ? 1.0f : 0.0f;
    }
}
