// ref_shim.cc — thin wrappers around fragments of the reference lifted at build
// time by oracle/build_ref.sh into a temporary directory (removed after the compile).  TEST INFRASTRUCTURE ONLY.
// This file holds no reference code: each `#include "x.inc"` below is replaced by
// the reference's own statements (with the repairs build_ref.sh lists) when the
// library is built in a container that has /root/reference.
#include <climits>
#include <cstring>
#include <random>
#include <string>
#include <vector>

// Stand-in for the protobuf-generated message (src/protos/serverless_learn.proto:81-83),
// exposing only the three accessors the hot path uses.
struct Update {
    std::vector<double> d;
    int delta_size() const { return (int)d.size(); }
    const std::vector<double>& delta() const { return d; }
    void add_delta(double v) { d.push_back(v); }
};

#include "file0_consts.inc"  // DUMMY_FILE_LENGTH, file, CHUNK_SIZE
#include "state.inc"         // model_state, old_state, LEARN_RATE

static void load_state(const double* model, const double* old, int n) {
    model_state.assign(model, model + n);
    old_state.assign(old, old + n);
}
static void store_state(double* model, double* old, int n) {
    std::memcpy(model, model_state.data(), sizeof(double) * (size_t)n);
    std::memcpy(old, old_state.data(), sizeof(double) * (size_t)n);
}

extern "C" {

long ref_file0_length() { return DUMMY_FILE_LENGTH; }
long ref_chunk_size() { return CHUNK_SIZE; }
double ref_learn_rate() { return LEARN_RATE; }

// Runs the generator loop once per process and copies min(cap, length) bytes.
long ref_file0(unsigned char* out, long cap) {
    if (file.empty()) {
#include "file0_body.inc"
    }
    long n = (long)file.size() < cap ? (long)file.size() : cap;
    std::memcpy(out, file.data(), (size_t)n);
    return (long)file.size();
}

// ExchangeUpdates handler body.  Returns the reply length.
int ref_merge(double* model, double* old, int n, const double* delta_in, int n_in, double* reply,
              int reply_cap) {
    load_state(model, old, n);
    Update in;
    in.d.assign(delta_in, delta_in + n_in);
    Update out;
    const Update* update = &in;
    Update* return_update = &out;
    {
#include "merge_body.inc"
    }
    if ((int)model_state.size() != n) return -1;  // grew: caller must size buffers first
    store_state(model, old, n);
    int m = out.delta_size() < reply_cap ? out.delta_size() : reply_cap;
    std::memcpy(reply, out.d.data(), sizeof(double) * (size_t)m);
    return out.delta_size();
}

// Initiator: delta build.
int ref_delta(const double* model, const double* old, int n, double* out_delta) {
    load_state(model, old, n);
#include "delta_body.inc"
    std::memcpy(out_delta, update.d.data(), sizeof(double) * (size_t)update.delta_size());
    return update.delta_size();
}

// Initiator: apply the reply, then `old_state = model_state`.
int ref_apply(double* model, double* old, int n, const double* back, int n_back) {
    load_state(model, old, n);
    Update update_back;
    update_back.d.assign(back, back + n_back);
    {
#include "apply_body.inc"
    }
    {
#include "oldcopy_body.inc"
    }
    if ((int)model_state.size() != n) return -1;
    store_state(model, old, n);
    return 0;
}

// One tick of simulate_training (without the sleep).
void ref_tick(double* model, int n) {
    std::vector<double> zero((size_t)n, 0.0);
    load_state(model, zero.data(), n);
#include "tick_body.inc"
    std::memcpy(model, model_state.data(), sizeof(double) * (size_t)n);
}

}  // extern "C"
