// sl_job — the reference's three roles in one C++ process on the B200 path (host code above the C ABI):
// a file server that holds one shard per worker, a master that registers the workers, pushes the shards
// (DoPush -> ReceiveFile into HBM) and sends the peer lists (CheckUp), W workers that train and gossip.
// No gRPC here (not in this image): the role cores of serverless_learn_b200/csrc/roles.h are what the
// gRPC shims of INTEGRATION.md would call.  Prints one JSON line.
//
//   sl_job [--workers 4] [--rows 4096] [--features 1024] [--batch 256] [--rounds 8] [--local-steps 16]
//          [--eta 0.1] [--mode async|sync]
//   async: the reference's structure (src/worker.cc:221-231 training threads, :194-219 gossip): every round
//          each worker takes local steps (all workers side by side in one kernel), then gossips once with a
//          rand()%N partner;  sync: the build-defined synchronous extension, one kernel per round.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <memory>
#include <string>
#include <vector>

#include "../serverless_learn_b200/csrc/roles.h"

int main(int argc, char** argv) {
    int W = 4, F = 1024, B = 256, rounds = 8, K = 16;
    long long rows = 4096;
    double eta = 0.1;
    std::string mode = "async";
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string k = argv[i];
        const char* v = argv[i + 1];
        if (k == "--workers") W = atoi(v);
        else if (k == "--rows") rows = atoll(v);
        else if (k == "--features") F = atoi(v);
        else if (k == "--batch") B = atoi(v);
        else if (k == "--rounds") rounds = atoi(v);
        else if (k == "--local-steps") K = atoi(v);
        else if (k == "--eta") eta = atof(v);
        else if (k == "--mode") mode = v;
        else { fprintf(stderr, "unknown option %s\n", k.c_str()); return 2; }
    }
    if (W < 1 || W > 8 || (mode != "async" && mode != "sync")) { fprintf(stderr, "bad arguments\n"); return 2; }

    slr::Directory dir;
    slr::FileServerCore fs;                                   // file 0 as the reference builds it
    slr::MasterCore master(0, F, B, eta, 0.5);
    if (!master.ok()) { fprintf(stderr, "master: %s\n", sl_last_error(nullptr)); return 1; }
    std::vector<std::unique_ptr<slr::WorkerCore>> workers;
    for (int k = 0; k < W; ++k) {
        const std::string addr = "worker" + std::to_string(k) + ":50051";
        workers.emplace_back(new slr::WorkerCore(0, addr, F, F, B, eta, 0.5));
        if (!workers.back()->ok()) { fprintf(stderr, "worker %d: %s\n", k, sl_last_error(nullptr)); return 1; }
        dir.add(addr, workers.back().get());
        master.RegisterBirth(addr);                                          // src/worker.cc:117-129
        if (!fs.AddSyntheticShard((uint32_t)k + 1, (uint64_t)k * rows, (uint64_t)rows, (uint32_t)F, 0x5EED, true))
            return 1;
    }
    if (master.RequestPushesSharded(fs, dir) != W) { fprintf(stderr, "a push failed\n"); return 1; }
    if (master.DoCheckups(fs, dir) != W) { fprintf(stderr, "a check-up failed\n"); return 1; }

    sl_handle* hs[8];
    for (int k = 0; k < W; ++k) hs[k] = workers[k]->handle();
    std::vector<double> loss((size_t)W * K), mean_loss;
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < rounds; ++r) {
        double m = 0.0;
        if (mode == "async") {
            if (sl_train_steps_multi(hs, W, K, loss.data()) != 0) { fprintf(stderr, "%s\n", sl_last_error(hs[0])); return 1; }
            for (int k = 0; k < W; ++k) m += loss[(size_t)k * K + K - 1] / W;
            for (int k = 0; k < W; ++k)
                if (workers[k]->GossipOnce(dir) < 0) { fprintf(stderr, "gossip failed\n"); return 1; }
        } else {
            if (sl_train_steps_group(hs, W, K, loss.data()) != 0) { fprintf(stderr, "%s\n", sl_last_error(hs[0])); return 1; }
            m = loss[(size_t)K - 1];
        }
        mean_loss.push_back(m);
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::vector<double> w0((size_t)F);
    sl_get_model(hs[0], w0.data(), F);
    printf("{\"mode\": \"%s\", \"workers\": %d, \"rows_per_worker\": %lld, \"F\": %d, \"B\": %d, \"rounds\": %d, "
           "\"local_steps\": %d, \"samples_per_s\": %.6g, \"seconds\": %.6g, \"w0_0\": %.17g, \"mean_loss\": [",
           mode.c_str(), W, rows, F, B, rounds, K, (double)W * B * K * rounds / secs, secs, w0[0]);
    for (size_t i = 0; i < mean_loss.size(); ++i) printf("%s%.17g", i ? ", " : "", mean_loss[i]);
    printf("]}\n");
    return 0;
}
