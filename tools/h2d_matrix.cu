// Pinned host -> device copy bandwidth of GPU subsets on one node: which GPUs share a host
// uplink, and in which order GPUs should be taken so that N concurrent copiers get the most
// out of the host.  One JSON line per measurement on stdout.
//
//   h2d_matrix [bytes_per_gpu_per_rep] [reps]
//
// Sections: "single" (each GPU alone), "pair" (every pair), "subset" (named subsets),
// "variant" (two copy streams per GPU, write-combined staging) and "greedy" (grow a set one GPU
// at a time, always adding the GPU that gives the largest measured total; the resulting order
// is what bench.py / dist.py use as the rank -> GPU map).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <vector>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e_));           \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

struct Gpu {
    uint8_t* host = nullptr;     // pinned, default flags
    uint8_t* host_wc = nullptr;  // pinned, write-combined
    uint8_t* dev = nullptr;
    cudaStream_t s[2] = {nullptr, nullptr};
    cudaEvent_t e0 = nullptr, e1[2] = {nullptr, nullptr};
};

static std::vector<Gpu> g;
static uint64_t g_bytes = 1ull << 30;
static int g_reps = 4;

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Copies g_bytes x g_reps to every GPU of `set` at once; returns total GB/s (all bytes over the
// wall time until the last GPU finished) and per-GPU GB/s from CUDA events.
static double run(const std::vector<int>& set, int streams, bool wc, std::vector<double>* per_gpu) {
    for (int d : set) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceSynchronize());
    }
    const double t0 = now();
    for (int d : set) {
        CK(cudaSetDevice(d));
        CK(cudaEventRecord(g[d].e0, g[d].s[0]));
        if (streams == 2) CK(cudaStreamWaitEvent(g[d].s[1], g[d].e0, 0));
    }
    const uint64_t part = g_bytes / streams;
    for (int r = 0; r < g_reps; ++r)
        for (int d : set) {
            CK(cudaSetDevice(d));
            const uint8_t* src = wc ? g[d].host_wc : g[d].host;
            for (int k = 0; k < streams; ++k)
                CK(cudaMemcpyAsync(g[d].dev + k * part, src + k * part, part, cudaMemcpyHostToDevice, g[d].s[k]));
        }
    for (int d : set) {
        CK(cudaSetDevice(d));
        for (int k = 0; k < streams; ++k) CK(cudaEventRecord(g[d].e1[k], g[d].s[k]));
    }
    for (int d : set) {
        CK(cudaSetDevice(d));
        for (int k = 0; k < streams; ++k) CK(cudaEventSynchronize(g[d].e1[k]));
    }
    const double dt = now() - t0;
    if (per_gpu) {
        per_gpu->clear();
        for (int d : set) {
            float ms = 0.f, worst = 0.f;
            for (int k = 0; k < streams; ++k) {
                CK(cudaEventElapsedTime(&ms, g[d].e0, g[d].e1[k]));
                worst = std::max(worst, ms);
            }
            per_gpu->push_back((double)g_bytes * g_reps / (worst * 1e-3) / 1e9);
        }
    }
    return (double)g_bytes * g_reps * set.size() / dt / 1e9;
}

static void report(const char* section, const std::vector<int>& set, int streams, bool wc) {
    std::vector<double> per;
    run(set, streams, wc, nullptr);  // warm-up
    const double total = run(set, streams, wc, &per);
    std::string ids, rates;
    for (size_t i = 0; i < set.size(); ++i) {
        ids += (i ? ", " : "") + std::to_string(set[i]);
        char b[32];
        snprintf(b, sizeof b, "%s%.1f", i ? ", " : "", per[i]);
        rates += b;
    }
    printf("{\"section\": \"%s\", \"gpus\": [%s], \"streams_per_gpu\": %d, \"write_combined\": %s, "
           "\"total_gb_s\": %.1f, \"per_gpu_gb_s\": [%s]}\n",
           section, ids.c_str(), streams, wc ? "true" : "false", total, rates.c_str());
    fflush(stdout);
}

int main(int argc, char** argv) {
    if (argc > 1) g_bytes = strtoull(argv[1], 0, 10);
    if (argc > 2) g_reps = atoi(argv[2]);
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    g.resize(n);
    for (int d = 0; d < n; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaHostAlloc((void**)&g[d].host, g_bytes, cudaHostAllocPortable));
        CK(cudaHostAlloc((void**)&g[d].host_wc, g_bytes, cudaHostAllocPortable | cudaHostAllocWriteCombined));
        memset(g[d].host, d + 1, g_bytes);
        memset(g[d].host_wc, d + 1, g_bytes);
        CK(cudaMalloc((void**)&g[d].dev, g_bytes));
        for (int k = 0; k < 2; ++k) {
            CK(cudaStreamCreateWithFlags(&g[d].s[k], cudaStreamNonBlocking));
            CK(cudaEventCreate(&g[d].e1[k]));
        }
        CK(cudaEventCreate(&g[d].e0));
    }
    printf("{\"section\": \"setup\", \"gpus\": %d, \"bytes_per_gpu_per_rep\": %llu, \"reps\": %d}\n", n,
           (unsigned long long)g_bytes, g_reps);
    for (int d = 0; d < n; ++d) report("single", {d}, 1, false);
    for (int a = 0; a < n; ++a)
        for (int b = a + 1; b < n; ++b) report("pair", {a, b}, 1, false);
    if (n >= 8) {
        report("subset", {0, 1, 2, 3}, 1, false);
        report("subset", {4, 5, 6, 7}, 1, false);
        report("subset", {0, 1, 4, 5}, 1, false);
        report("subset", {0, 2, 4, 6}, 1, false);
        report("subset", {0, 1, 2, 3, 4, 5, 6, 7}, 1, false);
    }
    std::vector<int> all;
    for (int d = 0; d < n; ++d) all.push_back(d);
    report("variant", {0}, 2, false);
    report("variant", {0}, 1, true);
    if (n > 1) {
        report("variant", all, 2, false);
        report("variant", all, 1, true);
    }
    // Greedy order: the first k GPUs of `order` are the best measured k-subset grown one at a time.
    std::vector<int> order = {0}, rest;
    for (int d = 1; d < n; ++d) rest.push_back(d);
    std::vector<double> totals = {0.0};
    {
        std::vector<double> per;
        run(order, 1, false, nullptr);
        totals[0] = run(order, 1, false, &per);
    }
    while (!rest.empty()) {
        int best = -1;
        double best_total = -1.0;
        for (int c : rest) {
            std::vector<int> trial = order;
            trial.push_back(c);
            run(trial, 1, false, nullptr);
            const double t = run(trial, 1, false, nullptr);
            if (t > best_total) { best_total = t; best = c; }
        }
        order.push_back(best);
        totals.push_back(best_total);
        rest.erase(std::find(rest.begin(), rest.end(), best));
    }
    std::string o, t;
    for (size_t i = 0; i < order.size(); ++i) {
        o += (i ? ", " : "") + std::to_string(order[i]);
        char b[32];
        snprintf(b, sizeof b, "%s%.1f", i ? ", " : "", totals[i]);
        t += b;
    }
    printf("{\"section\": \"greedy\", \"order\": [%s], \"total_gb_s_with_first_k\": [%s]}\n", o.c_str(), t.c_str());
    return 0;
}
