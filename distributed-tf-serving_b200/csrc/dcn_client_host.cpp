// dcn_client_host.cpp -- C++ host mirror of the reference's load driver (DCNClient.main,
// DCNClient.java:205-241) over the C ABI of the engine (dlopen, no link-time dependency).
//
//   DC:209-210  inputs are built ONCE and shared read-only by every driver thread
//   DC:213-224  concurrentNum (6) OS threads, each issuing requestNum sequential requests
//   DC:137-203  processARequest: here ONE ctr_predict_async + ctr_wait per request (the engine
//               fans out over its GPUs and merges in order), then the ascending sort (DC:195)
//   DC:141,198  the timed window: inputs exist -> merged, sorted scores on the host
//   DC:234-236  mean latency at exit (this driver also prints p50 / p99)
#include "../../include/ctr_engine.h"

#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {
struct Api {
    decltype(&ctr_engine_create) create;
    decltype(&ctr_engine_destroy) destroy;
    decltype(&ctr_engine_load_weights) load;
    decltype(&ctr_engine_model_info) info;
    decltype(&ctr_predict_async) predict_async;
    decltype(&ctr_wait) wait;
    decltype(&ctr_last_error) last_error;
};

template <typename T>
bool sym(void* h, const char* name, T* out) {
    *out = reinterpret_cast<T>(dlsym(h, name));
    if (!*out) fprintf(stderr, "missing symbol %s\n", name);
    return *out != nullptr;
}

// Float.compareTo order (DC:195 sorts boxed Floats): -0.0 < +0.0, NaN last
inline int32_t java_key(float f) {
    int32_t b;
    memcpy(&b, &f, 4);
    if (f != f) b = 0x7FC00000;
    return b < 0 ? (b ^ 0x7FFFFFFF) : b;
}
}  // namespace

int main(int argc, char** argv) {
    std::string lib = "libctr_b200.so", weights;
    int candidate_num = 1500, request_num = 1000, concurrent_num = 6, n_devices = 1;   // DC:29-31
    bool do_sort = true;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", what); exit(2); }
            return argv[++i];
        };
        if (a == "--help") {
            printf("usage: dcn_client_host --weights FILE [--lib libctr_b200.so] [--candidates 1500] [--requests 1000]\n"
                   "                       [--threads 6] [--devices 1] [--no-sort]\n"
                   "Mirror of DCNClient.main: N driver threads x M sequential requests over the C ABI.\n");
            return 0;
        } else if (a == "--weights") weights = next("--weights");
        else if (a == "--lib") lib = next("--lib");
        else if (a == "--candidates") candidate_num = atoi(next("--candidates"));
        else if (a == "--requests") request_num = atoi(next("--requests"));
        else if (a == "--threads") concurrent_num = atoi(next("--threads"));
        else if (a == "--devices") n_devices = atoi(next("--devices"));
        else if (a == "--no-sort") do_sort = false;
        else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    if (weights.empty()) { fprintf(stderr, "--weights is required\n"); return 2; }
    void* h = dlopen(lib.c_str(), RTLD_NOW);
    if (!h) { fprintf(stderr, "dlopen failed: %s\n", dlerror()); return 1; }
    Api api;
    if (!(sym(h, "ctr_engine_create", &api.create) && sym(h, "ctr_engine_destroy", &api.destroy) &&
          sym(h, "ctr_engine_load_weights", &api.load) && sym(h, "ctr_engine_model_info", &api.info) &&
          sym(h, "ctr_predict_async", &api.predict_async) && sym(h, "ctr_wait", &api.wait) &&
          sym(h, "ctr_last_error", &api.last_error)))
        return 1;

    ctr_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.struct_size = sizeof(cfg);
    cfg.n_devices = n_devices;                                       // hostsList.size(), DC:38
    for (int i = 0; i < n_devices; ++i) cfg.device_ids[i] = i;
    cfg.max_candidates = candidate_num;
    cfg.workers_per_device = 4;
    ctr_engine* e = nullptr;
    if (api.create(&cfg, &e) != CTR_OK) { fprintf(stderr, "ctr_engine_create: %s\n", api.last_error()); return 1; }
    if (api.load(e, weights.c_str()) != CTR_OK) { fprintf(stderr, "ctr_engine_load_weights: %s\n", api.last_error()); return 1; }
    ctr_model_info mi;
    api.info(e, &mi);
    const int F = mi.num_fields;                                     // FIELD_NUM, DC:25

    // DC:57-74: ids 1..F per candidate, weights 1.0f -- built once, shared by all threads (DC:209-210)
    std::vector<int64_t> ids((size_t)candidate_num * F);
    std::vector<float> wts((size_t)candidate_num * F, 1.0f);
    for (int c = 0; c < candidate_num; ++c)
        for (int f = 0; f < F; ++f) ids[(size_t)c * F + f] = f + 1;

    std::mutex mu;
    std::vector<double> times;                                       // timeLists, DC:44
    std::atomic<int> failures{0};
    auto driver = [&]() {
        std::vector<float> scores(candidate_num);
        std::vector<int32_t> keys(candidate_num);
        for (int j = 0; j < request_num; ++j) {
            auto t0 = std::chrono::steady_clock::now();              // DC:141
            ctr_ticket t;
            int rc = api.predict_async(e, ids.data(), wts.data(), candidate_num, scores.data(), nullptr, nullptr, &t);
            if (rc == CTR_OK) rc = api.wait(e, t, -1);               // join, DC:158-159
            if (rc != CTR_OK) { ++failures; fprintf(stderr, "request failed: %s\n", api.last_error()); return; }
            if (do_sort)                                             // DC:195
                std::sort(scores.begin(), scores.end(), [](float a, float b) { return java_key(a) < java_key(b); });
            double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();   // DC:198-199
            std::lock_guard<std::mutex> lk(mu);
            times.push_back(ms);
        }
    };
    std::vector<std::thread> threads;
    auto w0 = std::chrono::steady_clock::now();
    for (int i = 0; i < concurrent_num; ++i) threads.emplace_back(driver);   // DC:213-224
    for (auto& t : threads) t.join();                                        // DC:226-232
    double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
    if (times.empty() || failures) { api.destroy(e); return 1; }
    double sum = 0;
    for (double v : times) sum += v;
    std::sort(times.begin(), times.end());
    printf("Average time cost with %d is %.4f ms with %zu requests\n", candidate_num, sum / times.size(), times.size());   // DC:234-236
    printf("p50 %.4f ms  p99 %.4f ms  %.0f candidates/s over %d driver threads\n", times[times.size() / 2],
           times[std::min(times.size() - 1, (size_t)(times.size() * 0.99))], (double)times.size() * candidate_num / wall, concurrent_num);
    api.destroy(e);
    dlclose(h);
    return 0;
}
