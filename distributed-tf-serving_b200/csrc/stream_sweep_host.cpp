// stream_sweep_host.cpp -- BASELINE.json config 5 over the C ABI: a ranking request stream at request
// sizes 64 ... 65 536 candidates, measured two ways.
//
//   closed loop  the reference's own load shape (DCNClient.java:213-224): concurrentNum (6) driver
//                threads, each issuing sequential requests; window DC:141 -> DC:198 without the sort
//                (inputs exist -> ordered scores on the host)
//   open loop    Poisson arrivals at a stated fraction of the closed-loop capacity, submitted with
//                ctr_predict_async + completion callback (the CompletableFuture shape of DC:151-154);
//                latency = completion - SCHEDULED arrival, so queueing shows up in the tail
//
// Inputs are Criteo-shaped (13 dense + 26 sparse fields), generated
// here with a fixed-seed integer hash, written once into ctr_alloc_pinned buffers and shared read-only
// by all requests (DC:209-210 shares its input lists the same way).
#include "../../include/ctr_engine.h"

#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {
using Clock = std::chrono::steady_clock;

struct Api {
    decltype(&ctr_engine_create) create;
    decltype(&ctr_engine_destroy) destroy;
    decltype(&ctr_engine_load_weights) load;
    decltype(&ctr_engine_model_info) info;
    decltype(&ctr_predict_async) predict_async;
    decltype(&ctr_wait) wait;
    decltype(&ctr_alloc_pinned) alloc_pinned;
    decltype(&ctr_free_pinned) free_pinned;
    decltype(&ctr_last_error) last_error;
};

template <typename T>
bool sym(void* h, const char* name, T* out) {
    *out = reinterpret_cast<T>(dlsym(h, name));
    if (!*out) fprintf(stderr, "missing symbol %s\n", name);
    return *out != nullptr;
}

inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline double unit(uint64_t k) { return (double)(mix64(k) >> 11) * (1.0 / 9007199254740992.0); }

double pct(std::vector<double>& v, double p) {
    if (v.empty()) return 0;
    std::sort(v.begin(), v.end());
    return v[std::min(v.size() - 1, (size_t)(v.size() * p))];
}

// One slot of the open-loop ring: an output buffer plus the bookkeeping its callback needs.
struct Flight {
    float* out = nullptr;
    Clock::time_point scheduled;
    std::atomic<int> busy{0};
    std::vector<double>* sink = nullptr;
    std::mutex* sink_mu = nullptr;
    std::atomic<long>* done = nullptr;
    std::atomic<int>* failures = nullptr;
};

void on_done(void* user, ctr_ticket, int32_t status) {
    Flight* f = static_cast<Flight*>(user);
    const double ms = std::chrono::duration<double, std::milli>(Clock::now() - f->scheduled).count();
    if (status != CTR_OK) ++*f->failures;
    {
        std::lock_guard<std::mutex> lk(*f->sink_mu);
        f->sink->push_back(ms);
    }
    f->busy.store(0, std::memory_order_release);
    ++*f->done;
}
}  // namespace

int main(int argc, char** argv) {
    std::string lib = "libctr_b200.so", weights, out_path;
    int n_devices = 1, threads = 6, requests = 10000, warmup = 1000, max_size = 65536, min_size = 64, workers = 4;
    long rows_per_field = 1L << 22;
    double load = 0.7;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", what); exit(2); }
            return argv[++i];
        };
        if (a == "--help") {
            printf("usage: stream_sweep_host --weights FILE [--lib libctr_b200.so] [--devices 1] [--threads 6] [--workers 4]\n"
                   "         [--requests 10000] [--warmup 1000] [--min 64] [--max 65536] [--load 0.7]\n"
                   "         [--rows-per-field 4194304] [--out table.md]\n"
                   "Config 5: closed loop (6 driver threads) and open loop (Poisson arrivals at --load x the\n"
                   "closed-loop request rate) per request size; prints a markdown table.\n");
            return 0;
        } else if (a == "--weights") weights = next("--weights");
        else if (a == "--lib") lib = next("--lib");
        else if (a == "--devices") n_devices = atoi(next("--devices"));
        else if (a == "--threads") threads = atoi(next("--threads"));
        else if (a == "--workers") workers = atoi(next("--workers"));
        else if (a == "--requests") requests = atoi(next("--requests"));
        else if (a == "--warmup") warmup = atoi(next("--warmup"));
        else if (a == "--min") min_size = atoi(next("--min"));
        else if (a == "--max") max_size = atoi(next("--max"));
        else if (a == "--load") load = atof(next("--load"));
        else if (a == "--rows-per-field") rows_per_field = atol(next("--rows-per-field"));
        else if (a == "--out") out_path = next("--out");
        else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    if (weights.empty()) { fprintf(stderr, "--weights is required\n"); return 2; }
    void* h = dlopen(lib.c_str(), RTLD_NOW);
    if (!h) { fprintf(stderr, "dlopen failed: %s\n", dlerror()); return 1; }
    Api api;
    if (!(sym(h, "ctr_engine_create", &api.create) && sym(h, "ctr_engine_destroy", &api.destroy) &&
          sym(h, "ctr_engine_load_weights", &api.load) && sym(h, "ctr_engine_model_info", &api.info) &&
          sym(h, "ctr_predict_async", &api.predict_async) && sym(h, "ctr_wait", &api.wait) &&
          sym(h, "ctr_alloc_pinned", &api.alloc_pinned) && sym(h, "ctr_free_pinned", &api.free_pinned) &&
          sym(h, "ctr_last_error", &api.last_error)))
        return 1;

    ctr_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.struct_size = sizeof(cfg);
    cfg.n_devices = n_devices;
    for (int i = 0; i < n_devices; ++i) cfg.device_ids[i] = i;
    cfg.max_candidates = max_size;
    cfg.workers_per_device = workers;
    ctr_engine* e = nullptr;
    if (api.create(&cfg, &e) != CTR_OK) { fprintf(stderr, "ctr_engine_create: %s\n", api.last_error()); return 1; }
    if (api.load(e, weights.c_str()) != CTR_OK) { fprintf(stderr, "ctr_engine_load_weights: %s\n", api.last_error()); return 1; }
    ctr_model_info mi;
    api.info(e, &mi);
    const int F = mi.num_fields;
    const int n_dense = F > 13 ? 13 : 0;

    // shared read-only inputs: `threads` different slabs of max_size candidates in pinned memory
    const size_t slab = (size_t)max_size * F;
    int64_t* ids = nullptr;
    float* wts = nullptr;
    if (api.alloc_pinned(e, (int64_t)(slab * threads * sizeof(int64_t)), (void**)&ids) != CTR_OK ||
        api.alloc_pinned(e, (int64_t)(slab * threads * sizeof(float)), (void**)&wts) != CTR_OK) {
        fprintf(stderr, "ctr_alloc_pinned: %s\n", api.last_error());
        return 1;
    }
    for (size_t c = 0; c < (size_t)max_size * threads; ++c)
        for (int f = 0; f < F; ++f) {
            const uint64_t k = c * 64 + f;
            if (f < n_dense) {
                ids[c * F + f] = f + 1;
                wts[c * F + f] = (float)std::log1p(std::exp(2.0 * unit(k) - 1.0));
            } else {
                ids[c * F + f] = (n_dense + 1) + (int64_t)(f - n_dense) * rows_per_field + (int64_t)(mix64(k ^ 0xABCDEFull) % (uint64_t)rows_per_field);
                wts[c * F + f] = 1.0f;
            }
        }
    constexpr int kRing = 64;
    std::vector<float*> outs(std::max(threads, kRing));
    for (auto& o : outs)
        if (api.alloc_pinned(e, (int64_t)max_size * sizeof(float), (void**)&o) != CTR_OK) { fprintf(stderr, "ctr_alloc_pinned: %s\n", api.last_error()); return 1; }

    std::string table = "| candidates / request | closed p50 ms | closed p99 ms | closed requests/s | closed candidates/s | open-loop load | open requests/s | open p50 ms | open p99 ms |\n"
                        "|---:|---:|---:|---:|---:|---:|---:|---:|---:|\n";
    std::atomic<int> failures{0};
    for (int C = min_size; C <= max_size; C *= 2) {
        // bound the work per size (the large sizes move 30 MB per request)
        const int n_timed = (int)std::max<long>(threads * 50L, std::min<long>(requests, (1L << 28) / C));
        const int n_warm = (int)std::min<long>(warmup, std::max<long>(threads * 5L, n_timed / 10));

        // ---------------- closed loop ----------------
        std::vector<double> lat;
        std::mutex mu;
        const int per_thread = (n_timed + threads - 1) / threads, warm_per_thread = (n_warm + threads - 1) / threads;
        std::vector<Clock::time_point> t_begin(threads), t_end(threads);
        auto driver = [&](int t) {
            std::vector<double> mine;
            mine.reserve(per_thread);
            for (int j = 0; j < warm_per_thread + per_thread; ++j) {
                if (j == warm_per_thread) t_begin[t] = Clock::now();
                const auto t0 = Clock::now();
                ctr_ticket tk;
                int rc = api.predict_async(e, ids + t * slab, wts + t * slab, C, outs[t], nullptr, nullptr, &tk);
                if (rc == CTR_OK) rc = api.wait(e, tk, -1);
                if (rc != CTR_OK) { ++failures; fprintf(stderr, "request failed: %s\n", api.last_error()); return; }
                if (j >= warm_per_thread) mine.push_back(std::chrono::duration<double, std::milli>(Clock::now() - t0).count());
            }
            t_end[t] = Clock::now();
            std::lock_guard<std::mutex> lk(mu);
            lat.insert(lat.end(), mine.begin(), mine.end());
        };
        {
            std::vector<std::thread> th;
            for (int t = 0; t < threads; ++t) th.emplace_back(driver, t);
            for (auto& t : th) t.join();
        }
        if (failures) break;
        double span = 0;
        for (int t = 0; t < threads; ++t) span = std::max(span, std::chrono::duration<double>(t_end[t] - t_begin[t]).count());
        const double closed_rps = (double)per_thread * threads / span;
        const double c50 = pct(lat, 0.50), c99 = pct(lat, 0.99);

        // ---------------- open loop: Poisson arrivals at load x closed-loop capacity ----------------
        const double rate = load * closed_rps;
        std::vector<double> olat;
        std::mutex omu;
        std::atomic<long> done{0};
        std::vector<Flight> ring(kRing);
        for (int i = 0; i < kRing; ++i) { ring[i].out = outs[i]; ring[i].sink = &olat; ring[i].sink_mu = &omu; ring[i].done = &done; ring[i].failures = &failures; }
        const long n_open = n_timed + n_warm;
        auto t_next = Clock::now();
        Clock::time_point open_begin = t_next;
        uint64_t seed = 0x1234567ull + (uint64_t)C;
        for (long j = 0; j < n_open; ++j) {
            if (j == n_warm) { open_begin = t_next; std::lock_guard<std::mutex> lk(omu); olat.clear(); }
            const double gap = -std::log(1.0 - unit(seed++)) / rate;                          // exponential inter-arrival
            t_next += std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(gap));
            while (Clock::now() < t_next) { /* spin: sleeping is too coarse at 40k requests/s */ }
            Flight& f = ring[j % kRing];
            while (f.busy.load(std::memory_order_acquire)) { /* 64 requests already in flight: the wait counts, arrival time is fixed */ }
            f.busy.store(1, std::memory_order_relaxed);
            f.scheduled = t_next;
            ctr_ticket tk;
            const int t = (int)(j % threads);
            if (api.predict_async(e, ids + t * slab, wts + t * slab, C, f.out, on_done, &f, &tk) != CTR_OK) {
                ++failures;
                fprintf(stderr, "request failed: %s\n", api.last_error());
                break;
            }
        }
        if (failures) break;
        for (auto& f : ring) while (f.busy.load(std::memory_order_acquire)) std::this_thread::yield();
        const double open_span = std::chrono::duration<double>(Clock::now() - open_begin).count();
        std::vector<double> otimed;
        {
            std::lock_guard<std::mutex> lk(omu);
            otimed = olat;                       // completions after the warm-up mark (a few warm-up stragglers included)
        }
        const double o50 = pct(otimed, 0.50), o99 = pct(otimed, 0.99);
        char line[512];
        snprintf(line, sizeof(line), "| %d | %.3f | %.3f | %.0f | %.3g | %.0f %% | %.0f | %.3f | %.3f |\n", C, c50, c99, closed_rps, closed_rps * C,
                 load * 100, (double)(n_open - n_warm) / open_span, o50, o99);
        table += line;
        fputs(line, stdout);
        fflush(stdout);
    }
    if (!out_path.empty() && !failures) {
        FILE* fo = fopen(out_path.c_str(), "w");
        if (fo) {
            fprintf(fo, "# Request stream sweep (BASELINE config 5), %d x B200, %s\n\n", n_devices, n_devices > 1 ? "candidates sharded over the GPUs inside one engine" : "one engine");
            fprintf(fo, "Through the C ABI from pinned host buffers (reference layout, %d B per candidate over PCIe); window = inputs exist -> ordered\n"
                        "scores on the host. Closed loop: %d driver threads x sequential requests (DCNClient.java:213-224). Open loop: Poisson arrivals at\n"
                        "the stated fraction of the closed-loop request rate, latency from the SCHEDULED arrival to the completion callback.\n"
                        "Up to %d timed requests per size after %d warm-up (fewer at the largest sizes: at most 2^28 candidates per size).\n\n",
                    F * 12, threads, requests, warmup);
            fputs(table.c_str(), fo);
            fclose(fo);
        }
    }
    for (auto& o : outs) api.free_pinned(e, o);
    api.free_pinned(e, ids);
    api.free_pinned(e, wts);
    api.destroy(e);
    dlclose(h);
    return failures ? 1 : 0;
}
