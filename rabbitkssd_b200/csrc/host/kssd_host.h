// Host-side glue shared by the drop-in shims (sketch_gpu.cpp, dist_gpu.cpp).
//
// The shims keep the reference's five hot signatures (src/sketch.h:62-66, src/dist.h:35-37) and are compiled
// AGAINST THE REFERENCE'S OWN HEADERS (-I<RabbitKSSD>/src): a maintainer drops the two .cpp files next to
// sketch.cpp / dist.cpp and links libkssd_b200.so.  Everything device-side goes through the
// C-ABI in include/kssd_b200.h; there is no CPU implementation of the hot loops in here.
//
// Devices: one kssd_ctx per GPU.  Default: every visible device (CUDA_VISIBLE_DEVICES applies); KSSD_B200_DEVICES="0,2"
// selects a subset, KSSD_B200_DEVICE=<n> a single device.  Sketching deals the input files over the
// pool; alldist / dist shard the query rows over it with the index replicated, one host thread per GPU
// writing the reference's per-thread result sub-files (src/dist.cpp:152-168).
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "kssd_b200.h"

namespace kssd_host {

inline void check(int rc, const char *what) {
    if (rc != KSSD_OK) {
        fprintf(stderr, "ERROR: %s: %s\n", what, kssd_b200_last_error());
        exit(1);
    }
}

// The reference reports errors with a message on stderr and exit(1) (e.g. src/dist.cpp:41-43); the shims do the same with
// the library's text.
inline const std::vector<int> &device_ids(bool fatal = true) {
    static std::vector<int> ids;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!ids.empty()) return ids;
    const char *many = getenv("KSSD_B200_DEVICES");
    const char *one = getenv("KSSD_B200_DEVICE");
    if (one && *one && !(many && *many)) {
        ids.push_back(atoi(one));
        return ids;
    }
    int n = 0;
    if (kssd_b200_device_count(&n) != KSSD_OK) {
        if (fatal) check(KSSD_E_CUDA, "libkssd_b200");
        return ids;  // empty: no usable device (early, non-fatal probe)
    }
    if (many && *many && strcmp(many, "all")) {
        for (const char *p = many; *p;) {
            char *end = nullptr;
            const long v = strtol(p, &end, 10);
            if (end == p) break;
            ids.push_back((int)v);
            p = *end == ',' ? end + 1 : end;
        }
    }
    if (ids.empty())
        for (int i = 0; i < n; i++) ids.push_back(i);
    return ids;
}
inline size_t device_count() { return device_ids().size(); }
// contexts are created on first use: a job that needs one GPU initialises one.  Slots have their own locks, so several
// contexts can be created at the same time (contexts_up_to).
struct ContextSlot {
    std::mutex mu;
    kssd_ctx *ctx = nullptr;
};
inline kssd_ctx *context_at(size_t i, bool fatal = true) {
    static std::mutex mu;
    static std::vector<ContextSlot *> pool;
    const std::vector<int> &ids = device_ids(fatal);
    ContextSlot *slot = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        while (pool.size() < ids.size()) pool.push_back(new ContextSlot());
        if (i < pool.size()) slot = pool[i];
    }
    if (!slot) {
        if (!fatal) return nullptr;
        fprintf(stderr, "ERROR: libkssd_b200: device slot %zu of %zu\n", i, ids.size());
        exit(1);
    }
    std::lock_guard<std::mutex> lock(slot->mu);
    if (!slot->ctx && kssd_b200_init(&slot->ctx, ids[i]) != KSSD_OK) {
        slot->ctx = nullptr;
        if (!fatal) return nullptr;
        fprintf(stderr, "ERROR: libkssd_b200: %s\n", kssd_b200_last_error());
        exit(1);
    }
    return slot->ctx;
}
// the first G contexts, created concurrently (creating one takes ~1 s)
inline std::vector<kssd_ctx *> contexts_up_to(size_t G) {
    std::vector<kssd_ctx *> out(G, nullptr);
    std::vector<std::thread> th;
    for (size_t g = 1; g < G; g++) th.emplace_back([&out, g] { out[g] = context_at(g); });
    if (G) out[0] = context_at(0);
    for (std::thread &t : th) t.join();
    return out;
}
// How many GPUs a job is worth: bringing one in costs ~1 s (context) plus a copy of the index, so another GPU is used only
// per `per_gpu` units of work (bytes to sketch, cells to compare); `env` overrides per_gpu (tests force 2 GPUs with it).
inline size_t gpus_for(double work, double per_gpu, const char *env) {
    if (const char *e = getenv(env)) per_gpu = atof(e);
    if (!(per_gpu > 0)) per_gpu = 1;
    const double want = 1.0 + work / per_gpu;
    const size_t have = device_count();
    return want >= (double)have ? have : (size_t)want;
}
inline kssd_ctx *context() { return context_at(0); }

// CUDA context creation takes ~0.9 s on a B200 box -- as long as sketching a thousand bacterial genomes.  The shims start it
// on a background thread when the program is loaded (a static object in sketch_gpu.cpp), so that it overlaps with argument
// parsing, the read of the .shuf file and the stat() of the input list; context_at() then finds the context ready or waits
// for the same mutex.  KSSD_B200_LAZY_INIT=1 turns it off (e.g. `rabbit_kssd info`, which never needs a GPU, then pays nothing).
struct EarlyInit {
    std::thread t;
    EarlyInit() {
        if (getenv("KSSD_B200_LAZY_INIT")) return;
        t = std::thread([] { context_at(0, false); });  // a host without a GPU still runs shuffle / info / convert
    }
    ~EarlyInit() {
        if (t.joinable()) t.join();
    }
};

// ---- what transSketches left behind ------------------------------------------------------------
// `alldist -i <genome list>` and `dist -r <genome list>` run sketch -> transSketches -> index_tridist / index_dist in ONE
// process; the reference writes the 1 GiB dense .index, reads it back and prefix-sums 2^28 counters (src/dist.cpp:83-130).
// Here the CSR that transSketches built is still resident on device 0: the distance step uses it when it is asked for the
// very files transSketches wrote, and the dense .index is written by a background thread meanwhile (the format is frozen,
// so the files are still produced; the thread is joined before anyone reads them and at exit).
struct IndexCache {
    bool valid = false;
    bool use64 = false;
    std::string index_file, dict_file;
    size_t n_genomes = 0;
};
inline IndexCache &index_cache() {
    static IndexCache c;
    return c;
}

struct PendingWrites {
    std::mutex mu;
    std::vector<std::pair<std::string, std::thread>> jobs;
    void add(const std::string &path, std::thread &&t) {
        std::lock_guard<std::mutex> lock(mu);
        jobs.emplace_back(path, std::move(t));
    }
    void wait(const std::string *path = nullptr) {  // null: all
        std::lock_guard<std::mutex> lock(mu);
        for (auto &j : jobs)
            if ((!path || j.first == *path) && j.second.joinable()) j.second.join();
    }
};
inline PendingWrites &pending_writes() {
    static PendingWrites *p = [] {
        PendingWrites *q = new PendingWrites();
        atexit([] { pending_writes().wait(); });
        return q;
    }();
    return *p;
}

// Row boundaries b[0..parts] giving every part the same cost of the n x n triangle: row i costs its n-1-i cells
// (src/dist.cpp:207 scans j in (i, n)) plus row_cost cell-equivalents of per-row work (rabbitkssd_b200/multigpu.py).
inline std::vector<size_t> triangle_bounds(size_t n, size_t parts, double row_cost) {
    std::vector<size_t> b(1, 0);
    const double m = (double)n + row_cost;
    const double total = (double)n * m - (double)n * ((double)n + 1) / 2.0;
    for (size_t p = 1; p < parts; p++) {
        const double target = total * (double)p / (double)parts;
        const double disc = (2 * m - 1) * (2 * m - 1) - 8 * target;
        double r = ((2 * m - 1) - std::sqrt(disc > 0 ? disc : 0)) / 2.0;
        size_t ri = r <= 0 ? 0 : (size_t)std::llround(r);
        if (ri < b.back()) ri = b.back();
        if (ri > n) ri = n;
        b.push_back(ri);
    }
    b.push_back(n);
    return b;
}

inline std::vector<size_t> rect_bounds(size_t n, size_t parts) {
    std::vector<size_t> b(1, 0);
    for (size_t p = 0; p < parts; p++) b.push_back(b.back() + n / parts + (p < n % parts ? 1 : 0));
    return b;
}

}  // namespace kssd_host
