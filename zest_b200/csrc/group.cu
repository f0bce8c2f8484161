// Several GPUs behind one call: zg_group_* (declared in include/zest_gpu.h).
//
// zest is ONE process (/root/reference/src/xet_bridge.zig:231-264: one bridge reconstructs a file; src/parallel_download.zig:
// 91-134: one task group fans a file's terms out), so a drop-in that wants all GPUs of a box cannot ask the caller to
// start a process per GPU.  A group owns one context per device and one worker thread per device; a verify call
// partitions the batch over the devices (greedy longest-processing-time by compressed bytes, SURVEY.md 8e -- xorbs are
// independent, nothing is exchanged between GPUs, no collective), every worker runs zg_verify_xorbs on its part through
// its own context (its own lanes, streams and PCIe link), and verdicts / roots / chunk hashes come back in the
// caller's order.  Host code only: the kernels are the ones behind zg_verify_xorbs.
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/zest_gpu.h"
#include "kernels.h"

using zg::set_last_error;

struct zg_group {
    struct Worker {
        zg_ctx* ctx = nullptr;
        int device = -1;
        std::thread th;
        std::mutex mu;
        std::condition_variable cv;
        std::function<void()> job;  // set by the caller, cleared by the worker
        bool has_job = false, quit = false;
        int rc = ZG_OK;
        char err[512] = "";
    };
    std::vector<Worker*> w;
    std::mutex call_mu;  // one group call at a time (each context still serves other callers through its lanes)
    std::mutex done_mu;
    std::condition_variable done_cv;
    int pending = 0;
};

namespace {

void worker_main(zg_group* g, zg_group::Worker* w) {
    cudaSetDevice(w->device);
    for (;;) {
        std::function<void()> job;
        {
            std::unique_lock<std::mutex> lk(w->mu);
            w->cv.wait(lk, [&] { return w->has_job || w->quit; });
            if (w->quit) return;
            job = std::move(w->job);
            w->has_job = false;
        }
        job();
        {
            std::lock_guard<std::mutex> lk(g->done_mu);
            g->pending--;
        }
        g->done_cv.notify_all();
    }
}

// Run one job per worker (null = none) and wait for all of them.
void run_all(zg_group* g, std::vector<std::function<void()>>& jobs) {
    {
        std::lock_guard<std::mutex> lk(g->done_mu);
        g->pending = 0;
        for (auto& j : jobs) g->pending += j ? 1 : 0;
    }
    for (size_t i = 0; i < jobs.size(); i++) {
        if (!jobs[i]) continue;
        {
            std::lock_guard<std::mutex> lk(g->w[i]->mu);
            g->w[i]->job = std::move(jobs[i]);
            g->w[i]->has_job = true;
        }
        g->w[i]->cv.notify_one();
    }
    std::unique_lock<std::mutex> lk(g->done_mu);
    g->done_cv.wait(lk, [&] { return g->pending == 0; });
}

}  // namespace

extern "C" {

int zg_partition_xorbs(const size_t* lens, size_t n, int n_parts, int32_t* part) {
    if (n_parts < 1 || (n && (!lens || !part))) return set_last_error(ZG_ERR_INVALID_ARG, "zg_partition_xorbs: bad argument");
    std::vector<size_t> order(n);
    for (size_t i = 0; i < n; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return lens[a] > lens[b]; });  // ties: lower index first
    std::vector<unsigned long long> load((size_t)n_parts, 0);
    for (size_t i : order) {
        int best = 0;
        for (int p = 1; p < n_parts; p++)
            if (load[p] < load[best]) best = p;  // ties: lower part first
        part[i] = best;
        load[best] += lens[i];
    }
    return ZG_OK;
}

int zg_group_init(const int* devices, int n_devices, const zg_opts* opts, zg_group** out) {
    if (!out) return set_last_error(ZG_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return set_last_error(ZG_ERR_NO_DEVICE, "no CUDA device (%s); libzest_gpu has no CPU path", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    }
    if (n_devices <= 0) n_devices = count;
    if (n_devices > 64) return set_last_error(ZG_ERR_LIMIT, "at most 64 devices per group");
    zg_group* g = new (std::nothrow) zg_group();
    if (!g) return set_last_error(ZG_ERR_OOM, "out of host memory");
    int prev = -1;
    cudaGetDevice(&prev);
    for (int i = 0; i < n_devices; i++) {
        const int dev = devices ? devices[i] : i;
        zg_ctx* c = nullptr;
        int r = zg_init(dev, opts, &c);
        if (r != ZG_OK) {
            zg_group_shutdown(g);
            if (prev >= 0) cudaSetDevice(prev);
            return r;
        }
        auto* w = new zg_group::Worker();
        w->ctx = c;
        w->device = dev;
        g->w.push_back(w);
        w->th = std::thread(worker_main, g, w);
    }
    if (prev >= 0) cudaSetDevice(prev);
    *out = g;
    return ZG_OK;
}

void zg_group_shutdown(zg_group* g) {
    if (!g) return;
    for (auto* w : g->w) {
        {
            std::lock_guard<std::mutex> lk(w->mu);
            w->quit = true;
        }
        w->cv.notify_one();
        if (w->th.joinable()) w->th.join();
        zg_shutdown(w->ctx);
        delete w;
    }
    delete g;
}

int zg_group_size(const zg_group* g) { return g ? (int)g->w.size() : 0; }

zg_ctx* zg_group_ctx(zg_group* g, int i) { return (g && i >= 0 && (size_t)i < g->w.size()) ? g->w[(size_t)i]->ctx : nullptr; }

int zg_group_verify_xorbs(zg_group* g, const uint8_t* const* blobs, const size_t* lens, const uint8_t* expected, size_t n,
                          int8_t* verdict, uint8_t* roots, uint8_t* chunk_hashes, size_t chunk_hash_cap, uint64_t* bytes_verified) {
    if (!g || g->w.empty()) return set_last_error(ZG_ERR_INVALID_ARG, "null or empty group");
    if (n && (!blobs || !lens || !verdict || !roots)) return set_last_error(ZG_ERR_INVALID_ARG, "null argument");
    if (bytes_verified) *bytes_verified = 0;
    if (n == 0) return ZG_OK;
    std::lock_guard<std::mutex> call(g->call_mu);
    const int P = (int)g->w.size();
    std::vector<int32_t> part(n);
    zg_partition_xorbs(lens, n, P, part.data());
    std::vector<std::vector<size_t>> mine((size_t)P);
    for (size_t i = 0; i < n; i++) mine[(size_t)part[i]].push_back(i);  // ascending: each part keeps the caller's order

    // Chunk hashes come back in the caller's xorb order, so every xorb's place in that array must be known before the
    // parts run: count the chunks of each xorb first (header walk only, in the workers, in parallel).
    std::vector<uint32_t> n_chunks;
    std::vector<size_t> hash_at;
    if (chunk_hashes) {
        n_chunks.assign(n, 0);
        std::vector<std::function<void()>> jobs((size_t)P);
        for (int p = 0; p < P; p++) {
            if (mine[(size_t)p].empty()) continue;
            jobs[(size_t)p] = [&, p] {
                for (size_t i : mine[(size_t)p]) {
                    uint32_t c = 0;
                    if (zg_xorb_index(blobs[i], lens[i], 0, &c, nullptr, nullptr, nullptr, nullptr, nullptr) != ZG_OK) c = 0;
                    n_chunks[i] = c;  // a blob that cannot be walked contributes no hashes (as in zg_verify_xorbs)
                }
            };
        }
        run_all(g, jobs);
        hash_at.assign(n + 1, 0);
        for (size_t i = 0; i < n; i++) hash_at[i + 1] = hash_at[i] + n_chunks[i];
        if (hash_at[n] > chunk_hash_cap)
            return set_last_error(ZG_ERR_CAPACITY, "chunk_hashes holds %zu hashes, the batch has %zu", chunk_hash_cap, hash_at[n]);
    }

    std::vector<uint64_t> part_bytes((size_t)P, 0);
    std::vector<std::function<void()>> jobs((size_t)P);
    for (int p = 0; p < P; p++) {
        auto& idx = mine[(size_t)p];
        if (idx.empty()) continue;
        zg_group::Worker* w = g->w[(size_t)p];
        w->rc = ZG_OK;
        jobs[(size_t)p] = [&, p, w] {
            const auto& ix = mine[(size_t)p];
            const size_t m = ix.size();
            std::vector<const uint8_t*> b(m);
            std::vector<size_t> l(m);
            std::vector<uint8_t> ex(expected ? 32 * m : 0), rt(32 * m), ch;
            std::vector<int8_t> vd(m);
            size_t nh = 0;
            for (size_t k = 0; k < m; k++) {
                b[k] = blobs[ix[k]];
                l[k] = lens[ix[k]];
                if (expected) memcpy(ex.data() + 32 * k, expected + 32 * ix[k], 32);
                if (chunk_hashes) nh += n_chunks[ix[k]];
            }
            if (chunk_hashes) ch.resize(32 * nh + 32);
            uint64_t bv = 0;
            w->rc = zg_verify_xorbs(w->ctx, b.data(), l.data(), expected ? ex.data() : nullptr, m, vd.data(), rt.data(),
                                    chunk_hashes ? ch.data() : nullptr, nh, &bv);
            if (w->rc != ZG_OK) { strncpy(w->err, zg_last_error(), sizeof w->err - 1); w->err[sizeof w->err - 1] = 0; }
            part_bytes[(size_t)p] = bv;
            size_t hp = 0;
            for (size_t k = 0; k < m; k++) {
                verdict[ix[k]] = vd[k];
                memcpy(roots + 32 * ix[k], rt.data() + 32 * k, 32);
                if (chunk_hashes) {
                    memcpy(chunk_hashes + 32 * hash_at[ix[k]], ch.data() + 32 * hp, 32 * (size_t)n_chunks[ix[k]]);
                    hp += n_chunks[ix[k]];
                }
            }
        };
    }
    run_all(g, jobs);
    int rc = ZG_OK;
    uint64_t total = 0;
    for (int p = 0; p < P; p++) {
        total += part_bytes[(size_t)p];
        if (!mine[(size_t)p].empty() && g->w[(size_t)p]->rc != ZG_OK && rc == ZG_OK)
            rc = set_last_error(g->w[(size_t)p]->rc, "device %d: %s", g->w[(size_t)p]->device, g->w[(size_t)p]->err);
    }
    if (bytes_verified) *bytes_verified = total;
    return rc;
}

}  // extern "C"
