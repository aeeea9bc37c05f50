// lsc_sharded_index: a table whose rows are spread over several GPUs, driven from ONE host thread.
//
// This is the handle a Java datasource binds to: QueryStepDataSource.fetchData is called from a single JVM
// thread (langstream-agents/langstream-ai-agents/src/main/java/com/datastax/oss/streaming/ai/datasource/
// QueryStepDataSource.java:22-37) and has to answer from all GPUs of the box.  The library therefore keeps, per
// device, one worker thread with its own stream, an lsc_index shard and an lsc_exchange endpoint; an API call
// hands every worker its part and returns when all are done.  A search is one fused scan launch per GPU per pass:
// each GPU scans its shard, stores its candidates into the peers' receive buffers over NVLink and merges what the
// peers stored into its own (topk_fuse.cuh); worker 0 copies the merged result back.
// Write side: JdbcWriter semantics (UPDATE, INSERT when no row matched, DELETE on null value -
// langstream-agents/langstream-vector-agents/src/main/java/ai/langstream/agents/vector/jdbc/JdbcWriter.java:139-179)
// with new ids placed on the shard that holds the fewest rows.
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "exchange.h"

namespace lsc {
void index_set_sm_budget(lsc_index* ix, int sms);  // index.cu
int index_sm_count(const lsc_index* ix);
}  // namespace lsc

using namespace lsc;

namespace {

struct Worker {
    int device = 0;
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = false, stop = false;
    int rc = LSC_OK;
    std::string err;
    // per-device staging of a search
    cudaStream_t stream = nullptr;
    float* d_q = nullptr;
    int64_t cap_q = 0;  // floats
    int64_t* d_ids = nullptr;
    float* d_scores = nullptr;
    int64_t cap_out = 0;

    void loop() {
        cudaSetDevice(device);
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            cv.wait(lk, [&] { return has_job || stop; });
            if (stop) return;
            lk.unlock();
            const int r = job();
            std::string e = r != LSC_OK ? std::string(lsc_last_error()) : std::string();
            lk.lock();
            rc = r;
            err = std::move(e);
            has_job = false;
            done = true;
            cv.notify_all();
        }
    }
    void submit(std::function<int()> f) {
        std::lock_guard<std::mutex> lk(m);
        job = std::move(f);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return done; });
        done = false;
        return rc;
    }
};

template <typename T>
int grow(T** p, int64_t* cap, int64_t need, int64_t floor_) {
    if (need <= *cap) return LSC_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const int64_t c = std::max(need, floor_);
    LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(p), static_cast<size_t>(c) * sizeof(T)));
    *cap = c;
    return LSC_OK;
}

}  // namespace

struct lsc_sharded_index {
    int dim = 0;
    int ndev = 0;
    std::vector<lsc_index*> shards;
    std::vector<lsc_exchange*> xs;
    std::vector<Worker*> workers;
    std::unordered_map<int64_t, int> shard_of;  // id -> shard
    std::vector<int64_t> counts;                // rows per shard
    std::mutex mu;                              // one API call at a time

    // run f(shard index) on every worker, return the first failure
    int on_all(const std::function<int(int)>& f) {
        for (int i = 0; i < ndev; ++i) workers[i]->submit([&f, i] { return f(i); });
        int rc = LSC_OK;
        std::string msg;
        for (int i = 0; i < ndev; ++i) {
            const int r = workers[i]->wait();
            if (r != LSC_OK && rc == LSC_OK) {
                rc = r;
                msg = workers[i]->err;
            }
        }
        if (rc != LSC_OK) set_error(msg);
        return rc;
    }
};

extern "C" {

void lsc_sharded_index_destroy(lsc_sharded_index* s) {
    if (s == nullptr) return;
    for (Worker* w : s->workers) {
        if (w == nullptr) continue;
        if (w->th.joinable()) {
            {
                std::lock_guard<std::mutex> lk(w->m);
                w->stop = true;
                w->cv.notify_all();
            }
            w->th.join();
        }
        cudaSetDevice(w->device);
        if (w->stream) cudaStreamSynchronize(w->stream);
        if (w->d_q) cudaFree(w->d_q);
        if (w->d_ids) cudaFree(w->d_ids);
        if (w->d_scores) cudaFree(w->d_scores);
        if (w->stream) cudaStreamDestroy(w->stream);
        delete w;
    }
    for (lsc_index* ix : s->shards) lsc_index_destroy(ix);
    for (lsc_exchange* x : s->xs) lsc_exchange_destroy(x);
    delete s;
}

int lsc_sharded_index_create(int32_t dim, int64_t capacity_per_shard, const int* devices, int32_t ndev,
                             lsc_sharded_index** out) {
    if (out == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_create: out is null");
    *out = nullptr;
    if (devices == nullptr || ndev < 1 || ndev > TKX_MAX_WORLD)
        return fail(LSC_E_INVALID, "lsc_sharded_index_create: between 1 and 8 devices");
    lsc_sharded_index* s = new lsc_sharded_index();
    s->dim = dim;
    s->ndev = ndev;
    s->counts.assign(ndev, 0);
    auto body = [&]() -> int {
        for (int i = 0; i < ndev; ++i) {
            lsc_index* ix = nullptr;
            LSC_TRY(lsc_index_create(dim, capacity_per_shard, devices[i], &ix));
            s->shards.push_back(ix);
            lsc_exchange* x = nullptr;
            LSC_TRY(lsc_exchange_create(devices[i], i, ndev, 1024, &x));
            s->xs.push_back(x);
        }
        LSC_TRY(lsc_exchange_connect_local(s->xs.data(), ndev));
        // Shards that share a GPU (a handle built over one device, as the single-GPU tests do) must be able to
        // run their scans side by side - each waits for the other's candidates - so they split the SMs.
        for (int i = 0; i < ndev; ++i) {
            int same = 0;
            for (int j = 0; j < ndev; ++j) same += devices[j] == devices[i] ? 1 : 0;
            if (same > 1) index_set_sm_budget(s->shards[i], std::max(2, (index_sm_count(s->shards[i]) / same) & ~1));
        }
        for (int i = 0; i < ndev; ++i) {
            Worker* w = new Worker();
            w->device = devices[i];
            s->workers.push_back(w);
            LSC_CUDA(cudaSetDevice(devices[i]));
            LSC_CUDA(cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking));
            w->th = std::thread([w] { w->loop(); });
        }
        return LSC_OK;
    };
    const int rc = body();
    if (rc != LSC_OK) {
        const std::string msg = lsc_last_error();
        lsc_sharded_index_destroy(s);
        set_error(msg);
        return rc;
    }
    *out = s;
    return LSC_OK;
}

int32_t lsc_sharded_index_shards(const lsc_sharded_index* s) { return s ? s->ndev : 0; }

lsc_index* lsc_sharded_index_shard(lsc_sharded_index* s, int32_t i) {
    return (s != nullptr && i >= 0 && i < s->ndev) ? s->shards[i] : nullptr;
}

int64_t lsc_sharded_index_size(const lsc_sharded_index* s) {
    if (s == nullptr) return 0;
    int64_t n = 0;
    for (lsc_index* ix : s->shards) n += lsc_index_size(ix);
    return n;
}

int lsc_sharded_index_upsert(lsc_sharded_index* s, const int64_t* ids, const float* vecs, int32_t n) {
    if (s == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_upsert: handle is null");
    if (n < 0) return fail(LSC_E_INVALID, "lsc_sharded_index_upsert: negative n");
    if (n == 0) return LSC_OK;
    if (ids == nullptr || vecs == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_upsert: null pointer");
    std::lock_guard<std::mutex> lock(s->mu);
    // route: an id that exists stays on its shard, a new id goes to the shard with the fewest rows
    std::vector<std::vector<int64_t>> sid(s->ndev);
    std::vector<std::vector<int>> srow(s->ndev);
    std::vector<int64_t> counts = s->counts;
    std::unordered_map<int64_t, int> fresh;
    for (int i = 0; i < n; ++i) {
        int sh;
        auto it = s->shard_of.find(ids[i]);
        if (it != s->shard_of.end()) {
            sh = it->second;
        } else {
            auto f = fresh.find(ids[i]);
            if (f != fresh.end()) {
                sh = f->second;
            } else {
                sh = static_cast<int>(std::min_element(counts.begin(), counts.end()) - counts.begin());
                fresh[ids[i]] = sh;
                ++counts[sh];
            }
        }
        sid[sh].push_back(ids[i]);
        srow[sh].push_back(i);
    }
    const int dim = s->dim;
    const int rc = s->on_all([&](int sh) -> int {
        const int m = static_cast<int>(sid[sh].size());
        if (m == 0) return LSC_OK;
        std::vector<float> rows(static_cast<size_t>(m) * dim);
        for (int j = 0; j < m; ++j) memcpy(&rows[static_cast<size_t>(j) * dim], vecs + static_cast<size_t>(srow[sh][j]) * dim, dim * sizeof(float));
        return lsc_index_upsert(s->shards[sh], sid[sh].data(), rows.data(), m);
    });
    // the map follows what the shards now hold, also after a partial failure
    for (const auto& kv : fresh) s->shard_of[kv.first] = kv.second;
    for (int sh = 0; sh < s->ndev; ++sh) s->counts[sh] = lsc_index_size(s->shards[sh]);
    return rc;
}

int lsc_sharded_index_delete(lsc_sharded_index* s, const int64_t* ids, int32_t n) {
    if (s == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_delete: handle is null");
    if (n < 0) return fail(LSC_E_INVALID, "lsc_sharded_index_delete: negative n");
    if (n == 0) return LSC_OK;
    if (ids == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_delete: null pointer");
    std::lock_guard<std::mutex> lock(s->mu);
    std::vector<std::vector<int64_t>> sid(s->ndev);
    for (int i = 0; i < n; ++i) {
        auto it = s->shard_of.find(ids[i]);
        if (it == s->shard_of.end()) continue;  // unknown ids are ignored
        sid[it->second].push_back(ids[i]);
        s->shard_of.erase(it);
    }
    const int rc = s->on_all([&](int sh) -> int {
        if (sid[sh].empty()) return LSC_OK;
        return lsc_index_delete(s->shards[sh], sid[sh].data(), static_cast<int32_t>(sid[sh].size()));
    });
    for (int sh = 0; sh < s->ndev; ++sh) s->counts[sh] = lsc_index_size(s->shards[sh]);
    return rc;
}

int lsc_sharded_index_search(lsc_sharded_index* s, const float* queries, int32_t B, int32_t k, int64_t* out_ids,
                             float* out_scores) {
    if (s == nullptr) return fail(LSC_E_INVALID, "lsc_sharded_index_search: handle is null");
    if (B < 0) return fail(LSC_E_INVALID, "lsc_sharded_index_search: negative B");
    if (k < 1 || k > TKX_MAX_XK) return fail(LSC_E_UNSUPPORTED, "lsc_sharded_index_search: k must be in [1,32]");
    if (B == 0) return LSC_OK;
    if (queries == nullptr || out_ids == nullptr || out_scores == nullptr)
        return fail(LSC_E_INVALID, "lsc_sharded_index_search: null pointer");
    std::lock_guard<std::mutex> lock(s->mu);
    const int dim = s->dim;
    const int64_t nout = static_cast<int64_t>(B) * k;
    // Phase 1: every shard sizes its buffers.  Allocation synchronises the device; it must be over on ALL shards
    // before the first scan starts, because a running scan waits for its peers (shards may share a GPU).
    LSC_TRY(s->on_all([&](int sh) -> int {
        Worker* w = s->workers[sh];
        LSC_CUDA(cudaSetDevice(w->device));
        LSC_TRY(grow(&w->d_q, &w->cap_q, static_cast<int64_t>(B) * dim, 1 << 16));
        if (nout > w->cap_out) {
            int64_t c1 = w->cap_out, c2 = w->cap_out;
            LSC_TRY(grow(&w->d_ids, &c1, nout, 4096));
            LSC_TRY(grow(&w->d_scores, &c2, nout, 4096));
            w->cap_out = std::min(c1, c2);
        }
        return lsc_index_reserve(s->shards[sh], std::min<int32_t>(B, s->xs[sh]->xB), k);
    }));
    // Phase 2: upload, scan + exchange + merge (one fused launch per pass), read back.
    return s->on_all([&](int sh) -> int {
        Worker* w = s->workers[sh];
        LSC_CUDA(cudaSetDevice(w->device));
        LSC_CUDA(cudaMemcpyAsync(w->d_q, queries, static_cast<size_t>(B) * dim * sizeof(float), cudaMemcpyHostToDevice, w->stream));
        if (s->ndev == 1)
            LSC_TRY(lsc_index_search_dev(s->shards[sh], w->d_q, B, k, w->d_ids, w->d_scores, w->stream));
        else
            LSC_TRY(lsc_index_search_sharded_dev(s->shards[sh], s->xs[sh], w->d_q, B, k, w->d_ids, w->d_scores, w->stream));
        if (sh == 0) {  // every shard ends up with the merged result; one copy goes back
            LSC_CUDA(cudaMemcpyAsync(out_ids, w->d_ids, static_cast<size_t>(nout) * sizeof(int64_t), cudaMemcpyDeviceToHost, w->stream));
            LSC_CUDA(cudaMemcpyAsync(out_scores, w->d_scores, static_cast<size_t>(nout) * sizeof(float), cudaMemcpyDeviceToHost, w->stream));
        }
        return lsc_index_check(s->shards[sh], w->stream);  // synchronises the stream
    });
}

}  // extern "C"
