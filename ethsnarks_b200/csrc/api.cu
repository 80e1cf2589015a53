// C ABI of libb200snark.so (declared in include/b200snark.h): context, key ingest, and the prove orchestration.
// Host side only — every arithmetic operation runs in the kernels of msm_*.cu / ntt.cu; there is no CPU fallback.
#include "../../include/b200snark.h"
#include "internal.hpp"
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <algorithm>
#include <memory>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <nvtx3/nvToolsExt.h>      // header-only; the ranges are no-ops unless a profiler injects itself

using namespace b200;

// ------------------------------------------------------------------------------------------------ peer-device workers
struct b200_ctx::Worker::Impl {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<void()> job;
    bool has_job = false, busy = false, quit = false;
    std::exception_ptr err;
};
b200_ctx::Worker::Worker(int device) : impl(new Impl()) {
    impl->th = std::thread([this, device] {
        cudaSetDevice(device);
        std::unique_lock<std::mutex> lk(impl->m);
        for (;;) {
            impl->cv.wait(lk, [this] { return impl->has_job || impl->quit; });
            if (impl->quit) return;
            std::function<void()> job = std::move(impl->job);
            impl->has_job = false;
            lk.unlock();
            std::exception_ptr e;
            try { job(); } catch (...) { e = std::current_exception(); }
            lk.lock();
            impl->err = e;
            impl->busy = false;
            impl->cv.notify_all();
        }
    });
}
b200_ctx::Worker::~Worker() {
    { std::lock_guard<std::mutex> lk(impl->m); impl->quit = true; }
    impl->cv.notify_all();
    impl->th.join();
    delete impl;
}
void b200_ctx::Worker::start(std::function<void()> job) {
    std::lock_guard<std::mutex> lk(impl->m);
    impl->job = std::move(job);
    impl->has_job = true;
    impl->busy = true;
    impl->err = nullptr;
    impl->cv.notify_all();
}
void b200_ctx::Worker::wait() {
    std::unique_lock<std::mutex> lk(impl->m);
    impl->cv.wait(lk, [this] { return !impl->busy; });
    if (impl->err) { std::exception_ptr e = impl->err; impl->err = nullptr; std::rethrow_exception(e); }
}

// ------------------------------------------------------------------------------------------------ ctx plumbing
void* b200_ctx::scratch(const char* name, size_t bytes) {
    Slot& s = slots[lane ? std::string(name) + "#" + std::to_string(lane) : std::string(name)];
    if (s.cap < bytes) {
        if (s.p) { cudaDeviceSynchronize(); cudaFree(s.p); s.p = nullptr; s.cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&s.p, want);
        if (e != cudaSuccess) {
            e = cudaMalloc(&s.p, bytes);
            want = bytes;
        }
        if (e != cudaSuccess) {
            s.p = nullptr;
            (void)cudaGetLastError();   // clear the sticky-free allocation error
            throw std::runtime_error(std::string("cudaMalloc(") + name + ", " + std::to_string(bytes) + " B): " + cudaGetErrorString(e));
        }
        s.cap = want;
    }
    return s.p;
}
void b200_ctx::t_begin(const char* name) {
    if (!timing_enabled) return;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, cur());
    timing_events.push_back({name, {a, b}});
    nvtx_ids.push_back((unsigned long long)nvtxRangeStartA(name));
}
void b200_ctx::t_end() {
    if (!timing_enabled) return;
    // close the most recent event pair that has not been closed yet (phases nest like a stack)
    for (size_t i = timing_events.size(); i-- > 0;) {
        if (timing_events[i].first.empty() || timing_events[i].first[0] != '#') {
            cudaEventRecord(timing_events[i].second.second, cur());
            if (i < nvtx_ids.size()) nvtxRangeEnd((nvtxRangeId_t)nvtx_ids[i]);
            timing_events[i].first = "#" + timing_events[i].first;
            return;
        }
    }
}
void b200_ctx::t_collect() {
    timings.clear();
    timing_starts.clear();
    for (auto& ev : timing_events) {
        float ms = 0.f, at = 0.f;
        if (!ev.first.empty() && ev.first[0] == '#') cudaEventElapsedTime(&ms, ev.second.first, ev.second.second);
        cudaEventElapsedTime(&at, timing_events.front().second.first, ev.second.first);
        timing_starts.push_back(at);
        timings.push_back({ev.first.empty() ? ev.first : ev.first.substr(ev.first[0] == '#' ? 1 : 0), ms});
    }
    for (auto& ev : timing_events) { cudaEventDestroy(ev.second.first); cudaEventDestroy(ev.second.second); }
    timing_events.clear();
    nvtx_ids.clear();
}

namespace {

struct Guard {   // turns C++ exceptions into status codes at the ABI boundary
    b200_ctx* ctx;
    template <typename F>
    int run(F&& f) {
        try {
            if (ctx) B200_CUDA_CHECK(cudaSetDevice(ctx->device));
            return f();
        } catch (const std::invalid_argument& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_ARG;
        } catch (const std::domain_error& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_FORMAT;
        } catch (const std::exception& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_CUDA;
        }
    }
};

struct DevBuf {   // RAII device allocation
    void* p = nullptr;
    explicit DevBuf(size_t bytes) { B200_CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(bytes, 16))); }
    ~DevBuf() { cudaFree(p); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

inline size_t aff_bytes(int group) { return group == 1 ? 64 : 128; }
inline size_t xyzz_bytes(int group) { return group == 1 ? 128 : 256; }
inline size_t aff_limbs(int group) { return group == 1 ? 8 : 16; }

// host affine points + flags -> device array with (0,0) as the infinity encoding
void upload_points(b200_ctx* ctx, int group, const uint64_t* xy, const uint8_t* inf, size_t n, void* d_out) {
    const size_t ab = aff_bytes(group);
    if (!n) return;
    if (!inf) {
        B200_CUDA_CHECK(cudaMemcpyAsync(d_out, xy, n * ab, cudaMemcpyHostToDevice, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        return;
    }
    std::vector<uint8_t> tmp(n * ab);
    memcpy(tmp.data(), xy, n * ab);
    for (size_t i = 0; i < n; ++i)
        if (inf[i]) memset(tmp.data() + i * ab, 0, ab);
    B200_CUDA_CHECK(cudaMemcpy(d_out, tmp.data(), n * ab, cudaMemcpyHostToDevice));
}
void download_points(b200_ctx* ctx, int group, const void* d_in, size_t n, uint64_t* xy, uint8_t* inf) {
    const size_t ab = aff_bytes(group);
    if (!n) return;
    B200_CUDA_CHECK(cudaMemcpyAsync(xy, d_in, n * ab, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    const size_t nl = aff_limbs(group);
    for (size_t i = 0; i < n; ++i) {
        uint64_t o = 0;
        for (size_t l = 0; l < nl; ++l) o |= xy[i * nl + l];
        if (inf) inf[i] = (o == 0);
        if (o == 0 && group) {   // canonical output for infinity: (0, 1) in Montgomery form is not needed; keep (0,0)
        }
    }
}

// result of g_finish: affine point followed by a u32 flag
void fetch_finished(b200_ctx* ctx, int group, const void* d_fin, uint64_t* xy, uint8_t* inf) {
    uint8_t buf[128 + 16];
    const size_t ab = aff_bytes(group);
    B200_CUDA_CHECK(cudaMemcpyAsync(buf, d_fin, ab + 4, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    memcpy(xy, buf, ab);
    uint32_t f;
    memcpy(&f, buf + ab, 4);
    *inf = (uint8_t)(f != 0);
}

}  // namespace

struct b200_bases {
    b200_ctx* ctx = nullptr;
    int group = 1;
    size_t n = 0;
    void* d = nullptr;   // affine Montgomery, (0,0) = infinity
    void* table = nullptr;   // optional window table [W][n] (b200_bases_precompute)
    int table_c = 0;
};

extern "C" {

const char* b200_strerror(int code) {
    switch (code) {
        case B200_OK: return "ok";
        case B200_ERR_CUDA: return "CUDA error (no device, launch or allocation failure)";
        case B200_ERR_ARG: return "invalid argument";
        case B200_ERR_FORMAT: return "malformed proving key";
        case B200_ERR_STATE: return "call not valid in this state (e.g. prove before circuit_load)";
        default: return "unknown error";
    }
}

static int init_one(int dev, b200_ctx** out);
static void shutdown_one(b200_ctx* ctx);

// n_devices > 1: one context that owns several GPUs of the box ("one handle = one device set").  Keys loaded
// on it are split 1/N per device and b200_prove runs on all of them from this one process, the per-device partial sums
// travelling to the first device over peer memory (NVLink); everything else (standalone MSM, NTT, generator) runs on the
// first device.
int b200_init(const int* device_ids, int n_devices, b200_ctx** out) {
    if (!out) return B200_ERR_ARG;
    *out = nullptr;
    if (n_devices <= 1) return init_one((device_ids && n_devices == 1) ? device_ids[0] : 0, out);
    if (!device_ids || n_devices > 64) return B200_ERR_ARG;
    b200_ctx* root = nullptr;
    int rc = init_one(device_ids[0], &root);
    if (rc) return rc;
    cudaEventCreateWithFlags(&root->ev_group, cudaEventDisableTiming);
    for (int i = 1; i < n_devices; ++i) {
        b200_ctx* p = nullptr;
        rc = init_one(device_ids[i], &p);
        if (rc) { b200_shutdown(root); return rc; }
        cudaEventCreateWithFlags(&p->ev_group, cudaEventDisableTiming);
        root->peers.push_back(p);
        root->workers.push_back(new b200_ctx::Worker(device_ids[i]));
    }
    // peer access in both directions between every pair; a pair without it cannot join one context
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < n_devices; ++j) {
            if (device_ids[i] == device_ids[j]) continue;   // a device may be listed more than once (one-GPU tests of this path)
            int can = 0;
            cudaDeviceCanAccessPeer(&can, device_ids[i], device_ids[j]);
            if (!can) { root->last_error = "b200_init: no peer access between the devices"; b200_shutdown(root); return B200_ERR_CUDA; }
            cudaSetDevice(device_ids[i]);
            cudaError_t e = cudaDeviceEnablePeerAccess(device_ids[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { b200_shutdown(root); return B200_ERR_CUDA; }
            (void)cudaGetLastError();
        }
    cudaSetDevice(device_ids[0]);
    *out = root;
    return B200_OK;
}
static int init_one(int dev, b200_ctx** out) {
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { (void)cudaGetLastError(); return B200_ERR_CUDA; }
    if (dev < 0 || dev >= count) return B200_ERR_ARG;
    if (cudaSetDevice(dev) != cudaSuccess) return B200_ERR_CUDA;
    b200_ctx* c = new b200_ctx();
    c->device = dev;
    // Stream priorities order the tails: lane 1 (the G2 MSM, the longest chain) first, then the main stream (H chain and
    // H MSM), then the two witness-only G1 lanes, whose accumulate blocks fill the SMs the latency-bound reductions of
    // the higher-priority chains leave idle.
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);   // numerically lower = higher priority
    auto prio = [&](int level) { return std::max(prio_hi, std::min(prio_lo, prio_lo - level)); };
    int lv[4] = {1, 2, 0, 0};                               // main, G2 lane, A lane, L lane
    if (const char* e = getenv("B200_LANE_PRIO")) {         // tuning knob: four digits, e.g. "1211"
        for (int i = 0; i < 4 && e[i] >= '0' && e[i] <= '9'; ++i) lv[i] = e[i] - '0';
    }
    if (const char* e = getenv("B200_G2_ACC_SMEM_KB")) c->g2_acc_smem = atoi(e) * 1024;
    if (cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio(lv[0])) != cudaSuccess) { delete c; return B200_ERR_CUDA; }
    c->lanes[0] = c->stream;
    for (int i = 1; i < 4; ++i) cudaStreamCreateWithPriority(&c->lanes[i], cudaStreamNonBlocking, prio(lv[i]));
    for (int i = 0; i < 4; ++i) cudaEventCreateWithFlags(&c->lane_done[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_witness, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_shared, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_front, cudaEventDisableTiming);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    *out = c;
    return B200_OK;
}
void b200_shutdown(b200_ctx* ctx) {
    if (!ctx) return;
    for (auto* w : ctx->workers) delete w;
    for (auto* p : ctx->peers) { if (p->ev_group) { cudaSetDevice(p->device); cudaEventDestroy(p->ev_group); } shutdown_one(p); }
    if (ctx->ev_group) { cudaSetDevice(ctx->device); cudaEventDestroy(ctx->ev_group); }
    shutdown_one(ctx);
}
static void shutdown_one(b200_ctx* ctx) {
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (auto& kv : ctx->slots) cudaFree(kv.second.p);
    ctx->t_collect();
    for (int i = 1; i < 4; ++i) cudaStreamDestroy(ctx->lanes[i]);
    for (int i = 0; i < 4; ++i) cudaEventDestroy(ctx->lane_done[i]);
    cudaEventDestroy(ctx->ev_witness);
    cudaEventDestroy(ctx->ev_shared);
    cudaEventDestroy(ctx->ev_front);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}
const char* b200_last_error(b200_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }
void* b200_stream(b200_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
unsigned long long b200_launch_count(b200_ctx* ctx) {
    if (!ctx) return 0;
    unsigned long long n = ctx->launches;
    for (auto* p : ctx->peers) n += p->launches;
    return n;
}
int b200_device_count(b200_ctx* ctx) { return ctx ? 1 + (int)ctx->peers.size() : 0; }
int b200_set_option(b200_ctx* ctx, const char* name, long value) {
    if (!ctx || !name) return B200_ERR_ARG;
    for (auto* p : ctx->peers) b200_set_option(p, name, value);
    if (!strcmp(name, "window")) ctx->window_override = (int)value;
    else if (!strcmp(name, "chunk")) ctx->chunk_override = (int)value;
    else if (!strcmp(name, "timing")) ctx->timing_enabled = value != 0;
    else if (!strcmp(name, "precompute")) ctx->precompute = (int)value;
    else if (!strcmp(name, "overlap")) ctx->overlap = (int)value;
    else if (!strcmp(name, "reduce_1d")) ctx->reduce_1d = (int)value;
    else if (!strcmp(name, "g2_pair")) ctx->g2_pair = (int)value;
    else if (!strcmp(name, "ntt_radix4")) ctx->ntt_radix4 = (int)value;
    else if (!strcmp(name, "shared_sort")) ctx->shared_sort = (int)value;
    else if (!strcmp(name, "pair_reduce")) ctx->pair_reduce = (int)value;
    else if (!strcmp(name, "reduce_fold")) ctx->reduce_fold = (int)value;
    else if (!strcmp(name, "pair_tail")) ctx->pair_tail = (int)value;
    else if (!strcmp(name, "finalize_chunks")) ctx->finalize_chunks = (int)value;
    else if (!strcmp(name, "finalize_big_seg")) ctx->finalize_big_seg = (int)value;
    else if (!strcmp(name, "split_chains")) ctx->split_chains = (int)value;
    else return B200_ERR_ARG;
    return B200_OK;
}
int b200_timing_count(b200_ctx* ctx) { return ctx ? (int)ctx->timings.size() : 0; }
int b200_timing_get(b200_ctx* ctx, int i, char* name, size_t cap, float* ms) {
    if (!ctx || i < 0 || i >= (int)ctx->timings.size()) return B200_ERR_ARG;
    if (name && cap) { strncpy(name, ctx->timings[i].first.c_str(), cap - 1); name[cap - 1] = 0; }
    if (ms) *ms = ctx->timings[i].second;
    return B200_OK;
}

int b200_timing_start(b200_ctx* ctx, int i, float* ms) {
    if (!ctx || !ms || i < 0 || i >= (int)ctx->timing_starts.size()) return B200_ERR_ARG;
    *ms = ctx->timing_starts[i];
    return B200_OK;
}

int b200_imad_peak(b200_ctx* ctx, int reps, double* macs_per_s) {
    if (!ctx || !macs_per_s || reps < 1) return B200_ERR_ARG;
    return Guard{ctx}.run([&] { *macs_per_s = imad_peak(ctx, reps); return (int)B200_OK; });
}

// ------------------------------------------------------------------------------------------------ sub-operations
int b200_field_op(b200_ctx* ctx, int field, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
    if (!ctx || !a || !out || field < 0 || field > 1 || op < 0 || op > 11) return B200_ERR_ARG;
    if ((op <= 2 || op >= 9) && !b) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        DevBuf da(n * 32), db(n * 32), dout(n * 32);
        B200_CUDA_CHECK(cudaMemcpyAsync(da.p, a, n * 32, cudaMemcpyHostToDevice, ctx->stream));
        if (b) B200_CUDA_CHECK(cudaMemcpyAsync(db.p, b, n * 32, cudaMemcpyHostToDevice, ctx->stream));
        field_op(ctx, field, op, da.p, db.p, dout.p, n);
        B200_CUDA_CHECK(cudaMemcpyAsync(out, dout.p, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        return (int)B200_OK;
    });
}

static int group_add(b200_ctx* ctx, int group, const uint64_t* a_xy, const uint8_t* a_inf, const uint64_t* b_xy, const uint8_t* b_inf,
                     uint64_t* out_xy, uint8_t* out_inf, size_t n) {
    if (!ctx || !a_xy || !b_xy || !out_xy) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        const size_t ab = aff_bytes(group);
        DevBuf da(n * ab), db(n * ab), dout(n * ab);
        upload_points(ctx, group, a_xy, a_inf, n, da.p);
        upload_points(ctx, group, b_xy, b_inf, n, db.p);
        g_add_affine(ctx, group, da.p, db.p, dout.p, n);
        download_points(ctx, group, dout.p, n, out_xy, out_inf);
        return (int)B200_OK;
    });
}
int b200_g1_add(b200_ctx* ctx, const uint64_t* a_xy, const uint8_t* a_inf, const uint64_t* b_xy, const uint8_t* b_inf, uint64_t* out_xy,
                uint8_t* out_inf, size_t n) {
    return group_add(ctx, 1, a_xy, a_inf, b_xy, b_inf, out_xy, out_inf, n);
}
int b200_g2_add(b200_ctx* ctx, const uint64_t* a_xy, const uint8_t* a_inf, const uint64_t* b_xy, const uint8_t* b_inf, uint64_t* out_xy,
                uint8_t* out_inf, size_t n) {
    return group_add(ctx, 2, a_xy, a_inf, b_xy, b_inf, out_xy, out_inf, n);
}

// sum of n affine points in one kernel (the "reduce" of SURVEY 8e: the all-gathered per-GPU partial sums of a sharded MSM)
static int group_sum(b200_ctx* ctx, int group, const uint64_t* xy, const uint8_t* inf, size_t n, uint64_t* out_xy, uint8_t* out_inf) {
    if (!ctx || (!xy && n) || !out_xy || !out_inf || n > (1u << 20)) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        DevBuf d(std::max<size_t>(1, n) * aff_bytes(group)), fin(aff_bytes(group) + 256);
        upload_points(ctx, group, xy, inf, n, d.p);
        g_finish(ctx, group, nullptr, 0, d.p, (int)n, fin.p, false);
        fetch_finished(ctx, group, fin.p, out_xy, out_inf);
        return (int)B200_OK;
    });
}
int b200_g1_sum(b200_ctx* ctx, const uint64_t* xy, const uint8_t* inf, size_t n, uint64_t* out_xy, uint8_t* out_inf) {
    return group_sum(ctx, 1, xy, inf, n, out_xy, out_inf);
}
int b200_g2_sum(b200_ctx* ctx, const uint64_t* xy, const uint8_t* inf, size_t n, uint64_t* out_xy, uint8_t* out_inf) {
    return group_sum(ctx, 2, xy, inf, n, out_xy, out_inf);
}

static int fixed_base(b200_ctx* ctx, int group, const uint64_t* scalars, size_t n, uint64_t* out_xy, uint8_t* out_inf) {
    if (!ctx || !scalars || !out_xy) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        DevBuf ds(n * 32), dout(n * aff_bytes(group));
        B200_CUDA_CHECK(cudaMemcpyAsync(ds.p, scalars, n * 32, cudaMemcpyHostToDevice, ctx->stream));
        g_fixed_base(ctx, group, ds.p, n, dout.p);
        download_points(ctx, group, dout.p, n, out_xy, out_inf);
        return (int)B200_OK;
    });
}
int b200_g1_fixed_base(b200_ctx* ctx, const uint64_t* s, size_t n, uint64_t* out_xy, uint8_t* out_inf) { return fixed_base(ctx, 1, s, n, out_xy, out_inf); }
int b200_g2_fixed_base(b200_ctx* ctx, const uint64_t* s, size_t n, uint64_t* out_xy, uint8_t* out_inf) { return fixed_base(ctx, 2, s, n, out_xy, out_inf); }

int b200_ntt(b200_ctx* ctx, uint64_t* data, uint32_t log_m, int mode) {
    if (!ctx || !data || mode < 0 || mode > 3 || log_m > 28) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        const size_t m = (size_t)1 << log_m;
        Domain dom;
        domain_create(ctx, log_m, &dom);
        std::shared_ptr<Domain> guard(&dom, [](Domain* d) { domain_free(d); });
        DevBuf d(m * 32);
        B200_CUDA_CHECK(cudaMemcpyAsync(d.p, data, m * 32, cudaMemcpyHostToDevice, ctx->stream));
        ctx->t_begin("ntt");
        ntt_natural(ctx, dom, d.p, mode);
        ctx->t_end();
        B200_CUDA_CHECK(cudaMemcpyAsync(data, d.p, m * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        ctx->t_collect();
        return (int)B200_OK;
    });
}

// ------------------------------------------------------------------------------------------------ base sets + MSM
int b200_bases_upload(b200_ctx* ctx, int group, const uint64_t* xy, const uint8_t* inf, size_t n, b200_bases** out) {
    if (!ctx || !out || (group != 1 && group != 2) || (n && !xy)) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        std::unique_ptr<b200_bases> b(new b200_bases());
        b->ctx = ctx; b->group = group; b->n = n;
        B200_CUDA_CHECK(cudaMalloc(&b->d, std::max<size_t>(n * aff_bytes(group), 16)));
        try { upload_points(ctx, group, xy, inf, n, b->d); } catch (...) { cudaFree(b->d); throw; }
        *out = b.release();
        return (int)B200_OK;
    });
}
int b200_bases_generate(b200_ctx* ctx, int group, const uint64_t* scalars, size_t n, b200_bases** out) {
    if (!ctx || !out || (group != 1 && group != 2) || (n && !scalars)) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        std::unique_ptr<b200_bases> b(new b200_bases());
        b->ctx = ctx; b->group = group; b->n = n;
        DevBuf ds(n * 32);
        B200_CUDA_CHECK(cudaMemcpyAsync(ds.p, scalars, n * 32, cudaMemcpyHostToDevice, ctx->stream));
        B200_CUDA_CHECK(cudaMalloc(&b->d, std::max<size_t>(n * aff_bytes(group), 16)));
        try {
            g_fixed_base(ctx, group, ds.p, n, b->d);
            B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        } catch (...) { cudaFree(b->d); throw; }
        *out = b.release();
        return (int)B200_OK;
    });
}
int b200_bases_precompute(b200_bases* b, int window) {
    if (!b || window < 0 || window > 24) return B200_ERR_ARG;
    b200_ctx* ctx = b->ctx;
    return Guard{ctx}.run([&] {
        if (b->table) { cudaStreamSynchronize(ctx->stream); cudaFree(b->table); b->table = nullptr; b->table_c = 0; }
        if (!b->n) return (int)B200_OK;
        const int c = window ? std::max(4, window) : msm_table_window(b->n);
        const int W = (255 + c - 1) / c;
        B200_CUDA_CHECK(cudaMalloc(&b->table, (size_t)W * b->n * aff_bytes(b->group)));
        g_precompute(ctx, b->group, b->d, b->n, c, b->table);
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        b->table_c = c;
        return (int)B200_OK;
    });
}
int b200_bases_download(b200_bases* b, uint64_t* xy, uint8_t* inf) {
    if (!b || !xy) return B200_ERR_ARG;
    return Guard{b->ctx}.run([&] { download_points(b->ctx, b->group, b->d, b->n, xy, inf); return (int)B200_OK; });
}
size_t b200_bases_count(b200_bases* b) { return b ? b->n : 0; }
void b200_bases_free(b200_bases* b) {
    if (!b) return;
    cudaSetDevice(b->ctx->device);
    cudaStreamSynchronize(b->ctx->stream);
    cudaFree(b->d);
    cudaFree(b->table);
    delete b;
}

int b200_msm(b200_ctx* ctx, b200_bases* bases, const uint64_t* scalars, size_t n, int on_device, uint64_t* out_xy, uint8_t* out_inf) {
    if (!ctx || !bases || !out_xy || !out_inf || n > bases->n || (n && !scalars)) return B200_ERR_ARG;
    return Guard{ctx}.run([&] {
        const int g = bases->group;
        const void* d_scalars = scalars;
        ctx->t_begin("msm.total");
        if (!on_device) {
            void* ds = ctx->scratch("msm_scalars_in", n * 32);
            B200_CUDA_CHECK(cudaMemcpyAsync(ds, scalars, n * 32, cudaMemcpyHostToDevice, ctx->stream));
            d_scalars = ds;
        }
        void* acc = ctx->scratch("msm_result", xyzz_bytes(g) + 256);
        void* fin = (char*)acc + xyzz_bytes(g);
        if (bases->table) msm_run(ctx, g, bases->table, d_scalars, nullptr, n, true, acc, bases->table_c, bases->n);
        else msm_run(ctx, g, bases->d, d_scalars, nullptr, n, true, acc);
        g_finish(ctx, g, acc, 1, nullptr, 0, fin, false);
        ctx->t_end();
        fetch_finished(ctx, g, fin, out_xy, out_inf);
        ctx->t_collect();
        return (int)B200_OK;
    });
}

static int msm_host(b200_ctx* ctx, int group, const uint64_t* bases_xy, const uint8_t* inf, const uint64_t* scalars, size_t n,
                    uint64_t* out_xy, uint8_t* out_inf) {
    b200_bases* b = nullptr;
    int rc = b200_bases_upload(ctx, group, bases_xy, inf, n, &b);
    if (rc) return rc;
    rc = b200_msm(ctx, b, scalars, n, 0, out_xy, out_inf);
    b200_bases_free(b);
    return rc;
}
int b200_msm_g1(b200_ctx* ctx, const uint64_t* bases_xy, const uint8_t* inf, const uint64_t* scalars, size_t n, uint64_t out_xy[8], uint8_t* out_inf) {
    return msm_host(ctx, 1, bases_xy, inf, scalars, n, out_xy, out_inf);
}
int b200_msm_g2(b200_ctx* ctx, const uint64_t* bases_xy, const uint8_t* inf, const uint64_t* scalars, size_t n, uint64_t out_xy[16], uint8_t* out_inf) {
    return msm_host(ctx, 2, bases_xy, inf, scalars, n, out_xy, out_inf);
}

}  // extern "C"
