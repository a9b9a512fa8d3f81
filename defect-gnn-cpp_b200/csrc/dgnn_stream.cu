// dgnn_stream.cu -- dgnn_build_stream: a job of more structures than one edge arena holds (BASELINE config 5: 10^6
// structures per node, ~4 TB of graphs) built chunk by chunk through TWO arenas.  The reference's batch driver is a
// per-file loop that builds one structure's neighbour list at a time and drops it (src/preprocess/preprocess_betti.cpp:
// 37-79); here a producer thread pulls chunk c from the caller's source and builds it into arena c & 1 while the
// caller's sink consumes chunk c - 1 from the other arena -- upload, count and fill of the next chunk overlap whatever
// the consumer does with the current one (device work on its own stream, D2H, host reductions).  Host code only.
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>

#include "../../include/dgnn_cabi.h"

extern "C" dgnn_ctx* dgnn_internal_twin(dgnn_ctx* ctx);                    // dgnn_api.cu
extern "C" void dgnn_internal_set_error(dgnn_ctx* ctx, const char* msg);   // dgnn_api.cu

namespace {
struct Slot {
    dgnn_graphs g;
    dgnn_build_info bi;
    int64_t chunk = 0;
    int rc = DGNN_OK;
    bool full = false;
    bool end = false;  // the source is exhausted (or failed: rc)
};
}  // namespace

extern "C" int dgnn_build_stream(dgnn_ctx* ctx, dgnn_stream_source source, void* source_user, dgnn_stream_sink sink,
                                 void* sink_user, dgnn_stream_info* info) {
    if (!ctx || !source) return DGNN_ERR_INVALID_ARG;
    dgnn_stream_info I;
    memset(&I, 0, sizeof(I));
    dgnn_ctx* arena[2] = {ctx, dgnn_internal_twin(ctx)};
    if (!arena[1]) {
        dgnn_internal_set_error(ctx, "dgnn_build_stream: could not create the second arena");
        return DGNN_ERR_CUDA;
    }
    Slot slots[2];
    std::mutex mu;
    std::condition_variable cv;
    bool stop = false;  // the consumer gave up: the producer must not start another chunk
    const auto t0 = std::chrono::steady_clock::now();

    std::thread producer([&] {
        for (int64_t c = 0;; ++c) {
            Slot& s = slots[c & 1];
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return !s.full || stop; });
                if (stop) return;
            }
            dgnn_batch b;
            memset(&b, 0, sizeof(b));
            const int have = source(source_user, c, &b);
            int rc = DGNN_OK;
            dgnn_graphs g;
            dgnn_build_info bi;
            memset(&g, 0, sizeof(g));
            memset(&bi, 0, sizeof(bi));
            if (have > 0) {
                rc = dgnn_build(arena[c & 1], &b, &g);
                if (rc == DGNN_OK) dgnn_last_build_info(arena[c & 1], &bi);
                else if ((c & 1) == 1) dgnn_internal_set_error(ctx, dgnn_last_error(arena[1]));
            } else if (have < 0) {
                rc = DGNN_ERR_INVALID_ARG;
                dgnn_internal_set_error(ctx, "dgnn_build_stream: the source callback failed");
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                s.g = g;
                s.bi = bi;
                s.chunk = c;
                s.rc = rc;
                s.end = have <= 0 || rc != DGNN_OK;
                s.full = true;
            }
            cv.notify_all();
            if (have <= 0 || rc != DGNN_OK) return;
        }
    });

    int rc = DGNN_OK;
    for (int64_t c = 0;; ++c) {
        Slot& s = slots[c & 1];
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return s.full; });
        }
        if (s.end) {
            rc = s.rc;
            break;
        }
        const auto k0 = std::chrono::steady_clock::now();
        const int kr = sink ? sink(sink_user, s.chunk, &s.g, arena[c & 1]) : 0;
        I.sink_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - k0).count();
        I.n_chunks += 1;
        I.n_structs += s.g.n_structs;
        I.n_atoms += s.g.n_atoms;
        I.n_edges += s.g.n_edges;
        I.build_device_ms += s.bi.total_ms;
        I.h2d_ms += s.bi.h2d_ms;
        I.kernel_launches += s.bi.kernel_launches;
        {
            std::lock_guard<std::mutex> lk(mu);
            s.full = false;
            if (kr != 0) stop = true;
        }
        cv.notify_all();
        if (kr != 0) {
            rc = DGNN_ERR_INVALID_ARG;
            dgnn_internal_set_error(ctx, "dgnn_build_stream: the sink callback failed");
            break;
        }
    }
    {
        std::lock_guard<std::mutex> lk(mu);
        stop = true;
    }
    cv.notify_all();
    producer.join();
    I.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (info) *info = I;
    return rc;
}
