// C-ABI glue: error reporting, device queries and the engine dispatch of the loss.
#include "common.cuh"
#include "supcon_common.cuh"
#include "supcon_tc.cuh"

#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <utility>
#include <vector>

namespace scb {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool attr_done(const void* kern, int attr, int value) {
    // (kernel, attribute, device) -> last value applied; a few dozen entries, guarded by a mutex
    struct Key { const void* k; int a; int d; };
    static std::mutex mu;
    static std::vector<std::pair<Key, int>> seen;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    std::lock_guard<std::mutex> lock(mu);
    for (auto& e : seen)
        if (e.first.k == kern && e.first.a == attr && e.first.d == dev) {
            if (e.second == value) return true;
            e.second = value;
            return false;
        }
    seen.push_back({Key{kern, attr, dev}, value});
    return false;
}

bool pdl_enabled() {
    static const bool on = [] {
        // On unless SCB_PDL=0.  Worth ~1.6 us of a 70 us forward+backward on B200 when the step is replayed from a
        // CUDA graph (the dependent grid's prologue overlaps its predecessor's tail); nothing in eager mode.
        const char* e = getenv("SCB_PDL");
        return e == nullptr || e[0] != '0';
    }();
    return on;
}

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static int fp32_splits(int n_rows, int n_k) {
    const int row_tiles = (int)ceil_div(n_rows, 32), col_tiles = (int)ceil_div(n_k, 32);
    int s = (int)ceil_div(2 * sm_count(), row_tiles);
    if (s > col_tiles) s = col_tiles;
    if (s > 64) s = 64;
    if (s < 1) s = 1;
    return s;
}

static int check_problem(const scb_supcon_problem* p, int row_begin, int n_rows) {
    SCB_REQUIRE(p != nullptr && p->k != nullptr, "supcon: null problem / k");
    SCB_REQUIRE(p->n_views >= 1 && p->seg >= 1 && p->n_seg >= 1, "supcon: bad V/seg/n_seg");
    SCB_REQUIRE(p->row_layout == SCB_ROWS_VIEW_MAJOR || (p->row_layout == SCB_ROWS_SAMPLE_MAJOR && p->mode == SCB_MODE_ALL),
                "supcon: row_layout must be SCB_ROWS_VIEW_MAJOR, or SCB_ROWS_SAMPLE_MAJOR with mode `all`");
    SCB_REQUIRE((int64_t)p->n_k == (int64_t)p->n_views * p->seg * p->n_seg,
                "supcon: n_k (%d) != n_seg*n_views*seg", p->n_k);
    SCB_REQUIRE(p->dim >= 1 && p->ldk >= p->dim, "supcon: bad dim/ldk");
    SCB_REQUIRE(p->mode == SCB_MODE_ALL || p->mode == SCB_MODE_ONE, "Unknown mode: %d", p->mode);
    SCB_REQUIRE(p->mask_kind != SCB_MASK_LABELS || p->labels, "supcon: labels missing");
    SCB_REQUIRE(p->mask_kind != SCB_MASK_EXPLICIT || p->mask, "supcon: mask missing");
    SCB_REQUIRE(p->mask_kind >= 0 && p->mask_kind <= 2, "supcon: bad mask_kind");
    SCB_REQUIRE(p->scalar_dtype == SCB_F32 || p->scalar_dtype == SCB_BF16 || p->scalar_dtype == SCB_F16,
                "supcon: bad scalar_dtype %d", p->scalar_dtype);
    SCB_REQUIRE(row_begin >= 0 && n_rows >= 1 && row_begin + n_rows <= p->n_k, "supcon: bad row range");
    if (p->engine == SCB_ENGINE_FP32) {
        SCB_REQUIRE(p->k_dtype == SCB_F32, "supcon: ENGINE_FP32 needs an F32 contrast matrix");
    } else if (p->engine == SCB_ENGINE_BF16 || p->engine == SCB_ENGINE_BF16X3) {
        SCB_REQUIRE(p->k_dtype == SCB_BF16, "supcon: tensor-core engines need a BF16 contrast matrix");
        SCB_REQUIRE(p->engine != SCB_ENGINE_BF16X3 || p->k_lo, "supcon: ENGINE_BF16X3 needs k_lo");
    } else {
        set_error("supcon: unknown engine %d", p->engine);
        return SCB_ERR_ARG;
    }
    return SCB_OK;
}

}  // namespace scb

using namespace scb;

extern "C" const char* scb_last_error(void) { return g_err; }
extern "C" int scb_version(void) { return 100; }
extern "C" int64_t scb_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

extern "C" int scb_debug_set_trace(void* device_buffer) {
    tc_set_trace((long long*)device_buffer);
    return SCB_OK;
}

extern "C" int scb_debug_set_events(void* start, void* stop) {
    tc_set_events(start, stop);
    return SCB_OK;
}

extern "C" int64_t scb_stream_capture_id(void* stream_) {
    cudaStreamCaptureStatus status = cudaStreamCaptureStatusNone;
    unsigned long long id = 0;
    if (cudaStreamGetCaptureInfo((cudaStream_t)stream_, &status, &id) != cudaSuccess) { cudaGetLastError(); return -1; }
    return status == cudaStreamCaptureStatusActive ? (int64_t)id : 0;
}

extern "C" int scb_device_cc(void) {
    int dev = 0, major = 0, minor = 0;
    SCB_CUDA_OK(cudaGetDevice(&dev));
    SCB_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    SCB_CUDA_OK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
    return major * 10 + minor;
}

extern "C" int64_t scb_supcon_workspace_bytes(const scb_supcon_problem* p, int n_rows) {
    if (!p || n_rows < 1) return 0;
    int64_t fp32 = (int64_t)64 * 4 * n_rows * sizeof(float);
    int64_t tc = tc_workspace_bytes(p, n_rows);
    return (fp32 > tc ? fp32 : tc) + kFinScratchBytes + 256;
}

extern "C" int64_t scb_supcon_label_cache_bytes(const scb_supcon_problem* p) {
    if (!p || p->engine == SCB_ENGINE_FP32) return 0;
    return tc_label_cache_bytes(p);
}

extern "C" int scb_supcon_forward(const scb_supcon_problem* p, int row_begin, int n_rows,
                                  int n_anchor_total, float* out_lse, float* out_possum,
                                  float* out_npos, float* out_loss_sum, void* workspace,
                                  int64_t workspace_bytes, void* stream_) {
    if (int e = check_problem(p, row_begin, n_rows)) return e;
    SCB_REQUIRE(out_lse && out_possum && out_npos && out_loss_sum && workspace, "supcon_forward: null pointer");
    SCB_REQUIRE(n_anchor_total >= 1, "supcon_forward: n_anchor_total must be >= 1");
    SCB_REQUIRE(workspace_bytes >= scb_supcon_workspace_bytes(p, n_rows), "supcon_forward: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    // workspace = [finalize scratch (zeroed, self-cleaning) | engine scratch]
    float* part = (float*)((char*)workspace + kFinScratchBytes);
    int n_split = 1, flat_q = 0, flat_ct = 0;
    if (p->engine == SCB_ENGINE_FP32) {
        n_split = fp32_splits(n_rows, p->n_k);
        if (int e = fp32_forward(p, row_begin, n_rows, part, n_split, stream)) return e;
    } else {
        return tc_forward(p, row_begin, n_rows, n_anchor_total, out_lse, out_possum, out_npos, out_loss_sum, part, nullptr,
                          stream);
    }
    return launch_finalize(p, part, n_split, flat_q, flat_ct, row_begin, n_rows, 1.0 / (double)n_anchor_total, out_lse,
                           out_possum, out_npos, out_loss_sum, workspace, stream);
}

extern "C" int scb_supcon_forward_peers(const scb_supcon_problem* p, int row_begin, int n_rows, int n_anchor_total,
                                        float* out_possum, const scb_peer_channel* ch, void* workspace,
                                        int64_t workspace_bytes, void* stream_) {
    if (int e = check_problem(p, row_begin, n_rows)) return e;
    SCB_REQUIRE(out_possum && workspace && ch && ch->state, "supcon_forward_peers: null pointer");
    SCB_REQUIRE(ch->n_dst == p->n_seg && row_begin % n_rows == 0 && (int64_t)n_rows * p->n_seg == p->n_k,
                "supcon_forward_peers: one destination per segment, every rank holding n_k / n_seg rows");
    for (int w = 0; w < ch->n_dst; ++w)
        SCB_REQUIRE(ch->dst[w] && ch->flag[w], "supcon_forward_peers: destination %d is null", w);
    SCB_REQUIRE(n_anchor_total >= 1, "supcon_forward_peers: n_anchor_total must be >= 1");
    SCB_REQUIRE(workspace_bytes >= scb_supcon_workspace_bytes(p, n_rows), "supcon_forward_peers: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    float* part = (float*)((char*)workspace + kFinScratchBytes);
    int n_split = 1, flat_q = 0, flat_ct = 0;
    if (p->engine == SCB_ENGINE_FP32) {
        n_split = fp32_splits(n_rows, p->n_k);
        if (int e = fp32_forward(p, row_begin, n_rows, part, n_split, stream)) return e;
    } else {
        return tc_forward(p, row_begin, n_rows, n_anchor_total, nullptr, out_possum, nullptr, nullptr, part, ch, stream);
    }
    return launch_finalize(p, part, n_split, flat_q, flat_ct, row_begin, n_rows, 1.0 / (double)n_anchor_total, nullptr,
                           out_possum, nullptr, nullptr, workspace, stream, ch);
}

extern "C" int scb_supcon_backward(const scb_supcon_problem* p, int row_begin, int n_rows,
                                   const float* lse_all, const float* npos_all, const float* grad_out,
                                   int n_anchor_total, float* out_dk, int ld_dk, void* workspace,
                                   int64_t workspace_bytes, void* stream_) {
    if (int e = check_problem(p, row_begin, n_rows)) return e;
    SCB_REQUIRE(lse_all && npos_all && grad_out && out_dk, "supcon_backward: null pointer");
    SCB_REQUIRE(n_anchor_total >= 1 && ld_dk >= p->dim, "supcon_backward: bad n_anchor_total / ld_dk");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (p->engine == SCB_ENGINE_FP32)
        return fp32_backward(p, row_begin, n_rows, lse_all, npos_all, grad_out, n_anchor_total, out_dk,
                             ld_dk, stream);
    SCB_REQUIRE(workspace && workspace_bytes >= scb_supcon_workspace_bytes(p, n_rows),
                "supcon_backward: workspace too small");
    return tc_backward(p, row_begin, n_rows, lse_all, npos_all, grad_out, n_anchor_total, out_dk, ld_dk,
                       (char*)workspace + kFinScratchBytes, stream);
}

extern "C" int scb_supcon_backward_features(const scb_supcon_problem* p, int row_begin, int n_rows,
                                            const float* lse_all, const float* npos_all, const float* grad_out,
                                            int n_anchor_total, const void* features, int feat_dtype, int n_samples,
                                            int64_t stride_sample, int64_t stride_view, int normalize,
                                            const float* inv_norm, void* out_dfeatures, int64_t out_stride_sample,
                                            int64_t out_stride_view, void* workspace, int64_t workspace_bytes,
                                            void* stream_) {
    if (int e = check_problem(p, row_begin, n_rows)) return e;
    (void)npos_all;
    SCB_REQUIRE(lse_all && grad_out && features && out_dfeatures, "supcon_backward_features: null pointer");
    SCB_REQUIRE(!normalize || inv_norm, "supcon_backward_features: normalize needs inv_norm");
    SCB_REQUIRE(n_anchor_total >= 1, "supcon_backward_features: bad n_anchor_total");
    if (p->engine == SCB_ENGINE_FP32) {
        set_error("supcon_backward_features: the fp32 engine uses scb_supcon_backward + scb_supcon_finish");
        return SCB_ERR_UNSUPPORTED;
    }
    SCB_REQUIRE(workspace && workspace_bytes >= scb_supcon_workspace_bytes(p, n_rows),
                "supcon_backward_features: workspace too small");
    return tc_backward_features(p, row_begin, n_rows, lse_all, grad_out, n_anchor_total, features, feat_dtype,
                                n_samples, stride_sample, stride_view, normalize, inv_norm, out_dfeatures,
                                out_stride_sample, out_stride_view, (char*)workspace + kFinScratchBytes,
                                (cudaStream_t)stream_);
}
