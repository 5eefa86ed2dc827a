// Row bookkeeping shared by the loss engines.
#pragma once
#include "common.cuh"

namespace scb {

// view-major:   row = q*(V*seg) + v*seg + b     sample-major: row = q*(V*seg) + b*V + v
// sample id = q*seg + b in both                                  (include/simclr_b200.h, "Row order")
struct RowMap {
    int vseg;   // V * seg
    int seg;
    int nv;     // V
    int il;     // 1 = sample-major (views of a sample interleaved)
    __host__ __device__ static RowMap of(const scb_supcon_problem* p) {
        RowMap m; m.vseg = p->n_views * p->seg; m.seg = p->seg; m.nv = p->n_views;
        m.il = (p->row_layout == SCB_ROWS_SAMPLE_MAJOR) ? 1 : 0; return m;
    }
    __device__ __forceinline__ int sid(int row) const {
        const int q = row / vseg, rem = row - q * vseg;
        return q * seg + (il ? rem / nv : rem % seg);
    }
    __device__ __forceinline__ int view_of(int row) const {
        const int rem = row % vseg;
        return il ? rem % nv : rem / seg;
    }
    // the row that holds view v of sample `sample_id`
    __device__ __forceinline__ int row_of(int sample_id, int v) const {
        const int q = sample_id / seg, b = sample_id - q * seg;
        return q * vseg + (il ? b * nv + v : v * seg + b);
    }
    // the row that holds view vv of the same sample
    __device__ __forceinline__ int sibling(int row, int vv) const {
        const int v = view_of(row);
        return il ? row + (vv - v) : row + (vv - v) * seg;
    }
    // anchors: every row (`all`) or the view-0 rows (`one`)            losses.py:54-59
    __device__ __forceinline__ bool is_anchor(int row, int mode) const {
        return mode == SCB_MODE_ALL || view_of(row) == 0;
    }
};

// ---- peer-memory exchange (include/simclr_b200.h, scb_peer_channel)
struct PeerOut {
    char* dst[16];
    unsigned* flag[16];
    unsigned* state;     // [0] ticket, [1] exchange counter
    int n_dst;           // 0: no peers, plain local output
    __host__ static PeerOut none() { PeerOut o; o.n_dst = 0; o.state = nullptr; return o; }
    __host__ static PeerOut of(const scb_peer_channel* ch) {
        PeerOut o;
        o.n_dst = ch->n_dst;
        o.state = ch->state;
        for (int w = 0; w < 16; ++w) {
            o.dst[w] = w < ch->n_dst ? (char*)ch->dst[w] : nullptr;
            o.flag[w] = w < ch->n_dst ? ch->flag[w] : nullptr;
        }
        return o;
    }
};

// Called by every thread at the end of a producing kernel (after its last store to the peers).  The last
// CTA to arrive bumps the exchange counter and raises this rank's flag on every rank.  Ordering: each
// CTA fences its remote stores at system scope before taking the ticket; the flag stores are releases.
__device__ __forceinline__ void peer_signal(const PeerOut& po) {
    __syncthreads();                 // the CTA's stores happen-before thread 0's fence (fences are cumulative):
    if (threadIdx.x == 0) {          // one system-scope fence per CTA instead of one per thread
        __threadfence_system();
        const unsigned total = gridDim.x * gridDim.y * gridDim.z;
        if (atomicAdd(po.state, 1u) == total - 1) {
            po.state[0] = 0u;
            const unsigned e = po.state[1] + 1u;
            po.state[1] = e;
            __threadfence_system();
            for (int w = 0; w < po.n_dst; ++w)
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(po.flag[w]), "r"(e) : "memory");
        }
    }
}

// base mask m[sid_i, sid_j] before the self-contrast column is removed      losses.py:39-50
struct MaskSpec {
    const int64_t* labels;
    const float* mask;
    int kind;
    int n_samples;   // n_seg * seg
    __host__ __device__ static MaskSpec of(const scb_supcon_problem* p) {
        MaskSpec m; m.labels = p->labels; m.mask = p->mask; m.kind = p->mask_kind;
        m.n_samples = p->n_seg * p->seg; return m;
    }
    __device__ __forceinline__ float value(int sid_i, int sid_j) const {
        if (kind == SCB_MASK_SIMCLR) return sid_i == sid_j ? 1.f : 0.f;
        if (kind == SCB_MASK_LABELS) return labels[sid_i] == labels[sid_j] ? 1.f : 0.f;
        return mask[(int64_t)sid_i * n_samples + sid_j];
    }
};

// engines (supcon_generic.cu / supcon_tc.cu)
int fp32_forward(const scb_supcon_problem* p, int row_begin, int n_rows, float* part, int n_split,
                 cudaStream_t stream);
int fp32_backward(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                  const float* npos_all, const float* grad_out, int n_anchor_total, float* out_dk,
                  int ld_dk, cudaStream_t stream);
// flat_q > 0: tensor-core engine, row block rb (256 rows) owns slots of the CTAs overlapping its tiles
int launch_finalize(const scb_supcon_problem* p, const float* part, int n_split, int flat_q, int flat_ct,
                    int row_begin, int n_rows, double inv_anchors, float* out_lse, float* out_possum, float* out_npos,
                    float* out_loss_sum, void* scratch, cudaStream_t stream, const scb_peer_channel* ch = nullptr);
// Start of every loss workspace: [4096 bytes: ticket + per-CTA partial sums of the finalize kernel (fp32 engine)]
// [kTicketRegionBytes: row-block tickets of the tensor-core forward].  Both are zero on entry and re-arm themselves;
// nothing else is ever written there, whatever engine or shape used the workspace last.
constexpr int kTicketRegionBytes = 16384;
constexpr int kFinScratchBytes = 4096 + kTicketRegionBytes;

}  // namespace scb
