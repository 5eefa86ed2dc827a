// Tensor-core engine of the SupCon / NT-Xent loss for sm_100a: flash-style forward and
// recompute backward that never write a similarity tile to memory.
//
//   * K (the [n_k, ldk] bf16 contrast matrix) is staged by TMA (cp.async.bulk.tensor, 128-byte
//     swizzle) in [128 rows x 64 col] sub-tiles.  The forward walks 256-row blocks (two UMMA M = 128
//     halves, every B tile feeds two MMAs), the backward 128-row blocks; column tiles are 128 wide.
//   * S = K_i K_j^T runs on tcgen05.mma (kind::f16, M=128, N=128, K=16) with fp32 accumulators in
//     TMEM; one elected thread issues, completion is tracked with tcgen05.commit -> mbarrier.
//   * forward: four epilogue warpgroups pull S out of TMEM with tcgen05.ld and keep the online
//     (max, sum exp, positive sum) row state in registers.
//   * backward: the epilogue turns S into P = dL/dS, packs it to bf16 and writes it back into TMEM over the
//     S columns it came from; a second tcgen05.mma takes P as its A operand from TMEM and the very same K_j
//     tile, read MN-major, as B: dK_i accumulates in TMEM.  (Structure and the measurements behind it:
//     comment above supcon_tc_bwd_kernel and DESIGN.md section 3.2.)
//
// Reference being replaced: /root/reference/losses.py:64-91 and its autograd backward.
// Work split: the RB x CT tile grid is flattened row-block-major and cut into equal spans, one
// per CTA (persistent, one CTA per SM).  Forward: CTAs that share a row block leave (max, sum, possum, npos)
// partials; the last of them (ticket per row block) merges them in the same kernel and the last row block
// writes the scalar loss (fwd_merge_rows) -- no finalize kernel.  Backward: a row block is shared by at most two
// CTAs, launched as a cluster pair; the two dK halves meet over distributed shared memory and each CTA turns
// half of the rows into d features (finish_staged) -- no reduce kernel, no partial dK in global memory (spans
// that do not line up with row blocks fall back to per-CTA partial blocks + supcon_tc_reduce_finish).
// No float atomics anywhere: results are bitwise reproducible.
#include "supcon_tc.cuh"
#include "supcon_common.cuh"
#include "tc_ptx.cuh"

#include <stdlib.h>
#include <string.h>

#include <atomic>

namespace scb {
using namespace ptx;

constexpr int kBMH = 128;       // rows per half = one UMMA M
constexpr int kBM = 256;        // rows per CTA tile
constexpr int kBN = 128;        // columns per tile
constexpr int kSub = 16384;     // bytes of one [128 x 64] bf16 swizzled sub-tile
constexpr int kTcThreads = 576; // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..17 = four epilogue warpgroups
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kSafeSpread = 60.f;
#ifndef SCB_POLY_LANE
#define SCB_POLY_LANE (-1)
#endif
// Which of every 4 exponentials of the fast paths runs on the FMA pipe (ex2_fma) instead of MUFU; -1 = none.
// Measured with lane 3: 75.1 -> 74.3 us per step and 6e-6 instead of 1e-6 error in the log-sum-exps: the
// epilogues are bound by per-tile latency, not by the MUFU pipe, so it stays off.
constexpr int kPolyLane = SCB_POLY_LANE;

// Merge of the forward's per-CTA row partials inside the forward kernel itself: the last CTA to deliver its
// partials for a row block (ticket per row block) turns them into lse / possum / npos and the block's share of
// the loss; the last row block to finish adds the shares in block order (deterministic) and writes the scalar.
struct FwdOut {
    unsigned* rb_count;     // [RB] tickets, zero on entry, zero again on exit
    unsigned long long* acc;  // packed (fixed-point loss sum << 12 | row blocks merged), self re-arming
    float* out_lse;
    float* out_possum;
    float* out_npos;
    float* out_loss;        // [0] fp32 value, [1] value in scalar_dtype (low-order bytes)
    float loss_scale;
    double inv_anchors;
    int scalar_dtype;
    int n_k, my_slot;       // peers: layout of every rank's [lse (n_k) | npos (n_k) | loss shares] array
    PeerOut po;             // n_dst > 0: lse / npos / loss share go to every rank, then the flags are raised
};

// Everything the last step of the backward needs to turn a summed dK row into d features: scale by the upstream
// gradient, closed-form positive-pair term (fp32), un-stack, backward of the optional L2 normalisation, cast.
struct FinishParams {
    const void* feat;       // features [n_samples, n_views, dim] (only read when normalize != 0)
    void* out;              // d features, same dtype
    int feat_dtype, n_samples, normalize;
    int64_t stride_s, stride_v, ostride_s, ostride_v;
    const float* inv_norm;
    int grad_dtype;
    float coef_scale_invT;  // (T/Tb) / n_anchor_total / T
    const float* cls_all;   // labels: [n_samples_total][ldk] sum of the rows of the class (slot = representative sample)
    const float* cls_anc;   // labels: same over the anchor rows only (== cls_all in mode `all`)
};

struct TcParams {
    const __nv_bfloat16* k;
    const __nv_bfloat16* k_lo;          // SCB_ENGINE_BF16X3: low-order halves (k ~= k + k_lo), else nullptr
    int ldk, n_k, dim;
    int row_begin, n_rows, a_count;     // local rows; rows the forward treats as anchors
    int CT, RB, T, q;                   // column tiles, row blocks, tiles, tiles per CTA
    RowMap map;
    int n_views, mode, mask_kind;
    float c2, inv_T;                    // log2(e)/T, 1/T
    float* part;                        // forward partials [slot][4][n_rows]
    const int* rep_row;                 // labels: class representative (a sample id) per global row
    const float* npos_row;              // labels: number of positives per global row
    const float* nia_row;               // labels: [anchor]/npos per global row
    const float* lse_all;
    const float* grad_out;
    float coef_scale;                   // (T/Tb) / n_anchor_total
    float* out_dk;
    int ld_dk;
    float* dk_part;                     // backward partials [slot][RB*128][ldk] (one slot per CTA sharing a row block)
    int pair;                           // backward, direct finish: CTAs per row block (1, or 2 = a cluster pair)
    FwdOut fo;
    FinishParams fin;
    long long* trace;                   // optional (debug): per-role timestamps of CTA 0
};

// The timeline instrumentation below is compiled in only with -DSCB_TRACE=1 (tools/build_trace_lib.sh): even
// untaken, its tests cost a few dozen instructions per tile on paths where every instruction is latency.
#ifndef SCB_TRACE
#define SCB_TRACE 0
#endif
constexpr bool kTrace = SCB_TRACE != 0;

// wall clock in ns.  clock64() is NOT usable for these timelines: the per-sub-partition counters
// fall behind while all of a sub-partition's warps are parked on a barrier.
__device__ __forceinline__ long long global_ns() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// debug tracing: CTA 0, lane 0 of a warp appends (id, clock) pairs to its own 512-slot lane of the buffer.
// trace[4095] != 0 switches these per-event stamps off (the stores perturb the pipeline) and keeps only the
// per-CTA entry / exit stamps.
struct Tracer {
    long long* buf; int n; unsigned mask;
    __device__ Tracer(long long* base, int role) : buf(nullptr), n(0), mask(0xffffffffu) {
        if (kTrace && base != nullptr && blockIdx.x == 0 && (threadIdx.x & 31) == 0 && role >= 0 && base[4095] == 0) {
            buf = base + role * 1024;
            if (base[4094] != 0) mask = (unsigned)base[4094];    // only events with bit (id & 31) set (a stamp costs ~150 ns)
        }
    }
    __device__ __forceinline__ void operator()(int id) {
        if (kTrace && buf != nullptr && n < 510 && ((mask >> (id & 31)) & 1u)) {
            buf[2 * n] = id; buf[2 * n + 1] = global_ns(); ++n; buf[1022] = n;
        }
    }
};

// diagnostics: wall-clock (ns) entry / exit of every CTA at trace[3072 + 2*cta (+1)]
__device__ __forceinline__ void stamp_cta(long long* trace, int which) {
    if (kTrace && trace != nullptr && threadIdx.x == 0 && blockIdx.x < 512) trace[3072 + 2 * blockIdx.x + which] = global_ns();
}

__device__ __forceinline__ float bf16_dot_rows(const __nv_bfloat16* a, const __nv_bfloat16* b, int n) {
    float acc = 0.f;
    for (int d = 0; d < n; d += 8) {
        const uint4 ua = *reinterpret_cast<const uint4*>(a + d);
        const uint4 ub = *reinterpret_cast<const uint4*>(b + d);
        const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&ua);
        const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&ub);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 fa = __bfloat1622float2(pa[i]), fb = __bfloat1622float2(pb[i]);
            acc = fmaf(fa.x, fb.x, acc);
            acc = fmaf(fa.y, fb.y, acc);
        }
    }
    return acc;
}

__device__ __forceinline__ double warp_sum_f64(double v) {      // fixed xor tree: deterministic
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void store_loss_scalar(float* out, float loss, int scalar_dtype) {
    out[0] = loss;
    out[1] = 0.f;                         // [1]: the loss in the caller's scalar dtype (low-order bytes)
    if (scalar_dtype == SCB_BF16) *reinterpret_cast<__nv_bfloat16*>(out + 1) = __float2bfloat16_rn(loss);
    else if (scalar_dtype == SCB_F16) *reinterpret_cast<__half*>(out + 1) = __float2half_rn(loss);
    else out[1] = loss;
}

// Called by the 512 epilogue threads of the CTA that delivered the LAST partials of row block `rb` (losses.py:83-91
// for those 256 rows): merges the (max, sum, possum, npos) slots of every row, writes lse / possum / npos, and leaves
// the block's share of the loss in a packed 64-bit accumulator; whoever completes the last row block holds the total.
// Out of line: it runs once or twice per CTA and must not sit in the middle of the tile loop's code.
__device__ __noinline__ void fwd_merge_rows(const TcParams& p, int rb, int n_cta, double* s_red) {
    const int e = (int)threadIdx.x - 64;          // 0..511
    const FwdOut& o = p.fo;
    // (the acquire side of the row-block ticket is the ticket's own acq_rel atomic + the barrier behind it; the
    // partials are read with ld.global.cg, i.e. from the L2 where the other CTAs' released stores are)
    double contrib = 0.0;
    const int li = rb * kBM + e;
    if (e < kBM && li < p.a_count) {
        const int n_split = 2 * n_cta;
        float m = -INFINITY, sum = 0.f, ps = 0.f, np = 0.f;
        // partial maxima are in the log2 domain (as the tile loop keeps them); MUFU ex2 / lg2 keep this code short --
        // it runs once per row block, cold, so its size is its cost
        for (int s0 = 0; s0 < n_split; s0 += 12) {         // twelve slots at a time (a row block is shared by <= 6 CTAs): one round trip
            float vm[12], vs[12], vp[12], vn[12];
#pragma unroll
            for (int u = 0; u < 12; ++u) {
                const bool ok = s0 + u < n_split;
                const float* q = p.part + ((int64_t)(s0 + u) * 4) * p.n_rows + li;
                vm[u] = ok ? __ldcg(q) : -INFINITY;
                vs[u] = ok ? __ldcg(q + p.n_rows) : 0.f;
                vp[u] = ok ? __ldcg(q + 2 * (int64_t)p.n_rows) : 0.f;
                vn[u] = ok ? __ldcg(q + 3 * (int64_t)p.n_rows) : 0.f;
            }
            float cm = vm[0];
#pragma unroll
            for (int u = 1; u < 12; ++u) cm = fmaxf(cm, vm[u]);
            const float nm = fmaxf(m, cm);
            if (nm > -INFINITY) {
                float acc = sum * ex2(m - nm);                 // m = -inf: sum is 0 and ex2(-inf) = 0
#pragma unroll
                for (int u = 0; u < 12; ++u) acc = fmaf(vs[u], ex2(vm[u] - nm), acc);
                sum = acc;
                m = nm;
            }
#pragma unroll
            for (int u = 0; u < 12; ++u) { ps += vp[u]; np += vn[u]; }
        }
        const float lse = (m + __log2f(sum)) * kLn2;
        contrib = (double)(-o.loss_scale * (ps / np - lse));          // losses.py:87,90 (0/0 -> nan)
        o.out_possum[li] = ps;
        if (o.po.n_dst > 0) {
            for (int w = 0; w < o.po.n_dst; ++w) {
                float* d = reinterpret_cast<float*>(o.po.dst[w]);
                d[p.row_begin + li] = lse;
                d[o.n_k + p.row_begin + li] = np;
            }
        } else {
            o.out_lse[li] = lse;
            o.out_npos[li] = np;
        }
    }
    if (p.n_rows > p.a_count) {
        // local rows that are not anchors (mode `one`, views > 0): lse = 0, possum = 0, npos = 1; every merging CTA
        // fills its share of them
        const int extra = p.n_rows - p.a_count, chunk = (extra + p.RB - 1) / p.RB;
        const int x1 = min(extra, (rb + 1) * chunk);
        for (int x = rb * chunk + e; x < x1; x += 512) {
            const int r = p.a_count + x;
            o.out_possum[r] = 0.f;
            if (o.po.n_dst > 0) {
                for (int w = 0; w < o.po.n_dst; ++w) {
                    float* d = reinterpret_cast<float*>(o.po.dst[w]);
                    d[p.row_begin + r] = 0.f;
                    d[o.n_k + p.row_begin + r] = 1.f;
                }
            } else {
                o.out_lse[r] = 0.f;
                o.out_npos[r] = 1.f;
            }
        }
    }
    contrib = warp_sum_f64(contrib);
    if ((e & 31) == 0) s_red[e >> 5] = contrib;
    named_bar_sync(1, 512);                       // also orders every thread's stores before thread 0's atomic below
    if (e < 32) {
        double t = (e < 16) ? s_red[e] : 0.0;
        t = warp_sum_f64(t);
        if (e == 0) {
            o.rb_count[rb] = 0u;                  // re-arm the ticket for the next launch
            // The block's share joins the others in ONE 64-bit atomic: bits 12.. carry the sum in fixed point
            // (2^-20 units: integer addition is exact and order-independent, so the loss is deterministic), bits
            // 0..11 count the row blocks.  Whoever brings the count to RB holds the exact total -- no second
            // ticket, no re-read of the shares.  A non-finite share (an anchor without positives, losses.py:87)
            // is sent as a huge value that the reader turns back into NaN.
            const bool finite = fabs(t) < 1e12;
            const long long fx = finite ? __double2ll_rn(t * 1048576.0) : (1ll << 50);
            const unsigned long long add = ((unsigned long long)fx << 12) + 1ull;
            // peers read lse / npos after this rank's flag: order them at system scope; on one GPU nothing in this
            // kernel reads them, the kernel boundary publishes them
            if (o.po.n_dst > 0) __threadfence_system();
            const unsigned long long now = atomicAdd(o.acc, add) + add;
            if ((int)(now & 4095ull) == p.RB) {
                *o.acc = 0ull;
                const long long tot_fx = (long long)now >> 12;
                float loss = (float)((double)tot_fx * (1.0 / 1048576.0) * o.inv_anchors);
                if (tot_fx >= (1ll << 49) || tot_fx <= -(1ll << 49)) loss = __int_as_float(0x7fc00000);
                if (o.po.n_dst > 0) {
                    // this rank's share of the global mean goes to every rank, then the rank's flag is raised there
                    __threadfence_system();
                    for (int w = 0; w < o.po.n_dst; ++w)
                        reinterpret_cast<float*>(o.po.dst[w])[2 * o.n_k + 4 * o.my_slot] = loss;
                    __threadfence_system();
                    const unsigned ex = o.po.state[1] + 1u;
                    o.po.state[1] = ex;
                    for (int w = 0; w < o.po.n_dst; ++w)
                        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(o.po.flag[w]), "r"(ex) : "memory");
                } else {
                    store_loss_scalar(o.out_loss, loss, o.scalar_dtype);
                }
            }
        }
    }
    named_bar_sync(1, 512);                       // s_red may be reused by this CTA's next merge
}

// =================================================================================== forward
// smem: A[NSA][A_BYTES] | B[NSB][B_BYTES] | barriers | tmem slot | merge scratch
//
// kX3 (SCB_ENGINE_BF16X3): every operand is a pair of bf16 matrices, k ~= hi + lo with lo = bf16(k - hi), and a tile
// is three MMA groups into the same fp32 accumulator, S = hi hi^T + hi lo^T + lo hi^T (the lo lo^T term is below
// 2^-16 of the result): fp32-class logits on the bf16 tensor-core path.  The staged tiles are twice as large, so A
// and B are single-buffered (the tile's MMA time triples anyway).
template <int KCH, bool kLabels, bool kX3>
__global__ void __launch_bounds__(kTcThreads, 1)
supcon_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_lo,
                     const __grid_constant__ TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int A_PART = 2 * KCH * kSub;           // 256 rows of one bf16 matrix
    constexpr int B_PART = KCH * kSub;
    constexpr int A_BYTES = (kX3 ? 2 : 1) * A_PART;  // hi | lo
    constexpr int B_BYTES = (kX3 ? 2 : 1) * B_PART;
    constexpr int NSA = kX3 ? 1 : 2;
    constexpr int NSB = kX3 ? 1 : 2;
    uint8_t* a_smem = smem;
    uint8_t* b_smem = smem + NSA * A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(b_smem + NSB * B_BYTES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
    double* s_red = reinterpret_cast<double*>(bars + 18);                  // [16] merge scratch
    volatile uint32_t* s_ticket = reinterpret_cast<volatile uint32_t*>(bars + 34);
    // shared-memory copy of the parameters for the out-of-line tail code: through a reference, kernel parameters
    // are read with generic loads from the parameter window (an L2 round trip per dependent field; measured 7 us
    // per tail), from shared memory they cost an LDS
    TcParams* sp = reinterpret_cast<TcParams*>(bars + 40);
    for (int i = threadIdx.x; i < (int)(sizeof(TcParams) / 4); i += kTcThreads)
        reinterpret_cast<uint32_t*>(sp)[i] = reinterpret_cast<const uint32_t*>(&p)[i];
    const uint32_t bar0 = smem_u32(bars);
    auto a_full = [&](int b) { return bar0 + 8u * (0 + b); };
    auto a_empty = [&](int b) { return bar0 + 8u * (2 + b); };
    auto b_full = [&](int s) { return bar0 + 8u * (4 + s); };
    auto b_empty = [&](int s) { return bar0 + 8u * (6 + s); };
    auto s_full = [&](int h, int b) { return bar0 + 8u * (8 + h * 2 + b); };
    auto s_empty = [&](int h, int b) { return bar0 + 8u * (12 + h * 2 + b); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // trace builds: p.trace[4094] selects the CTA whose phases are stamped (0 by default)
    const unsigned trace_cta = (kTrace && p.trace != nullptr) ? (unsigned)p.trace[4094] : 0u;
    auto stamp = [&](int slot) { if (kTrace && p.trace != nullptr && blockIdx.x == trace_cta) { p.trace[slot] = clock64(); p.trace[slot + 16] = global_ns(); } };
    if (threadIdx.x == 0) stamp(0);
    stamp_cta(p.trace, 0);
    if ((smem_u32(smem) & 1023u) != 0u) __trap();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 12; ++i) mbar_init(bar0 + 8u * i, 1);
        for (int i = 12; i < 16; ++i) mbar_init(bar0 + 8u * i, 8);     // s_empty: one arrival per warp of the row half
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 512);
    if (warp == 0 && lane == 0) prefetch_tensormap(&tmap);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) stamp(1);
    pdl_launch_dependents();
    pdl_wait();                                  // everything above touched no global memory

    const int t0 = blockIdx.x * p.q;
    const int t1 = min(p.T, t0 + p.q);

    if (warp == 0) {
        // ---------------------------------------------------- TMA producer (warp-uniform, one lane issues)
        int n_a = 0, n_b = 0, cur_rb = -1;
        for (int tau = t0; tau < t1; ++tau) {
            const int rb = tau / p.CT, ct = tau - rb * p.CT;
            if (rb != cur_rb) {
                cur_rb = rb;
                const int buf = n_a % NSA;
                mbar_wait(a_empty(buf), ((n_a / NSA) & 1) ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(a_full(buf), A_BYTES);
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int kc = 0; kc < KCH; ++kc) {
                            tma_load_2d(smem_u32(a_smem + buf * A_BYTES + (h * KCH + kc) * kSub), &tmap, kc * 64,
                                        p.row_begin + rb * kBM + h * kBMH, a_full(buf));
                            if (kX3)
                                tma_load_2d(smem_u32(a_smem + buf * A_BYTES + A_PART + (h * KCH + kc) * kSub), &tmap_lo,
                                            kc * 64, p.row_begin + rb * kBM + h * kBMH, a_full(buf));
                        }
                }
                __syncwarp();
                ++n_a;
            }
            const int st = n_b % NSB;
            mbar_wait(b_empty(st), ((n_b / NSB) & 1) ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(b_full(st), B_BYTES);
#pragma unroll
                for (int kc = 0; kc < KCH; ++kc) {
                    tma_load_2d(smem_u32(b_smem + st * B_BYTES + kc * kSub), &tmap, kc * 64, ct * kBN, b_full(st));
                    if (kX3)
                        tma_load_2d(smem_u32(b_smem + st * B_BYTES + B_PART + kc * kSub), &tmap_lo, kc * 64, ct * kBN,
                                    b_full(st));
                }
            }
            __syncwarp();
            ++n_b;
        }
    } else if (warp == 1) {
        // ---------------------------------------------------- MMA issuer (warp-uniform, one lane issues)
        constexpr uint32_t idesc = idesc_bf16(128, 128, 0);
        int n_a = 0, n_b = 0, idx = 0, cur_rb = -1, abuf = 0;
        for (int tau = t0; tau < t1; ++tau) {
            const int rb = tau / p.CT;
            if (rb != cur_rb) {
                if (cur_rb >= 0) {
                    if (elect_one()) umma_commit(a_empty(abuf));
                    __syncwarp();
                }
                cur_rb = rb;
                abuf = n_a % NSA;
                mbar_wait(a_full(abuf), (n_a / NSA) & 1);
                if (n_a == 0 && lane == 0) stamp(2);
                ++n_a;
            }
            const int st = n_b % NSB;
            mbar_wait(b_full(st), (n_b / NSB) & 1);
            if (n_b == 0 && lane == 0) stamp(3);
            ++n_b;
            const int sbuf = idx & 1;
            const uint32_t sphase = (idx >> 1) & 1;
            const uint32_t b_lo = desc_lo(smem_u32(b_smem + st * B_BYTES), 16);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                mbar_wait(s_empty(h, sbuf), sphase ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_base + sbuf * 256 + h * 128;
                const uint32_t a_lo = desc_lo(smem_u32(a_smem + abuf * A_BYTES + h * KCH * kSub), 16);
                if (elect_one()) {
#pragma unroll
                    for (int grp = 0; grp < (kX3 ? 3 : 1); ++grp) {       // hi hi^T, hi lo^T, lo hi^T
                        const uint32_t ag = a_lo + ((grp == 2 ? A_PART : 0) >> 4);
                        const uint32_t bg = b_lo + ((grp == 1 ? B_PART : 0) >> 4);
#pragma unroll
                        for (int k16 = 0; k16 < KCH * 4; ++k16) {
                            const uint32_t off = ((k16 >> 2) * kSub + (k16 & 3) * 32) >> 4;
                            if (grp == 0 && k16 == 0) umma_bf16_lo<false>(d, ag + off, bg + off, idesc);
                            else umma_bf16_lo<true>(d, ag + off, bg + off, idesc);
                        }
                    }
                    umma_commit(s_full(h, sbuf));
                    if (h == 1) umma_commit(b_empty(st));
                }
                __syncwarp();
            }
            ++idx;
        }
    } else {
        // ---------------------------------------------------- epilogue warpgroups
        const int w4 = (warp - 2) >> 2;                 // warpgroup = (row half wg, column half ca)
        const int wg = w4 >> 1, ca = w4 & 1;
        const int q4 = warp & 3;                        // TMEM lane quarter this warp may touch
        const int r = q4 * 32 + lane;                   // row within the half
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        float m = -INFINITY, sum = 0.f, possum = 0.f;
        int cur_rb = -1, idx = 0, li = 0, gi = 0, rep_i = 0;

        // SimCLR positives with two views are read out of S itself: row i's single positive is the
        // column of its other view (losses.py:42-43).  pj = that column; [pj_lo, pj_hi] = the range the
        // warp's 32 rows cover, so the test per 32-column chunk is warp-uniform.
        const bool capture = !kLabels && p.n_views == 2;
        int pj = -1, pj_lo = 0, pj_hi = -1;

        auto flush = [&](int rb) {
            if (threadIdx.x == 64) stamp(5);              // (trace) tiles of the row block done
            const int first_cta = (rb * p.CT) / p.q;      // CTAs [first_cta, last_cta] overlap this row block
            if (li < p.a_count) {
                // one slot per (CTA overlapping the row block, column half)
                const int slot = 2 * (blockIdx.x - first_cta) + ca;
                float ps = 0.f, np = 0.f;
                if (kLabels) {
                    ps = possum * p.inv_T;
                    if (slot == 0) np = p.npos_row[gi];
                } else if (capture) {
                    ps = possum * p.inv_T;
                    if (slot == 0) np = 1.f;
                } else if (slot == 0) {
                    // more than two views: the other views of the same sample, as dot products
                    const int v = p.map.view_of(gi);
                    for (int vv = 0; vv < p.n_views; ++vv)
                        if (vv != v)
                        {
                            const int64_t oi = (int64_t)gi * p.ldk, oj = (int64_t)p.map.sibling(gi, vv) * p.ldk;
                            ps += bf16_dot_rows(p.k + oi, p.k + oj, p.ldk);
                            if (kX3) ps += bf16_dot_rows(p.k + oi, p.k_lo + oj, p.ldk) + bf16_dot_rows(p.k_lo + oi, p.k + oj, p.ldk);
                        }
                    ps *= p.inv_T;
                    np = (float)(p.n_views - 1);
                }
                float* dst = p.part + (int64_t)slot * 4 * p.n_rows + li;
                dst[0] = m;                               // log2 domain
                dst[p.n_rows] = sum;
                dst[2 * (int64_t)p.n_rows] = ps;
                dst[3 * (int64_t)p.n_rows] = np;
            }
            // row-block ticket: the CTA that delivers the last partials of the block merges them (fwd_merge_rows)
            const int n_cta = ((rb + 1) * p.CT - 1) / p.q - first_cta + 1;
            named_bar_sync(1, 512);                       // all partial stores of this CTA are issued ...
            if (threadIdx.x == 64) {
                // ... and released by the ticket itself (acq_rel at gpu scope: the barrier above makes the release
                // cumulative over the whole CTA's stores, the acquire side covers the merge below) -- one L2 round
                // trip instead of a full fence plus the atomic
                unsigned tk;
                asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], %2;" : "=r"(tk) : "l"(p.fo.rb_count + rb), "r"(1u) : "memory");
                *s_ticket = tk;
            }
            named_bar_sync(1, 512);
            if (threadIdx.x == 64) stamp(6);              // (trace) partials out, ticket taken
            if (*s_ticket == (unsigned)n_cta - 1u) {
                fwd_merge_rows(*sp, rb, n_cta, s_red);
                if (threadIdx.x == 64) stamp(8);          // (trace) row block merged
            }
        };

        // incremental tile bookkeeping (no division per tile); one copy of the flush and of the chunk code:
        // the code between two tiles is latency-bound scalar work and has to stay short and cache-resident
        int rb = t0 / p.CT, ct = t0 - rb * p.CT;
        for (int tau = t0;; ++tau) {
            const bool more = tau < t1;
            if (!more || rb != cur_rb) {
                if (cur_rb >= 0) flush(cur_rb);
                if (!more) break;
                cur_rb = rb;
                m = -INFINITY; sum = 0.f; possum = 0.f;
                li = rb * kBM + wg * kBMH + r;
                gi = p.row_begin + li;
                if (kLabels) rep_i = (gi < p.n_k) ? p.rep_row[gi] : -1;
                if (capture) {
                    pj = -1;
                    if (gi < p.n_k) pj = p.map.sibling(gi, 1 - p.map.view_of(gi));
                    int lo = pj < 0 ? 0x7fffffff : pj, hi = pj;
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) {
                        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
                    }
                    pj_lo = lo; pj_hi = hi;
                }
            }
            const int sbuf = idx & 1;
            mbar_wait(s_full(wg, sbuf), (idx >> 1) & 1);
            if (idx == 0 && threadIdx.x == 64) stamp(4);
            tc_fence_after();
            const int j0 = ct * kBN + ca * 64;            // this warpgroup's 64 columns of the tile
            const int g_lo = p.row_begin + rb * kBM + wg * kBMH;
            const bool pred = (j0 + 64 > p.n_k) || (g_lo < j0 + 64 && j0 < g_lo + kBMH);
            const uint32_t taddr = tmem_base + lane_addr + sbuf * 256 + wg * 128 + ca * 64;
            // one 32-column chunk of this thread's row: online (max, sum) update + positive sum
            auto process = [&](uint32_t (&v)[32], int cc) {
                const int jc = j0 + cc * 32;
                if (capture && jc <= pj_hi && jc + 32 > pj_lo) {
                    const int c_hit = pj - jc;
#pragma unroll
                    for (int c = 0; c < 32; ++c) possum += (c == c_hit) ? __uint_as_float(v[c]) : 0.f;
                }
                if (!pred) {
                    float m8[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        m8[i] = fmaxf(fmaxf(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1])),
                                      fmaxf(__uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3])));
                    const float cmax = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])),
                                             fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
                    const float nm = fmaxf(m, cmax * p.c2);
                    sum *= ex2(m - nm);
                    m = nm;
                    if (kPolyLane >= 0) {
                        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            const float a = fmaf(__uint_as_float(v[c]), p.c2, -m);
                            acc[c & 3] += ((c & 3) == kPolyLane) ? ex2_fma(a) : ex2(a);     // one in four on the FMA pipe
                        }
                        sum += (acc[0] + acc[1]) + (acc[2] + acc[3]);
                    } else {
                        // two elements per FFMA2 / FADD2: the scale-and-shift and the accumulation take half the issue
                        // slots, the exponentials stay one MUFU each
                        const uint64_t c2p = pack2(p.c2, p.c2), mneg = pack2(-m, -m);
                        uint64_t acc01 = pack2(0.f, 0.f), acc23 = acc01;
#pragma unroll
                        for (int c = 0; c < 32; c += 4) {
                            const uint64_t x01 = ffma2(pack2(__uint_as_float(v[c]), __uint_as_float(v[c + 1])), c2p, mneg);
                            const uint64_t x23 = ffma2(pack2(__uint_as_float(v[c + 2]), __uint_as_float(v[c + 3])), c2p, mneg);
                            float x0, x1, x2, x3;
                            unpack2(x01, x0, x1);
                            unpack2(x23, x2, x3);
                            acc01 = fadd2(acc01, pack2(ex2(x0), ex2(x1)));
                            acc23 = fadd2(acc23, pack2(ex2(x2), ex2(x3)));
                        }
                        float s0, s1, s2, s3;
                        unpack2(acc01, s0, s1);
                        unpack2(acc23, s2, s3);
                        sum += (s0 + s1) + (s2 + s3);
                    }
                    if (kLabels) {
                        float pacc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int c4 = 0; c4 < 8; ++c4) {
                            const int4 rj = __ldcg(reinterpret_cast<const int4*>(p.rep_row + jc) + c4);
                            pacc[0] += (rj.x == rep_i) ? __uint_as_float(v[c4 * 4 + 0]) : 0.f;
                            pacc[1] += (rj.y == rep_i) ? __uint_as_float(v[c4 * 4 + 1]) : 0.f;
                            pacc[2] += (rj.z == rep_i) ? __uint_as_float(v[c4 * 4 + 2]) : 0.f;
                            pacc[3] += (rj.w == rep_i) ? __uint_as_float(v[c4 * 4 + 3]) : 0.f;
                        }
                        possum += (pacc[0] + pacc[1]) + (pacc[2] + pacc[3]);
                    }
                } else {
                    // tiles that contain the self column (losses.py:74-79) or run past n_k
                    float cmax = -INFINITY;
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const int j = jc + c;
                        const bool ok = (j < p.n_k) && (j != gi);
                        const float x = ok ? __uint_as_float(v[c]) : -INFINITY;
                        v[c] = __float_as_uint(x);
                        cmax = fmaxf(cmax, x);
                    }
                    const float nm = fmaxf(m, cmax * p.c2);
                    if (nm > -INFINITY) {
                        sum *= ex2(m - nm);
                        m = nm;
#pragma unroll
                        for (int c = 0; c < 32; ++c) sum += ex2(fmaf(__uint_as_float(v[c]), p.c2, -m));
                    }
                    if (kLabels) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            const int j = jc + c;
                            const float x = __uint_as_float(v[c]);
                            if (x > -INFINITY && p.rep_row[j] == rep_i) possum += x;
                        }
                    }
                }
            };
#pragma unroll 1
            for (int cc = 0; cc < 2; ++cc) {
                uint32_t v[32];
                tmem_ld32(taddr + cc * 32, v);
                tmem_ld_wait_dep(v);
                if (cc == 1) {
                    // every column of S[wg][sbuf] this thread needs is in registers: hand the buffer back
                    // (one arrival per warp: 32 per-thread arrivals on one barrier word serialise)
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(s_empty(wg, sbuf));
                }
                process(v, cc);
            }
            ++idx;
            if (++ct == p.CT) { ct = 0; ++rb; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) stamp(7);
    stamp_cta(p.trace, 1);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// In-register transpose of the 32 x 32 block a warp holds (v[i] of lane l = element (l, i)):
// afterwards v[i] of lane l = element (i, l), so a store of v[i] writes 32 consecutive columns of row i.
__device__ __forceinline__ void warp_transpose32(uint32_t (&v)[32], int lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool hi = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if ((i & s) == 0) {
                const uint32_t send = hi ? v[i] : v[i + s];
                const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, s);
                if (hi) v[i] = recv; else v[i + s] = recv;
            }
        }
    }
}

// One warp's [32 rows x 32 columns] block of dK: TMEM -> registers -> transposed, coalesced stores.  Out of line
// (scalar arguments only): it runs once or twice per CTA and its ~600 instructions would otherwise sit in the
// middle of the epilogue's loop code.
__device__ __noinline__ void drain_block(uint32_t taddr, float* dst, int ldk, int lane) {
    uint32_t v[32];
    tmem_ld32(taddr, v);
    tmem_ld_wait_dep(v);
    warp_transpose32(v, lane);
#pragma unroll
    for (int rr = 0; rr < 32; ++rr) dst[(int64_t)rr * ldk] = __uint_as_float(v[rr]);
}

// pack pv[off .. off+15] to bf16 and write them as 8 consecutive 32-bit TMEM columns of this thread's lane
// (UMMA A operand in TMEM: two K-consecutive bf16 per column, even index in the low half)
template <int N>
__device__ __forceinline__ void store_p16(uint32_t taddr, const float (&pv)[N], int off) {
    uint32_t wds[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h2 = __floats2bfloat162_rn(pv[off + 2 * i], pv[off + 2 * i + 1]);
        wds[i] = *reinterpret_cast<uint32_t*>(&h2);
    }
    tmem_st8(taddr, wds);
}

// ------------------------------------------------------------------ finishing a summed dK row
// The tiles only carry the softmax part of the gradient, P_ij + P_ji (rounded to bf16 for the second MMA).  The
// positive-pair part, -(1/T) sum_{j != i} (M_ij [i anchor]/npos_i + M_ji [j anchor]/npos_j) k_j, has a closed form
// per row and is added here in fp32 (Appendix A of SURVEY.md):
//   SimCLR: the other views of the same sample, weight ([anchor_i] + [anchor_j]) / (V - 1);
//   labels: with S_all / S_anc the sums of the rows / anchor rows of the row's class and c = 1 / npos,
//           nia_i (S_all - k_i) + c (S_anc - [anchor_i] k_i)            (label_rows_kernel builds the sums).
__device__ __forceinline__ void load_k4(const __nv_bfloat16* row, int d, float (&o)[4]) {
    const uint2 raw = *reinterpret_cast<const uint2*>(row + d);
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
    const float2 f0 = __bfloat1622float2(h2[0]), f1 = __bfloat1622float2(h2[1]);
    o[0] = f0.x; o[1] = f0.y; o[2] = f1.x; o[3] = f1.y;
}

// the positive-pair term of global row gi at features d .. d+3 (to be ADDED to the softmax part)
__device__ __forceinline__ void positive_term(const TcParams& p, int gi, int d, float (&pos)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) pos[i] = 0.f;
    if (p.mask_kind == SCB_MASK_SIMCLR && p.n_views >= 2) {
        const int v = p.map.view_of(gi);
        const float ai = p.map.is_anchor(gi, p.mode) ? 1.f : 0.f;
        const float cpos = -1.f / (float)(p.n_views - 1);
        for (int vv = 0; vv < p.n_views; ++vv) {
            if (vv == v) continue;
            const int j = p.map.sibling(gi, vv);
            const float wgt = cpos * (ai + (p.map.is_anchor(j, p.mode) ? 1.f : 0.f));
            float kj[4];
            load_k4(p.k + (int64_t)j * p.ldk, d, kj);
            if (p.k_lo != nullptr) {
                float kl[4];
                load_k4(p.k_lo + (int64_t)j * p.ldk, d, kl);
#pragma unroll
                for (int i = 0; i < 4; ++i) kj[i] += kl[i];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) pos[i] = fmaf(wgt, kj[i], pos[i]);
        }
    } else if (p.mask_kind == SCB_MASK_LABELS) {
        const int rep = p.rep_row[gi];
        const float nia_i = p.nia_row[gi];
        const float c = 1.f / p.npos_row[gi];
        const float ai = p.map.is_anchor(gi, p.mode) ? 1.f : 0.f;
        const float4 sa = *reinterpret_cast<const float4*>(p.fin.cls_all + (int64_t)rep * p.ldk + d);
        const float4 sn = *reinterpret_cast<const float4*>(p.fin.cls_anc + (int64_t)rep * p.ldk + d);
        float ki[4];
        load_k4(p.k + (int64_t)gi * p.ldk, d, ki);
        if (p.k_lo != nullptr) {
            float kl[4];
            load_k4(p.k_lo + (int64_t)gi * p.ldk, d, kl);
#pragma unroll
            for (int i = 0; i < 4; ++i) ki[i] += kl[i];
        }
        pos[0] = -(nia_i * (sa.x - ki[0]) + c * (sn.x - ai * ki[0]));
        pos[1] = -(nia_i * (sa.y - ki[1]) + c * (sn.y - ai * ki[1]));
        pos[2] = -(nia_i * (sa.z - ki[2]) + c * (sn.z - ai * ki[2]));
        pos[3] = -(nia_i * (sa.w - ki[3]) + c * (sn.w - ai * ki[3]));
    }
}

__device__ __forceinline__ float load_elem(const void* base, int dtype, int64_t idx) {
    if (dtype == SCB_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
    if (dtype == SCB_F16) return __half2float(reinterpret_cast<const __half*>(base)[idx]);
    return reinterpret_cast<const float*>(base)[idx];
}
__device__ __forceinline__ void store_elem(void* base, int dtype, int64_t idx, float v) {
    if (dtype == SCB_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
    else if (dtype == SCB_F16) reinterpret_cast<__half*>(base)[idx] = __float2half_rn(v);
    else reinterpret_cast<float*>(base)[idx] = v;
}

// NB local rows per warp at a time (lane = features lane*4 .. +3): acc[b] holds the summed softmax part of row
// lrow[b] (-1 = no row).  All global loads of the batch are issued before the first dependent use, so the batch
// costs one memory round trip, not NB.  Writes d features[b, v, :] (losses.py:52-53 un-stacked,
// main_supcon.py:283 backward).
template <int NB>
__device__ __forceinline__ void finish_rows(const TcParams& p, float (&acc)[NB][4], const int (&lrow)[NB], int lane,
                                            float coef) {
    const FinishParams& f = p.fin;
    const int d = lane * 4;
    const bool on = d < p.ldk;
    float pos[NB][4], z[NB][4], inv[NB];
    int64_t src_off[NB], dst_off[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        const bool ok = lrow[b] >= 0;
        const int lr = ok ? lrow[b] : 0;
        // local row = v*n_samples + b (view-major) or b*n_views + v (sample-major)  ->  features[b, v, :]
        const int v_l = p.map.il ? lr % p.n_views : lr / f.n_samples;
        const int b_l = p.map.il ? lr / p.n_views : lr - v_l * f.n_samples;
        src_off[b] = (int64_t)b_l * f.stride_s + (int64_t)v_l * f.stride_v;
        dst_off[b] = (int64_t)b_l * f.ostride_s + (int64_t)v_l * f.ostride_v;
#pragma unroll
        for (int i = 0; i < 4; ++i) { pos[b][i] = 0.f; z[b][i] = 0.f; }
        inv[b] = 1.f;
        if (ok && on) positive_term(p, p.row_begin + lr, d, pos[b]);
        if (ok && f.normalize) {
            inv[b] = f.inv_norm[lr];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (d + i < p.dim) z[b][i] = load_elem(f.feat, f.feat_dtype, src_off[b] + d + i);
        }
    }
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        if (lrow[b] < 0) continue;                 // warp-uniform
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = on ? (acc[b][i] + pos[b][i]) * coef : 0.f;
        if (f.normalize) {
            float dot = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) { z[b][i] *= inv[b]; dot = fmaf(a[i], z[b][i], dot); }
            dot = warp_sum(dot);
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = (a[i] - z[b][i] * dot) * inv[b];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (d + i < p.dim) store_elem(f.out, f.feat_dtype, dst_off[b] + d + i, a[i]);
    }
}

// Direct finish of the backward (one or two CTAs per row block): a warp's [32 rows x 32 columns] block of dK goes
// from TMEM into a [rows][pitch] fp32 staging array in shared memory -- this CTA's, or the partner CTA's of the
// cluster pair (shared::cluster address) for the rows the partner finishes.  Out of line, scalar arguments.
__device__ __noinline__ void stage_block(uint32_t taddr, uint32_t dst, int remote) {
    uint32_t v[32];
    tmem_ld32(taddr, v);
    tmem_ld_wait_dep(v);
    if (remote) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16u * i), "r"(v[4 * i]),
                         "r"(v[4 * i + 1]), "r"(v[4 * i + 2]), "r"(v[4 * i + 3]) : "memory");
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16u * i), "r"(v[4 * i]),
                         "r"(v[4 * i + 1]), "r"(v[4 * i + 2]), "r"(v[4 * i + 3]) : "memory");
    }
}

// ---- compact 8-wide row finish of the direct backward.  This code runs ONCE per CTA, cold: every kilobyte of it is
// an instruction-cache miss at the tail of the kernel, so it is written for size -- one copy of everything, rolled
// loops, 16-byte typed vector accesses (the host only takes this path for 16-byte aligned features with dim % 8 == 0).
__device__ __forceinline__ void load_k8(const __nv_bfloat16* row, int d, float (&o)[8]) {
    const uint4 raw = *reinterpret_cast<const uint4*>(row + d);
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h2[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
}
__device__ __noinline__ void load8_typed(const void* base, int dtype, int64_t idx, float (&o)[8]) {
    if (dtype == SCB_F32) {
        const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + idx);
        const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + idx + 4);
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    } else if (dtype == SCB_BF16) {
        load_k8(reinterpret_cast<const __nv_bfloat16*>(base) + idx, 0, o);
    } else {
        const uint4 raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(base) + idx);
        const __half2* h2 = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(h2[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
    }
}
__device__ __noinline__ void store8_typed(void* base, int dtype, int64_t idx, const float (&v)[8]) {
    if (dtype == SCB_F32) {
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + idx) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + idx + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else {
        uint4 raw;
        if (dtype == SCB_BF16) {
            __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
            for (int i = 0; i < 4; ++i) h2[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(base) + idx) = raw;
        } else {
            __half2* h2 = reinterpret_cast<__half2*>(&raw);
#pragma unroll
            for (int i = 0; i < 4; ++i) h2[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
            *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(base) + idx) = raw;
        }
    }
}

// The sixteen epilogue warps turn the staged rows into d features.  A row is spread over ldk / 8 lanes (8 features
// each) and a lane group works on kFinRows rows at once, so that their global loads share one round trip: with
// ldk = 128 the 64 rows of a CTA pair's half are done in a single pass.  own / recv: [n_keep][pitch] fp32 (recv =
// the partner CTA's contribution, or nullptr).
constexpr int kFinRows = 1;       // 2 measured slower: the tail is bound by the amount of (cold) code, not by load latency
__device__ __noinline__ void finish_staged(const TcParams& p, const float* own, const float* recv, int pitch,
                                           int lrow0, int n_keep, int ew, int lane) {
    const FinishParams& f = p.fin;
    const float coef = f.coef_scale_invT;                    // already times the upstream gradient (kernel prologue)
    const int lpr = p.ldk >> 3, rpw = 32 / lpr;              // lanes per row, rows per warp pass
    const int sub = lane / lpr, d = (lane - sub * lpr) * 8;
    const bool in_dim = d < p.dim;
#pragma unroll 1
    for (int r0 = ew * rpw * kFinRows; r0 < n_keep; r0 += 16 * rpw * kFinRows) {
        float a[kFinRows][8], z[kFinRows][8], inv[kFinRows];
        int64_t src_off[kFinRows], dst_off[kFinRows];
        bool ok[kFinRows];
        // ---- phase 1: everything that comes from memory, for all rows of the group
#pragma unroll
        for (int q = 0; q < kFinRows; ++q) {
            const int rr = r0 + q * rpw + sub;
            ok[q] = rr < n_keep && lrow0 + rr < p.n_rows;
            const int rs = ok[q] ? rr : 0;
            const int lr = lrow0 + rs;
            const int gi = p.row_begin + lr;
            const float4 a0 = *reinterpret_cast<const float4*>(own + rs * pitch + d);
            const float4 a1 = *reinterpret_cast<const float4*>(own + rs * pitch + d + 4);
            a[q][0] = a0.x; a[q][1] = a0.y; a[q][2] = a0.z; a[q][3] = a0.w;
            a[q][4] = a1.x; a[q][5] = a1.y; a[q][6] = a1.z; a[q][7] = a1.w;
            if (recv != nullptr) {
                const float4 c0 = *reinterpret_cast<const float4*>(recv + rs * pitch + d);
                const float4 c1 = *reinterpret_cast<const float4*>(recv + rs * pitch + d + 4);
                a[q][0] += c0.x; a[q][1] += c0.y; a[q][2] += c0.z; a[q][3] += c0.w;
                a[q][4] += c1.x; a[q][5] += c1.y; a[q][6] += c1.z; a[q][7] += c1.w;
            }
            // local row = v*n_samples + b (view-major) or b*n_views + v (sample-major)  ->  features[b, v, :]
            // (ONE division; view, siblings and anchor flags follow from it: the local rows are one segment)
            const int qd = (unsigned)lr / (unsigned)(p.map.il ? p.n_views : f.n_samples);
            const int v_l = p.map.il ? lr - qd * p.n_views : qd;
            const int b_l = p.map.il ? qd : lr - qd * f.n_samples;
            src_off[q] = (int64_t)b_l * f.stride_s + (int64_t)v_l * f.stride_v + d;
            dst_off[q] = (int64_t)b_l * f.ostride_s + (int64_t)v_l * f.ostride_v + d;
            const bool all = p.mode == SCB_MODE_ALL;
            // positive-pair term (see positive_term): one weighted row per other view, or the class sums
            if (p.mask_kind == SCB_MASK_SIMCLR && p.n_views >= 2) {
                const float ai = (all || v_l == 0) ? 1.f : 0.f;
                const float cpos = -1.f / (float)(p.n_views - 1);
                const int vstep = p.map.il ? 1 : p.map.seg;
#pragma unroll 1
                for (int vv = 0; vv < p.n_views; ++vv) {
                    if (vv == v_l) continue;
                    const int j = gi + (vv - v_l) * vstep;
                    const float wgt = cpos * (ai + ((all || vv == 0) ? 1.f : 0.f));
                    float kj[8];
                    load_k8(p.k + (int64_t)j * p.ldk, d, kj);
#pragma unroll
                    for (int i = 0; i < 8; ++i) a[q][i] = fmaf(wgt, kj[i], a[q][i]);
                    if (p.k_lo != nullptr) {                    // bf16x3: k = hi + lo
                        load_k8(p.k_lo + (int64_t)j * p.ldk, d, kj);
#pragma unroll
                        for (int i = 0; i < 8; ++i) a[q][i] = fmaf(wgt, kj[i], a[q][i]);
                    }
                }
            } else if (p.mask_kind == SCB_MASK_LABELS) {
                const int rep = p.rep_row[gi];
                const float nia_i = p.nia_row[gi];
                const float c = 1.f / p.npos_row[gi];
                const float ai = (all || v_l == 0) ? 1.f : 0.f;
                float ki[8], sa[8], sn[8];
                load_k8(p.k + (int64_t)gi * p.ldk, d, ki);
                if (p.k_lo != nullptr) {
                    float kl[8];
                    load_k8(p.k_lo + (int64_t)gi * p.ldk, d, kl);
#pragma unroll
                    for (int i = 0; i < 8; ++i) ki[i] += kl[i];
                }
                load8_typed(f.cls_all, SCB_F32, (int64_t)rep * p.ldk + d, sa);
                load8_typed(f.cls_anc, SCB_F32, (int64_t)rep * p.ldk + d, sn);
#pragma unroll
                for (int i = 0; i < 8; ++i) a[q][i] -= nia_i * (sa[i] - ki[i]) + c * (sn[i] - ai * ki[i]);
            }
            inv[q] = 1.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) z[q][i] = 0.f;
            if (f.normalize) {                               // backward of main_supcon.py:283
                inv[q] = f.inv_norm[lr];
                if (in_dim) load8_typed(f.feat, f.feat_dtype, src_off[q], z[q]);
            }
        }
        // ---- phase 2: arithmetic and the stores
#pragma unroll
        for (int q = 0; q < kFinRows; ++q) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[q][i] *= coef;
            if (f.normalize) {
                float dot = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) { z[q][i] *= inv[q]; dot = fmaf(a[q][i], z[q][i], dot); }
                for (int o = lpr >> 1; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
#pragma unroll
                for (int i = 0; i < 8; ++i) a[q][i] = (a[q][i] - z[q][i] * dot) * inv[q];
            }
            if (ok[q] && in_dim) store8_typed(f.out, f.feat_dtype, dst_off[q], a[q]);
        }
    }
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {     // every thread of every CTA of the cluster
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// bf16x3: pv[off .. off+15] as bf16 hi (8 words at taddr) and bf16 lo = bf16(pv - hi) (8 words at taddr + 8)
template <int N>
__device__ __forceinline__ void store_p16_split(uint32_t taddr, const float (&pv)[N], int off) {
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float a = pv[off + 2 * i], b = pv[off + 2 * i + 1];
        const __nv_bfloat162 h2 = __floats2bfloat162_rn(a, b);
        const float2 hf = __bfloat1622float2(h2);
        const __nv_bfloat162 l2 = __floats2bfloat162_rn(a - hf.x, b - hf.y);
        hi[i] = *reinterpret_cast<const uint32_t*>(&h2);
        lo[i] = *reinterpret_cast<const uint32_t*>(&l2);
    }
    tmem_st8(taddr, hi);
    tmem_st8(taddr + 8, lo);
}

// =================================================================================== backward
// 18 warps: warp 0 = TMA producer + per-tile column factors, warp 1 = MMA issuer + TMEM owner,
// warps 2..17 = four epilogue warpgroups.
//
// A CTA owns a 128-row block (one UMMA M) at a time and walks its span of 128-column tiles.  S is
// TRIPLE-buffered in TMEM and all sixteen epilogue warps work on the same tile (warpgroup g takes columns
// g*32 .. g*32+31 of every row), so
//   * the MUFU pipe always has four warps per sub-partition to draw from (two were not enough to keep it
//     busy: a warp's exps share six scoreboards), and
//   * the tensor pipe runs two tiles ahead: while the warps turn S(t) into P(t), S(t+1) and S(t+2) are
//     already waiting and dK += P(t-1) K_j(t-1) is in flight.  The chain MMA1 -> tcgen05.ld -> exp ->
//     tcgen05.st -> MMA2 is as long as before, but three tiles are inside it instead of one per row half.
// P never touches shared memory: each thread packs its 32 values to bf16 and writes the 16 words with
// tcgen05.st over the first half of the S columns it has just read; the second MMA takes its A operand
// from TMEM.  (With P in shared memory the two MMAs plus the P stores need ~160 B/clk of shared memory.)
// No ordering barrier is needed between MMA2(t) (reads P in S buffer t%3) and MMA1(t+3) (overwrites it):
// both are issued by the same thread and the tensor pipe executes in issue order.
//
// Column data of a tile (1-exp factors C_j, log-sum-exps, "unsafe" flag) is staged next to its B tile by
// the otherwise idle producer warp, and becomes visible through the same b_full barrier.
//
// smem: A[KCH*kSub] | B[5][KCH*kSub] | cols[5][2][128] f32 | flags[5][4] u32 | barriers | tmem slot
// TMEM: S[b] at columns b*128 (b = 0..2), dK at 384 .. 384 + 64*KCH.

constexpr int kBMB = 128;       // rows per row block of the backward
constexpr int kBwdThreads = 608;     // the forward's 18 warps + one more issuing warp
constexpr int kBwdMma1Warp = 18;

// kX3 (SCB_ENGINE_BF16X3): operands split into bf16 hi + lo (see the forward kernel).  S takes three MMA groups; P is
// split the same way -- of every 16 columns a thread has read, the first 8 TMEM words get bf16(P), the next 8 get
// bf16(P - bf16(P)) -- and dK += P_hi K_hi + P_hi K_lo + P_lo K_hi.  Tiles are twice as large: two B stages.
template <int KCH, bool kDirect, bool kX3>
__global__ void __launch_bounds__(kBwdThreads, 1)
supcon_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_lo,
                     const __grid_constant__ TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int A_PART = KCH * kSub;
    constexpr int B_PART = KCH * kSub;
    constexpr int A_BYTES = (kX3 ? 2 : 1) * A_PART;
    constexpr int B_BYTES = (kX3 ? 2 : 1) * B_PART;
    constexpr int NSB = kX3 ? 2 : 5;
    constexpr int NS = 3;                          // S buffers
    constexpr int ND = KCH * 64;                   // dK columns
    constexpr uint32_t kDkCol = 384;
    uint8_t* a_smem = smem;
    uint8_t* b_smem = a_smem + A_BYTES;
    float* cols = reinterpret_cast<float*>(b_smem + NSB * B_BYTES);       // [NSB][2][128]
    uint32_t* flags = reinterpret_cast<uint32_t*>(cols + NSB * 256);      // [NSB][4]
    uint64_t* bars = reinterpret_cast<uint64_t*>(flags + NSB * 4 + 4);    // 23 used
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
    TcParams* sp = reinterpret_cast<TcParams*>(bars + 28);          // see the forward kernel
    if (kDirect)
        for (int i = threadIdx.x; i < (int)(sizeof(TcParams) / 4); i += kBwdThreads)
            reinterpret_cast<uint32_t*>(sp)[i] = reinterpret_cast<const uint32_t*>(&p)[i];
    const uint32_t bar0 = smem_u32(bars);
    const uint32_t a_full = bar0, a_empty = bar0 + 8u;
    auto b_full = [&](int s) { return bar0 + 8u * (2 + s); };
    auto b_empty = [&](int s) { return bar0 + 8u * (7 + s); };
    auto s_full = [&](int b) { return bar0 + 8u * (12 + b); };
    auto p_full = [&](int b) { return bar0 + 8u * (15 + b); };
    const uint32_t dz_full = bar0 + 8u * 18, dz_empty = bar0 + 8u * 19;
    auto s_free = [&](int b) { return bar0 + 8u * (20 + b); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (kTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) p.trace[1023] = global_ns();
    stamp_cta(p.trace, 0);
    if ((smem_u32(smem) & 1023u) != 0u) __trap();
    if (threadIdx.x == 0) {
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int s = 0; s < NSB; ++s) { mbar_init(b_full(s), 2); mbar_init(b_empty(s), 1); }
        for (int b = 0; b < NS; ++b) { mbar_init(s_full(b), 1); mbar_init(p_full(b), 8); mbar_init(s_free(b), 1); }    // p_full: one arrival per warp of a team
        mbar_init(dz_full, 1);
        mbar_init(dz_empty, 16);                          // one arrival per epilogue warp
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 512);
    if (warp == 0 && lane == 0) prefetch_tensormap(&tmap);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                                  // everything above touched no global memory
    if (kTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) p.trace[2047] = global_ns();
    // the tail's scale factor, upstream gradient included: fetched now, so that the finish does not start with a
    // global-memory round trip
    if (kDirect && threadIdx.x == 32)
        sp->fin.coef_scale_invT = load_scalar(p.grad_out, p.fin.grad_dtype) * p.fin.coef_scale_invT;

    const int t0 = blockIdx.x * p.q;
    const int n_tiles = min(p.T, t0 + p.q) - t0;
    const int rb0 = t0 / p.CT, ct0 = t0 - rb0 * p.CT;
    // reference point of the 1-exp path: P_ij = 2^(s - l_i) (1 + R_i C_j), R_i = 2^(l_i - m0), C_j = 2^(m0 - l_j);
    // any m0 works as long as every |l - m0| stays below kSafeSpread, so one global value does
    const float m0 = __ldcg(p.lse_all) * kLog2e;     // coherent loads: see pdl_wait() in common.cuh

    if (warp == 0) {
        // ---------------------------------------------------- TMA producer + column factors
        auto load_lse4 = [&](int ct_) {
            const int j = ct_ * kBN + lane * 4;
            float4 l;
            l.x = (j + 0 < p.n_k) ? __ldcg(p.lse_all + j + 0) : 0.f;
            l.y = (j + 1 < p.n_k) ? __ldcg(p.lse_all + j + 1) : 0.f;
            l.z = (j + 2 < p.n_k) ? __ldcg(p.lse_all + j + 2) : 0.f;
            l.w = (j + 3 < p.n_k) ? __ldcg(p.lse_all + j + 3) : 0.f;
            return l;
        };
        // the log-sum-exps of a tile's columns are fetched three tiles ahead: a load issued and consumed in
        // consecutive iterations would put one L2 round trip into every iteration of this loop
        int rb = rb0, ct = ct0, n_a = 0, ct_pf = ct0;
        auto next_pf = [&](int t_pf) {
            float4 l = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t_pf < n_tiles) l = load_lse4(ct_pf);
            if (++ct_pf == p.CT) ct_pf = 0;
            return l;
        };
        float4 l_cur = next_pf(0), l_n1 = next_pf(1), l_n2 = next_pf(2);
        for (int t = 0; t < n_tiles; ++t) {
            const int st = t % NSB;
            if (t == 0 || ct == 0) {
                mbar_wait(a_empty, (n_a & 1) ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(a_full, A_BYTES);
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc) {
                        tma_load_2d(smem_u32(a_smem + kc * kSub), &tmap, kc * 64, p.row_begin + rb * kBMB, a_full);
                        if (kX3)
                            tma_load_2d(smem_u32(a_smem + A_PART + kc * kSub), &tmap_lo, kc * 64, p.row_begin + rb * kBMB,
                                        a_full);
                    }
                }
                __syncwarp();
                ++n_a;
            }
            mbar_wait(b_empty(st), ((t / NSB) & 1) ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(b_full(st), B_BYTES);
#pragma unroll
                for (int kc = 0; kc < KCH; ++kc) {
                    tma_load_2d(smem_u32(b_smem + st * B_BYTES + kc * kSub), &tmap, kc * 64, ct * kBN, b_full(st));
                    if (kX3)
                        tma_load_2d(smem_u32(b_smem + st * B_BYTES + B_PART + kc * kSub), &tmap_lo, kc * 64, ct * kBN,
                                    b_full(st));
                }
            }
            __syncwarp();
            // cj[c] = [anchor_j] 2^(m0 - l_j)  (1-exp path),  lj[c] = l_j or +inf  (2-exp path)
            const float lv[4] = {l_cur.x, l_cur.y, l_cur.z, l_cur.w};
            float cv[4], l2v[4];
            bool unsafe = false;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = ct * kBN + lane * 4 + u;
                const bool okj = (j < p.n_k) && p.map.is_anchor(j, p.mode);
                const float l2 = lv[u] * kLog2e;
                const float d = m0 - l2;
                const bool bad = okj && !(fabsf(d) <= kSafeSpread);
                unsafe = unsafe || bad;
                cv[u] = (okj && !bad) ? ex2(d) : 0.f;
                l2v[u] = okj ? l2 : INFINITY;
            }
            float* cst = cols + st * 256;
            *reinterpret_cast<float4*>(cst + lane * 4) = make_float4(cv[0], cv[1], cv[2], cv[3]);
            *reinterpret_cast<float4*>(cst + 128 + lane * 4) = make_float4(l2v[0], l2v[1], l2v[2], l2v[3]);
            const unsigned any = __any_sync(0xffffffffu, unsafe) ? 1u : 0u;
            if (lane == 0) flags[st * 4] = any;
            __syncwarp();
            if (lane == 0) mbar_arrive(b_full(st));          // second arrival: the column data is in place
            if (++ct == p.CT) { ct = 0; ++rb; }
            l_cur = l_n1;
            l_n1 = l_n2;
            l_n2 = next_pf(t + 3);
        }
    } else if (warp == kBwdMma1Warp) {
        // ---------------------------------------------------- MMA1 issuer: S[t%3] = A (128 rows) x B^T (128 columns)
        // Two issuing warps, because one warp that alternates "wait for P(t), issue MMA2(t), issue MMA1(t+3)"
        // puts two barrier wake-ups and two blocking issue bursts in series on every tile and leaves the tensor
        // pipe idle in between.  tcgen05 ops of different threads are not ordered, so the reuse of an S buffer
        // is handed over explicitly: s_free[b] completes when MMA2(t), which reads P out of it, has finished.
        constexpr uint32_t idesc1 = idesc_bf16(128, 128, 0);
        Tracer tr(p.trace, 0);
        int n_a = 0, ct1 = ct0;
        const uint32_t a_lo = desc_lo(smem_u32(a_smem), 16);
        for (int t = 0; t < n_tiles; ++t) {
            const bool first_of_rb = (t == 0) || (ct1 == 0);
            const bool last_of_rb = (t == n_tiles - 1) || (ct1 == p.CT - 1);
            const int st = t % NSB, buf = t % NS;
            tr(100);
            if (first_of_rb) { mbar_wait(a_full, n_a & 1); ++n_a; }
            mbar_wait(b_full(st), (t / NSB) & 1);
            mbar_wait(s_free(buf), ((t / NS) & 1) ^ 1);
            tc_fence_after();
            tr(101);
            const uint32_t d = tmem_base + buf * 128;
            const uint32_t b_lo = desc_lo(smem_u32(b_smem + st * B_BYTES), 16);
            if (elect_one()) {
#pragma unroll
                for (int grp = 0; grp < (kX3 ? 3 : 1); ++grp) {           // hi hi^T, hi lo^T, lo hi^T
                    const uint32_t ag = a_lo + ((grp == 2 ? A_PART : 0) >> 4);
                    const uint32_t bg = b_lo + ((grp == 1 ? B_PART : 0) >> 4);
#pragma unroll
                    for (int k16 = 0; k16 < KCH * 4; ++k16) {
                        const uint32_t off = ((k16 >> 2) * kSub + (k16 & 3) * 32) >> 4;
                        if (grp == 0 && k16 == 0) umma_bf16_lo<false>(d, ag + off, bg + off, idesc1);
                        else umma_bf16_lo<true>(d, ag + off, bg + off, idesc1);
                    }
                }
                umma_commit(s_full(buf));
                if (last_of_rb) umma_commit(a_empty);
            }
            __syncwarp();
            tr(102);
            if (++ct1 == p.CT) ct1 = 0;
        }
    } else if (warp == 1) {
        // ---------------------------------------------------- MMA2 issuer: dK += P(t) x K_j
        // P(t): 128 x 128 bf16 held in TMEM over S[t%3]; K_j: the tile's 128 rows read MN-major
        constexpr uint32_t idesc2 = idesc_bf16(128, ND, 1);
        int n_dz = 0, ct2 = ct0;
        for (int t = 0; t < n_tiles; ++t) {
            const bool first_of_rb = (t == 0) || (ct2 == 0);
            const bool last_of_rb = (t == n_tiles - 1) || (ct2 == p.CT - 1);
            const int st = t % NSB, buf = t % NS;
            mbar_wait(p_full(buf), (t / NS) & 1);
            if (first_of_rb) mbar_wait(dz_empty, (n_dz & 1) ^ 1);
            tc_fence_after();
            const uint32_t d = tmem_base + kDkCol;
            const uint32_t pa = tmem_base + buf * 128;
            const uint32_t b_lo = desc_lo(smem_u32(b_smem + st * B_BYTES), kSub);
            if (elect_one()) {
                // 16 columns of P = 8 TMEM words; the warpgroup of column half h keeps its 64 columns in words
                // h*64 .. h*64+31
                if (!kX3) {
                    umma_bf16_ts_dyn(d, pa, b_lo, idesc2, first_of_rb ? 0u : 1u);
#pragma unroll
                    for (int j16 = 1; j16 < 8; ++j16)
                        umma_bf16_ts<true>(d, pa + (j16 >> 2) * 64 + (j16 & 3) * 8, b_lo + ((j16 * 2048) >> 4), idesc2);
                } else {
                    // P_hi K_hi, P_hi K_lo, P_lo K_hi: of the 16 words of a 16-column group, hi = first 8, lo = next 8
                    umma_bf16_ts_dyn(d, pa, b_lo, idesc2, first_of_rb ? 0u : 1u);
#pragma unroll
                    for (int grp = 0; grp < 3; ++grp) {
                        const uint32_t bg = b_lo + ((grp == 1 ? B_PART : 0) >> 4);
                        const uint32_t pg = pa + (grp == 2 ? 8u : 0u);
#pragma unroll
                        for (int j16 = (grp == 0 ? 1 : 0); j16 < 8; ++j16)
                            umma_bf16_ts<true>(d, pg + (j16 >> 2) * 64 + (j16 & 3) * 16, bg + ((j16 * 2048) >> 4), idesc2);
                    }
                }
                umma_commit(s_free(buf));
                umma_commit(b_empty(st));
                if (last_of_rb) umma_commit(dz_full);
            }
            __syncwarp();
            if (last_of_rb) ++n_dz;
            if (++ct2 == p.CT) ct2 = 0;
        }
    } else {
        // ---------------------------------------------------- epilogue warpgroups
        // Two teams of two warpgroups take alternate tiles (team = tile parity); inside a team warpgroup
        // `half` owns columns half*64 .. half*64+63 of the tile.  Every hand-over in a tile's chain
        // (barrier wake-up, tcgen05.ld, tcgen05.st + wait, barrier arrive) costs 50..150 ns; with one team
        // they add up on every tile, with two the other team's exps fill the gaps.
        const int g = (warp - 2) >> 2;                  // warpgroup
        const int team = g >> 1, half = g & 1;
        const int q4 = warp & 3;                        // TMEM lane quarter this warp may touch
        const int r = q4 * 32 + lane;                   // row within the block
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;

        int rb = rb0, ct = ct0, cur_rb = -1, li = 0, gi = 0, n_dz = 0;
        bool rows_fast = false;
        float lse2_i = INFINITY, R_i = 0.f;
        Tracer tr(p.trace, (warp == 2) ? 1 : (warp == 10 ? 2 : -1));   // lane quarter 2 of teams 0 and 1

        // dK of a finished row block leaves TMEM as plain vector stores into this CTA's slot of the
        // partial buffer (the reduce kernel adds the slots); no float atomics on the hot path.
        // All four warpgroups take part (32 columns each), whichever team computed the block's last tile.
        auto drain = [&](int rb_) {
            tr(8);
            mbar_wait(dz_full, n_dz & 1);
            tr(9);
            ++n_dz;
            tc_fence_after();
            if (g * 32 < ND) {
                const int slot = blockIdx.x - (rb_ * p.CT) / p.q;
                // rows of this warp: q4*32 .. +32; after the transpose a store covers 128 contiguous bytes
                float* dst = p.dk_part + ((int64_t)slot * p.RB * kBMB + (int64_t)rb_ * kBMB + q4 * 32) * p.ldk +
                             g * 32 + lane;
                drain_block(tmem_base + lane_addr + kDkCol + g * 32, dst, p.ldk, lane);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dz_empty);
            tr(10);
        };

        if (kTrace && p.trace != nullptr && warp == 2 && lane == 0 && blockIdx.x < 256) p.trace[4096 + 2 * blockIdx.x] = global_ns();
        // The team walks its own tiles (t = team, team + 2, ...).  All bookkeeping is incremental: the code
        // between two tiles is a chain of dependent scalar instructions, and every one of them is latency.
        int st = team, bpar = 0, buf = team, spar = 0;      // B stage / S buffer of tile t and their phase parities
        ct = ct0 + team;
        while (ct >= p.CT) { ct -= p.CT; ++rb; }
        int next_drain = rb0;                               // row blocks are drained in order, by all sixteen warps
        const int rb_end = (t0 + n_tiles - 1) / p.CT + 1;
        for (int t = team;; t += 2) {
            const bool more = t < n_tiles;
            if (!kDirect) {
                const int rb_target = more ? rb : rb_end;   // every block before this tile's (at the end: all of them)
                while (next_drain < rb_target) drain(next_drain++);
            }
            if (!more) break;
            if (kTrace && t >= n_tiles - 2 && p.trace != nullptr && warp == 2 && lane == 0 && blockIdx.x < 256)
                p.trace[4096 + 2 * blockIdx.x + 1] = global_ns();      // diagnostics: last tile reached
            {
                if (rb != cur_rb) {
                    cur_rb = rb;
                    li = rb * kBMB + r;
                    gi = p.row_begin + li;
                    const bool row_ok = (li < p.n_rows);
                    const bool anchor_i = row_ok && p.map.is_anchor(gi, p.mode);
                    lse2_i = anchor_i ? p.lse_all[gi] * kLog2e : INFINITY;
                    const float d = lse2_i - m0;
                    const bool row_safe = anchor_i && (fabsf(d) <= kSafeSpread);
                    R_i = row_safe ? ex2(d) : 0.f;
                    // the 1-exp path needs every row of the warp to be an anchor with a tame log-sum-exp
                    rows_fast = __all_sync(0xffffffffu, row_safe);
                }
                const int j0 = ct * kBN + half * 64;          // first column of this warpgroup's slice
                const float* cj = cols + st * 256 + half * 64;   // 1-exp column factors
                const float* lj = cj + 128;                   // column log-sum-exps (log2 domain) or +inf
                const uint32_t taddr = tmem_base + lane_addr + buf * 128 + half * 64;
                tr(1);
                // S(t) ready implies the tile's column data is in place: MMA1(t) was issued after its b_full
                // completed (B tile + the producer warp's column stores), so no second wait here
                mbar_wait(s_full(buf), spar);
                tr(3);
                tc_fence_after();
                const bool fast = rows_fast && (flags[st * 4] == 0u) && (j0 + 64 <= p.n_k);
                if (fast) {
                    // ---------------- common case: one exp per element; two 32-column chunks, one copy of the code
                    const int g_lo = p.row_begin + rb * kBMB + q4 * 32;   // first row of this warp
#pragma unroll 1
                    for (int cc = 0; cc < 2; ++cc) {
                        uint32_t v[32];
                        tmem_ld32(taddr + cc * 32, v);
                        tmem_ld_wait_dep(v);
                        if (cc == 0) tr(4);
                        // exp, weight, (label term), (self column), bf16 pack, 16 words back into TMEM
                        float pv[32];
                        if (kPolyLane >= 0) {
#pragma unroll
                            for (int c4 = 0; c4 < 8; ++c4) {
                                const float4 c4v = *reinterpret_cast<const float4*>(cj + cc * 32 + c4 * 4);
                                const float cvv[4] = {c4v.x, c4v.y, c4v.z, c4v.w};
#pragma unroll
                                for (int u = 0; u < 4; ++u) {
                                    const float a = fmaf(__uint_as_float(v[c4 * 4 + u]), p.c2, -lse2_i);
                                    const float e = (u == kPolyLane) ? ex2_fma(a) : ex2(a);   // one in four on the FMA pipe
                                    pv[c4 * 4 + u] = e * fmaf(R_i, cvv[u], 1.f);
                                }
                            }
                        } else {
                            // the same arithmetic on packed pairs (FFMA2 / FMUL2): 2.5 instead of 4 issue slots per
                            // element around the MUFU, bitwise identical results
                            const uint64_t c2p = pack2(p.c2, p.c2), lneg = pack2(-lse2_i, -lse2_i);
                            const uint64_t Rp = pack2(R_i, R_i), onep = pack2(1.f, 1.f);
#pragma unroll
                            for (int c4 = 0; c4 < 8; ++c4) {
                                const float4 c4v = *reinterpret_cast<const float4*>(cj + cc * 32 + c4 * 4);
                                const uint64_t a01 = ffma2(pack2(__uint_as_float(v[c4 * 4]), __uint_as_float(v[c4 * 4 + 1])), c2p, lneg);
                                const uint64_t a23 = ffma2(pack2(__uint_as_float(v[c4 * 4 + 2]), __uint_as_float(v[c4 * 4 + 3])), c2p, lneg);
                                float a0, a1, a2, a3;
                                unpack2(a01, a0, a1);
                                unpack2(a23, a2, a3);
                                const uint64_t w01 = ffma2(Rp, pack2(c4v.x, c4v.y), onep);
                                const uint64_t w23 = ffma2(Rp, pack2(c4v.z, c4v.w), onep);
                                const uint64_t p01 = fmul2(pack2(ex2(a0), ex2(a1)), w01);
                                const uint64_t p23 = fmul2(pack2(ex2(a2), ex2(a3)), w23);
                                unpack2(p01, pv[c4 * 4], pv[c4 * 4 + 1]);
                                unpack2(p23, pv[c4 * 4 + 2], pv[c4 * 4 + 3]);
                            }
                        }
                        const int jc = j0 + cc * 32;
                        // the self column (losses.py:74-79) only exists in slices on the diagonal: warp-uniform test
                        if (g_lo < jc + 32 && jc < g_lo + 32) {
                            const int c_self = gi - jc;
#pragma unroll
                            for (int c = 0; c < 32; ++c) pv[c] = (c == c_self) ? 0.f : pv[c];
                        }
                        if (!kX3) {
                            store_p16(taddr + cc * 16, pv, 0);
                            store_p16(taddr + cc * 16 + 8, pv, 16);
                        } else {
                            store_p16_split(taddr + cc * 32, pv, 0);
                            store_p16_split(taddr + cc * 32 + 16, pv, 16);
                        }
                    }
                } else {
                    // ---------------- rare case (ragged tiles, non-anchor rows, wild LSE ranges): two exps,
                    // sixteen columns at a time straight from TMEM, compact code
#pragma unroll 1
                    for (int g16 = 0; g16 < 4; ++g16) {
                        uint32_t v8a[8], v8b[8];
                        tmem_ld8(taddr + g16 * 16, v8a);
                        tmem_ld8(taddr + g16 * 16 + 8, v8b);
                        tmem_ld_wait_dep8(v8a);
                        tmem_ld_wait_dep8(v8b);
                        float pv[16];
#pragma unroll
                        for (int u = 0; u < 16; ++u) {
                            const int c = g16 * 16 + u;
                            const int j = j0 + c;
                            const float x = __uint_as_float(u < 8 ? v8a[u & 7] : v8b[u & 7]);
                            const float e = ex2(fmaf(x, p.c2, -lse2_i)) + ex2(fmaf(x, p.c2, -lj[c]));   // +inf -> 0
                            pv[u] = (j >= p.n_k || j == gi) ? 0.f : e;
                        }
                        if (!kX3) store_p16(taddr + g16 * 8, pv, 0);     // words [8g, 8g+8) lie inside columns already read
                        else store_p16_split(taddr + g16 * 16, pv, 0); // hi | lo words over the 16 columns just read
                    }
                }
                tr(6);
                // publishing P also hands the S buffer back: every TMEM access of this thread is complete
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_full(buf));    // one arrival per warp: per-thread arrivals on one word serialise
                tr(7);
            }
            ct += 2;
            while (ct >= p.CT) { ct -= p.CT; ++rb; }
            st += 2;
            if (st >= NSB) { st -= NSB; bpar ^= 1; }
            buf += 2;
            if (buf >= NS) { buf -= NS; spar ^= 1; }
        }
        (void)bpar; (void)next_drain; (void)rb_end;
        if (kDirect) {
            // the CTA's single row block is complete once the last MMA2 has landed in TMEM
            mbar_wait(dz_full, 0);
            tc_fence_after();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (kDirect) {
        // ---------------------------------------------------- direct finish (no partial buffer, no reduce kernel)
        // The CTA walked (part of) ONE row block.  With two CTAs per block (a cluster pair, p.pair == 2) CTA c of
        // the pair finishes rows c*64 .. c*64+63: each CTA stages its own contribution to the rows it keeps in
        // `own` and ships the other 64 rows into the partner's `recv` over distributed shared memory.  The B
        // stages are free by now (all MMAs have completed), the staging arrays live there.
        const int pitch = p.ldk + 4;                       // floats; 16-byte rows, conflict-free float4 columns
        float* own = reinterpret_cast<float*>(b_smem);
        float* recv = own + 128 * pitch;
        const int rb_d = rb0;
        const uint32_t crank = (p.pair == 2) ? cluster_ctarank() : 0u;
        if (p.pair == 2) cluster_sync_all();               // the partner, too, is done with its B stages
        if (warp >= 2 && warp < 18) {
            const int g = (warp - 2) >> 2, q4 = warp & 3;
            tc_fence_after();
            if (g * 32 < ND) {
                const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + kDkCol + g * 32;
                if (p.pair == 2) {
                    const bool keep = (uint32_t)(q4 >> 1) == crank;
                    const int rr = (q4 & 1) * 32 + lane;                     // row within the kept half
                    const uint32_t local = smem_u32((keep ? own : recv) + rr * pitch + g * 32);
                    stage_block(taddr, keep ? local : mapa_shared(local, crank ^ 1u), keep ? 0 : 1);
                } else {
                    stage_block(taddr, smem_u32(own + (q4 * 32 + lane) * pitch + g * 32), 0);
                }
            }
        }
        if (p.pair == 2) cluster_sync_all(); else __syncthreads();
        if (warp >= 2 && warp < 18) {
            if (p.pair == 2) finish_staged(*sp, own, recv, pitch, rb_d * kBMB + (int)crank * 64, 64, warp - 2, lane);
            else finish_staged(*sp, own, nullptr, pitch, rb_d * kBMB, 128, warp - 2, lane);
        }
        tc_fence_before();
        __syncthreads();
    }
    if (kTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) p.trace[3071] = global_ns();
    stamp_cta(p.trace, 1);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------ label bookkeeping
// One warp per sample s: scans the labels once (ballots), which gives the class representative rep = smallest
// sample id with the same label (exact equality classes for arbitrary int64 labels) and the class size; fills
//   rep_row[j] = rep, npos_row[j] = V*count - 1, nia_row[j] = [anchor_j] / npos_row[j]      for the V rows j of s.
// With cls_all != nullptr the warp of a representative then adds up the rows of its class in a fixed order:
//   cls_all[rep] = sum of all rows of the class, cls_anc[rep] = sum of its anchor rows (mode `one`: view 0)
// -- the fp32 closed form of the positive-pair part of the gradient (positive_term above).
__global__ void __launch_bounds__(256)
label_rows_kernel(const int64_t* __restrict__ labels, int n_samples, RowMap map, int n_views, int mode,
                  const __nv_bfloat16* k, const __nv_bfloat16* k_lo, int ldk,
                  int* __restrict__ rep_row,
                  float* __restrict__ npos_row, float* __restrict__ nia_row, float* __restrict__ cls_all,
                  float* __restrict__ cls_anc) {
    const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    pdl_launch_dependents();
    pdl_wait();
    if (s >= n_samples) return;
    const int64_t mine = labels[s];
    int cnt = 0, rep = -1;
    for (int t0 = 0; t0 < n_samples; t0 += 32) {
        const int t = t0 + lane;
        const bool eq = t < n_samples && labels[t] == mine;
        const unsigned bits = __ballot_sync(0xffffffffu, eq);
        cnt += __popc(bits);
        if (bits != 0u && rep < 0) rep = t0 + __ffs(bits) - 1;
    }
    const float np = (float)(n_views * cnt - 1);
    for (int v = lane; v < n_views; v += 32) {
        const int j = map.row_of(s, v);
        rep_row[j] = rep;
        npos_row[j] = np;
        nia_row[j] = map.is_anchor(j, mode) ? 1.f / np : 0.f;
    }
    if (cls_all == nullptr || rep != s) return;
    const int d = lane * 4;
    const bool on = d < ldk;
    float sa[4] = {0.f, 0.f, 0.f, 0.f}, sn[4] = {0.f, 0.f, 0.f, 0.f};
    for (int t0 = 0; t0 < n_samples; t0 += 32) {
        const int t = t0 + lane;
        const bool eq = t < n_samples && labels[t] == mine;
        unsigned bits = __ballot_sync(0xffffffffu, eq);
        while (bits != 0u) {
            const int tt = t0 + __ffs(bits) - 1;
            bits &= bits - 1u;
            for (int v = 0; v < n_views; ++v) {
                const int j = map.row_of(tt, v);
                if (on) {
                    float kj[4];
                    load_k4(k + (int64_t)j * ldk, d, kj);
                    if (k_lo != nullptr) {
                        float kl[4];
                        load_k4(k_lo + (int64_t)j * ldk, d, kl);
#pragma unroll
                        for (int i = 0; i < 4; ++i) kj[i] += kl[i];
                    }
                    const bool anc = map.is_anchor(j, mode);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { sa[i] += kj[i]; sn[i] += anc ? kj[i] : 0.f; }
                }
            }
        }
    }
    if (on) {
        *reinterpret_cast<float4*>(cls_all + (int64_t)s * ldk + d) = make_float4(sa[0], sa[1], sa[2], sa[3]);
        if (cls_anc != cls_all)
            *reinterpret_cast<float4*>(cls_anc + (int64_t)s * ldk + d) = make_float4(sn[0], sn[1], sn[2], sn[3]);
    }
}

// ------------------------------------------------------------------ dK reduction (general geometry)
// Used when a row block is shared by more than two CTAs or a CTA walks several row blocks (the backward main
// kernel then leaves per-slot partial dK blocks in the workspace).  One warp per local row:
//   sum over the slots of the row block + closed-form positive-pair term, times coef/T.
__device__ __forceinline__ void sum_dk_slots(const TcParams& p, int lrow, int d, float (&acc)[4]) {
    const int rb = lrow >> 7;                      // 128-row blocks of the backward
    const int n_slots = ((rb + 1) * p.CT - 1) / p.q - (rb * p.CT) / p.q + 1;
    const int slot_rows = p.RB * kBMB;
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = 0.f;
    for (int s0 = 0; s0 < n_slots; s0 += 8) {          // eight slots at a time: all loads in flight together
        float4 t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            t[u] = (s0 + u < n_slots)
                       ? __ldcs(reinterpret_cast<const float4*>(p.dk_part + ((int64_t)(s0 + u) * slot_rows + lrow) * p.ldk + d))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 8; ++u) { acc[0] += t[u].x; acc[1] += t[u].y; acc[2] += t[u].z; acc[3] += t[u].w; }
    }
}

// fp32 dK rows out (scb_supcon_backward)
__global__ void __launch_bounds__(256)
supcon_tc_reduce_dk_kernel(const __grid_constant__ TcParams p) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    pdl_launch_dependents();
    pdl_wait();
    if (warp >= p.n_rows) return;
    const float coef = load_scalar(p.grad_out, p.fin.grad_dtype) * p.fin.coef_scale_invT;
    float* dst = p.out_dk + (int64_t)warp * p.ld_dk;
    const int d = lane * 4;
    if (d >= p.ldk) return;
    float acc[4], pos[4];
    sum_dk_slots(p, warp, d, acc);
    positive_term(p, p.row_begin + warp, d, pos);
    *reinterpret_cast<float4*>(dst + d) = make_float4(coef * (acc[0] + pos[0]), coef * (acc[1] + pos[1]),
                                                      coef * (acc[2] + pos[2]), coef * (acc[3] + pos[3]));
}

// d features out (scb_supcon_backward_features): same reduction + un-stack, normalise-backward, cast
__global__ void __launch_bounds__(256)
supcon_tc_reduce_finish_kernel(const __grid_constant__ TcParams p) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    pdl_launch_dependents();
    pdl_wait();
    if (warp >= p.n_rows) return;
    const float coef = load_scalar(p.grad_out, p.fin.grad_dtype) * p.fin.coef_scale_invT;
    float acc[1][4];
    const int lrow[1] = {warp};
    if (lane * 4 < p.ldk) sum_dk_slots(p, warp, lane * 4, acc[0]);
    else { acc[0][0] = acc[0][1] = acc[0][2] = acc[0][3] = 0.f; }
    finish_rows<1>(p, acc, lrow, lane, coef);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || ptr == nullptr)
        return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(ptr);
    return fn;
}

static bool thread_has_context() {
    typedef CUresult (*GetCurrentFn)(CUcontext*);
    static GetCurrentFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &ptr, cudaEnableDefault, &qres) != cudaSuccess || !ptr) return false;
        fn = reinterpret_cast<GetCurrentFn>(ptr);
    }
    CUcontext ctx = nullptr;
    return fn(&ctx) == CUDA_SUCCESS && ctx != nullptr;
}

static int make_tmap_ptr(const scb_supcon_problem* p, const void* kptr, CUtensorMap* out);
static int make_tmap(const scb_supcon_problem* p, CUtensorMap* out) { return make_tmap_ptr(p, p->k, out); }
static int make_tmap_ptr(const scb_supcon_problem* p, const void* kptr, CUtensorMap* out) {
    // The encoder is a driver-API entry point and needs a context current on THIS thread; autograd calls
    // the backward from its own thread, where nothing may have touched the runtime yet.
    if (!thread_has_context()) SCB_CUDA_OK(cudaFree(nullptr));     // binds the primary context (never during capture:
                                                                    // a thread that captures has used the runtime before)
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return SCB_ERR_CUDA; }
    SCB_REQUIRE((reinterpret_cast<uintptr_t>(kptr) & 15u) == 0, "supcon: k must be 16-byte aligned");
    // the descriptor depends on (address, rows, row stride) only: forward and backward of a step, and every
    // step of a training loop that re-uses its buffers, hit this small per-thread cache instead of the driver
    struct Entry { const void* k; int n_k, ldk; CUtensorMap map; };
    static thread_local Entry cache[8] = {};
    static thread_local int next = 0;
    for (const Entry& e : cache)
        if (e.k == kptr && e.n_k == p->n_k && e.ldk == p->ldk) { *out = e.map; return SCB_OK; }
    cuuint64_t gdim[2] = {(cuuint64_t)p->ldk, (cuuint64_t)p->n_k};
    cuuint64_t gstride[1] = {(cuuint64_t)p->ldk * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(kptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with code %d", (int)r); return SCB_ERR_CUDA; }
    cache[next] = Entry{kptr, p->n_k, p->ldk, *out};
    next = (next + 1) & 7;
    return SCB_OK;
}

// Generic [rows, cols] bf16 matrix with row stride `ld` (elements), box [128 rows x 64 columns], 128-byte swizzle: the
// operand layout of every tcgen05 kernel in this library (used by head_tc.cu).  Out-of-range rows / columns read zero.
int tc_make_tmap_2d(const void* ptr, int64_t rows, int64_t cols, int64_t ld, CUtensorMap* out) {
    if (!thread_has_context()) SCB_CUDA_OK(cudaFree(nullptr));
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return SCB_ERR_CUDA; }
    SCB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15u) == 0 && (ld * 2) % 16 == 0,
                "tensor map: base and row stride must be 16-byte aligned");
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with code %d", (int)r); return SCB_ERR_CUDA; }
    return SCB_OK;
}

static constexpr int kMaxSlots = 32;
// diagnostics hooks: set from the caller's thread, read by whichever thread launches (autograd's, for the backward)
static std::atomic<long long*> g_trace{nullptr};   // set through scb_debug_set_trace
void tc_set_trace(long long* p) { g_trace.store(p, std::memory_order_release); }
// measurement aid (scb_debug_set_events): events recorded right before / after the backward main kernel
static std::atomic<cudaEvent_t> g_ev_start{nullptr}, g_ev_stop{nullptr};
void tc_set_events(void* a, void* b) {
    g_ev_start.store((cudaEvent_t)a, std::memory_order_release);
    g_ev_stop.store((cudaEvent_t)b, std::memory_order_release);
}

static_assert(sizeof(TcParams) % 4 == 0 && sizeof(TcParams) <= 1024, "TcParams copy in shared memory");
struct TcGeom { int CT, RB, T, q, grid, n_split, direct; };

static TcGeom tc_geom(int n_k, int rows) {
    TcGeom g;
    g.CT = (int)ceil_div(n_k, kBN);
    g.RB = (int)ceil_div(rows, kBM);
    g.T = g.CT * g.RB;
    int q = (int)ceil_div(g.T, sm_count());
    const int qmin = (int)ceil_div(g.CT, kMaxSlots / 2 - 1);
    if (q < qmin) q = qmin;
    g.q = q;
    g.grid = (int)ceil_div(g.T, q);
    g.n_split = 2 * ((int)ceil_div(g.CT, q) + 1);      // (CTA overlapping the row block) x (column half)
    if (g.n_split > kMaxSlots) g.n_split = kMaxSlots;
    g.direct = 0;
    return g;
}

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

// backward: 128-row blocks; n_split = slots of a row block = CTAs whose tile span overlaps it
static constexpr int kMaxDkSlots = 16;
static TcGeom tc_geom_bwd(int n_k, int rows) {
    TcGeom g;
    g.CT = (int)ceil_div(n_k, kBN);
    g.RB = (int)ceil_div(rows, kBMB);
    g.T = g.CT * g.RB;
    int q = (int)ceil_div(g.T, sm_count());
    const int qmin = (int)ceil_div(g.CT, kMaxDkSlots - 1);
    if (q < qmin) q = qmin;
    // A span that stays inside one row block saves its CTA a mid-kernel drain (3.3 us) and the reduce kernel a
    // slot: take the next q that divides CT (or is a multiple of it) when that costs at most 15 % more tiles
    // per CTA.
    bool aligned = false;
    for (int qa = q; qa * 100 <= q * 115; ++qa)
        if (g.CT % qa == 0 || qa % g.CT == 0) { q = qa; aligned = true; break; }
    g.q = q;
    g.grid = (int)ceil_div(g.T, q);
    g.n_split = aligned ? (int)ceil_div(g.CT, q) : (int)ceil_div(g.CT, q) + 1;
    if (g.n_split < 1) g.n_split = 1;
    if (g.n_split > kMaxDkSlots) g.n_split = kMaxDkSlots;
    // one or two CTAs per row block, each inside it: the kernel finishes the rows itself (cluster pair, no partials)
    g.direct = (aligned && q <= g.CT && g.CT % q == 0 && g.CT / q <= 2) ? g.CT / q : 0;
    return g;
}

static size_t dk_part_bytes(const scb_supcon_problem* p, int n_rows) {
    const TcGeom g = tc_geom_bwd(p->n_k, n_rows);
    return align256((size_t)g.n_split * g.RB * kBMB * p->ldk * sizeof(float));
}

// label block: rep_row | npos_row | nia_row [n_k] | cls_all | cls_anc [n_samples][ldk]
static size_t label_block_bytes(const scb_supcon_problem* p) {
    if (p->mask_kind != SCB_MASK_LABELS) return 0;
    return 3 * align256((size_t)p->n_k * 4) + 2 * align256((size_t)p->n_seg * p->seg * p->ldk * sizeof(float));
}
int64_t tc_label_cache_bytes(const scb_supcon_problem* p) { return (int64_t)label_block_bytes(p); }

static constexpr int kMaxFwdRowBlocks = 4090;                 // 256-row blocks of the forward: up to ~1M local rows
static_assert((size_t)(kMaxFwdRowBlocks + 4) * sizeof(unsigned) <= (size_t)kTicketRegionBytes, "ticket region");
struct TcWs { float* part; char* labels; unsigned* rb_count; float* dk_part; size_t total; };
static TcWs tc_ws_carve(const scb_supcon_problem* p, int n_rows, void* base) {
    TcWs w;
    char* ws = (char*)base;
    // The tickets must find zeros: they sit at a FIXED place in front of the engine scratch (supcon_common.cuh),
    // where nothing else is ever written whatever the previous call on this workspace was, and re-arm themselves.
    w.rb_count = base ? (unsigned*)(ws - kTicketRegionBytes) : nullptr;       // [kMaxFwdRowBlocks] tickets | done counter
    w.part = (float*)ws; ws += align256((size_t)kMaxSlots * 4 * n_rows * sizeof(float));
    w.labels = ws; ws += label_block_bytes(p);
    w.dk_part = (float*)ws; ws += dk_part_bytes(p, n_rows);
    w.total = (size_t)(ws - (char*)base);
    return w;
}

int64_t tc_workspace_bytes(const scb_supcon_problem* p, int n_rows) {
    if (p->engine == SCB_ENGINE_FP32) return 0;
    return (int64_t)tc_ws_carve(p, n_rows, nullptr).total;
}

static int tc_supported(const scb_supcon_problem* p) {
    if (p->engine == SCB_ENGINE_BF16X3 && p->k_lo == nullptr) { set_error("the bf16x3 engine needs k_lo"); return SCB_ERR_ARG; }
    if (p->mask_kind == SCB_MASK_EXPLICIT) { set_error("tensor-core engine: explicit masks run on ENGINE_FP32"); return SCB_ERR_UNSUPPORTED; }
    if ((int64_t)p->n_k > (int64_t)kMaxFwdRowBlocks * kBM) { set_error("tensor-core engine: more than %d rows", kMaxFwdRowBlocks * kBM); return SCB_ERR_UNSUPPORTED; }
    if (p->ldk != 64 && p->ldk != 128) { set_error("tensor-core engine: ldk must be 64 or 128 (dim <= 128), got %d", p->ldk); return SCB_ERR_UNSUPPORTED; }
    return SCB_OK;
}

static void fill_params(const scb_supcon_problem* p, int row_begin, int n_rows, int a_count, const TcGeom& g,
                        void* workspace, TcParams* t) {
    memset(t, 0, sizeof(*t));
    t->k = (const __nv_bfloat16*)p->k;
    t->k_lo = (p->engine == SCB_ENGINE_BF16X3) ? (const __nv_bfloat16*)p->k_lo : nullptr;
    t->ldk = p->ldk; t->n_k = p->n_k; t->dim = p->dim;
    t->row_begin = row_begin; t->n_rows = n_rows; t->a_count = a_count;
    t->CT = g.CT; t->RB = g.RB; t->T = g.T; t->q = g.q;
    t->map = RowMap::of(p);
    t->n_views = p->n_views; t->mode = p->mode; t->mask_kind = p->mask_kind;
    t->c2 = p->inv_temperature * kLog2e;
    t->inv_T = p->inv_temperature;
    const TcWs w = tc_ws_carve(p, n_rows, workspace);
    t->part = w.part;
    if (p->mask_kind == SCB_MASK_LABELS) {
        char* lb = p->label_cache ? (char*)p->label_cache : w.labels;
        t->rep_row = (const int*)lb; lb += align256((size_t)p->n_k * 4);
        t->npos_row = (const float*)lb; lb += align256((size_t)p->n_k * 4);
        t->nia_row = (const float*)lb; lb += align256((size_t)p->n_k * 4);
        t->fin.cls_all = (const float*)lb; lb += align256((size_t)p->n_seg * p->seg * p->ldk * sizeof(float));
        t->fin.cls_anc = (p->mode == SCB_MODE_ALL) ? t->fin.cls_all : (const float*)lb;
    }
    t->fo.rb_count = w.rb_count;
    t->fo.acc = reinterpret_cast<unsigned long long*>(w.rb_count + kMaxFwdRowBlocks);
    t->fo.po = PeerOut::none();
    t->dk_part = w.dk_part;
    t->pair = 0;
    t->trace = g_trace.load(std::memory_order_acquire);
}

static int launch_label_rows(const scb_supcon_problem* p, const TcParams& t, bool with_sums, cudaStream_t stream) {
    if (p->mask_kind != SCB_MASK_LABELS) return SCB_OK;
    prefer_max_smem_carveout(label_rows_kernel);
    const int n_samples = p->n_seg * p->seg;
    SCB_CUDA_OK(launch_pdl(label_rows_kernel, dim3((unsigned)ceil_div((int64_t)n_samples * 32, 256)), dim3(256), 0, stream,
                           p->labels, n_samples, t.map, p->n_views, p->mode, t.k, t.k_lo, p->ldk, (int*)t.rep_row,
                           (float*)t.npos_row, (float*)t.nia_row, with_sums ? (float*)t.fin.cls_all : (float*)nullptr,
                           with_sums ? (float*)t.fin.cls_anc : (float*)nullptr));
    count_launch();
    return SCB_OK;
}

template <typename Kern>
static int set_smem(Kern kern, size_t bytes) {
    SCB_CUDA_OK(set_max_dynamic_smem(kern, (int)bytes));
    return SCB_OK;
}

// programmatic stream serialisation + (optionally) a thread-block cluster of `cluster_x` consecutive CTAs
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                      int cluster_x, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    if (cluster_x > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = (unsigned)cluster_x;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Forward for the anchors among the local rows: ONE launch (plus the label bookkeeping kernel for label masks);
// the per-CTA partials are merged inside the kernel (fwd_merge_rows).  ch != nullptr: the per-row results and the
// rank's loss share go to every rank's statistics array and the rank's flag is raised (scb_supcon_forward_peers).
int tc_forward(const scb_supcon_problem* p, int row_begin, int n_rows, int n_anchor_total, float* out_lse,
               float* out_possum, float* out_npos, float* out_loss, void* workspace, const scb_peer_channel* ch,
               cudaStream_t stream) {
    if (int e = tc_supported(p)) return e;
    CUtensorMap tmap;
    if (int e = make_tmap(p, &tmap)) return e;
    const int a_count = (p->mode == SCB_MODE_ALL) ? n_rows : (n_rows < p->seg ? n_rows : p->seg);
    const TcGeom g = tc_geom(p->n_k, a_count);
    TcParams t;
    fill_params(p, row_begin, n_rows, a_count, g, workspace, &t);
    t.fo.out_lse = out_lse; t.fo.out_possum = out_possum; t.fo.out_npos = out_npos; t.fo.out_loss = out_loss;
    t.fo.loss_scale = p->loss_scale;
    t.fo.inv_anchors = 1.0 / (double)n_anchor_total;
    t.fo.scalar_dtype = p->scalar_dtype;
    t.fo.n_k = p->n_k;
    t.fo.my_slot = row_begin / n_rows;               // rank: every rank holds n_rows consecutive rows
    if (ch) t.fo.po = PeerOut::of(ch);
    // with a label cache the class sums of the backward are built here, once per step
    if (int e = launch_label_rows(p, t, p->label_cache != nullptr, stream)) return e;
    const int kch = p->ldk / 64;
    const bool x3 = p->engine == SCB_ENGINE_BF16X3;
    CUtensorMap tmap_lo = tmap;
    if (x3) { if (int e = make_tmap_ptr(p, p->k_lo, &tmap_lo)) return e; }
    // bf16: A[2] + B[2]; bf16x3: (hi | lo) tiles, single-buffered -- the same 192 KiB either way at ldk = 128
    const size_t smem = (size_t)2 * (2 * kch * kSub) + 2 * (kch * kSub) + 40 * 8 + 1024;
    const bool lab = p->mask_kind == SCB_MASK_LABELS;
#define SCB_LAUNCH_FWD(KCH_, LAB_, X3_)                                                                 \
    do {                                                                                                \
        if (int e = set_smem(supcon_tc_fwd_kernel<KCH_, LAB_, X3_>, smem)) return e;                    \
        SCB_CUDA_OK(launch_pdl(supcon_tc_fwd_kernel<KCH_, LAB_, X3_>, dim3(g.grid), dim3(kTcThreads), smem, stream, tmap, \
                               tmap_lo, t));                                                            \
    } while (0)
    if (x3) {
        if (kch == 2) { if (lab) SCB_LAUNCH_FWD(2, true, true); else SCB_LAUNCH_FWD(2, false, true); }
        else { if (lab) SCB_LAUNCH_FWD(1, true, true); else SCB_LAUNCH_FWD(1, false, true); }
    } else {
        if (kch == 2) { if (lab) SCB_LAUNCH_FWD(2, true, false); else SCB_LAUNCH_FWD(2, false, false); }
        else { if (lab) SCB_LAUNCH_FWD(1, true, false); else SCB_LAUNCH_FWD(1, false, false); }
    }
#undef SCB_LAUNCH_FWD
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

// main kernel of the backward.  fin == nullptr: fills the per-slot partial dK blocks in the workspace (a reduce
// kernel follows).  fin != nullptr and a geometry with one or two CTAs per row block: the kernel writes d features
// itself (g_out->direct != 0) and nothing follows.
static int tc_backward_main(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                            const float* grad_out, int n_anchor_total, void* workspace, const FinishParams* fin,
                            TcGeom* g_out, TcParams* t_out, cudaStream_t stream) {
    if (int e = tc_supported(p)) return e;
    CUtensorMap tmap;
    if (int e = make_tmap(p, &tmap)) return e;
    TcGeom g = tc_geom_bwd(p->n_k, n_rows);
    static const bool no_direct = [] { const char* e = getenv("SCB_BWD_DIRECT"); return e != nullptr && e[0] == '0'; }();
    if (fin == nullptr || no_direct) g.direct = 0;
    if (fin != nullptr && g.direct) {
        // the in-kernel finish uses 16-byte vector accesses on 8 consecutive features
        const int es = fin->feat_dtype == SCB_F32 ? 4 : 2;
        const bool vec_ok = p->dim % 8 == 0 && (reinterpret_cast<uintptr_t>(fin->out) & 15u) == 0 &&
                            fin->ostride_s % 8 == 0 && fin->ostride_v % 8 == 0 &&
                            (!fin->normalize || ((reinterpret_cast<uintptr_t>(fin->feat) & 15u) == 0 &&
                                                 fin->stride_s % 8 == 0 && fin->stride_v % 8 == 0));
        (void)es;
        if (!vec_ok) g.direct = 0;
    }
    TcParams t;
    fill_params(p, row_begin, n_rows, n_rows, g, workspace, &t);
    t.lse_all = lse_all; t.grad_out = grad_out;
    t.coef_scale = p->loss_scale / (float)n_anchor_total;
    {
        const float* ca = t.fin.cls_all; const float* cn = t.fin.cls_anc;
        if (fin) t.fin = *fin;
        t.fin.cls_all = ca; t.fin.cls_anc = cn;
        t.fin.grad_dtype = p->scalar_dtype;
        t.fin.coef_scale_invT = t.coef_scale * p->inv_temperature;
    }
    t.pair = g.direct;
    // class sums of the positive-pair term: built by the forward when the caller passed a label cache
    if (p->label_cache == nullptr)
        if (int e = launch_label_rows(p, t, true, stream)) return e;
    const int kch = p->ldk / 64;
    const bool x3 = p->engine == SCB_ENGINE_BF16X3;
    CUtensorMap tmap_lo = tmap;
    if (x3) { if (int e = make_tmap_ptr(p, p->k_lo, &tmap_lo)) return e; }
    const int nsb = x3 ? 2 : 5, parts = x3 ? 2 : 1;
    const size_t smem = (size_t)parts * (kch * kSub) + (size_t)nsb * parts * (kch * kSub) + nsb * 256 * 4 +
                        (nsb * 4 + 4) * 4 + 28 * 8 + 1024;
#define SCB_LAUNCH_BWD(KCH_, DIR_, X3_)                                                                 \
    do {                                                                                                \
        if (int e = set_smem(supcon_tc_bwd_kernel<KCH_, DIR_, X3_>, smem)) return e;                    \
        SCB_CUDA_OK(launch_pdl_cluster(supcon_tc_bwd_kernel<KCH_, DIR_, X3_>, dim3(g.grid), dim3(kBwdThreads), smem, stream, \
                                       DIR_ ? g.direct : 1, tmap, tmap_lo, t));                         \
    } while (0)
    if (cudaEvent_t ev = g_ev_start.load(std::memory_order_acquire)) SCB_CUDA_OK(cudaEventRecord(ev, stream));
    if (x3) {
        if (kch == 2) { if (g.direct) SCB_LAUNCH_BWD(2, true, true); else SCB_LAUNCH_BWD(2, false, true); }
        else { if (g.direct) SCB_LAUNCH_BWD(1, true, true); else SCB_LAUNCH_BWD(1, false, true); }
    } else {
        if (kch == 2) { if (g.direct) SCB_LAUNCH_BWD(2, true, false); else SCB_LAUNCH_BWD(2, false, false); }
        else { if (g.direct) SCB_LAUNCH_BWD(1, true, false); else SCB_LAUNCH_BWD(1, false, false); }
    }
#undef SCB_LAUNCH_BWD
    if (cudaEvent_t ev = g_ev_stop.load(std::memory_order_acquire)) SCB_CUDA_OK(cudaEventRecord(ev, stream));
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    *g_out = g;
    *t_out = t;
    return SCB_OK;
}

int tc_backward(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all, const float* npos_all,
                const float* grad_out, int n_anchor_total, float* out_dk, int ld_dk, void* workspace,
                cudaStream_t stream) {
    (void)npos_all;
    SCB_REQUIRE(ld_dk >= p->ldk && ld_dk % 4 == 0 && (reinterpret_cast<uintptr_t>(out_dk) & 15u) == 0,
                "supcon_backward (tensor-core engine): out_dk must be 16-byte aligned with ld_dk >= ldk (%d)", p->ldk);
    TcGeom g;
    TcParams t;
    if (int e = tc_backward_main(p, row_begin, n_rows, lse_all, grad_out, n_anchor_total, workspace, nullptr, &g, &t,
                                 stream))
        return e;
    t.out_dk = out_dk; t.ld_dk = ld_dk;
    prefer_max_smem_carveout(supcon_tc_reduce_dk_kernel);
    SCB_CUDA_OK(launch_pdl(supcon_tc_reduce_dk_kernel, dim3((unsigned)ceil_div((int64_t)n_rows * 32, 256)), dim3(256), 0,
                           stream, t));
    count_launch();
    return SCB_OK;
}

int tc_backward_features(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                         const float* grad_out, int n_anchor_total, const void* features, int feat_dtype,
                         int n_samples, int64_t stride_s, int64_t stride_v, int normalize, const float* inv_norm,
                         void* out, int64_t ostride_s, int64_t ostride_v, void* workspace, cudaStream_t stream) {
    SCB_REQUIRE(n_rows == n_samples * p->n_views, "supcon_backward_features: n_rows must be n_samples * n_views");
    SCB_REQUIRE(feat_dtype == SCB_F32 || feat_dtype == SCB_BF16 || feat_dtype == SCB_F16,
                "supcon_backward_features: unsupported feature dtype %d", feat_dtype);
    FinishParams fin;
    memset(&fin, 0, sizeof(fin));
    fin.feat = features; fin.out = out; fin.feat_dtype = feat_dtype; fin.n_samples = n_samples;
    fin.normalize = normalize; fin.stride_s = stride_s; fin.stride_v = stride_v;
    fin.ostride_s = ostride_s; fin.ostride_v = ostride_v; fin.inv_norm = inv_norm;
    TcGeom g;
    TcParams t;
    if (int e = tc_backward_main(p, row_begin, n_rows, lse_all, grad_out, n_anchor_total, workspace, &fin, &g, &t, stream))
        return e;
    if (g.direct) return SCB_OK;
    const unsigned blocks = (unsigned)ceil_div((int64_t)n_rows * 32, 256);
    prefer_max_smem_carveout(supcon_tc_reduce_finish_kernel);
    SCB_CUDA_OK(launch_pdl(supcon_tc_reduce_finish_kernel, dim3(blocks), dim3(256), 0, stream, t));
    count_launch();
    return SCB_OK;
}

}  // namespace scb
