// Decode-phase kernels of the generation backend (sm_100a).
//
// None of these exist in the reference: /root/reference/src/adapters/local-llm.ts:159 hands the
// prompt to an external server.  They are the "beyond the boundary" list of SURVEY.md 8(a-K):
// embed gather, RMSNorm, decode projections (GEMV), RoPE + KV append, paged GQA attention with a
// shared-prefix (cascade) phase, SwiGLU, lm_head + sampling.  All are HBM-bound streams: weights and
// KV pages are read once with 16-byte coalesced loads, reductions are warp shuffles / fixed-order
// shared-memory sums (deterministic: no atomics anywhere on the data path).
#pragma once
#include <type_traits>

#include "common.cuh"
#include "gemv.cuh"

// a projection input leaves its producer as fp16 (prefill: row-major [rows][K] for the tcgen05 GEMMs) or, for the decode
// projection kernel, as the bf16 hi + lo of that fp16 value in the XF layout (common.cuh): xf = 1; xf = 2: the XF layout
// holding the fp16 value itself (XF16 form, for projections over the fp16 weight twins)
__device__ __forceinline__ void store_act4(void* base, int xf, int K, int row, int k, const float (&y)[4]) {
    if (xf) {
        xf_store4(reinterpret_cast<bf16*>(base), K, row, k, y, xf == 2);
    } else {
        uint2 o;
        o.x = pack_f16(y[0], y[1]); o.y = pack_f16(y[2], y[3]);
        *reinterpret_cast<uint2*>(reinterpret_cast<f16*>(base) + (size_t)row * K + k) = o;
    }
}
// Rows of a prefill activation buffer -> slots of the decode buffer, with their RMSNorm scale: how the last prompt tokens
// reach the decode lm_head kernel.  One CTA per row.
struct GatherRowsArgs {
    const f16* src;        // [S][H]
    const float* src_r;    // [S]
    bf16* dst;             // XF layout over H columns
    float* dst_r;          // [32]
    int H;
    int raw;               // 1: XF16 form (common.cuh xf_store4)
    int rows[32];
};
static __global__ void __launch_bounds__(256) gather_rows_kernel(const GatherRowsArgs a) {
    const int t = blockIdx.x, r = a.rows[t];
    for (int c = threadIdx.x * 4; c < a.H; c += blockDim.x * 4) {
        const uint2 v = *reinterpret_cast<const uint2*>(a.src + (size_t)r * a.H + c);
        const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), f1 = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
        const float y[4] = {f0.x, f0.y, f1.x, f1.y};
        xf_store4(a.dst, a.H, t, c, y, a.raw);
    }
    if (threadIdx.x == 0) a.dst_r[t] = a.src_r[r];
}

// =================================================================================================
// Embedding gather from the CM-layout table (gemv.cuh):  x[t][:] = float(E[tok[t]][:]), and the first layer's
// RMSNorm in its split form (gemv.cuh): xw = fp16(x * w) (row-major fp16, or XF layout with xf != 0), per-16-element partial sums of x^2 (decode consumers) and
// r = rsqrt(mean x^2 + eps) (prefill / tensor-parallel consumers).  One CTA per slot, one thread per 16 elements.
// =================================================================================================
static __global__ void __launch_bounds__(256) embed_norm_kernel(const bf16* __restrict__ E, const int* __restrict__ tok,
                                                                float* __restrict__ x, const bf16* __restrict__ norm_w,
                                                                void* __restrict__ xw, int xf, float* __restrict__ ss_part,
                                                                float* __restrict__ inv_rms, int H, int vocab, float eps) {
    griddep_launch();
    griddep_wait();
    __shared__ float red[8];
    const int t = blockIdx.x;
    int id = tok[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const int nkb = H >> 5;
    const int tile = id >> 4, h = (id >> 3) & 1, r = id & 7;
    float tot = 0.f;
    for (int part = threadIdx.x; part < H / 16; part += blockDim.x) {
        float ss = 0.f;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int k = part * 16 + c * 8;
            const int kb = k >> 5, q = (k & 31) >> 3;
            const uint4 v = *reinterpret_cast<const uint4*>(E + ((((size_t)tile * nkb + kb) * 2 + h) * 4 + q) * 64 + r * 8);
            const uint4 wv = *reinterpret_cast<const uint4*>(norm_w + k);
            const float4 a = make_float4(bf16lo(v.x), bf16hi(v.x), bf16lo(v.y), bf16hi(v.y));
            const float4 b = make_float4(bf16lo(v.z), bf16hi(v.z), bf16lo(v.w), bf16hi(v.w));
            float4* o = reinterpret_cast<float4*>(x + (size_t)t * H + k);
            o[0] = a;
            o[1] = b;
            const float y0[4] = {a.x * bf16lo(wv.x), a.y * bf16hi(wv.x), a.z * bf16lo(wv.y), a.w * bf16hi(wv.y)};
            const float y1[4] = {b.x * bf16lo(wv.z), b.y * bf16hi(wv.z), b.z * bf16lo(wv.w), b.w * bf16hi(wv.w)};
            store_act4(xw, xf, H, t, k, y0);
            store_act4(xw, xf, H, t, k + 4, y1);
            ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
        }
        if (ss_part) ss_part[part * 32 + (t & 31)] = ss;
        tot += ss;
    }
    if (inv_rms) {
        tot = warp_sum(tot);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = tot;
        __syncthreads();
        if (threadIdx.x == 0) {
            float a2 = 0.f;
            for (int i = 0; i < (int)(blockDim.x >> 5); ++i) a2 += red[i];
            inv_rms[t] = 1.0f / sqrtf(a2 / (float)H + eps);
        }
    }
}

// fp16 twin of a bf16 weight matrix (same element order: the CM layout is kept), for the tcgen05 prefill GEMMs whose two
// operands must share one 16-bit format (gemm_tc5.cuh).  bf16 has 8 significant bits, fp16 11: exact for |w| >= 2^-14.
static __global__ void bf16_to_f16_kernel(const bf16* __restrict__ in, f16* __restrict__ out, size_t n8) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = reinterpret_cast<const uint4*>(in)[i];
        uint4 o;
        o.x = pack_f16(bf16lo(v.x), bf16hi(v.x)); o.y = pack_f16(bf16lo(v.y), bf16hi(v.y));
        o.z = pack_f16(bf16lo(v.z), bf16hi(v.z)); o.w = pack_f16(bf16lo(v.w), bf16hi(v.w));
        reinterpret_cast<uint4*>(out)[i] = o;
    }
}

// Split RMSNorm for the paths that have a whole row in one place (prefill, tensor parallel): x (+= delta) stays fp32,
// xw = fp16(x * w), inv_rms[t] = rsqrt(mean x^2 + eps).  One CTA per token.
template <typename TD>
static __global__ void __launch_bounds__(256) rms_xw_kernel(float* __restrict__ x, const TD* __restrict__ delta,
                                                            const bf16* __restrict__ w, void* __restrict__ xw, int xf,
                                                            float* __restrict__ inv_rms, int H, float eps) {
    const int t = blockIdx.x;
    float* xr = x + (size_t)t * H;
    const TD* dr = delta ? delta + (size_t)t * H : nullptr;
    float ss = 0.f;
    for (int c = threadIdx.x * 4; c < H; c += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(xr + c);
        if (dr) {
            float4 d;
            if (sizeof(TD) == 4) {
                d = *reinterpret_cast<const float4*>(dr + c);
            } else {                                    // fp16 partial sums (tensor-parallel prefill)
                const uint2 dh = *reinterpret_cast<const uint2*>(dr + c);
                const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&dh.x)), d1 = __half22float2(*reinterpret_cast<const __half2*>(&dh.y));
                d = make_float4(d0.x, d0.y, d1.x, d1.y);
            }
            v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            *reinterpret_cast<float4*>(xr + c) = v;
        }
        const uint2 wv = *reinterpret_cast<const uint2*>(w + c);
        const float y[4] = {v.x * bf16lo(wv.x), v.y * bf16hi(wv.x), v.z * bf16lo(wv.y), v.w * bf16hi(wv.y)};
        store_act4(xw, xf, H, t, c, y);
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    __shared__ float red[8];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) tot += red[i];
        inv_rms[t] = 1.0f / sqrtf(tot / (float)H + eps);
    }
}

// =================================================================================================
// RMSNorm:  y[t] = bf16( x[t] * rsqrt(mean(x[t]^2) + eps) * w )      (optionally x[t] += delta[t] first)
// one CTA per token slot; the fp32 residual stream stays in x.
// =================================================================================================
static __global__ void __launch_bounds__(256) rmsnorm_kernel(float* __restrict__ x, const float* __restrict__ delta,
                                                      const bf16* __restrict__ w, bf16* __restrict__ y, int H,
                                                      int ldy, float eps) {
    const int t = blockIdx.x;
    float* xr = x + (size_t)t * H;
    const float* dr = delta ? delta + (size_t)t * H : nullptr;
    float ss = 0.f;
    for (int c = threadIdx.x * 4; c < H; c += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(xr + c);
        if (dr) {
            const float4 d = *reinterpret_cast<const float4*>(dr + c);
            v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            *reinterpret_cast<float4*>(xr + c) = v;
        }
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    __shared__ float red[8];
    __shared__ float s_r;
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) tot += red[i];
        s_r = 1.0f / sqrtf(tot / (float)H + eps);
    }
    __syncthreads();
    const float r = s_r;
    for (int c = threadIdx.x * 4; c < H; c += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4*>(xr + c);
        const uint2 wv = *reinterpret_cast<const uint2*>(w + c);
        uint2 o;
        o.x = pack_bf16(v.x * r * bf16lo(wv.x), v.y * r * bf16hi(wv.x));
        o.y = pack_bf16(v.z * r * bf16lo(wv.y), v.w * r * bf16hi(wv.y));
        *reinterpret_cast<uint2*>(y + (size_t)t * ldy + c) = o;
    }
}

// final arg-max over the per-CTA candidates; advances the decode state on the device so that a
// decode step can be replayed as a CUDA graph without host involvement.
struct SampleArgs {
    const float* cta_max;
    const int* cta_idx;
    int n_cta, T;
    int* next_tok;      // [T] written always
    int* tok;           // nullable: tok[t] = next (decode: feeds the next step)
    int* pos;           // nullable: pos[t] += 1
    int* out_tokens;    // nullable: out_tokens[t*max_out + *step] = next
    int max_out;
    int* step_ptr;      // nullable: incremented once per launch
};

static __global__ void __launch_bounds__(256) sample_final_kernel(const SampleArgs a) {
    griddep_launch();
    griddep_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int step = a.step_ptr ? *a.step_ptr : 0;
    for (int t = warp; t < a.T; t += 8) {
        float best = -INFINITY;
        int bi = 0x7fffffff;
        for (int c = lane; c < a.n_cta; c += 32) {
            const float v = a.cta_max[c * 32 + t];
            const int i = a.cta_idx[c * 32 + t];
            if (v > best || (v == best && i < bi)) { best = v; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float v = __shfl_xor_sync(0xffffffffu, best, o);
            const int i = __shfl_xor_sync(0xffffffffu, bi, o);
            if (v > best || (v == best && i < bi)) { best = v; bi = i; }
        }
        if (lane == 0) {
            if (bi == 0x7fffffff) bi = 0;
            a.next_tok[t] = bi;
            if (a.tok) a.tok[t] = bi;
            if (a.pos) a.pos[t] += 1;
            if (a.out_tokens && step < a.max_out) a.out_tokens[(size_t)t * a.max_out + step] = bi;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && a.step_ptr) *a.step_ptr = step + 1;
}

// =================================================================================================
// Paged GQA decode attention, split over the KV length, with a shared-prefix (cascade) phase.
//
// Work item = one WARP: 16 query rows x one chunk of `chunk_pages` KV pages of one kv head.
//   * per-slot items (blockIdx.z < T): rows = the G q-heads of slot z; keys [shared, pos+1)
//   * shared items   (blockIdx.z >= T): rows = 16 consecutive (slot, q-head) pairs of ALL slots;
//     keys [0, shared) -- read from HBM once for every knight instead of once per knight.
// Each warp streams its pages through a private 4-stage cp.async ring in shared memory (XOR-swizzled
// 16-byte pieces, conflict-free ldmatrix), computes S = QK^T and O += PV with mma.sync m16n8k16 and an
// online softmax in the exp2 domain, and writes an un-normalised partial (O, m, l).  attn_merge_kernel
// combines the partials in fixed order (attn_merge2_kernel in the engine: one CTA per (slot, head), two L2 round trips).
// =================================================================================================
struct AttnArgs {
    const bf16* Q;
    int ldq;
    const bf16* Kp;
    const bf16* Vp;
    const int* page_table;
    int pt_stride;
    const int* slot_seq;
    const int* slot_pos;
    const int* shared_len_ptr;  // nullable => 0
    int T, n_q, n_kv;
    int chunk_pages;            // KV pages per work item
    int n_shared_tiles;         // 0 => cascade off
    float* part_o;              // [T][n_q][max_chunks][hd]
    float* part_ml;             // [T][n_q][max_chunks][2]
    int max_chunks;
    L2Prefetch pf[2];           // later kernels' weights to pull into L2 while this kernel runs (common.cuh)
    int early;                  // 1 = index setup and old KV pages before the PDL wait (needs a producer that releases
                                // this kernel only after its own wait: the engine's QKV projection does)
    float scale_log2;           // log2(e)/sqrt(hd)
    int max_len;                // > 0: positions are clamped to max_len - 1 (callers whose positions the host cannot check:
                                // a bad position then gives a truncated result, never a partial outside the scratch)
};

#ifndef RT_ATTN_STAGES
#define RT_ATTN_STAGES 3
#endif
constexpr int ATTN_STAGES = RT_ATTN_STAGES;   // cp.async ring depth per warp (measured best on B200: 3 stages, 128-key chunks at head_dim 128)
constexpr int ATTN_WARPS = 4;

// XOR swizzle of the 16-byte piece index inside a shared-memory row of HD bf16, chosen so that the 8 rows an
// ldmatrix touches land in 8 different bank groups: rows are 64 B (HD=32), 128 B (HD=64) or 256 B (HD=128).
template <int HD>
__device__ __forceinline__ int attn_swz(int row) { return HD >= 64 ? (row & 7) : ((row >> 1) & 3); }

template <int HD>
__device__ __forceinline__ void attn_issue_page(uint32_t sbase, const bf16* __restrict__ Kp, const bf16* __restrict__ Vp,
                                                size_t page_elem_off, int lane) {
    constexpr int CPR = HD / 8;               // 16-byte pieces per row
    constexpr int PIECES = KV_PAGE * CPR;     // per K (or V) page
#pragma unroll
    for (int j = 0; j < PIECES / 32; ++j) {
        const int ci = lane + 32 * j;
        const int row = ci / CPR, c = ci % CPR;
        const uint32_t dst = (uint32_t)(row * (HD * 2) + ((c ^ attn_swz<HD>(row)) << 4));
        cp_async16(sbase + dst, Kp + page_elem_off + row * HD + c * 8);
        cp_async16(sbase + KV_PAGE * HD * 2 + dst, Vp + page_elem_off + row * HD + c * 8);
    }
}

template <int HD>
__global__ void __launch_bounds__(ATTN_WARPS * 32, HD == 128 ? 2 : 4) attn_decode_kernel(const AttnArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int STAGE_BYTES = 2 * KV_PAGE * HD * 2;
    constexpr int KT = HD / 16;  // k-steps for QK^T; pairs of 8-wide d tiles for PV

    RT_TR(50);
    griddep_launch();
    if (threadIdx.x == 0) {     // weights do not depend on the previous kernel: before the PDL wait
        const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z), n_cta = gridDim.x * gridDim.y * gridDim.z;
        l2_prefetch_slice(a.pf[0], cta, n_cta);
        l2_prefetch_slice(a.pf[1], cta, n_cta);
    }
    // ---- everything that does not depend on the QKV projection of THIS layer runs before the PDL wait: the index
    // setup and the cp.async of KV pages that hold only OLDER tokens (their HBM latency hides behind that projection).
    // Safe because the QKV kernel releases this one only after its own wait (gemv.cuh), i.e. with every earlier kernel
    // -- the previous step's sampler, which advanced slot_pos, included -- complete.  `a.early` = 0 turns it off.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const int h = blockIdx.y;
    const int G = a.n_q / a.n_kv;
    const int c = blockIdx.x * ATTN_WARPS + warp;
    const int C = a.chunk_pages * KV_PAGE;
    if (!a.early) griddep_wait();
    const int sh = (a.n_shared_tiles > 0 && a.shared_len_ptr) ? *a.shared_len_ptr : 0;
    const int nsh_chunks = (sh + C - 1) / C;
    // keys below the lowest position any slot holds in this step were written by earlier steps
    int min_pos = lane < a.T ? a.slot_pos[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) min_pos = min(min_pos, __shfl_xor_sync(0xffffffffu, min_pos, o));

    int key0, key1, pidx, seq;
    int row_slot[2], row_head[2];
    bool row_ok[2];
    if ((int)blockIdx.z < a.T) {
        const int slot = blockIdx.z;
        seq = a.slot_seq[slot];
        int len = a.slot_pos[slot] + 1;
        if (a.max_len > 0) len = min(len, a.max_len);
        key0 = sh + c * C;
        key1 = min(key0 + C, len);
        pidx = nsh_chunks + c;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int R = g + 8 * i;
            row_ok[i] = R < G;
            row_slot[i] = slot;
            row_head[i] = h * G + (R < G ? R : 0);
        }
    } else {
        const int tile = blockIdx.z - a.T;
        seq = a.slot_seq[0];
        key0 = c * C;
        key1 = min(key0 + C, sh);
        pidx = c;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int R = tile * 16 + g + 8 * i;
            row_ok[i] = R < a.T * G;
            row_slot[i] = row_ok[i] ? R / G : 0;
            row_head[i] = h * G + (row_ok[i] ? R % G : 0);
        }
    }
    if (key0 >= key1 || pidx >= a.max_chunks) return;  // warp-uniform

    const uint32_t sbase = smem_u32(smem_raw) + warp * (ATTN_STAGES * STAGE_BYTES);
    const int* pt = a.page_table + (size_t)seq * a.pt_stride;
    const int page0 = key0 / KV_PAGE;
    const int npages = (key1 - key0 + KV_PAGE - 1) / KV_PAGE;
    const size_t page_stride = (size_t)a.n_kv * KV_PAGE * HD;
    const size_t head_off = (size_t)h * KV_PAGE * HD;

    // ring prologue, old pages first (uncommitted: they join the first group committed below, which is waited for no
    // later than they are needed)
    unsigned early_mask = 0u;
    if (a.early) {
#pragma unroll
        for (int s = 0; s < ATTN_STAGES - 1; ++s) {
            if (s < npages && (page0 + s + 1) * KV_PAGE <= min_pos) {
                attn_issue_page<HD>(sbase + s * STAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + s] * page_stride + head_off, lane);
                early_mask |= 1u << s;
            }
        }
        griddep_wait();
    }
    RT_TR(51);

    // Q fragments (A operand), loaded once
    uint32_t qa[KT][4];
    {
        const bf16* q0 = a.Q + (size_t)row_slot[0] * a.ldq + (size_t)row_head[0] * HD;
        const bf16* q1 = a.Q + (size_t)row_slot[1] * a.ldq + (size_t)row_head[1] * HD;
#pragma unroll
        for (int t = 0; t < KT; ++t) {
            qa[t][0] = row_ok[0] ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 2 * q) : 0u;
            qa[t][1] = row_ok[1] ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 2 * q) : 0u;
            qa[t][2] = row_ok[0] ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 8 + 2 * q) : 0u;
            qa[t][3] = row_ok[1] ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 8 + 2 * q) : 0u;
        }
    }

#pragma unroll
    for (int s = 0; s < ATTN_STAGES - 1; ++s) {
        if (s < npages && !(early_mask & (1u << s)))
            attn_issue_page<HD>(sbase + s * STAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + s] * page_stride + head_off, lane);
        cp_async_commit();
    }

    float o[HD / 8][4];
#pragma unroll
    for (int d = 0; d < HD / 8; ++d)
#pragma unroll
        for (int i = 0; i < 4; ++i) o[d][i] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    for (int p = 0; p < npages; ++p) {
        {
            const int pn = p + ATTN_STAGES - 1;
            if (pn < npages)
                attn_issue_page<HD>(sbase + (pn % ATTN_STAGES) * STAGE_BYTES, a.Kp, a.Vp,
                                    (size_t)pt[page0 + pn] * page_stride + head_off, lane);
            cp_async_commit();
        }
        cp_async_wait<ATTN_STAGES - 1>();
        __syncwarp();
        const uint32_t sK = sbase + (p % ATTN_STAGES) * STAGE_BYTES;
        const uint32_t sV = sK + KV_PAGE * HD * 2;

        float s[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) s[j][i] = 0.f;
        {
            const int m = lane >> 3, i8 = lane & 7;
            const int key = (m >> 1) * 8 + i8;
#pragma unroll
            for (int t = 0; t < KT; ++t) {
                const int chunk = 2 * t + (m & 1);
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4(r0, r1, r2, r3, sK + key * (HD * 2) + ((chunk ^ attn_swz<HD>(key)) << 4));
                mma_bf16_16816(s[0], qa[t][0], qa[t][1], qa[t][2], qa[t][3], r0, r1);
                mma_bf16_16816(s[1], qa[t][0], qa[t][1], qa[t][2], qa[t][3], r2, r3);
            }
        }
        // scale, mask, online softmax (rows g -> idx 0,1 ; rows g+8 -> idx 2,3)
        const int kbase = key0 + p * KV_PAGE;
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int key = kbase + j * 8 + 2 * q + (i & 1);
                const float v = key < key1 ? s[j][i] * a.scale_log2 : -INFINITY;
                s[j][i] = v;
                mx[i >> 1] = fmaxf(mx[i >> 1], v);
            }
        float alpha[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            const float mn = fmaxf(m_run[r], mx[r]);
            alpha[r] = exp2f(m_run[r] - mn);   // m_run = -inf -> 0 ; every page holds >= 1 valid key so mn is finite
            m_run[r] = mn;
            l_run[r] *= alpha[r];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float pv = exp2f(s[j][i] - m_run[i >> 1]);
                s[j][i] = pv;
                l_run[i >> 1] += pv;
            }
#pragma unroll
        for (int d = 0; d < HD / 8; ++d) {
            o[d][0] *= alpha[0]; o[d][1] *= alpha[0];
            o[d][2] *= alpha[1]; o[d][3] *= alpha[1];
        }
        // P as bf16 hi + bf16 lo (two MMAs): the decode kernel is HBM-bound, so the second MMA is free and
        // the softmax weights keep ~16 bits instead of 8
        const uint32_t pa0 = pack_bf16(s[0][0], s[0][1]), pa1 = pack_bf16(s[0][2], s[0][3]);
        const uint32_t pa2 = pack_bf16(s[1][0], s[1][1]), pa3 = pack_bf16(s[1][2], s[1][3]);
        const uint32_t pl0 = pack_bf16(s[0][0] - bf16lo(pa0), s[0][1] - bf16hi(pa0));
        const uint32_t pl1 = pack_bf16(s[0][2] - bf16lo(pa1), s[0][3] - bf16hi(pa1));
        const uint32_t pl2 = pack_bf16(s[1][0] - bf16lo(pa2), s[1][1] - bf16hi(pa2));
        const uint32_t pl3 = pack_bf16(s[1][2] - bf16lo(pa3), s[1][3] - bf16hi(pa3));
        {
            const int m = lane >> 3, i8 = lane & 7;
            const int key = (m & 1) * 8 + i8;
#pragma unroll
            for (int dp = 0; dp < KT; ++dp) {
                const int chunk = 2 * dp + (m >> 1);
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4_trans(r0, r1, r2, r3, sV + key * (HD * 2) + ((chunk ^ attn_swz<HD>(key)) << 4));
                mma_bf16_16816(o[2 * dp], pa0, pa1, pa2, pa3, r0, r1);
                mma_bf16_16816(o[2 * dp + 1], pa0, pa1, pa2, pa3, r2, r3);
#ifndef RT_ATTN_NO_LO
                mma_bf16_16816(o[2 * dp], pl0, pl1, pl2, pl3, r0, r1);
                mma_bf16_16816(o[2 * dp + 1], pl0, pl1, pl2, pl3, r2, r3);
#endif
            }
        }
        __syncwarp();
    }

#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        if (!row_ok[r]) continue;
        const size_t base = ((size_t)row_slot[r] * a.n_q + row_head[r]) * a.max_chunks + pidx;
        float* po = a.part_o + base * HD;
#pragma unroll
        for (int d = 0; d < HD / 8; ++d)
            *reinterpret_cast<float2*>(po + d * 8 + 2 * q) = make_float2(o[d][2 * r], o[d][2 * r + 1]);
        if (q == 0) *reinterpret_cast<float2*>(a.part_ml + base * 2) = make_float2(m_run[r], l_run[r]);
    }
    RT_TR(54);
}

struct MergeArgs {
    const float* part_o;
    const float* part_ml;
    int max_chunks;
    const int* slot_pos;
    const int* shared_len_ptr;
    int cascade;       // n_shared_tiles > 0
    int n_sh_parts;    // >= 0: the shared prefix contributes exactly this many partials (attn_decode_tc5.cuh: its key ranges are
                       // not multiples of chunk_keys); -1: ceil(shared / chunk_keys) as written by attn_decode_kernel
    int max_len;       // > 0: clamp of slot_pos + 1 (see AttnArgs)
    int chunk_keys;    // C
    int n_q, hd;
    bf16* out_bf;      // nullable: the input of the O-projection, XF layout over ldo columns
    float* out_f32;    // nullable  [T][ldo]
    int ldo;
    int raw;           // 1: out_bf in the XF16 form (common.cuh xf_store1)
    int early;         // slot_pos / shared_len read before the PDL wait (see AttnArgs.early)
};

static __global__ void attn_merge_kernel(const MergeArgs a) {
    const int t = blockIdx.y, head = blockIdx.x, d = threadIdx.x;
    int len = a.slot_pos[t] + 1;
    if (a.max_len > 0) len = min(len, a.max_len);
    const int sh = (a.cascade && a.shared_len_ptr) ? *a.shared_len_ptr : 0;
    const int C = a.chunk_keys;
    const int n = (sh + C - 1) / C + (len - sh + C - 1) / C;
    const size_t base = ((size_t)t * a.n_q + head) * a.max_chunks;
    float M = -INFINITY;
    for (int i = 0; i < n; ++i) M = fmaxf(M, a.part_ml[(base + i) * 2]);
    float L = 0.f, acc = 0.f;
    for (int i = 0; i < n; ++i) {
        const float w = exp2f(a.part_ml[(base + i) * 2] - M);
        L += a.part_ml[(base + i) * 2 + 1] * w;
        acc += a.part_o[(base + i) * a.hd + d] * w;
    }
    const float r = acc / L;
    if (a.out_bf) xf_store1(a.out_bf, a.ldo, t, head * a.hd + d, r, a.raw);
    if (a.out_f32) a.out_f32[(size_t)t * a.ldo + head * a.hd + d] = r;
}

// Merge for the engine's decode step: one CTA of NWM warps per (slot, q head).  The stage is a chain of dependent round trips
// on a few KB (index -> (m, l) pairs -> partial outputs), so it is laid out to have as few of them as possible after the
// PDL wait: the number of partials comes from values that do not change within a step and is read BEFORE the wait when the
// caller says so (`early`: inside the decode graph); after it every warp requests its partial outputs (chunks i = warp,
// warp + NWM, ...: with 16 warps ALL <= 136 partials of a row are in flight at once, 9 per warp) and, right behind them, all
// (m, l) pairs; it derives the global max and the weights while the outputs travel, accumulates, and the warp sums are added
// in fixed order.  Measured in the in-kernel timeline of tensor-parallel shards (profiles/r2_summary.md): 4 warps walking
// 130 partials in three dependent batches of 16 took 5.8 us per layer.
template <int HD, int NWM>
static __global__ void __launch_bounds__(NWM * 32) attn_merge2_kernel(const MergeArgs a) {
    RT_TR(60);
    griddep_launch();
    constexpr int VD = HD / 32;
    constexpr int MAXC = 136;               // MAX_CHUNKS (engine.cuh)
    constexpr int MAXB = (MAXC + 31) / 32;
    // partials a warp requests at once: 4 warps x 18 cover the 69 partials of a row of config 3 (64 shared + 5 own chunks) in
    // ONE round of loads; 16 warps x 9 cover all 136
    constexpr int PW = (MAXC + NWM - 1) / NWM > 18 ? 18 : (MAXC + NWM - 1) / NWM;
    __shared__ float sacc[NWM][HD];
    __shared__ float sL;
    const int t = blockIdx.y, head = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int len = 0, sh = 0;
    if (a.early) {
        len = a.slot_pos[t] + 1;
        sh = (a.cascade && a.shared_len_ptr) ? *a.shared_len_ptr : 0;
    }
    griddep_wait();
    RT_TR(61);
    if (!a.early) {
        len = a.slot_pos[t] + 1;
        sh = (a.cascade && a.shared_len_ptr) ? *a.shared_len_ptr : 0;
    }
    if (a.max_len > 0) len = min(len, a.max_len);
    const int C = a.chunk_keys;
    const int n = min((a.n_sh_parts >= 0 ? a.n_sh_parts : (sh + C - 1) / C) + (len - sh + C - 1) / C, a.max_chunks);
    const size_t base = ((size_t)t * a.n_q + head) * a.max_chunks;
    const float* src = a.part_o + base * HD + lane * VD;
    // one vector load per lane and partial (a partial row is HD contiguous floats, 16-byte aligned: HD * 4 bytes per row)
    using VecT = typename std::conditional<VD == 4, float4, typename std::conditional<VD == 2, float2, float>::type>::type;
    float acc[VD];
#pragma unroll
    for (int v = 0; v < VD; ++v) acc[v] = 0.f;
    float2 ml[MAXB];
    float M = -INFINITY, L = 0.f;
    for (int i0 = warp; i0 < n; i0 += NWM * PW) {
        VecT pv[PW];
#pragma unroll
        for (int j = 0; j < PW; ++j) {
            const int i = i0 + NWM * j;
            pv[j] = VecT{};
            if (i < n) pv[j] = *reinterpret_cast<const VecT*>(src + (size_t)i * HD);
        }
        if (i0 == warp) {                   // first (usually only) round: the (m, l) pairs travel together with the outputs
#pragma unroll
            for (int b = 0; b < MAXB; ++b) {
                const int i = b * 32 + lane;
                ml[b] = i < n ? *reinterpret_cast<const float2*>(a.part_ml + (base + i) * 2) : make_float2(-INFINITY, 0.f);
                M = fmaxf(M, ml[b].x);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
#pragma unroll
            for (int b = 0; b < MAXB; ++b) {
                ml[b].x = exp2f(ml[b].x - M);   // weight of chunk b*32+lane (0 for absent chunks)
                L += ml[b].y * ml[b].x;
            }
            L = warp_sum(L);
        }
#pragma unroll
        for (int j = 0; j < PW; ++j) {
            const int i = i0 + NWM * j;
            float w = 0.f;
#pragma unroll
            for (int b = 0; b < MAXB; ++b) {
                const float wb = __shfl_sync(0xffffffffu, ml[b].x, i & 31);
                if ((i >> 5) == b) w = wb;
            }
            const float* tf = reinterpret_cast<const float*>(&pv[j]);
#pragma unroll
            for (int v = 0; v < VD; ++v) acc[v] += tf[v] * w;
        }
    }
#pragma unroll
    for (int v = 0; v < VD; ++v) sacc[warp][lane * VD + v] = acc[v];
    if (threadIdx.x == 0) sL = L;           // warp 0 always has a partial (n >= 1)
    __syncthreads();
    if (threadIdx.x < HD) {
        float r = 0.f;
#pragma unroll
        for (int w = 0; w < NWM; ++w) r += sacc[w][threadIdx.x];
        const float o = r / sL;
        if (a.out_bf) xf_store1(a.out_bf, a.ldo, t, head * HD + threadIdx.x, o, a.raw);
        if (a.out_f32) a.out_f32[(size_t)t * a.ldo + head * HD + threadIdx.x] = o;
    }
    RT_TR(64);
}

// Tensor parallel: every rank holds the local winners (score, token id) of all ranks after an all-gather; pick the
// global arg-max (lowest id wins ties) and advance the decode state -- identically on every rank.
struct TpPickArgs {
    const float* all_val;   // [tp][64 words]: 32 scores then 32 ids per rank
    int tp, T;
    int* next_tok;
    int* tok;
    int* pos;
    int* out_tokens;
    int max_out;
    int* step_w;
};
static __global__ void tp_pick_kernel(const TpPickArgs a) {
    griddep_launch();
    griddep_wait();
    const int t = threadIdx.x;
    const int step = a.step_w ? *a.step_w : 0;
    if (t < a.T) {
        float best = -INFINITY;
        int bi = 0x7fffffff;
        for (int r = 0; r < a.tp; ++r) {
            const float v = a.all_val[r * 64 + t];
            const int i = reinterpret_cast<const int*>(a.all_val)[r * 64 + 32 + t];
            if (v > best || (v == best && i < bi)) { best = v; bi = i; }
        }
        if (bi == 0x7fffffff) bi = 0;
        a.next_tok[t] = bi;
        if (a.tok) a.tok[t] = bi;
        if (a.pos) a.pos[t] += 1;
        if (a.out_tokens && step < a.max_out) a.out_tokens[(size_t)t * a.max_out + step] = bi;
    }
    __syncthreads();
    if (t == 0 && a.step_w) *a.step_w = step + 1;
}
