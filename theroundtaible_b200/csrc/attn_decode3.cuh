// Decode attention, version 3: one thread-block CLUSTER per (slot, kv head) computes the whole context and merges its
// split-KV partials through distributed shared memory -- no partials in global memory, no merge kernel.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external server);
// SURVEY.md 8(a-K) row "paged decode attention (GQA)".
//
// Why: on short contexts (a few thousand tokens: BASELINE config 2, SEQ-mode turns) the attention stage is a latency
// chain, not an HBM stream; the measured timeline (RT_TRACE) shows attention + boundary + merge kernel + boundary taking
// ~10 us of a 36 us layer with < 3 MB of KV to read.  Removing the second kernel and its launch boundary is the lever.
// Cost: no cascade (every slot reads the shared prefix itself -- from L2 after the first), so the engine uses this
// version only while the context is short; long shared prefixes keep version 1 + attn_merge2_kernel.
//
//   grid (CS, n_kv, T), cluster (CS, 1, 1), CS in {1, 2, 4, 8}; CTA = A3_WARPS warps.
//   Global warp gw = rank * A3_WARPS + warp owns pages [gw*ppw, (gw+1)*ppw) of the slot's context (ppw = ceil(pages /
//   (CS * A3_WARPS))), streams them through its private cp.async ring (same fragment code as version 1: QK^T and PV on
//   mma.sync, online softmax in the exp2 domain, P as bf16 hi + lo) and ends with an un-normalised (O[16][HD], m, l).
//   Level 1: the warps of a CTA leave their partials in local shared memory; the CTA reduces them in fixed warp order.
//   Level 2: every CTA stores its reduced partial into the LEADER's shared memory (mapa + st.shared::cluster), one
//   barrier.cluster, and the leader reduces the CS partials in rank order and writes the normalised bf16 output row.
//   Fixed orders everywhere => bit-reproducible.
#pragma once
#include "kernels.cuh"

constexpr int A3_WARPS = 8;
// cp.async ring depth per warp: 3 pages in flight at head_dim <= 64, 2 at 128 (the rings of 8 warps must leave room for
// the level-2 area inside 227 KB)
template <int HD>
__host__ __device__ constexpr int a3_stages() { return HD >= 128 ? 2 : 3; }

struct Attn3Args {
    const bf16* Q;
    int ldq;
    const bf16* Kp;
    const bf16* Vp;
    const int* page_table;
    int pt_stride;
    const int* slot_seq;
    const int* slot_pos;
    int T, n_q, n_kv;
    float scale_log2;
    bf16* out;                  // normalised output = the O-projection's input: XF layout over ldo columns (common.cuh)
    int ldo;
    int early;                  // index setup and old KV pages before the PDL wait (see AttnArgs.early)
    int raw;                    // 1: `out` in the XF16 form (common.cuh xf_store4)
};

template <int HD>
__host__ __device__ constexpr int a3_ring_bytes() { return A3_WARPS * a3_stages<HD>() * 2 * KV_PAGE * HD * 2; }
// level-1 area: A3_WARPS x (16 rows x HD + 32) floats, overlays the rings once every warp has left its loop;
// level-2 area (leader only, written remotely): 8 x (16 x HD + 32) floats, separate so that no remote store can land in a ring
template <int HD>
__host__ __device__ constexpr int a3_part_floats() { return 16 * HD + 32; }
template <int HD>
__host__ __device__ constexpr int a3_smem_bytes() { return a3_ring_bytes<HD>() + 8 * a3_part_floats<HD>() * 4; }

__device__ __forceinline__ uint32_t a3_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t a3_cluster_size() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t a3_mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void a3_st_remote_f4(uint32_t addr, float4 v) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void a3_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int HD>
__global__ void __launch_bounds__(A3_WARPS * 32, 1) attn_decode3_kernel(const Attn3Args a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int STAGE_BYTES = 2 * KV_PAGE * HD * 2;
    constexpr int KT = HD / 16;
    constexpr int A3_STAGES = a3_stages<HD>();
    constexpr int PF = a3_part_floats<HD>();
    float* l1 = reinterpret_cast<float*>(smem_raw);                                   // level 1 (overlays the rings)
    float* l2 = reinterpret_cast<float*>(smem_raw + a3_ring_bytes<HD>());             // level 2 (leader's is the target)

    griddep_launch();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const int rank = (int)a3_cluster_rank(), CS = (int)a3_cluster_size();
    const int h = blockIdx.y, slot = blockIdx.z;
    const int G = a.n_q / a.n_kv;
    if (!a.early) griddep_wait();

    // keys below the lowest position any slot holds in this step were written by earlier steps
    int min_pos = lane < a.T ? a.slot_pos[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) min_pos = min(min_pos, __shfl_xor_sync(0xffffffffu, min_pos, o));
    const int seq = a.slot_seq[slot];
    const int len = a.slot_pos[slot] + 1;
    const int n_pages_all = (len + KV_PAGE - 1) / KV_PAGE;
    const int GW = CS * A3_WARPS;
    const int ppw = (n_pages_all + GW - 1) / GW;
    const int gw = rank * A3_WARPS + warp;
    const int page0 = gw * ppw;
    const int npages = max(0, min(ppw, n_pages_all - page0));
    const int key0 = page0 * KV_PAGE, key1 = min(key0 + npages * KV_PAGE, len);

    const bool row_ok0 = g < G, row_ok1 = g + 8 < G;
    const int head0 = h * G + (row_ok0 ? g : 0), head1 = h * G + (row_ok1 ? g + 8 : 0);

    const uint32_t sbase = smem_u32(smem_raw) + warp * (A3_STAGES * STAGE_BYTES);
    const int* pt = a.page_table + (size_t)seq * a.pt_stride;
    const size_t page_stride = (size_t)a.n_kv * KV_PAGE * HD;
    const size_t head_off = (size_t)h * KV_PAGE * HD;

    unsigned early_mask = 0u;
    if (a.early) {
#pragma unroll
        for (int s = 0; s < A3_STAGES - 1; ++s) {
            if (s < npages && (page0 + s + 1) * KV_PAGE <= min_pos) {
                attn_issue_page<HD>(sbase + s * STAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + s] * page_stride + head_off, lane);
                early_mask |= 1u << s;
            }
        }
        griddep_wait();
    }

    uint32_t qa[KT][4];
    {
        const bf16* q0 = a.Q + (size_t)slot * a.ldq + (size_t)head0 * HD;
        const bf16* q1 = a.Q + (size_t)slot * a.ldq + (size_t)head1 * HD;
#pragma unroll
        for (int t = 0; t < KT; ++t) {
            qa[t][0] = row_ok0 ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 2 * q) : 0u;
            qa[t][1] = row_ok1 ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 2 * q) : 0u;
            qa[t][2] = row_ok0 ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 8 + 2 * q) : 0u;
            qa[t][3] = row_ok1 ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 8 + 2 * q) : 0u;
        }
    }
#pragma unroll
    for (int s = 0; s < A3_STAGES - 1; ++s) {
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
            const int pn = p + A3_STAGES - 1;
            if (pn < npages)
                attn_issue_page<HD>(sbase + (pn % A3_STAGES) * STAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + pn] * page_stride + head_off, lane);
            cp_async_commit();
        }
        cp_async_wait<A3_STAGES - 1>();
        __syncwarp();
        const uint32_t sK = sbase + (p % A3_STAGES) * STAGE_BYTES;
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
            alpha[r] = exp2f(m_run[r] - mn);   // every page of a warp holds >= 1 valid key, so mn is finite
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
                mma_bf16_16816(o[2 * dp], pl0, pl1, pl2, pl3, r0, r1);
                mma_bf16_16816(o[2 * dp + 1], pl0, pl1, pl2, pl3, r2, r3);
            }
        }
        __syncwarp();
    }
    cp_async_wait<0>();
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    __syncthreads();                       // every warp has left its ring: level 1 may overlay it

    // ---- level 1: warp partials -> local shared memory: [warp][row 16][HD] then [warp][row][m, l] ----
    {
        float* mine = l1 + (size_t)warp * PF;
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int row = g + 8 * r;
#pragma unroll
            for (int d = 0; d < HD / 8; ++d)
                *reinterpret_cast<float2*>(mine + row * HD + d * 8 + 2 * q) = make_float2(o[d][2 * r], o[d][2 * r + 1]);
            if (q == 0) *reinterpret_cast<float2*>(mine + 16 * HD + row * 2) = make_float2(m_run[r], l_run[r]);
        }
    }
    __syncthreads();
    // reduce the A3_WARPS partials in warp order: thread -> (row, 4 consecutive dims); 16*HD/4 items
    const uint32_t l2_leader = a3_mapa(smem_u32(l2), 0);
    for (int it = threadIdx.x; it < 16 * HD / 4; it += A3_WARPS * 32) {
        const int row = it / (HD / 4), d4 = (it % (HD / 4)) * 4;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < A3_WARPS; ++w) M = fmaxf(M, l1[(size_t)w * PF + 16 * HD + row * 2]);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float L = 0.f;
#pragma unroll
        for (int w = 0; w < A3_WARPS; ++w) {
            const float mw = l1[(size_t)w * PF + 16 * HD + row * 2];
            const float sc = mw == -INFINITY ? 0.f : exp2f(mw - M);
            const float4 v = *reinterpret_cast<const float4*>(l1 + (size_t)w * PF + row * HD + d4);
            acc.x += v.x * sc; acc.y += v.y * sc; acc.z += v.z * sc; acc.w += v.w * sc;
            L += l1[(size_t)w * PF + 16 * HD + row * 2 + 1] * sc;
        }
        // ---- level 2: this CTA's partial -> the leader's shared memory, slot `rank` ----
        const uint32_t dst = l2_leader + (uint32_t)(((size_t)rank * PF + row * HD + d4) * 4);
        a3_st_remote_f4(dst, acc);
        if (d4 == 0) {
            const uint32_t dml = l2_leader + (uint32_t)(((size_t)rank * PF + 16 * HD + row * 2) * 4);
            asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(dml), "f"(M), "f"(L) : "memory");
        }
    }
    a3_cluster_sync();
    if (rank != 0) return;
    for (int it = threadIdx.x; it < G * HD / 4; it += A3_WARPS * 32) {
        const int row = it / (HD / 4), d4 = (it % (HD / 4)) * 4;
        float M = -INFINITY;
        for (int c = 0; c < CS; ++c) M = fmaxf(M, l2[(size_t)c * PF + 16 * HD + row * 2]);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float L = 0.f;
        for (int c = 0; c < CS; ++c) {
            const float mc = l2[(size_t)c * PF + 16 * HD + row * 2];
            const float sc = mc == -INFINITY ? 0.f : exp2f(mc - M);
            const float4 v = *reinterpret_cast<const float4*>(l2 + (size_t)c * PF + row * HD + d4);
            acc.x += v.x * sc; acc.y += v.y * sc; acc.z += v.z * sc; acc.w += v.w * sc;
            L += l2[(size_t)c * PF + 16 * HD + row * 2 + 1] * sc;
        }
        const float inv = 1.0f / L;
        const float y[4] = {acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv};
        xf_store4(a.out, a.ldo, slot, (h * G + row) * HD + d4, y, a.raw);
    }
}
