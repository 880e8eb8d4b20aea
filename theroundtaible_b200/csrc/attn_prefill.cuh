// Causal prefill attention over the paged KV cache (flash-attention style, tensor cores via mma.sync m16n8k16).
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external
// server); SURVEY.md 8(a-K) row "prefill causal attention".
//
// The prompt tokens of a chunk are a flat array; a SEGMENT is a run of consecutive tokens of one sequence with
// consecutive positions.  A work item (one CTA of 4 warps) = (segment piece of <= 64/G tokens, kv head): its 64 rows
// are (token, q-head-of-the-group) pairs, 16 per warp.  The CTA walks the keys 0 .. last position of its tokens in
// tiles of 4 pages (64 keys); every warp fetches one page of the tile with cp.async into a swizzled double buffer
// shared by the CTA, then each warp runs S = QK^T, causal mask, online softmax (exp2 domain), O += PV per page with
// the same fragment code as the decode kernel.  The chunk's own K/V are already in the cache (the QKV GEMM epilogue
// wrote them), so there is one code path for old and new keys.
#pragma once
#include "kernels.cuh"
#include "prefill.cuh"

// softmax weights P as bf16 hi + lo (two MMAs, as the decode kernel does) or as one bf16 (-DPFA_P_HILO=0: what most
// flash-attention kernels feed the tensor core).  Measured on config 3: 261 vs 251 ms of prefill; with one bf16 the
// typical logit error of the tiny hd-128 model rises from < 3e-3 to 3.4e-3 of max|logit| (bar: 1e-2), so hi + lo stays.
#ifndef PFA_P_HILO
#define PFA_P_HILO 1
#endif
#ifndef PFA_MIN_CTAS
#define PFA_MIN_CTAS 3
#endif
constexpr int PFA_STAGES = 2;
constexpr int PFA_TILE_PAGES = 4;

template <int HD>
__global__ void __launch_bounds__(128, PFA_MIN_CTAS) attn_prefill_kernel(const PfAttnArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int PAGE_BYTES = 2 * KV_PAGE * HD * 2;              // K + V of one page of one head
    constexpr int STAGE_BYTES = PFA_TILE_PAGES * PAGE_BYTES;
    constexpr int KT = HD / 16;
    griddep_launch();
    griddep_wait();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const int4 wk = a.work[blockIdx.x];
    const int tok0 = wk.x, ntok = wk.y, seq = wk.z, h = wk.w;
    const int G = a.n_q / a.n_kv;
    const int pos0 = a.tok_pos[tok0];
    const int ldq = a.n_q * HD;

    // rows of this lane: R = warp*16 + g (+8) -> (token R/G, head R%G)
    int row_tok[2], row_head[2], row_limit[2];
    bool row_ok[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int R = warp * 16 + g + 8 * i;
        const int tl = R / G;
        row_ok[i] = tl < ntok;
        row_tok[i] = tok0 + (row_ok[i] ? tl : 0);
        row_head[i] = h * G + (row_ok[i] ? R % G : 0);
        row_limit[i] = row_ok[i] ? pos0 + tl + 1 : 1;         // keys [0, limit) are visible
    }
    // the warp's largest limit (rows of a warp are ordered by token): last valid row of the warp
    int warp_limit;
    {
        const int Rl = min(warp * 16 + 15, ntok * G - 1);
        warp_limit = Rl >= warp * 16 ? pos0 + Rl / G + 1 : 1;
    }

    uint32_t qa[KT][4];
    {
        const bf16* q0 = a.Q + (size_t)row_tok[0] * ldq + (size_t)row_head[0] * HD;
        const bf16* q1 = a.Q + (size_t)row_tok[1] * ldq + (size_t)row_head[1] * HD;
#pragma unroll
        for (int t = 0; t < KT; ++t) {
            qa[t][0] = row_ok[0] ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 2 * q) : 0u;
            qa[t][1] = row_ok[1] ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 2 * q) : 0u;
            qa[t][2] = row_ok[0] ? *reinterpret_cast<const uint32_t*>(q0 + 16 * t + 8 + 2 * q) : 0u;
            qa[t][3] = row_ok[1] ? *reinterpret_cast<const uint32_t*>(q1 + 16 * t + 8 + 2 * q) : 0u;
        }
    }

    const uint32_t sbase = smem_u32(smem_raw);
    const int* pt = a.page_table + (size_t)seq * a.pt_stride;
    const int n_keys = pos0 + ntok;                                  // keys any row of this CTA may see
    const int n_pages = (n_keys + KV_PAGE - 1) / KV_PAGE;
    const int n_tiles = (n_pages + PFA_TILE_PAGES - 1) / PFA_TILE_PAGES;
    const size_t page_stride = (size_t)a.n_kv * KV_PAGE * HD;
    const size_t head_off = (size_t)h * KV_PAGE * HD;

    auto issue_tile = [&](int tile) {
        const int page = tile * PFA_TILE_PAGES + warp;               // warp w fetches page w of the tile
        if (page < n_pages)
            attn_issue_page<HD>(sbase + (tile % PFA_STAGES) * STAGE_BYTES + warp * PAGE_BYTES, a.Kp, a.Vp,
                                (size_t)pt[page] * page_stride + head_off, lane);
        cp_async_commit();
    };

    float o[HD / 8][4];
#pragma unroll
    for (int d = 0; d < HD / 8; ++d)
#pragma unroll
        for (int i = 0; i < 4; ++i) o[d][i] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    issue_tile(0);
    for (int tile = 0; tile < n_tiles; ++tile) {
        if (tile + 1 < n_tiles) issue_tile(tile + 1);
        else cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const uint32_t stage = sbase + (tile % PFA_STAGES) * STAGE_BYTES;
        // One online-softmax step per TILE (64 keys), not per page: the running max / rescale of the 16 x HD output
        // fragment is amortised over 4 pages.  Pages at or beyond the warp's causal limit are skipped (warp-uniform).
        const int kb_tile = tile * PFA_TILE_PAGES * KV_PAGE;
        const int npg = kb_tile >= warp_limit ? 0 : min(PFA_TILE_PAGES, (warp_limit - kb_tile + KV_PAGE - 1) / KV_PAGE);
        if (npg > 0) {
            float s[PFA_TILE_PAGES][2][4];
#pragma unroll
            for (int pp = 0; pp < PFA_TILE_PAGES; ++pp) {
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) s[pp][j][i] = 0.f;
                if (pp < npg) {
                    const uint32_t sK = stage + pp * PAGE_BYTES;
                    const int m = lane >> 3, i8 = lane & 7;
                    const int key = (m >> 1) * 8 + i8;
#pragma unroll
                    for (int t = 0; t < KT; ++t) {
                        const int chunk = 2 * t + (m & 1);
                        uint32_t r0, r1, r2, r3;
                        ldmatrix_x4(r0, r1, r2, r3, sK + key * (HD * 2) + ((chunk ^ attn_swz<HD>(key)) << 4));
                        mma_bf16_16816(s[pp][0], qa[t][0], qa[t][1], qa[t][2], qa[t][3], r0, r1);
                        mma_bf16_16816(s[pp][1], qa[t][0], qa[t][1], qa[t][2], qa[t][3], r2, r3);
                    }
                }
            }
            float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int pp = 0; pp < PFA_TILE_PAGES; ++pp)
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int key = kb_tile + pp * KV_PAGE + j * 8 + 2 * q + (i & 1);
                        const float v = (pp < npg && key < row_limit[i >> 1]) ? s[pp][j][i] * a.scale_log2 : -INFINITY;
                        s[pp][j][i] = v;
                        mx[i >> 1] = fmaxf(mx[i >> 1], v);
                    }
            float alpha[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
                mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
                const float mn = fmaxf(m_run[r], mx[r]);   // finite from tile 0 on: key 0 is visible to every row
                alpha[r] = exp2f(m_run[r] - mn);
                m_run[r] = mn;
                l_run[r] *= alpha[r];
            }
#pragma unroll
            for (int pp = 0; pp < PFA_TILE_PAGES; ++pp)
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float pv = exp2f(s[pp][j][i] - m_run[i >> 1]);
                        s[pp][j][i] = pv;
                        l_run[i >> 1] += pv;
                    }
#pragma unroll
            for (int d = 0; d < HD / 8; ++d) {
                o[d][0] *= alpha[0]; o[d][1] *= alpha[0];
                o[d][2] *= alpha[1]; o[d][3] *= alpha[1];
            }
#pragma unroll
            for (int pp = 0; pp < PFA_TILE_PAGES; ++pp) {
                if (pp >= npg) continue;
                const uint32_t sV = stage + pp * PAGE_BYTES + KV_PAGE * HD * 2;
                const uint32_t pa0 = pack_bf16(s[pp][0][0], s[pp][0][1]), pa1 = pack_bf16(s[pp][0][2], s[pp][0][3]);
                const uint32_t pa2 = pack_bf16(s[pp][1][0], s[pp][1][1]), pa3 = pack_bf16(s[pp][1][2], s[pp][1][3]);
#if PFA_P_HILO
                const uint32_t pl0 = pack_bf16(s[pp][0][0] - bf16lo(pa0), s[pp][0][1] - bf16hi(pa0));
                const uint32_t pl1 = pack_bf16(s[pp][0][2] - bf16lo(pa1), s[pp][0][3] - bf16hi(pa1));
                const uint32_t pl2 = pack_bf16(s[pp][1][0] - bf16lo(pa2), s[pp][1][1] - bf16hi(pa2));
                const uint32_t pl3 = pack_bf16(s[pp][1][2] - bf16lo(pa3), s[pp][1][3] - bf16hi(pa3));
#endif
                const int m = lane >> 3, i8 = lane & 7;
                const int key = (m & 1) * 8 + i8;
#pragma unroll
                for (int dp = 0; dp < KT; ++dp) {
                    const int chunk = 2 * dp + (m >> 1);
                    uint32_t r0, r1, r2, r3;
                    ldmatrix_x4_trans(r0, r1, r2, r3, sV + key * (HD * 2) + ((chunk ^ attn_swz<HD>(key)) << 4));
                    mma_bf16_16816(o[2 * dp], pa0, pa1, pa2, pa3, r0, r1);
                    mma_bf16_16816(o[2 * dp + 1], pa0, pa1, pa2, pa3, r2, r3);
#if PFA_P_HILO
                    mma_bf16_16816(o[2 * dp], pl0, pl1, pl2, pl3, r0, r1);
                    mma_bf16_16816(o[2 * dp + 1], pl0, pl1, pl2, pl3, r2, r3);
#endif
                }
            }
        }
        __syncthreads();   // everyone is done with this stage before it is refilled
    }

#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        if (!row_ok[r]) continue;
        const float inv = 1.0f / l_run[r];
        f16* dst = a.out + (size_t)row_tok[r] * ldq + (size_t)row_head[r] * HD;
#pragma unroll
        for (int d = 0; d < HD / 8; ++d)
            *reinterpret_cast<uint32_t*>(dst + d * 8 + 2 * q) = pack_f16(o[d][2 * r] * inv, o[d][2 * r + 1] * inv);
    }
}

constexpr size_t pfa_smem_bytes(int hd) { return (size_t)PFA_STAGES * PFA_TILE_PAGES * 2 * KV_PAGE * hd * 2; }
