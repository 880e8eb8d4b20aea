// Paged GQA decode attention, version 2: one CTA per (kv head, KV chunk) whose WARPS are the 16-row query tiles and
// which share ONE cp.async ring of K/V pages.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external
// server); SURVEY.md 8(a-K) rows "paged decode attention (GQA)" and "shared-prefix (cascade) attention".
//
// Version 1 (kernels.cuh) gives every warp a private ring, so the 2..4 row tiles of a shared-prefix item each fetch
// the same pages and only 2-3 pages per warp are in flight.  Here the tiles of an item sit in one CTA:
//   shared item  -- rows = (slot, q-head) pairs of ALL knights, warp w = row tile w, keys [0, shared)
//   own item of one slot -- one active warp, rows = its G q-heads, keys [shared, pos+1)
// A page (K + V of one kv head, 2-8 KB) is fetched ONCE per CTA, cooperatively, into a 4-stage XOR-swizzled ring
// (one named barrier per page), so 4 pages per CTA and ~6 CTAs per SM are in flight.  The per-page math is the same
// mma.sync m16n8k16 code as version 1 (online softmax in the exp2 domain, softmax weights as bf16 hi + lo).
// Output: un-normalised partials (O, m, l) per (slot, head, chunk), merged by attn_merge2_kernel.
#pragma once
#include "kernels.cuh"

constexpr int AD2_STAGES = 4;

template <int HD>
__device__ __forceinline__ void ad2_issue_page(uint32_t sdst, const bf16* __restrict__ Kp, const bf16* __restrict__ Vp,
                                               size_t page_elem_off, int tid, int nthr) {
    constexpr int CPR = HD / 8;
    constexpr int PIECES = KV_PAGE * CPR;
    for (int ci = tid; ci < PIECES; ci += nthr) {
        const int row = ci / CPR, c = ci % CPR;
        const uint32_t dst = (uint32_t)(row * (HD * 2) + ((c ^ attn_swz<HD>(row)) << 4));
        cp_async16(sdst + dst, Kp + page_elem_off + row * HD + c * 8);
        cp_async16(sdst + KV_PAGE * HD * 2 + dst, Vp + page_elem_off + row * HD + c * 8);
    }
}

// one page (16 keys) for one warp's 16 query rows; keys >= key_limit are masked
template <int HD>
__device__ __forceinline__ void ad2_page(uint32_t sK, uint32_t sV, const uint32_t (&qa)[HD / 16][4], int kbase, int key_limit,
                                         float scale_log2, float (&o)[HD / 8][4], float (&m_run)[2], float (&l_run)[2], int lane) {
    constexpr int KT = HD / 16;
    const int q = lane & 3;
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
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int key = kbase + j * 8 + 2 * q + (i & 1);
            const float v = key < key_limit ? s[j][i] * scale_log2 : -INFINITY;
            s[j][i] = v;
            mx[i >> 1] = fmaxf(mx[i >> 1], v);
        }
    float alpha[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
        const float mn = fmaxf(m_run[r], mx[r]);   // every page of an item holds >= 1 visible key, so mn is finite
        alpha[r] = exp2f(m_run[r] - mn);
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
}

template <int HD, int NWARP>
__global__ void __launch_bounds__(NWARP * 32) attn_decode2_kernel(const AttnArgs a) {
    extern __shared__ __align__(128) unsigned char ad2_smem[];
    constexpr int PAGE_BYTES = 2 * KV_PAGE * HD * 2;
    constexpr int KT = HD / 16;
    griddep_launch();
    griddep_wait();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    // 1-D grid: blockIdx.x = h * per_head + item; item < grid_nsh -> shared chunk, else (slot, own chunk)
    const int per_head = a.grid_nsh + a.T * a.grid_nown;
    const int h = blockIdx.x / per_head, item = blockIdx.x % per_head;
    const bool is_shared = item < a.grid_nsh;
    const int own_slot = is_shared ? 0 : (item - a.grid_nsh) / a.grid_nown;
    const int c = is_shared ? item : (item - a.grid_nsh) % a.grid_nown;
    const int G = a.n_q / a.n_kv;
    const int C = a.chunk_pages * KV_PAGE;
    const bool cascade = a.n_shared_tiles > 0;
    const int sh = (cascade && a.shared_len_ptr) ? *a.shared_len_ptr : 0;
    const int nsh_chunks = (sh + C - 1) / C;

    // ---- which item is this CTA, and which rows does this warp own ----
    int key0, key1, pidx, seq, nw;
    int row_slot[2], row_head[2];
    bool row_ok[2];
    if (is_shared) {
        if (!cascade) return;
        nw = min(a.n_shared_tiles, NWARP);
        seq = a.slot_seq[0];
        key0 = c * C;
        key1 = min(key0 + C, sh);
        pidx = c;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int R = warp * 16 + g + 8 * i;
            row_ok[i] = R < a.T * G;
            row_slot[i] = row_ok[i] ? R / G : 0;
            row_head[i] = h * G + (row_ok[i] ? R % G : 0);
        }
    } else {
        const int slot = own_slot;
        nw = 1;
        seq = a.slot_seq[slot];
        const int len = a.slot_pos[slot] + 1;
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
    }
    if (key0 >= key1 || warp >= nw) return;       // CTA-uniform / warp-uniform; nobody left behind at a barrier below
    const int nthr = nw * 32;
    const int tid = threadIdx.x;                  // < nthr for the active warps

    const uint32_t sbase = smem_u32(ad2_smem);
    const int* pt = a.page_table + (size_t)seq * a.pt_stride;
    const int page0 = key0 / KV_PAGE;
    const int npages = (key1 - key0 + KV_PAGE - 1) / KV_PAGE;
    const size_t page_stride = (size_t)a.n_kv * KV_PAGE * HD;
    const size_t head_off = (size_t)h * KV_PAGE * HD;

    // pages first (the long-latency part), then the query fragments
#pragma unroll
    for (int s = 0; s < AD2_STAGES - 1; ++s) {
        if (s < npages) ad2_issue_page<HD>(sbase + s * PAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + s] * page_stride + head_off, tid, nthr);
        cp_async_commit();
    }
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
    float o[HD / 8][4];
#pragma unroll
    for (int d = 0; d < HD / 8; ++d)
#pragma unroll
        for (int i = 0; i < 4; ++i) o[d][i] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    for (int p = 0; p < npages; ++p) {
        cp_async_wait<AD2_STAGES - 2>();          // this thread's pieces of page p have landed
        asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory");   // everybody's pieces landed; everybody finished page p-1
        {
            const int pn = p + AD2_STAGES - 1;     // refill the stage page p-1 occupied
            if (pn < npages)
                ad2_issue_page<HD>(sbase + (pn % AD2_STAGES) * PAGE_BYTES, a.Kp, a.Vp, (size_t)pt[page0 + pn] * page_stride + head_off, tid, nthr);
            cp_async_commit();
        }
        const uint32_t sK = sbase + (p % AD2_STAGES) * PAGE_BYTES;
        ad2_page<HD>(sK, sK + KV_PAGE * HD * 2, qa, key0 + p * KV_PAGE, key1, a.scale_log2, o, m_run, l_run, lane);
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
        for (int d = 0; d < HD / 8; ++d) *reinterpret_cast<float2*>(po + d * 8 + 2 * q) = make_float2(o[d][2 * r], o[d][2 * r + 1]);
        if (q == 0) *reinterpret_cast<float2*>(a.part_ml + base * 2) = make_float2(m_run[r], l_run[r]);
    }
}

constexpr size_t ad2_smem_bytes(int hd) { return (size_t)AD2_STAGES * 2 * KV_PAGE * hd * 2; }
