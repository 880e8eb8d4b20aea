// Causal prefill attention on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), head_dim 128 or 64, sm_100a only.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external server);
// SURVEY.md 8(a-K) row "prefill causal attention".  Default for head_dim 128 (RT_PFA_TC5=0: the mma.sync kernel of
// attn_prefill.cuh, which serves head_dim 32 / 64; RT_PFA_TC5=2 also takes head_dim 64 here -- correct, no faster).
//
// One CTA = up to 128 query rows of one kv head: BT = 128 / G consecutive tokens of a segment x the G q-heads of the group
// (row = token * G + head-in-group), walking the keys 0 .. position of its last token in tiles of 128 keys (8 KV pages).
//   warp 0   TMA: Q once (3-D box: 64 dims x G heads x BT tokens, two halves of the head dimension, SWIZZLE_128B), then
//            per key tile the K and V pages of the tile (2-D boxes of 64 dims x 16 keys straight out of the paged pool)
//            into a 2-stage ring.
//   warp 1   MMA issuer: S = Q K^T (M=128, N=128, K=128: 8 tcgen05.mma, both operands K-major) into TMEM columns 0..127;
//            after the softmax warps have written P: O_tile = P V (M=128, N=128, K=128 keys: A = P K-major from shared
//            memory, B = V MN-major -- V lies [key][dim], contiguous along N) into TMEM columns 128..255.
//   warps 2-9  softmax: TMEM lane = query row, the two warps of a lane quadrant take half a row each (64 keys, 64 dims):
//            read S once, causal mask, online softmax with a LAZY maximum (O in TMEM is rescaled only when the maximum
//            grew by more than 2^8), P as bf16 hi + lo into shared memory in the 128-byte-swizzled K-major layout;
//            at the end O / l as 128-byte row stores.
// Measured (config 3, 8B): 457 us per launch against ~740 us for the mma.sync kernel on the same chunk; prefill 222 -> 191 ms.
// In-kernel timing shows where the rest is: the softmax pass (TMEM read at 64 B/clk, one FFMA + ex2 + hi/lo packing per
// score) takes ~3.4k clocks per 128 x 128 tile and is serialised with the PV product by the single P buffer; the next step
// is 64-key tiles with a third K/V stage and two P buffers, so that softmax(j+1) overlaps PV(j).
// mbarriers: q_full, kv_full[2], kv_empty[2], s_full, s_free, p_full, p_free, o_full, o_free.
#pragma once
#include "gemm_tc5.cuh"

constexpr int P5_KEYS = 128;                         // keys per tile
constexpr int P5_HALF_BYTES = 128 * 128;             // 128 rows x 64 bf16 (one 128-byte swizzle row each)
constexpr int P5_P_BYTES = 2 * P5_HALF_BYTES;        // 32 KB (128 rows x 128 keys)
// head_dim 128: Q 32 KB, a K/V stage 64 KB;  head_dim 64: Q 16 KB, a stage 32 KB
constexpr int P5_STAGES = 2;
constexpr int P5_THREADS = 320;                       // TMA warp, MMA warp, 8 softmax warps
constexpr size_t p5_smem_bytes(int hd) { return (size_t)(hd / 64) * P5_HALF_BYTES * (1 + 2 * P5_STAGES) + 2 * P5_P_BYTES + 1024 + 256 + 1024; }   // P hi + P lo

struct Pf5Args {
    const int* page_table;
    int pt_stride;
    const int* tok_pos;     // [S]
    const int4* work;       // {first token (flat), n tokens (<= 128/G), sequence, kv head}
    int n_q, n_kv, G, BT;
    int n_pages_pool;       // pages per layer in the pool
    int layer;              // layer index (row offset into the pool-wide K/V tensor maps)
    float scale_log2;
    bf16* out;              // [S][n_q*128]
};

__device__ __forceinline__ void p5_tma_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
                 "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
// polling wait (mbarrier.test_wait never suspends the thread): the hand-offs of this kernel sit on the critical path of
// every key tile, and try_wait's suspend / wake-up granularity showed up as microseconds of bubbles per tile
__device__ __forceinline__ void p5_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred P1;\n\t"
            "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, P1;\n\t"
            "}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void p5_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc5_st32(uint32_t taddr, const float (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
        "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
        "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
        "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
        "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void p5_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int HD>
__global__ void __launch_bounds__(P5_THREADS, 1) attn_prefill_tc5_kernel(const __grid_constant__ CUtensorMap map_q,
                                                                  const __grid_constant__ CUtensorMap map_k,
                                                                  const __grid_constant__ CUtensorMap map_v, const Pf5Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    constexpr int NH = HD / 64;                          // 64-dim halves of the head dimension (one 128-byte swizzle row each)
    constexpr int P5_Q_BYTES = NH * P5_HALF_BYTES, P5_KV_STAGE = 2 * NH * P5_HALF_BYTES;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t s_q = sbase;
    const uint32_t s_kv = s_q + P5_Q_BYTES;
    const uint32_t s_p = s_kv + P5_STAGES * P5_KV_STAGE;
    const uint32_t bars = s_p + 2 * P5_P_BYTES;
    const uint32_t b_q = bars, b_kvf = bars + 8, b_kve = bars + 24, b_sf = bars + 40, b_sfree = bars + 48, b_pf = bars + 56,
                   b_pfree = bars + 64, b_of = bars + 72, b_ofree = bars + 80, tmem_slot = bars + 88, xchg = bars + 128;
    unsigned char* gen = smem + (sbase - smem_u32(smem));           // generic pointer to the aligned base
    volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - sbase));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int4 wk = a.work[blockIdx.x];
    const int tok0 = wk.x, ntok = wk.y, seq = wk.z, h = wk.w;

    if (threadIdx.x == 0) {
        mbar_init(b_q, 1);
        for (int s = 0; s < P5_STAGES; ++s) { mbar_init(b_kvf + 8 * s, 1); mbar_init(b_kve + 8 * s, 1); }
        mbar_init(b_sf, 1); mbar_init(b_sfree, 8); mbar_init(b_pf, 8); mbar_init(b_pfree, 1); mbar_init(b_of, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // stale shared memory must never reach the PV product as NaN: pages a tile does not load stay whatever they were
    for (int i = threadIdx.x; i < (P5_Q_BYTES + P5_STAGES * P5_KV_STAGE) / 16; i += P5_THREADS)
        reinterpret_cast<uint4*>(gen)[i] = make_uint4(0u, 0u, 0u, 0u);
    p5_fence_async_smem();
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(256));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    const uint32_t tmem = *tmem_slot_p;

    griddep_launch();
    griddep_wait();                                   // Q and this chunk's K/V come from the QKV GEMM

    const int pos0 = a.tok_pos[tok0];
    const int n_keys = pos0 + ntok;
    const int n_pages = (n_keys + KV_PAGE - 1) / KV_PAGE;
    const int n_tiles = (n_keys + P5_KEYS - 1) / P5_KEYS;

    if (warp == 0) {
        if (lane == 0) {
            const int* pt = a.page_table + (size_t)seq * a.pt_stride;
            const uint32_t q_bytes = (uint32_t)NH * 64u * (uint32_t)a.G * (uint32_t)a.BT * 2u;
            mbar_expect_tx(b_q, q_bytes);
            for (int c = 0; c < NH; ++c) p5_tma_3d(s_q + c * P5_HALF_BYTES, &map_q, b_q, c * 64, h * a.G, tok0);
            for (int j = 0; j < n_tiles; ++j) {
                const int s = j & 1;
                p5_wait(b_kve + 8 * s, ((j >> 1) & 1) ^ 1);
                const int npg = min(P5_KEYS / KV_PAGE, n_pages - j * (P5_KEYS / KV_PAGE));
                mbar_expect_tx(b_kvf + 8 * s, (uint32_t)npg * (uint32_t)(NH * 4096));
                const uint32_t st = s_kv + s * P5_KV_STAGE;
                for (int p = 0; p < npg; ++p) {
                    const int page = pt[j * (P5_KEYS / KV_PAGE) + p];
                    const int row = ((a.layer * a.n_pages_pool + page) * a.n_kv + h) * KV_PAGE;
                    for (int c = 0; c < NH; ++c) {
                        tma_load_2d(st + c * P5_HALF_BYTES + p * (KV_PAGE * 128), &map_k, b_kvf + 8 * s, c * 64, row);
                        tma_load_2d(st + (NH + c) * P5_HALF_BYTES + p * (KV_PAGE * 128), &map_v, b_kvf + 8 * s, c * 64, row);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptors: D = f32, A = B = bf16, M = 128, N = 128; bit 16 = B is MN-major (V is [key][dim])
            const uint32_t idesc_qk = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const uint32_t idesc_pv = (idesc_qk & ~(0x3Fu << 17)) | ((uint32_t)(HD >> 3) << 17) | (1u << 16);   // N = head_dim
            p5_wait(b_q, 0);
            // software pipeline: S(j) = Q K_j^T is issued BEFORE O += P(j-1) V_(j-1), so the tensor pipe works on the next
            // scores while the softmax warps turn the previous ones into P
            for (int j = 0; j <= n_tiles; ++j) {
                if (j < n_tiles) {
                    const int s = j & 1;
                    const uint32_t st = s_kv + s * P5_KV_STAGE;
                    p5_wait(b_kvf + 8 * s, (j >> 1) & 1);
                    p5_wait(b_sfree, (j & 1) ^ 1);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 4 * NH; ++k) {    // halves of the head dimension x 4 k16 steps
                        const uint64_t adesc = tc5_smem_desc(s_q + (k >> 2) * P5_HALF_BYTES + (k & 3) * 32, 16, 1024, 2);
                        const uint64_t bdesc = tc5_smem_desc(st + (k >> 2) * P5_HALF_BYTES + (k & 3) * 32, 16, 1024, 2);
                        tc5_mma_bf16(tmem, adesc, bdesc, idesc_qk, k > 0 ? 1u : 0u);
                    }
                    tc5_commit(b_sf);
                }
                if (j > 0) {
                    const int i = j - 1, s = i & 1;
                    const uint32_t st = s_kv + s * P5_KV_STAGE;
                    p5_wait(b_pf, i & 1);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 16; ++k) {        // P hi then P lo; 128 keys = 8 k16 steps each
                        const int kk = k & 7;
                        const uint64_t adesc = tc5_smem_desc(s_p + (k >> 3) * P5_P_BYTES + (kk >> 2) * P5_HALF_BYTES + (kk & 3) * 32, 16, 1024, 2);
                        // V, MN-major: 64-dim blocks P5_HALF_BYTES apart (LBO), 8-key groups 1024 B apart (SBO), 16 keys per step
                        const uint64_t bdesc = tc5_smem_desc(st + NH * P5_HALF_BYTES + kk * (16 * 128), P5_HALF_BYTES, 1024, 2);
                        tc5_mma_bf16(tmem + 128, adesc, bdesc, idesc_pv, (i > 0 || k > 0) ? 1u : 0u);
                    }
                    tc5_commit(b_of);                     // O includes tile i
                    tc5_commit(b_kve + 8 * s);
                    tc5_commit(b_pfree);
                }
            }
        }
    } else {
        // 8 softmax warps: TMEM lane quadrant = warp % 4 (hardware rule), and the two warps of a quadrant split a row's
        // 128 keys (and the 128 output dims) in halves; they agree on the row maximum through shared memory once per tile
        const int quad = warp & 3, half = (warp - 2) >> 2;
        const int r = quad * 32 + lane;                    // query row = TMEM lane
        const int t = r / a.G, g = r % a.G;
        const bool row_ok = t < ntok;
        const int limit = row_ok ? pos0 + t + 1 : 1;       // keys [0, limit) are visible
        int warp_min_limit = limit;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) warp_min_limit = min(warp_min_limit, __shfl_xor_sync(0xffffffffu, warp_min_limit, o));
        const uint32_t trow = tmem + ((uint32_t)(quad * 32) << 16);
        unsigned char* p_row = gen + (s_p - sbase) + r * 128;
        float* xch = reinterpret_cast<float*>(gen + (xchg - sbase));          // [2][128] exchange of half-row values
        // m_used: the maximum the stored O and l are expressed in.  It follows the running maximum lazily: O is rescaled
        // (TMEM load, multiply, store) only when the maximum has grown by more than 2^8 since, which after the first tiles
        // is rare; until then P simply carries a factor <= 256, exact in fp32 and harmless in bf16.
        float m_used = -INFINITY, l_run = 0.f;
        for (int j = 0; j < n_tiles; ++j) {
            const int kb = j * P5_KEYS + half * 64;
            p5_wait(b_sf, j & 1);
            tc5_fence_after();
            // Two passes over this warp's 64 score columns in chunks of 32 (one tcgen05.ld each), with the chunk loops NOT
            // unrolled: fully unrolled the tile body was ~25 KB of code and the single warp per scheduler stalled on
            // instruction fetch (ncu: 31 % of its cycles).
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                float v[32];
                tc5_ld32(trow + half * 64 + c * 32, v);
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (kb + c * 32 + i < limit) mx = fmaxf(mx, v[i] * a.scale_log2);
            }
            xch[half * 128 + r] = mx;
            asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
            mx = fmaxf(mx, xch[(half ^ 1) * 128 + r]);
            asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");     // both have read before the next tile writes
            const bool grow = mx > m_used + 8.0f;            // also true on the first tile (m_used = -inf)
            if (__any_sync(0xffffffffu, grow)) {             // same decision in both warps of the quadrant: same rows, same mx
                const float m_new = grow ? mx : m_used;
                const float f = (j > 0) ? exp2f(m_used - m_new) : 1.0f;   // 1 for rows that keep their maximum
                if (j > 0) {
                    p5_wait(b_of, (j - 1) & 1);              // every product issued so far has landed in O
                    tc5_fence_after();
#pragma unroll 1
                    for (int c = 0; c < NH; ++c) {
                        float v[32];
                        tc5_ld32(trow + 128 + half * (HD / 2) + c * 32, v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] *= f;
                        tc5_st32(trow + 128 + half * (HD / 2) + c * 32, v);
                    }
                    tc5_fence_before();
                    l_run *= f;
                }
                m_used = m_new;
            }
            p5_wait(b_pfree, (j & 1) ^ 1);                   // the previous tile's PV product has read P
            float lsum = 0.f;
            const float neg_m = -m_used;
            const bool diag = kb + 64 > warp_min_limit;      // some key of this warp's 64 is hidden from some row (warp-uniform)
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                float v[32];
                tc5_ld32(trow + half * 64 + c * 32, v);
                if (c == 1) {                                // S is consumed: the next QK^T may overwrite it
                    tc5_fence_before();
                    __syncwarp();
                    if (lane == 0) p5_mbar_arrive(b_sfree);
                }
                if (diag) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = (kb + c * 32 + i < limit) ? v[i] : -INFINITY;
                }
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {             // 8 keys = one 16-byte chunk of the row
                    uint32_t wh[4], wl[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int i0 = ch * 8 + 2 * e;
                        const float p0 = exp2f(fmaf(v[i0], a.scale_log2, neg_m)), p1 = exp2f(fmaf(v[i0 + 1], a.scale_log2, neg_m));
                        wh[e] = pack_bf16(p0, p1);
                        wl[e] = pack_bf16(p0 - bf16lo(wh[e]), p1 - bf16hi(wh[e]));
                        lsum += p0 + p1;
                    }
                    const int off = half * P5_HALF_BYTES + (((c * 4 + ch) ^ (r & 7)) << 4);
                    *reinterpret_cast<uint4*>(p_row + off) = make_uint4(wh[0], wh[1], wh[2], wh[3]);
                    *reinterpret_cast<uint4*>(p_row + P5_P_BYTES + off) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
                }
            }
            l_run += lsum;
            p5_fence_async_smem();
            __syncwarp();
            if (lane == 0) p5_mbar_arrive(b_pf);
        }
        // row sum = the two halves' partial sums
        xch[half * 128 + r] = l_run;
        asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
        l_run += xch[(half ^ 1) * 128 + r];
        p5_wait(b_of, (n_tiles - 1) & 1);
        tc5_fence_after();
        const float inv = 1.0f / l_run;
        bf16* dst = a.out + (size_t)(tok0 + t) * ((size_t)a.n_q * HD) + (size_t)(h * a.G + g) * HD + half * (HD / 2);
#pragma unroll 1
        for (int c = 0; c < NH; ++c) {
            float v[32];
            tc5_ld32(trow + 128 + half * (HD / 2) + c * 32, v);
            if (row_ok) {
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
                    uint4 pk;
                    pk.x = pack_bf16(v[i] * inv, v[i + 1] * inv); pk.y = pack_bf16(v[i + 2] * inv, v[i + 3] * inv);
                    pk.z = pack_bf16(v[i + 4] * inv, v[i + 5] * inv); pk.w = pack_bf16(v[i + 6] * inv, v[i + 7] * inv);
                    *reinterpret_cast<uint4*>(dst + c * 32 + i) = pk;
                }
            }
        }
        tc5_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc5_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
    }
}
