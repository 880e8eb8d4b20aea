// tcgen05 prefill attention, pipelined variant: 64-key tiles, a 3-stage K/V ring, TWO score buffers in TMEM and TWO P
// buffers in shared memory, so that the softmax of tile j+1 runs while the tensor pipe multiplies P(j) V(j).
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external server);
// SURVEY.md 8(a-K) row "prefill causal attention".  Same work decomposition, operand layouts and numerics as
// attn_prefill_tc5.cuh (read that header first); what changes is the schedule.  In-kernel timing of the 128-key version
// showed a serial chain per tile -- PV(j-1) done -> softmax(j) writes the single P buffer (~3.4k clocks) -> PV(j) -- with
// the tensor pipe idle two thirds of the time.  Here:
//   shared memory  Q 32 KB | K/V ring 3 x 32 KB (4 pages: K halves, V halves) | P 2 x (hi 16 KB + lo 16 KB)   = 192 KB
//   TMEM           S0 cols 0..63, S1 cols 64..127, O cols 128..255
//   MMA warp       iteration j: S[j&1] = Q K_j^T (needs s_free[j&1]); then O += P[(j-1)&1] V_(j-1) (needs p_full[(j-1)&1])
//   softmax warps  tile j: S[j&1] -> (max, lazy rescale) -> P[j&1] (needs p_free[j&1], i.e. PV(j-2) done, long past)
// so the only wait left on the softmax path is for the scores themselves.
#pragma once
#include "attn_prefill_tc5.cuh"

constexpr int P6_KEYS = 64;
constexpr int P6_STAGES = 3;
constexpr int P6_TILE_BYTES = 64 * 128;              // 64 rows (keys) x 64 bf16
constexpr int P6_PBUF = 2 * P5_HALF_BYTES;           // P hi (128 rows x 64 keys = 16 KB) + P lo
constexpr size_t p6_smem_bytes(int hd) {
    return (size_t)(hd / 64) * P5_HALF_BYTES + (size_t)P6_STAGES * 2 * (hd / 64) * P6_TILE_BYTES + 2 * P6_PBUF + 1024 + 256 + 1024 + 512;   // + page-id cache
}

// one MUFU instruction (exp2f without fast-math expands to range handling around it; inputs here are <= 8, and anything
// below -126 may flush to zero)
__device__ __forceinline__ float p6_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int HD>
__global__ void __launch_bounds__(P5_THREADS, 1) attn_prefill_tc5b_kernel(const __grid_constant__ CUtensorMap map_q,
                                                                          const __grid_constant__ CUtensorMap map_k,
                                                                          const __grid_constant__ CUtensorMap map_v, const Pf5Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    constexpr int NH = HD / 64;
    constexpr int Q_BYTES = NH * P5_HALF_BYTES, STAGE = 2 * NH * P6_TILE_BYTES;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t s_q = sbase;
    const uint32_t s_kv = s_q + Q_BYTES;
    const uint32_t s_p = s_kv + P6_STAGES * STAGE;
    const uint32_t bars = s_p + 2 * P6_PBUF;
    // barriers (8 bytes each): q, kvf[3], kve[3], sf[2], sfree[2], pf[2], pfree[2], of
    const uint32_t b_q = bars, b_kvf = bars + 8, b_kve = bars + 32, b_sf = bars + 56, b_sfree = bars + 72, b_pf = bars + 88,
                   b_pfree = bars + 104, b_of = bars + 120, tmem_slot = bars + 128, xchg = bars + 256;
    unsigned char* gen = smem + (sbase - smem_u32(smem));
    volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - sbase));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int4 wk = a.work[blockIdx.x];
    const int tok0 = wk.x, ntok = wk.y, seq = wk.z, h = wk.w;

    if (threadIdx.x == 0) {
        mbar_init(b_q, 1);
        for (int s = 0; s < P6_STAGES; ++s) { mbar_init(b_kvf + 8 * s, 1); mbar_init(b_kve + 8 * s, 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(b_sf + 8 * s, 1); mbar_init(b_sfree + 8 * s, 8); mbar_init(b_pf + 8 * s, 8); mbar_init(b_pfree + 8 * s, 1);
        }
        mbar_init(b_of, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // stale shared memory must never reach the PV product as NaN: pages a tile does not load stay whatever they were
    for (int i = threadIdx.x; i < (Q_BYTES + P6_STAGES * STAGE) / 16; i += P5_THREADS)
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
    const int n_tiles = (n_keys + P6_KEYS - 1) / P6_KEYS;
    constexpr int PPT = P6_KEYS / KV_PAGE;             // pages per tile

    if (warp == 0) {
        {
            const int* pt = a.page_table + (size_t)seq * a.pt_stride;
            // the page ids of the next 32 tiles are fetched by the whole warp at once into shared memory: read one by one by
            // the issuing lane, every id is a dependent global load in front of its TMA (four per tile)
            int* spt = reinterpret_cast<int*>(gen + (xchg - sbase) + 1024);
            if (lane == 0) {
                mbar_expect_tx(b_q, (uint32_t)NH * 64u * (uint32_t)a.G * (uint32_t)a.BT * 2u);
                for (int c = 0; c < NH; ++c) p5_tma_3d(s_q + c * P5_HALF_BYTES, &map_q, b_q, c * 64, h * a.G, tok0);
            }
            for (int j = 0; j < n_tiles; ++j) {
                if ((j & 31) == 0) {
                    __syncwarp();
                    for (int i = lane; i < 32 * PPT; i += 32) spt[i] = (j * PPT + i < n_pages) ? pt[j * PPT + i] : 0;
                    __syncwarp();
                }
                // one box per lane: 16 K/V boxes of a tile issued by 16 lanes at once.  Issued one after the other by a single
                // lane they took ~1.2 us per tile (in-kernel trace) and paced the whole pipeline
                const int s = j % P6_STAGES;
                if (lane == 0) {
                    p5_wait(b_kve + 8 * s, ((j / P6_STAGES) & 1) ^ 1);
                    mbar_expect_tx(b_kvf + 8 * s, (uint32_t)min(PPT, n_pages - j * PPT) * (uint32_t)(NH * 4096));
                }
                __syncwarp();
                const int npg = min(PPT, n_pages - j * PPT);
                const uint32_t st = s_kv + s * STAGE;
                if (lane < npg * 2 * NH) {
                    const int p = lane / (2 * NH), c = (lane % (2 * NH)) >> 1, is_v = lane & 1;
                    const int page = spt[(j & 31) * PPT + p];
                    const int row = ((a.layer * a.n_pages_pool + page) * a.n_kv + h) * KV_PAGE;
                    tma_load_2d(st + (is_v * NH + c) * P6_TILE_BYTES + p * (KV_PAGE * 128), is_v ? &map_v : &map_k, b_kvf + 8 * s, c * 64, row);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // D = f32, A = B = bf16, M = 128; QK^T: N = 64 keys, both K-major; PV: N = head_dim, B (V) MN-major (bit 16)
            const uint32_t idesc_base = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 4) << 24);
            const uint32_t idesc_qk = idesc_base | ((uint32_t)(P6_KEYS >> 3) << 17);
            const uint32_t idesc_pv = idesc_base | ((uint32_t)(HD >> 3) << 17) | (1u << 16);
            p5_wait(b_q, 0);
            for (int j = 0; j <= n_tiles; ++j) {
                if (j < n_tiles) {
                    const int s = j % P6_STAGES, b = j & 1;
                    const uint32_t st = s_kv + s * STAGE;
                    p5_wait(b_kvf + 8 * s, (j / P6_STAGES) & 1);
                    p5_wait(b_sfree + 8 * b, ((j >> 1) & 1) ^ 1);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 4 * NH; ++k) {    // halves of the head dimension x 4 k16 steps
                        const uint64_t adesc = tc5_smem_desc(s_q + (k >> 2) * P5_HALF_BYTES + (k & 3) * 32, 16, 1024, 2);
                        const uint64_t bdesc = tc5_smem_desc(st + (k >> 2) * P6_TILE_BYTES + (k & 3) * 32, 16, 1024, 2);
                        tc5_mma_bf16(tmem + b * P6_KEYS, adesc, bdesc, idesc_qk, k > 0 ? 1u : 0u);
                    }
                    tc5_commit(b_sf + 8 * b);
                }
                if (j > 0) {
                    const int i = j - 1, s = i % P6_STAGES, b = i & 1;
                    const uint32_t st = s_kv + s * STAGE;
                    p5_wait(b_pf + 8 * b, (i >> 1) & 1);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 8; ++k) {         // P hi then P lo; 64 keys = 4 k16 steps each
                        const int kk = k & 3;
                        const uint64_t adesc = tc5_smem_desc(s_p + b * P6_PBUF + (k >> 2) * P5_HALF_BYTES + kk * 32, 16, 1024, 2);
                        // V, MN-major: 64-dim blocks one tile apart (LBO), 8-key groups 1024 B apart (SBO), 16 keys per step
                        const uint64_t bdesc = tc5_smem_desc(st + NH * P6_TILE_BYTES + kk * (16 * 128), P6_TILE_BYTES, 1024, 2);
                        tc5_mma_bf16(tmem + 128, adesc, bdesc, idesc_pv, (i > 0 || k > 0) ? 1u : 0u);
                    }
                    tc5_commit(b_of);                     // O includes tile i
                    tc5_commit(b_kve + 8 * s);
                    tc5_commit(b_pfree + 8 * b);
                }
            }
        }
    } else {
        // 8 softmax warps: TMEM lane quadrant = warp % 4; the two warps of a quadrant take 32 keys (and half the output dims) each
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
        float m_used = -INFINITY, l_run = 0.f;             // lazy maximum, see attn_prefill_tc5.cuh
        for (int j = 0; j < n_tiles; ++j) {
            const int b = j & 1;
            const int kb = j * P6_KEYS + half * 32;
            p5_wait(b_sf + 8 * b, (j >> 1) & 1);
            tc5_fence_after();
            float v[32];
            tc5_ld32(trow + b * P6_KEYS + half * 32, v);
            tc5_fence_before();
            __syncwarp();
            if (lane == 0) p5_mbar_arrive(b_sfree + 8 * b);    // the scores are in registers
            if (kb + 32 > warp_min_limit) {                     // this warp's 32 keys touch the diagonal of some row
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = (kb + i < limit) ? v[i] : -INFINITY;
            }
            float mx = -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; ++i) mx = fmaxf(mx, v[i]);
            mx *= a.scale_log2;                                  // scale > 0: max commutes
            xch[half * 128 + r] = mx;
            asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
            mx = fmaxf(mx, xch[(half ^ 1) * 128 + r]);
            asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");     // both have read before the next tile writes
            const bool grow = mx > m_used + 8.0f;                // also true on the first tile (m_used = -inf)
            if (__any_sync(0xffffffffu, grow)) {
                const float m_new = grow ? mx : m_used;
                const float f = (j > 0) ? exp2f(m_used - m_new) : 1.0f;
                if (j > 0) {
                    p5_wait(b_of, (j - 1) & 1);                  // every product issued so far has landed in O
                    tc5_fence_after();
#pragma unroll 1
                    for (int c = 0; c < NH; ++c) {
                        float w[32];
                        tc5_ld32(trow + 128 + half * (HD / 2) + c * 32, w);
#pragma unroll
                        for (int i = 0; i < 32; ++i) w[i] *= f;
                        tc5_st32(trow + 128 + half * (HD / 2) + c * 32, w);
                    }
                    tc5_fence_before();
                    l_run *= f;
                }
                m_used = m_new;
            }
            const float neg_m = -m_used;
            uint32_t wh[16], wl[16];
            float lsum = 0.f;
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                const float p0 = p6_ex2(fmaf(v[2 * e], a.scale_log2, neg_m)), p1 = p6_ex2(fmaf(v[2 * e + 1], a.scale_log2, neg_m));
                wh[e] = pack_bf16(p0, p1);
                wl[e] = pack_bf16(p0 - bf16lo(wh[e]), p1 - bf16hi(wh[e]));
                lsum += p0 + p1;
            }
            l_run += lsum;
            p5_wait(b_pfree + 8 * b, ((j >> 1) & 1) ^ 1);        // PV(j-2) has read this P buffer (long past)
            unsigned char* pb = p_row + b * P6_PBUF;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {                     // 8 keys = one 16-byte chunk; chunks half*4 .. half*4+3 of the row
                const int off = ((half * 4 + ch) ^ (r & 7)) << 4;
                *reinterpret_cast<uint4*>(pb + off) = make_uint4(wh[4 * ch], wh[4 * ch + 1], wh[4 * ch + 2], wh[4 * ch + 3]);
                *reinterpret_cast<uint4*>(pb + P5_HALF_BYTES + off) = make_uint4(wl[4 * ch], wl[4 * ch + 1], wl[4 * ch + 2], wl[4 * ch + 3]);
            }
            p5_fence_async_smem();
            __syncwarp();
            if (lane == 0) p5_mbar_arrive(b_pf + 8 * b);
        }
        xch[half * 128 + r] = l_run;                             // row sum = the two halves' partial sums
        asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
        l_run += xch[(half ^ 1) * 128 + r];
        p5_wait(b_of, (n_tiles - 1) & 1);
        tc5_fence_after();
        const float inv = 1.0f / l_run;
        f16* dst = a.out + (size_t)(tok0 + t) * ((size_t)a.n_q * HD) + (size_t)(h * a.G + g) * HD + half * (HD / 2);
#pragma unroll 1
        for (int c = 0; c < NH; ++c) {
            float w[32];
            tc5_ld32(trow + 128 + half * (HD / 2) + c * 32, w);
            if (row_ok) {
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
                    uint4 pk;
                    pk.x = pack_f16(w[i] * inv, w[i + 1] * inv); pk.y = pack_f16(w[i + 2] * inv, w[i + 3] * inv);
                    pk.z = pack_f16(w[i + 4] * inv, w[i + 5] * inv); pk.w = pack_f16(w[i + 6] * inv, w[i + 7] * inv);
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
