// Decode attention for LONG contexts on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), sm_100a only.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external server);
// SURVEY.md 8(a-K) rows "paged decode attention (GQA)" and "shared-prefix (cascade) attention".
//
// Why tensor memory for a decode kernel: with a long SHARED prefix the rows that attend the same keys are all knights x
// all q-heads of a kv group (config 3: 5 x 4 = 20 rows, config 5: 8 x 8 = 64), and the mma.sync kernel (kernels.cuh) keeps
// one 16-row tile of output per warp in registers: the prefix is streamed once per tile, every warp walks its pages in a
// dependent chain, and the stage ran at 1.8 TB/s (18 us + an 8 us merge of 69 partials per layer on config 3).  Here the
// output of up to 128 rows lives in TMEM, K/V pages arrive by TMA straight from the paged pool, and one CTA per SM works a
// key RANGE for all rows at once -- the schedule of attn_prefill_tc5b.cuh (64-key tiles, 3-stage K/V ring, two score
// buffers in TMEM, two P buffers) with the query rows being (slot, q-head) pairs instead of (token, q-head) pairs.
//
// Work items (one CTA each, <= one wave; the engine sizes the key ranges so that all items have about the same number of
// tiles): a SHARED item = (kv head, key range of the shared prefix), valid rows = every slot; an OWN item = (slot, kv head,
// key range of that sequence's own keys), valid rows = the G rows of that slot.  Positions are read on the device (the
// decode graph is replayed unchanged while sequences grow): an item whose range starts beyond its sequence ends at once.
// Each item writes an un-normalised partial (O, m, l) per valid row; attn_merge2_kernel adds the <= ~25 partials of a row.
// Softmax warps whose 32 TMEM lanes hold no valid row only keep the barrier protocol going.
// The softmax of a tile is a serial chain per thread (load scores, max, exp2, pack P hi + lo, store) and sets the pace of
// the whole pipeline, and a decode step has few rows: so a row's 64 keys are spread over FOUR threads (16 softmax warps,
// four per TMEM lane quadrant) instead of the prefill kernel's two.
#pragma once
#include "attn_prefill_tc5b.cuh"

#ifndef RT_D5_PARTS
#define RT_D5_PARTS 2
#endif
constexpr int D5_PARTS = RT_D5_PARTS;                      // threads (warps per lane quadrant) sharing a row's keys: 2 or 4
constexpr int D5_THREADS = 64 + D5_PARTS * 128;            // TMA warp, MMA warp, 4 x D5_PARTS softmax warps

template <int N>
__device__ __forceinline__ void d5_ld(uint32_t taddr, float (&v)[N]);
template <>
__device__ __forceinline__ void d5_ld<32>(uint32_t taddr, float (&v)[32]) { tc5_ld32(taddr, v); }
template <>
__device__ __forceinline__ void d5_ld<16>(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
template <int N>
__device__ __forceinline__ void d5_st(uint32_t taddr, const float (&v)[N]);
template <>
__device__ __forceinline__ void d5_st<32>(uint32_t taddr, const float (&v)[32]) { tc5_st32(taddr, v); }
template <>
__device__ __forceinline__ void d5_st<16>(uint32_t taddr, const float (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

struct Dec5Item {
    int slot;      // -1: shared-prefix item (all slots); >= 0: own keys of that slot
    int h;         // kv head (local)
    int key0;      // SHARED: first key; OWN: first key RELATIVE to the shared length
    int span;      // keys (multiple of 64)
    int pidx;      // partial index of this item within its rows: shared items 0 .. n_sh-1, own items n_sh + chunk
    int seq;       // sequence whose page table the item walks (decode slot t of a generate call is sequence t; shared items: 0)
    int pad[2];
};
// The work items travel as a kernel PARAMETER (constant bank): the first thing every CTA does is read its item, and as a
// global load that was a whole L2 round trip in front of everything else (the prologue only hides under the QKV projection for
// the CTAs that find a free SM early, profiles/r2_summary.md 7).  They are a function of the plan alone, so a captured graph
// may keep them.
constexpr int D5_MAX_ITEMS = 296;
struct Dec5ItemList {
    Dec5Item it[D5_MAX_ITEMS];
};

struct Dec5Args {
    const int* page_table;
    int pt_stride;
    const int* slot_seq;
    const int* slot_pos;
    const int* shared_len_ptr;
    int T, n_q, n_kv, G, BT;
    int n_pages_pool;       // pages per layer in the pool
    int layer;
    float scale_log2;
    float* part_o;          // [T][n_q][max_chunks][hd]
    float* part_ml;         // [T][n_q][max_chunks][2]
    int max_chunks;
    int early;              // key ranges and old K/V tiles before the PDL wait (see AttnArgs.early)
};

template <int HD>
__global__ void __launch_bounds__(D5_THREADS, 1) attn_decode_tc5_kernel(const __grid_constant__ CUtensorMap map_q,
                                                                        const __grid_constant__ CUtensorMap map_k,
                                                                        const __grid_constant__ CUtensorMap map_v, const Dec5Args a,
                                                                        const __grid_constant__ Dec5ItemList items) {
    extern __shared__ __align__(1024) unsigned char smem[];
    constexpr int NH = HD / 64;
    constexpr int Q_BYTES = NH * P5_HALF_BYTES, STAGE = 2 * NH * P6_TILE_BYTES;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t s_q = sbase;
    const uint32_t s_kv = s_q + Q_BYTES;
    const uint32_t s_p = s_kv + P6_STAGES * STAGE;
    const uint32_t bars = s_p + 2 * P6_PBUF;
    const uint32_t b_q = bars, b_kvf = bars + 8, b_kve = bars + 32, b_sf = bars + 56, b_sfree = bars + 72, b_pf = bars + 88,
                   b_pfree = bars + 104, b_of = bars + 120, tmem_slot = bars + 128, xchg = bars + 256;
    unsigned char* gen = smem + (sbase - smem_u32(smem));
    volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - sbase));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Dec5Item it = items.it[blockIdx.x];
    RT_TR(80);

    if (threadIdx.x == 0) {
        mbar_init(b_q, 1);
        for (int s = 0; s < P6_STAGES; ++s) { mbar_init(b_kvf + 8 * s, 1); mbar_init(b_kve + 8 * s, 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(b_sf + 8 * s, 1); mbar_init(b_sfree + 8 * s, 4 * D5_PARTS); mbar_init(b_pf + 8 * s, 4 * D5_PARTS); mbar_init(b_pfree + 8 * s, 1);
        }
        mbar_init(b_of, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Stale shared memory must never reach the PV product as NaN through a page a tile does not load: the V half of every
    // stage starts zeroed (its P is zero there, and 0 x NaN is NaN).  Nothing else needs it: scores of keys beyond the range
    // are REPLACED by -inf, Q is written whole by its TMA box, and the rows of a product depend on their own P row only
    // (rows without a valid query are never stored).
    for (int s = 0; s < P6_STAGES; ++s)
        for (int i = threadIdx.x; i < NH * P6_TILE_BYTES / 16; i += D5_THREADS)
            reinterpret_cast<uint4*>(gen + Q_BYTES + s * STAGE + NH * P6_TILE_BYTES)[i] = make_uint4(0u, 0u, 0u, 0u);
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
    RT_TR(85);

    // The key range of the item.  `early` (the caller says: inside the decode graph, where the positions only change between
    // steps): worked out BEFORE the PDL wait, and the K/V tiles that cannot hold the key this step appends -- all of a shared
    // item, all but the last tile of the own item that reaches the current position -- are requested there too, so that
    // only Q (and that one tile) is fetched behind the QKV projection.
    int shared = 0, seq = 0, key0 = 0, key1 = 0, n_safe = 0;
    auto ranges = [&]() {
        shared = a.shared_len_ptr ? *a.shared_len_ptr : 0;
        seq = it.seq;
        if (it.slot < 0) {
            key0 = it.key0;
            key1 = min(key0 + it.span, shared);
            n_safe = 1 << 30;
        } else {
            const int pos = a.slot_pos[it.slot];
            key0 = shared + it.key0;
            key1 = min(key0 + it.span, pos + 1);
            n_safe = key0 + it.span > pos ? (max(key1 - key0, 0) + P6_KEYS - 1) / P6_KEYS - 1 : 1 << 30;
        }
    };
    constexpr int PPT = P6_KEYS / KV_PAGE;
    int* spt = reinterpret_cast<int*>(gen + (xchg - sbase) + D5_PARTS * 512);
    // a shared item's pages are known from the item alone: their ids are requested together with the positions (one round
    // trip instead of two in front of the first K/V tile)
    bool spt_ready = false;
    if (a.early && warp == 0 && it.slot < 0) {
        const int* pt0 = a.page_table + (size_t)it.seq * a.pt_stride + it.key0 / KV_PAGE;
        const int np_max = it.span / KV_PAGE;
        int ids[PPT];
#pragma unroll
        for (int i = 0; i < PPT; ++i) ids[i] = (lane + 32 * i < np_max) ? pt0[lane + 32 * i] : 0;
        ranges();
#pragma unroll
        for (int i = 0; i < PPT; ++i) spt[lane + 32 * i] = ids[i];
        __syncwarp();
        spt_ready = true;
    } else if (a.early) {
        ranges();
    }
    // tile j of the item -> its ring stage (one K/V box per lane: 16 boxes of a tile issued by 16 lanes at once.  Issued one
    // after the other by a single lane they took ~1.2 us per tile in the in-kernel trace and paced the whole pipeline)
    auto issue_tile = [&](int j, int n_pages, const int* pt) {
        if ((j & 31) == 0 && !(j == 0 && spt_ready)) {
            // the page ids of the next 32 tiles are fetched by the whole warp at once into shared memory: read one by one by
            // the issuing lane, every id is a dependent global load in front of its TMA (four per tile, ~2 us a tile)
            __syncwarp();
            for (int i = lane; i < 32 * PPT; i += 32) spt[i] = (j * PPT + i < n_pages) ? pt[j * PPT + i] : 0;
            __syncwarp();
        }
        const int s = j % P6_STAGES;
        if (lane == 0) {
            p5_wait(b_kve + 8 * s, ((j / P6_STAGES) & 1) ^ 1);
            mbar_expect_tx(b_kvf + 8 * s, (uint32_t)min(PPT, n_pages - j * PPT) * (uint32_t)(NH * 4096));
            RT_TR1(82);
        }
        __syncwarp();
        const int npg = min(PPT, n_pages - j * PPT);
        const uint32_t st = s_kv + s * STAGE;
        if (lane < npg * 2 * NH) {
            const int p = lane / (2 * NH), c = (lane % (2 * NH)) >> 1, is_v = lane & 1;
            const int page = spt[(j & 31) * PPT + p];
            const int row = ((a.layer * a.n_pages_pool + page) * a.n_kv + it.h) * KV_PAGE;
            tma_load_2d(st + (is_v * NH + c) * P6_TILE_BYTES + p * (KV_PAGE * 128), is_v ? &map_v : &map_k, b_kvf + 8 * s, c * 64, row);
        }
    };
    int n_pre = 0;
    if (a.early && warp == 0) {
        const int nk = max(key1 - key0, 0), npg = (nk + KV_PAGE - 1) / KV_PAGE, nt = (nk + P6_KEYS - 1) / P6_KEYS;
        const int* pt = a.page_table + (size_t)seq * a.pt_stride + key0 / KV_PAGE;
        n_pre = max(min(min(nt, n_safe), P6_STAGES), 0);
        for (int j = 0; j < n_pre; ++j) issue_tile(j, npg, pt);
    }
    griddep_wait();                                   // Q and the newest K/V come from the QKV projection of this layer
    RT_TR(81);
    if (!a.early) ranges();

    const int n_keys = max(key1 - key0, 0);           // 0: nothing to do (the sequence has not reached this range)
    const int n_pages = (n_keys + KV_PAGE - 1) / KV_PAGE;
    const int n_tiles = (n_keys + P6_KEYS - 1) / P6_KEYS;
    const int page0 = key0 / KV_PAGE;                 // key0 is page aligned (the shared length and the spans are)

    if (warp == 0) {
        if (n_tiles > 0) {
            const int* pt = a.page_table + (size_t)seq * a.pt_stride + page0;
            if (lane == 0) {
                mbar_expect_tx(b_q, (uint32_t)NH * 64u * (uint32_t)a.G * (uint32_t)a.BT * 2u);
                for (int c = 0; c < NH; ++c) p5_tma_3d(s_q + c * P5_HALF_BYTES, &map_q, b_q, c * 64, it.h * a.G, 0);
            }
            for (int j = n_pre; j < n_tiles; ++j) {
                issue_tile(j, n_pages, pt);
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && n_tiles > 0) {
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
                    RT_TR1(87);
                    p5_wait(b_sfree + 8 * b, ((j >> 1) & 1) ^ 1);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 4 * NH; ++k) {
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
                    RT_TR1(88);
                    tc5_fence_after();
#pragma unroll
                    for (int k = 0; k < 8; ++k) {         // P hi then P lo; 64 keys = 4 k16 steps each
                        const int kk = k & 3;
                        const uint64_t adesc = tc5_smem_desc(s_p + b * P6_PBUF + (k >> 2) * P5_HALF_BYTES + kk * 32, 16, 1024, 2);
                        const uint64_t bdesc = tc5_smem_desc(st + NH * P6_TILE_BYTES + kk * (16 * 128), P6_TILE_BYTES, 1024, 2);
                        tc5_mma_bf16(tmem + 128, adesc, bdesc, idesc_pv, (i > 0 || k > 0) ? 1u : 0u);
                    }
                    tc5_commit(b_of);
                    tc5_commit(b_kve + 8 * s);
                    tc5_commit(b_pfree + 8 * b);
                }
            }
        }
    } else {
        // 4 x D5_PARTS softmax warps: TMEM lane quadrant = warp % 4; the D5_PARTS warps of a quadrant take KPP keys of a tile
        // (and DPP of the output dims) each
        constexpr int KPP = P6_KEYS / D5_PARTS, DPP = HD / D5_PARTS, DCH = DPP > 32 ? 32 : DPP;   // TMEM chunks of <= 32 columns
        const int quad = warp & 3, part = (warp - 2) >> 2;
        const int r = quad * 32 + lane;                    // query row = TMEM lane = slot * G + head-in-group
        const int t = r / a.G, g = r % a.G;
        const bool row_ok = t < a.T && (it.slot < 0 || t == it.slot);
        const bool warp_ok = __any_sync(0xffffffffu, row_ok);    // the same for all warps of a quadrant (same rows)
        const uint32_t trow = tmem + ((uint32_t)(quad * 32) << 16);
        unsigned char* p_row = gen + (s_p - sbase) + r * 128;
        float* xch = reinterpret_cast<float*>(gen + (xchg - sbase));          // [D5_PARTS][128] exchange of part-row values
        float m_used = -INFINITY, l_run = 0.f;             // lazy maximum (attn_prefill_tc5.cuh)
        for (int j = 0; j < n_tiles; ++j) {
            const int b = j & 1;
            const int kb = j * P6_KEYS + part * KPP;           // first key of this warp's part of the tile, relative to key0
            p5_wait(b_sf + 8 * b, (j >> 1) & 1);
            if (warp == 4 && lane == 0) RT_TR1(83);
            tc5_fence_after();
            if (!warp_ok) {
                // no valid row in these 32 lanes: keep the hand-shakes going, leave this warp's P rows zero
                if (lane == 0) p5_mbar_arrive(b_sfree + 8 * b);
                p5_wait(b_pfree + 8 * b, ((j >> 1) & 1) ^ 1);
                if (lane == 0) p5_mbar_arrive(b_pf + 8 * b);
                continue;
            }
            float v[KPP];
            d5_ld<KPP>(trow + b * P6_KEYS + part * KPP, v);
            tc5_fence_before();
            __syncwarp();
            if (lane == 0) p5_mbar_arrive(b_sfree + 8 * b);    // the scores are in registers
            if (kb + KPP > n_keys) {                            // the range ends inside (or before) this part of the tile
#pragma unroll
                for (int i = 0; i < KPP; ++i) v[i] = (kb + i < n_keys) ? v[i] : -INFINITY;
            }
            float mx = -INFINITY;
#pragma unroll
            for (int i = 0; i < KPP; ++i) mx = fmaxf(mx, v[i]);
            mx *= a.scale_log2;                                  // scale > 0: max commutes
            xch[part * 128 + r] = mx;
            asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(32 * D5_PARTS) : "memory");
#pragma unroll
            for (int p2 = 0; p2 < D5_PARTS; ++p2) mx = fmaxf(mx, xch[p2 * 128 + r]);
            asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(32 * D5_PARTS) : "memory");   // all have read before the next tile writes
            const bool grow = mx > m_used + 8.0f;                // also true on the first tile (m_used = -inf); a tile always
                                                                 // holds >= 1 key of the range in its first part, so mx is finite
            if (__any_sync(0xffffffffu, grow)) {
                const float m_new = grow ? mx : m_used;
                const float f = (j > 0) ? exp2f(m_used - m_new) : 1.0f;
                if (j > 0) {
                    p5_wait(b_of, (j - 1) & 1);                  // every product issued so far has landed in O
                    tc5_fence_after();
#pragma unroll 1
                    for (int c = 0; c < DPP / DCH; ++c) {
                        float w[DCH];
                        d5_ld<DCH>(trow + 128 + part * DPP + c * DCH, w);
#pragma unroll
                        for (int i = 0; i < DCH; ++i) w[i] *= f;
                        d5_st<DCH>(trow + 128 + part * DPP + c * DCH, w);
                    }
                    tc5_fence_before();
                    l_run *= f;
                }
                m_used = m_new;
            }
            const float neg_m = -m_used;
            uint32_t wh[KPP / 2], wl[KPP / 2];
            float lsum = 0.f;
#pragma unroll
            for (int e = 0; e < KPP / 2; ++e) {
                const float p0 = p6_ex2(fmaf(v[2 * e], a.scale_log2, neg_m)), p1 = p6_ex2(fmaf(v[2 * e + 1], a.scale_log2, neg_m));
                wh[e] = pack_bf16(p0, p1);
                wl[e] = pack_bf16(p0 - bf16lo(wh[e]), p1 - bf16hi(wh[e]));
                lsum += p0 + p1;
            }
            l_run += lsum;
            p5_wait(b_pfree + 8 * b, ((j >> 1) & 1) ^ 1);        // PV(j-2) has read this P buffer
            unsigned char* pb = p_row + b * P6_PBUF;
#pragma unroll
            for (int ch = 0; ch < KPP / 8; ++ch) {               // 8 keys = one 16-byte chunk of the 128-byte row
                const int off = ((part * (KPP / 8) + ch) ^ (r & 7)) << 4;
                *reinterpret_cast<uint4*>(pb + off) = make_uint4(wh[4 * ch], wh[4 * ch + 1], wh[4 * ch + 2], wh[4 * ch + 3]);
                *reinterpret_cast<uint4*>(pb + P5_HALF_BYTES + off) = make_uint4(wl[4 * ch], wl[4 * ch + 1], wl[4 * ch + 2], wl[4 * ch + 3]);
            }
            p5_fence_async_smem();
            __syncwarp();
            if (lane == 0) p5_mbar_arrive(b_pf + 8 * b);
            if (warp == 4 && lane == 0) RT_TR1(84);
        }
        if (n_tiles > 0 && warp_ok) {
            xch[part * 128 + r] = l_run;                         // row sum = the parts' partial sums
            asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(32 * D5_PARTS) : "memory");
            l_run = 0.f;
#pragma unroll
            for (int p2 = 0; p2 < D5_PARTS; ++p2) l_run += xch[p2 * 128 + r];
            p5_wait(b_of, (n_tiles - 1) & 1);
            tc5_fence_after();
            // un-normalised partial of this key range: O relative to m_used, the row sum l, for attn_merge2_kernel
            const size_t base = ((size_t)(row_ok ? t : 0) * a.n_q + (size_t)(it.h * a.G + g)) * a.max_chunks + it.pidx;
            float* po = a.part_o + base * HD + part * DPP;
#pragma unroll 1
            for (int c = 0; c < DPP / DCH; ++c) {
                float w[DCH];
                d5_ld<DCH>(trow + 128 + part * DPP + c * DCH, w);
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < DCH; i += 4) *reinterpret_cast<float4*>(po + c * DCH + i) = make_float4(w[i], w[i + 1], w[i + 2], w[i + 3]);
                }
            }
            if (row_ok && part == 0) *reinterpret_cast<float2*>(a.part_ml + base * 2) = make_float2(m_used, l_run);
            tc5_fence_before();
        } else if (n_tiles > 0) {
            p5_wait(b_of, (n_tiles - 1) & 1);                    // do not leave before the tensor pipe is done with O
        }
    }
    __syncthreads();
    RT_TR(86);
    if (warp == 1) {
        tc5_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
    }
}
