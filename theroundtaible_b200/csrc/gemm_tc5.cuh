// Prefill projection GEMM on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), sm_100a only.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external
// server); SURVEY.md 8(a-K) row "prefill GEMM".
//
//   C[S][N] = A[S][K] * W[N][K]^T      A: fp16 activations, row-major, S tokens;  W: fp16 twin of the bf16 weights, CM layout
//
// A persistent CTA (one per SM) walks 128 x BN tiles (tokens x weight rows), token blocks fastest, with two accumulators in TMEM
// so that a tile's epilogue runs under the next tile's MMAs.  Per tile:
//   warp 0   TMA producer: per 64-wide k-block one 2-D tensor-map load of the A tile (SWIZZLE_128B, K-major) and
//            one 4-D tensor-map load of the W tile.  The CM layout of gemv.cuh stores 8-row x 16-byte core matrices
//            contiguously, and the 4-D box (256 elems, 2 k-blocks, 2 halves, BN/16 tiles) lands them in shared memory
//            as the canonical NO-SWIZZLE K-major operand: 8-row groups 1024 B apart (SBO), k-chunks 128 B apart (LBO).
//            The decode path streams the very same bytes; the weights exist once in HBM.
//   warp 1   tcgen05.mma issuer (one elected lane): 4 x (M=128, N=BN, K=16) MMAs per k-block into a TMEM accumulator;
//            tcgen05.commit releases the shared-memory stage / signals the epilogue.  Also owns TMEM alloc/dealloc.
//   warps 2-5 epilogue: tcgen05.ld 32x32b.x32 (thread = token row, 32 consecutive weight rows), fused
//            bias + RoPE + KV-page write | residual add | SwiGLU, 16-byte stores.
// 4-stage mbarrier pipeline (full/empty), 48 KB per stage at BN = 256; it keeps running across tile boundaries.  Accumulator
// hand-over: tmem_full[2] (tcgen05.commit) / tmem_free[2] (one arrival per epilogue warp).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "gemv.cuh"
#include "prefill.cuh"

constexpr int G5_BM = 128, G5_BK = 64, G5_STAGES = 4;
// ---- PTX wrappers -----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
                 "l"(map), "r"(bar), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tc5_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc5_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc5_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc5_mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc5_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout type [61,64): 0 = no swizzle, 2 = 128-byte swizzle
__device__ __forceinline__ uint64_t tc5_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46) | ((uint64_t)layout << 61);
}

template <int BN, int EPI>
__global__ void __launch_bounds__(192, 1) gemm_tc5_kernel(const __grid_constant__ CUtensorMap map_a,
                                                          const __grid_constant__ CUtensorMap map_w, const Gemm5Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    constexpr int A_BYTES = G5_BM * G5_BK * 2;   // 16 KB
    constexpr int B_BYTES = BN * G5_BK * 2;      // 32 KB at BN = 256
    constexpr int STAGE = A_BYTES + B_BYTES;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t bar_full = sbase + G5_STAGES * STAGE;        // 8 bytes each
    const uint32_t bar_empty = bar_full + 8 * G5_STAGES;
    const uint32_t bar_tmem = bar_empty + 8 * G5_STAGES;          // accumulator full [2]
    const uint32_t bar_tfree = bar_tmem + 16;                     // accumulator drained by the epilogue [2]
    const uint32_t tmem_slot = bar_tfree + 16;
    volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - smem_u32(smem)));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ... in token-block-fastest order, with TWO accumulators in
    // TMEM (2 x BN columns) so that the epilogue of a tile (TMEM loads, fused math, global stores) runs under the MMAs of the
    // next one, and the TMA ring never drains between tiles.
    const int m_blocks = (a.S + G5_BM - 1) / G5_BM, n_blocks = (a.N + BN - 1) / BN;
    const int n_tiles = m_blocks * n_blocks;
    const int num_kb = a.K / G5_BK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < G5_STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < 2; ++s) { mbar_init(bar_tmem + 8 * s, 1); mbar_init(bar_tfree + 8 * s, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(2 * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    const uint32_t tmem_d = *tmem_slot_p;

    griddep_launch();
    griddep_wait();      // activations (A) come from the previous kernel

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % G5_STAGES;
                    const uint32_t ph = (it / G5_STAGES) & 1;
                    mbar_wait(bar_empty + 8 * s, ph ^ 1);
                    mbar_expect_tx(bar_full + 8 * s, STAGE);
                    tma_load_2d(sbase + s * STAGE, &map_a, bar_full + 8 * s, kb * G5_BK, m_blk * G5_BM);
                    tma_load_4d(sbase + s * STAGE + A_BYTES, &map_w, bar_full + 8 * s, 0, kb * 2, 0, n_blk * (BN / 16));
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D=f32 [4,6)=1, A=bf16 [7,10)=1, B=bf16 [10,13)=1, K-major both, N>>3 [17,23), M>>4 [24,29)
            // A/B format 0 = f16, 1 = bf16.  kind::f16 wants ONE format for both operands (an f16 A with a bf16 B raised an
            // illegal-instruction fault on B200), so the engine multiplies fp16 activations by the fp16 twin of the weights
            const uint32_t fmt = a.ab_f16 ? 0u : 1u;
            const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(G5_BM >> 4) << 24);
            int it = 0, lt = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
                const int acc = lt & 1;
                mbar_wait(bar_tfree + 8 * acc, ((lt >> 1) & 1) ^ 1);          // the epilogue has drained this accumulator
                tc5_fence_after();
                const uint32_t d = tmem_d + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % G5_STAGES;
                    const uint32_t ph = (it / G5_STAGES) & 1;
                    mbar_wait(bar_full + 8 * s, ph);
                    tc5_fence_after();
                    const uint32_t sa = sbase + s * STAGE, sb = sa + A_BYTES;
#pragma unroll
                    for (int j = 0; j < G5_BK / 16; ++j) {
                        const uint64_t adesc = tc5_smem_desc(sa + j * 32, 16, 1024, 2);      // 128B swizzle, K-major
                        const uint64_t bdesc = tc5_smem_desc(sb + j * 256, 128, 1024, 0);    // no swizzle, K-major (CM)
                        tc5_mma_bf16(d, adesc, bdesc, idesc, (kb > 0 || j > 0) ? 1u : 0u);
                    }
                    tc5_commit(bar_empty + 8 * s);      // frees the stage once these MMAs have read it
                }
                tc5_commit(bar_tmem + 8 * acc);         // accumulator complete
            }
        }
    } else {
        // ---- epilogue: warps 2..5 own TMEM lane quadrants (warp % 4) ----
        const int quad = warp & 3;
        int lt = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
        const int m_blk = tile % m_blocks, n_blk = tile / m_blocks;
        const int acc = lt & 1;
        const int m = m_blk * G5_BM + quad * 32 + lane;
        mbar_wait(bar_tmem + 8 * acc, (lt >> 1) & 1);
        tc5_fence_after();
        const bool row_ok = m < a.S;
        int pos = 0, seq = 0;
        if (EPI == G5_EPI_QKV && row_ok) {
            pos = a.tok_pos[m];
            seq = a.tok_seq[m];
        }
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            float v[32];
            tc5_ld32(tmem_d + ((uint32_t)(quad * 32) << 16) + acc * BN + c * 32, v);
            const int n0 = n_blk * BN + c * 32;
            if (!row_ok || n0 >= a.N) continue;
            if (a.row_scale) {
                const float rs = a.row_scale[m];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] *= rs;
            }
            if (EPI == G5_EPI_RESID || EPI == G5_EPI_STORE) {
                float4* o = reinterpret_cast<float4*>(a.out + (size_t)m * a.ldo + n0);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float4 x = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                    if (EPI == G5_EPI_RESID) {
                        const float4 p = o[i];
                        x.x += p.x; x.y += p.y; x.z += p.z; x.w += p.w;
                    }
                    o[i] = x;
                }
            } else if (EPI == G5_EPI_STORE16) {
                uint4* o = reinterpret_cast<uint4*>(a.out_bf + (size_t)m * a.ldob + n0);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 p;
                    p.x = pack_f16(v[8 * i], v[8 * i + 1]); p.y = pack_f16(v[8 * i + 2], v[8 * i + 3]);
                    p.z = pack_f16(v[8 * i + 4], v[8 * i + 5]); p.w = pack_f16(v[8 * i + 6], v[8 * i + 7]);
                    o[i] = p;
                }
            } else if (EPI == G5_EPI_SWIGLU) {
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    float act[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const float gt = v[t * 16 + r], up = v[t * 16 + 8 + r];
                        act[r] = gt / (1.0f + expf(-gt)) * up;
                    }
                    uint4 p;
                    p.x = pack_f16(act[0], act[1]); p.y = pack_f16(act[2], act[3]);
                    p.z = pack_f16(act[4], act[5]); p.w = pack_f16(act[6], act[7]);
                    *reinterpret_cast<uint4*>(a.out_bf + (size_t)m * a.ldob + (n0 >> 1) + t * 8) = p;
                }
            } else {  // G5_EPI_QKV
                const int half = a.hd >> 1;
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    const int n = n0 + t * 16;
                    const int head = n / a.hd;
                    const int d0 = ((n % a.hd) >> 4) * 8;
                    float y1[8], y2[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        float x1 = v[t * 16 + r], x2 = v[t * 16 + 8 + r];
                        if (a.bias) {
                            x1 += a.bias[n + r];
                            x2 += a.bias[n + 8 + r];
                        }
                        if (head < a.n_q + a.n_kv) {
                            const float cs = a.cos_t[(size_t)pos * half + d0 + r], sn = a.sin_t[(size_t)pos * half + d0 + r];
                            y1[r] = x1 * cs - x2 * sn;
                            y2[r] = x2 * cs + x1 * sn;
                        } else {
                            y1[r] = x1;
                            y2[r] = x2;
                        }
                    }
                    bf16* dst;
                    if (head < a.n_q) {
                        dst = a.Q + (size_t)m * (a.n_q * a.hd) + (size_t)head * a.hd;
                    } else {
                        const bool is_k = head < a.n_q + a.n_kv;
                        const int kvh = head - a.n_q - (is_k ? 0 : a.n_kv);
                        const int page = a.page_table[(size_t)seq * a.pt_stride + pos / KV_PAGE];
                        dst = (is_k ? a.Kp : a.Vp) + (((size_t)page * a.n_kv + kvh) * KV_PAGE + (pos % KV_PAGE)) * a.hd;
                    }
                    uint4 p1, p2;
                    p1.x = pack_bf16(y1[0], y1[1]); p1.y = pack_bf16(y1[2], y1[3]);
                    p1.z = pack_bf16(y1[4], y1[5]); p1.w = pack_bf16(y1[6], y1[7]);
                    p2.x = pack_bf16(y2[0], y2[1]); p2.y = pack_bf16(y2[2], y2[3]);
                    p2.z = pack_bf16(y2[4], y2[5]); p2.w = pack_bf16(y2[6], y2[7]);
                    *reinterpret_cast<uint4*>(dst + d0) = p1;
                    *reinterpret_cast<uint4*>(dst + d0 + half) = p2;
                }
            }
        }
        tc5_fence_before();
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_tfree + 8 * acc) : "memory");
        }   // tiles
    }
    __syncthreads();
    if (warp == 1) {
        tc5_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(2 * BN));
    }
}

constexpr size_t gemm5_smem_bytes(int BN) { return (size_t)G5_STAGES * (G5_BM * G5_BK * 2 + BN * G5_BK * 2) + 1024 + 256; }
