// Shared pieces of the tcgen05 prefill attention (attn_prefill_tc5b.cuh): argument block, TMA / mbarrier / TMEM wrappers.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external server);
// SURVEY.md 8(a-K) row "prefill causal attention".  The first kernel built on these (128-key tiles, one P buffer:
// 457 us per config-3 launch against ~740 us for the mma.sync kernel, softmax serialised with the PV product) was
// superseded by the pipelined 64-key version and is no longer compiled (round-1 history has it).
//
// One CTA = up to 128 query rows of one kv head: BT = 128 / G consecutive tokens of a segment x the G q-heads of the group
// (row = token * G + head-in-group), walking the keys 0 .. position of its last token.  TMEM lane = query row; softmax with
// a LAZY maximum (O in TMEM is rescaled only when the row maximum grew by more than 2^8); P as bf16 hi + lo in shared
// memory in the 128-byte-swizzled K-major layout; V is the MN-major B operand as it lies in the paged pool.
#pragma once
#include "gemm_tc5.cuh"

constexpr int P5_KEYS = 128;                         // keys per tile
constexpr int P5_HALF_BYTES = 128 * 128;             // 128 rows x 64 bf16 (one 128-byte swizzle row each)
constexpr int P5_P_BYTES = 2 * P5_HALF_BYTES;        // 32 KB (128 rows x 128 keys)
// head_dim 128: Q 32 KB, a K/V stage 64 KB;  head_dim 64: Q 16 KB, a stage 32 KB
constexpr int P5_THREADS = 320;                       // TMA warp, MMA warp, 8 softmax warps

struct Pf5Args {
    const int* page_table;
    int pt_stride;
    const int* tok_pos;     // [S]
    const int4* work;       // {first token (flat), n tokens (<= 128/G), sequence, kv head}
    int n_q, n_kv, G, BT;
    int n_pages_pool;       // pages per layer in the pool
    int layer;              // layer index (row offset into the pool-wide K/V tensor maps)
    float scale_log2;
    f16* out;               // [S][n_q*128] fp16
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

