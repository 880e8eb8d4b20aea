// Shared helpers for the sm_100a kernels of the generation backend.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdexcept>
#include <string>

struct RtError : public std::runtime_error {
    int code;
    RtError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

#define RT_CUDA(expr)                                                                             \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            throw RtError(-2, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " +     \
                                  __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");    \
    } while (0)

#define RT_REQUIRE(cond, msg)                                                                     \
    do {                                                                                          \
        if (!(cond)) throw RtError(-1, std::string("invalid argument: ") + (msg));                \
    } while (0)

typedef __nv_bfloat16 bf16;
typedef __half f16;

__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- device helpers ----------------------------------------------------------------------------
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 t = __floats2bfloat162_rn(a, b);  // .x = a (low half), .y = b (high half)
    return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ uint16_t f2bf_bits(float a) {
    bf16 t = __float2bfloat16_rn(a);
    return *reinterpret_cast<uint16_t*>(&t);
}

// ---- the activation contract -----------------------------------------------------------------------------------
// Every projection input (x * norm weight, attention output, SwiGLU output) is rounded to fp16, saturating at
// +-65504, in BOTH phases; weights, Q, K and V are bf16; accumulation, residual stream and softmax are fp32
// (DESIGN.md 3 "numerics"; oracle/model.py `_rf16`).  With bf16 inputs two correct implementations drift 1.2e-2 of
// max |logit| apart on the 16-layer 1B model (rounding flips decorrelate), above the 1e-2 parity bar; fp16 gives 2.5e-3.
__device__ __forceinline__ float sat_f16(float v) { return v > 65504.f ? 65504.f : (v < -65504.f ? -65504.f : v); }
__device__ __forceinline__ uint32_t pack_f16(float a, float b) {
    __half2 t = __floats2half2_rn(sat_f16(a), sat_f16(b));   // .x = a (low half)
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ f16 to_f16(float a) { return __float2half_rn(sat_f16(a)); }
// The decode kernels multiply with mma.sync, whose operands must share one 16-bit format, and the weights are bf16:
// an fp16 value (11 significant bits) is fed as bf16 hi (8 bits, RNE) + bf16 lo (the remainder: <= 3 significant bits,
// exact), so the product with a bf16 weight is exact in the fp16 input.  The PRODUCER of an activation splits it once and
// stores both parts in the decode kernels' "XF" layout, which is the order the projection kernel's lanes fetch their
// MMA B fragments in (gemv.cuh): [slot group of 8][k-block of 32][hi | lo][slot in group][32 k] bf16, so the 8 slots x 32 k
// of one part are 512 contiguous bytes -- 4 full cache lines per warp load instead of 8 half lines of a row-major matrix.
__host__ __device__ inline size_t xf_index(int slot, int k, int K) {   // the hi part; the lo part is 256 elements further
    return (size_t)(slot >> 3) * ((size_t)K * 16) + (size_t)(k >> 5) * 512 + (size_t)((slot & 7) * 32 + (k & 31));
}
__host__ __device__ inline size_t xf_elems(int K) { return (size_t)K * 64; }      // 32 slots x 2 parts
__device__ __forceinline__ void split_f16x2(float a, float b, uint32_t& hi, uint32_t& lo) {
    const float ha = __half2float(__float2half_rn(sat_f16(a))), hb = __half2float(__float2half_rn(sat_f16(b)));
    hi = pack_bf16(ha, hb);
    lo = pack_bf16(ha - bf16lo(hi), hb - bf16hi(hi));
}
// `raw` (XF16 form): the fp16 value itself goes into the hi position and the lo position is not written -- what the
// projections read when their weights are the fp16 twins (gemv.cuh mode 2: one f16 x f16 MMA per fragment)
__device__ __forceinline__ void xf_store1(bf16* base, int K, int slot, int k, float v, int raw = 0) {
    const size_t i = xf_index(slot, k, K);
    if (raw) {
        reinterpret_cast<f16*>(base)[i] = to_f16(v);
        return;
    }
    const float h = __half2float(__float2half_rn(sat_f16(v)));
    const bf16 hi = __float2bfloat16_rn(h);
    base[i] = hi;
    base[i + 256] = __float2bfloat16_rn(h - __bfloat162float(hi));
}
__device__ __forceinline__ void xf_store4(bf16* base, int K, int slot, int k, const float (&y)[4], int raw = 0) {   // k % 4 == 0
    const size_t i = xf_index(slot, k, K);
    if (raw) {
        *reinterpret_cast<uint2*>(base + i) = make_uint2(pack_f16(y[0], y[1]), pack_f16(y[2], y[3]));
        return;
    }
    uint2 h, l;
    split_f16x2(y[0], y[1], h.x, l.x);
    split_f16x2(y[2], y[3], h.y, l.y);
    *reinterpret_cast<uint2*>(base + i) = h;
    *reinterpret_cast<uint2*>(base + i + 256) = l;
}

// streaming 16-byte load that does not pollute L1 (weights are read exactly once per step)
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// ---- optional in-kernel timeline (build with RT_NVCC_EXTRA=-DRT_TRACE; scripts/trace_timeline.py reads the dump) ----
#ifdef RT_TRACE
struct RtTraceRec { unsigned long long t; int tag; int cta; };
static __device__ RtTraceRec* g_rt_trace = nullptr;
static __device__ unsigned g_rt_trace_n = 0;
constexpr unsigned RT_TRACE_CAP = 1u << 22;
__device__ __forceinline__ void rt_trace(int tag) {
    if (threadIdx.x == 0 && g_rt_trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const unsigned i = atomicAdd(&g_rt_trace_n, 1u);
        if (i < RT_TRACE_CAP) g_rt_trace[i] = {t, tag, (int)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z))};
    }
}
#define RT_TR(tag) rt_trace(tag)
// a record from whichever single thread calls it (warp-specialised kernels: the role's leader lane)
__device__ __forceinline__ void rt_trace_one(int tag) {
    if (g_rt_trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const unsigned i = atomicAdd(&g_rt_trace_n, 1u);
        if (i < RT_TRACE_CAP) g_rt_trace[i] = {t, tag, (int)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z))};
    }
}
#define RT_TR1(tag) rt_trace_one(tag)
#else
#define RT_TR(tag)
#define RT_TR1(tag)
#endif

// ---- weight stream pipelining through L2 -----------------------------------------------------------------------
// A decode step is a chain of short HBM-bound kernels; between and inside some of them (attention on a short
// context, the split-KV merge, every kernel's tail) HBM idles.  A kernel may therefore be handed a region of a LATER
// kernel's weights: its CTAs each ask the memory system to pull one slice of it into the 126 MB L2
// (cp.async.bulk.prefetch.L2: fire and forget, no destination, no completion to wait on), so that the weight stream
// keeps flowing and the later kernel starts on L2 hits.
struct L2Prefetch {
    const void* ptr;
    unsigned long long bytes;      // 0 = none
};
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// called by ONE thread of CTA `cta` of `n_cta`
__device__ __forceinline__ void l2_prefetch_slice(const L2Prefetch& pf, int cta, int n_cta) {
    if (pf.bytes == 0ull) return;
    const unsigned long long per = ((pf.bytes + n_cta - 1) / n_cta + 127ull) & ~127ull;
    const unsigned long long off = per * (unsigned long long)cta;
    if (off >= pf.bytes) return;
    unsigned long long n = pf.bytes - off < per ? pf.bytes - off : per;
    n &= ~15ull;
    const char* base = reinterpret_cast<const char*>(pf.ptr) + off;
    for (unsigned long long o = 0; o < n; o += 32768ull) l2_prefetch_bulk(base + o, (unsigned)(n - o < 32768ull ? n - o : 32768ull));
}

// mma.sync m16n8k16 bf16 x bf16 -> fp32 (legacy tensor path; used only where the batch dimension is
// 8..32 wide and the kernel is HBM-bound: decode GEMV and decode attention)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// the same shape with fp16 operands (decode projections over the fp16 weight twins)
__device__ __forceinline__ void mma_f16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                              uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(saddr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(saddr));
}

__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- thread-block clusters: rank, distributed shared memory, cluster barrier ----------------------------------------
__device__ __forceinline__ uint32_t cl_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cl_mapa(uint32_t addr, uint32_t rank) {      // this CTA's shared address in CTA `rank`
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cl_st_f32(uint32_t addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void cl_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
