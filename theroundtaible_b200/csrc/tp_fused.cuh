// Fused tensor-parallel all-reduce + residual add + RMSNorm over NVLink peer memory (decode path).
//
// No reference counterpart (no collective exists in /root/reference); SURVEY.md 8(a-K) last row: the two
// per-layer all-reduces of [T][H] are LATENCY-bound (128-256 KB), so they are done one-shot inside the kernel that also
// needs the result: every rank pushes nothing, it just publishes "my partial is ready" and then PULLS all partials.
//
//   rank r, CTA t (one per token slot):
//     e = ++epoch[t]                                  (private to this CTA: graph replays need no host-side counter)
//     peers[p].flags[t][r] = e          for all p     (st.release.sys over NVLink)
//     wait local flags[t][p] >= e       for all p     (ld.acquire.sys)
//     x[t][:] += sum_p peers[p].partial[t][:]         (P2P loads, summed in rank order => identical on every rank)
//     xw[t][:] = bf16(x * w),  inv_rms[t] = rsqrt(mean x^2 + eps)     (split RMSNorm: the scale is applied by the
//                                                                       next projection's epilogue, gemv.cuh)
//
// The partial buffers are per call site (O-projection / down-projection), so a buffer is rewritten one full layer
// later; by then every peer has passed the next all-reduce's wait, i.e. finished reading it -- no trailing barrier.
// Each rank runs this kernel on its own GPU (never two ranks on one device).
#pragma once
#include "gemv.cuh"

constexpr int TPF_MAX = 8;

struct TpFusedArgs {
    const float* partial[TPF_MAX];   // every rank's partial-sum buffer for this call site (peer-mapped), [T][H]
    int* flags[TPF_MAX];             // every rank's flag array [32 slots][TPF_MAX]
    int* epoch;                      // local [32]
    int rank, tp, H;
    float* x;                        // local residual stream [T][H]
    const bf16* norm_w;
    bf16* xn;                        // local [T][H]: xw = bf16(x * w) (split RMSNorm, gemv.cuh)
    float* inv_rms;                  // local [T]: rsqrt(mean x^2 + eps) for the consuming projection's epilogue
    float eps;
};

__device__ __forceinline__ void tpf_st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int tpf_ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int NTHR>
__global__ void __launch_bounds__(NTHR) tp_fused_allreduce_norm_kernel(const TpFusedArgs a) {
    __shared__ float sred[NTHR / 32];
    griddep_launch();
    griddep_wait();                         // this rank's partial comes from the previous kernel
    const int t = blockIdx.x;
    int e = 0;
    if (threadIdx.x == 0) {
        e = a.epoch[t] + 1;
        a.epoch[t] = e;
    }
    e = __shfl_sync(0xffffffffu, e, 0);     // warp 0 only; the other warps pick it up after the barrier below
    if (threadIdx.x < a.tp) {
        __threadfence_system();
        tpf_st_release_sys(a.flags[threadIdx.x] + t * TPF_MAX + a.rank, e);
        while (tpf_ld_acquire_sys(a.flags[a.rank] + t * TPF_MAX + threadIdx.x) < e) {}
    }
    __syncthreads();

    const int n4 = a.H >> 2;
    constexpr int PER = 4;                  // float4 per thread held in registers (H <= 16 * NTHR)
    float4 v[PER];
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int c4 = threadIdx.x + k * NTHR;
        if (c4 < n4) {
            float4 acc = *(reinterpret_cast<const float4*>(a.x + (size_t)t * a.H) + c4);
            for (int p = 0; p < a.tp; ++p) {
                const float4 d = __ldcv(reinterpret_cast<const float4*>(a.partial[p] + (size_t)t * a.H) + c4);
                acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
            }
            v[k] = acc;
            *(reinterpret_cast<float4*>(a.x + (size_t)t * a.H) + c4) = acc;
            ss += acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
        } else {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int w = 0; w < NTHR / 32; ++w) tot += sred[w];
        a.inv_rms[t] = 1.0f / sqrtf(tot / (float)a.H + a.eps);
    }
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int c4 = threadIdx.x + k * NTHR;
        if (c4 < n4) {
            const uint2 wv = *reinterpret_cast<const uint2*>(a.norm_w + c4 * 4);
            uint2 o;
            o.x = pack_bf16(v[k].x * bf16lo(wv.x), v[k].y * bf16hi(wv.x));
            o.y = pack_bf16(v[k].z * bf16lo(wv.y), v[k].w * bf16hi(wv.y));
            *reinterpret_cast<uint2*>(a.xn + (size_t)t * a.H + c4 * 4) = o;
        }
    }
}
