// Fused tensor-parallel all-reduce + residual add + RMSNorm over NVLink peer memory (decode path).  Two forms: the PULL form
// below (any slot count; used for steps of more than 8 slots and with RT_TP_PUSH=0) and the PUSH form at the end of this
// file (steps of <= 8 slots: the projection delivers its partial sums itself, the receiver reads local memory only).
//
// No reference counterpart (no collective exists in /root/reference); SURVEY.md 8(a-K) last row: the two
// per-layer all-reduces of [T][H] are LATENCY-bound (128-256 KB), so they are done one-shot inside the kernel that also
// needs the result: every rank pushes nothing, it just publishes "my partial is ready" and then PULLS all partials.
//
//   rank r, CTA (c, t): column block c (NTHR float4 = 4*NTHR columns) of token slot t
//     e = ++epoch[t][c]                               (private to this CTA: graph replays need no host-side counter)
//     peers[p].flags[t][c][r] = e       for all p     (st.release.sys over NVLink)
//     wait local flags[t][c][p] >= e    for all p     (ld.acquire.sys)
//     x[t][cols] += sum_p peers[p].partial[t][cols]   (ONE float4 per thread and peer, all peers' P2P loads in flight at
//                                                      once, summed in rank order => identical on every rank)
//     xw[t][cols] = fp16(x * w);  ss_part[c][t] = sum of x^2 over the block     (split RMSNorm: the consuming projection
//                                                      adds the blocks and scales its accumulators, gemv.cuh)
//   Several CTAs per slot and one load per thread and peer keep the exchange at ~one NVLink round trip; the first
//   version (one CTA per slot, a serial loop over peers and four column groups) cost tp x 4 of them.
//
// The partial buffers are per call site (O-projection / down-projection), so a buffer is rewritten one full layer
// later; by then every peer has passed the next all-reduce's wait, i.e. finished reading it -- no trailing barrier.
// Each rank runs this kernel on its own GPU (never two ranks on one device).
#pragma once
#include "kernels.cuh"

constexpr int TPF_MAX = 8;
constexpr int TPF_CMAX = 16;         // column blocks per slot (H <= 4 * TPF_THREADS * TPF_CMAX)
constexpr int TPF_THREADS = 256;

struct TpFusedArgs {
    const float* partial[TPF_MAX];   // every rank's partial-sum buffer for this call site (peer-mapped), [T][H]
    int* flags[TPF_MAX];             // every rank's flag array [32 slots][TPF_CMAX][TPF_MAX]
    int* epoch;                      // local [32][TPF_CMAX]
    int rank, tp, H;
    float* x;                        // local residual stream [T][H]
    const bf16* norm_w;
    bf16* xn;                        // local, XF layout over H columns: xw = fp16(x * w) as bf16 hi + lo (gemv.cuh)
    float* ss_part;                  // local [gridDim.x][32]: per-column-block sums of x^2 for the consuming projection
    int raw;                         // 1: xn in the XF16 form (common.cuh xf_store4)
};

__device__ __forceinline__ void tpf_st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void tpf_st_relaxed_sys(int* p, int v) { asm volatile("st.relaxed.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int tpf_ld_relaxed_sys(const int* p) {
    int v;
    asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int tpf_ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(TPF_THREADS) tp_fused_allreduce_norm_kernel(const TpFusedArgs a) {
    __shared__ float sred[TPF_THREADS / 32];
    RT_TR(70);
    griddep_launch();
    griddep_wait();                         // this rank's partial comes from the previous kernel
    RT_TR(71);
    const int c = blockIdx.x, t = blockIdx.y;
    const int cell = t * TPF_CMAX + c;
    if (threadIdx.x < 32) {
        int e = 0;
        if (threadIdx.x == 0) {
            e = a.epoch[cell] + 1;
            a.epoch[cell] = e;
        }
        e = __shfl_sync(0xffffffffu, e, 0);
        if ((int)threadIdx.x < a.tp) {
            // "my partial is ready": it was written by the PREVIOUS kernel on this GPU (complete and visible in this GPU's
            // L2 once griddepcontrol.wait has returned), and the peers PULL it through that L2 -- so the flag needs no
            // system-scope release fence in front of it (the fence + release store + acquire polls cost ~4 us of the
            // 6-8 us this wait took in the in-kernel timeline, profiles/r2_summary.md).  The poll is relaxed; one acquire
            // load after it orders the pulls below behind the flag.
            tpf_st_relaxed_sys(a.flags[threadIdx.x] + cell * TPF_MAX + a.rank, e);
            const int* mine = a.flags[a.rank] + cell * TPF_MAX + threadIdx.x;
            // bounded: a peer that never arrives (a rank that died or left the step) ends this kernel with a trap -- an
            // error on this rank's stream -- instead of a GPU that spins for ever
            long long spins = 0;
            while (tpf_ld_relaxed_sys(mine) < e) {
                if (++spins > (1ll << 28)) {
                    printf("tp_fused_allreduce_norm: rank %d still waits for rank %d at epoch %d\n", a.rank, (int)threadIdx.x, e);
                    __trap();
                }
            }
            (void)tpf_ld_acquire_sys(mine);
        }
    }
    __syncthreads();
    RT_TR(72);                              // every peer's flag seen

    const int n4 = a.H >> 2;
    const int c4 = c * TPF_THREADS + threadIdx.x;
    float ss = 0.f;
    if (c4 < n4) {
        float4 d[TPF_MAX];
#pragma unroll
        for (int p = 0; p < TPF_MAX; ++p)
            if (p < a.tp) d[p] = __ldcv(reinterpret_cast<const float4*>(a.partial[p] + (size_t)t * a.H) + c4);
        float4 acc = *(reinterpret_cast<const float4*>(a.x + (size_t)t * a.H) + c4);
        const uint2 wv = *reinterpret_cast<const uint2*>(a.norm_w + c4 * 4);
#pragma unroll
        for (int p = 0; p < TPF_MAX; ++p)
            if (p < a.tp) { acc.x += d[p].x; acc.y += d[p].y; acc.z += d[p].z; acc.w += d[p].w; }
        *(reinterpret_cast<float4*>(a.x + (size_t)t * a.H) + c4) = acc;
        ss = acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
        const float y[4] = {acc.x * bf16lo(wv.x), acc.y * bf16hi(wv.x), acc.z * bf16lo(wv.y), acc.w * bf16hi(wv.y)};
        xf_store4(a.xn, a.H, t, c4 * 4, y, a.raw);
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int w = 0; w < TPF_THREADS / 32; ++w) tot += sred[w];
        a.ss_part[c * 32 + t] = tot;
    }
    RT_TR(74);
}

// =====================================================================================================================
// Push form of the same exchange (<= 8 slots): no flag, no pull.
//
// The row-parallel projection (gemv.cuh EPI_PUSH) stores its partial sums straight into EVERY rank's receive buffer
// recv[src rank][slot][H] -- remote stores over NVLink, fire and forget -- as 16-byte packets whose four fp32 words carry
// the parity of the slot's epoch in their lowest mantissa bit (a relative error of 2^-24, far below the fp16 rounding of the
// next projection's input).  The receiver reads only its OWN GPU's memory and takes a packet the moment its four parities
// match: behind the slowest rank's last tile there is one one-way NVLink flight and a local read, instead of a flag flight
// plus a pull round trip.
//
// Epochs.  A receiver CTA counts its own runs (private cell, like the pull kernel).  The SENDER's epoch of site s and slot t
// is advanced by CTA (0, t) of the OTHER site's receiver at its end: every kernel of the step waits for its predecessor
// (griddepcontrol.wait), so by then all CTAs of site s's last projection have read the old value, and site s's next projection
// reads the new one behind its own wait.  The two sites alternate strictly (one O-projection and one down-projection per
// layer with the same slots), which keeps both counts in step; site 0 starts at 1, site 1 at 0 (advanced to 1 before its
// first use); the buffers start zeroed = parity 0 = "not yet".
// Reuse: a rank can be ahead of a peer by less than one exchange of the same site (it needs the peer's packets of the site in
// between), so a packet is never overwritten before it was read -- the argument of the pull kernel.
// The receiver does wait for its own projection: a receiver that started early and spun beside it would also have to be
// ordered against the previous receiver of the same site, which nothing guarantees once the kernels of several layers fit on
// the GPU at once (seen with the 2-layer test models: both instances read the same run count).
// A rank that never delivers ends the spin with a trap.
struct TpLLArgs {
    const float* recv;               // this rank's receive buffer of the site: [tp][32][H]
    int* cell_epoch;                 // private [32][TPF_CMAX] run counters of this site's receiver CTAs
    int* next_site_epoch;            // the OTHER site's sender epochs [32]: CTA (0, t) advances entry t at its end
    int rank, tp, H;
    float* x;
    const bf16* norm_w;
    bf16* xn;
    float* ss_part;
    int raw;
};

__device__ __forceinline__ uint4 tpf_ld_relaxed_sys_v4(const float* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(TPF_THREADS) tp_ll_reduce_norm_kernel(const TpLLArgs a) {
    __shared__ float sred[TPF_THREADS / 32];
    __shared__ int s_par;
    RT_TR(70);
    griddep_launch();
    griddep_wait();
    const int c = blockIdx.x, t = blockIdx.y;
    const int cell = t * TPF_CMAX + c;
    if (threadIdx.x == 0) {
        const int e = a.cell_epoch[cell] + 1;
        a.cell_epoch[cell] = e;
        s_par = e & 1;
    }
    __syncthreads();
    const uint32_t par = (uint32_t)s_par;
    const int n4 = a.H >> 2;
    const int c4 = c * TPF_THREADS + threadIdx.x;
    const bool active = c4 < n4;
    const float* mine = a.recv + (size_t)t * a.H + (size_t)c4 * 4;        // + src * 32 * H
    const size_t src_stride = (size_t)32 * a.H;
    auto ok = [par](const uint4& v) { return ((v.x & 1u) == par) & ((v.y & 1u) == par) & ((v.z & 1u) == par) & ((v.w & 1u) == par); };

    RT_TR(71);

    float ss = 0.f;
    if (active) {
        uint4 d[TPF_MAX];
        uint32_t pending = (1u << a.tp) - 1u;
        long long spins = 0;
        while (pending) {
#pragma unroll
            for (int p = 0; p < TPF_MAX; ++p)          // all loads of a round in flight together
                if (p < a.tp && (pending >> p & 1u)) d[p] = tpf_ld_relaxed_sys_v4(mine + (size_t)p * src_stride);
#pragma unroll
            for (int p = 0; p < TPF_MAX; ++p)
                if (p < a.tp && (pending >> p & 1u) && ok(d[p])) pending &= ~(1u << p);
            if (pending && ++spins > (1ll << 28)) {
                printf("tp_ll_reduce_norm: rank %d misses packets (mask %x, slot %d, block %d)\n", a.rank, pending, t, c);
                __trap();
            }
        }
        float4 acc = *(reinterpret_cast<const float4*>(a.x + (size_t)t * a.H) + c4);
        const uint2 wv = *reinterpret_cast<const uint2*>(a.norm_w + c4 * 4);
#pragma unroll
        for (int p = 0; p < TPF_MAX; ++p)          // rank order: identical sums on every rank
            if (p < a.tp) {
                acc.x += __uint_as_float(d[p].x); acc.y += __uint_as_float(d[p].y);
                acc.z += __uint_as_float(d[p].z); acc.w += __uint_as_float(d[p].w);
            }
        *(reinterpret_cast<float4*>(a.x + (size_t)t * a.H) + c4) = acc;
        ss = acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
        const float y[4] = {acc.x * bf16lo(wv.x), acc.y * bf16hi(wv.x), acc.z * bf16lo(wv.y), acc.w * bf16hi(wv.y)};
        xf_store4(a.xn, a.H, t, c4 * 4, y, a.raw);
    }
    RT_TR(72);
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int w = 0; w < TPF_THREADS / 32; ++w) tot += sred[w];
        a.ss_part[c * 32 + t] = tot;
        if (c == 0) a.next_site_epoch[t] += 1;      // the other site's senders read it behind their own wait
    }
    RT_TR(74);
}
