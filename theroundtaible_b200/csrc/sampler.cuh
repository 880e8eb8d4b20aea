// top-k / top-p (nucleus) sampling over stored logits: rt_sampling.top_k / top_p of include/rt_engine.h.
// No counterpart in the reference: its request carries no sampling field at all (src/adapters/local-llm.ts:162-165)
// and the external server applies its own defaults; SURVEY.md 8(b) lists both knobs in rt_sampling.
//
// Semantics (those of Hugging Face's TopKLogitsWarper followed by TopPLogitsWarper, with ties made order-free):
//   z_i = logit_i / temperature
//   top-k : tau_k = the k-th largest z; keep {z_i >= tau_k}                       (all ties at tau_k are kept)
//   top-p : over the kept set, the SMALLEST set of highest-probability tokens whose probability mass reaches
//           top_p of the kept mass; tau_p = its lowest z; keep {z_i >= tau_p}     (ties at tau_p kept)
//   draw  : arg-max over the kept i of z_i + Gumbel(seed, step, slot, i) -- the SAME noise stream the unfiltered
//           temperature sampler in gemv.cuh uses, so filters that keep everything reproduce its tokens exactly.
// One CTA per token slot; the logits row (<= 152k floats, L2-resident: the lm_head just wrote it) is walked once per
// pass: max, 4 radix passes for tau_k, 1 mass pass + 4 weighted radix passes for tau_p, 1 arg-max pass.  Probability
// mass is accumulated as 2^-40 fixed point in 64-bit integers, so every sum is exact and order-independent
// (bit-reproducible run to run, no float atomics).
#pragma once
#include "gemv.cuh"

struct TopkpArgs {
    const float* logits;      // [T][ld]
    int ld, V, T;
    float inv_temp;           // > 0
    int top_k;                // <= 0 or >= V: off
    float top_p;              // <= 0 or >= 1: off
    unsigned long long seed;
    const int* step_ptr;      // nullable => step_const
    int step_const;
    int* next_tok;            // [T]
    int* tok;                 // nullable: tok[t] = next
    int* pos;                 // nullable: pos[t] += 1
    int* out_tokens;          // nullable
    int max_out;
    int* step_w;              // nullable: incremented once per launch by the last CTA
    int* done_counter;
    int* kept_out;            // nullable [T]: size of the kept set (parity hook)
    float* thresh_out;        // nullable [T]: the final threshold as a logit (parity hook)
};

constexpr int TOPKP_THREADS = 1024;

// order-preserving map float -> uint32 (larger float <=> larger key)
__device__ __forceinline__ unsigned topkp_key(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float topkp_unkey(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Radix select from the top: the key whose cumulative weight (scanning keys downwards) first reaches `target`, among
// the elements with key >= floor_key.  weight = 1 (WEIGHTED false: the target-th largest) or the fixed-point mass.
template <bool WEIGHTED>
__device__ unsigned topkp_select(const float* __restrict__ row, int V, float m, float inv_temp, unsigned floor_key,
                                 unsigned long long target, unsigned long long* hist, unsigned* s_bcast) {
    unsigned prefix = 0, mask = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int b = threadIdx.x; b < 256; b += TOPKP_THREADS) hist[b] = 0ull;
        __syncthreads();
        for (int i = threadIdx.x; i < V; i += TOPKP_THREADS) {
            const float l = row[i];
            const unsigned key = topkp_key(l);
            if (key < floor_key || (key & mask) != prefix) continue;
            unsigned long long wgt = 1ull;
            if (WEIGHTED) wgt = (unsigned long long)(expf((l - m) * inv_temp) * 1099511627776.0f);
            atomicAdd(&hist[(key >> shift) & 255u], wgt);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long acc = 0ull;
            int b = 255;
            for (; b > 0; --b) {
                if (acc + hist[b] >= target) break;
                acc += hist[b];
            }
            target -= acc;                      // what is still to be covered inside bin b
            *s_bcast = (unsigned)b;
            reinterpret_cast<unsigned long long*>(s_bcast + 2)[0] = target;
        }
        __syncthreads();
        prefix |= *s_bcast << shift;
        mask |= 0xFFu << shift;
        target = reinterpret_cast<unsigned long long*>(s_bcast + 2)[0];
        __syncthreads();
    }
    return prefix;
}

static __global__ void __launch_bounds__(TOPKP_THREADS) topkp_sample_kernel(const TopkpArgs a) {
    griddep_launch();
    griddep_wait();
    __shared__ unsigned long long hist[256];
    __shared__ __align__(8) unsigned s_bcast[4];
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    __shared__ unsigned long long s_sum[32];
    const int t = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* row = a.logits + (size_t)t * a.ld;
    const int step = a.step_ptr ? *a.step_ptr : a.step_const;

    // ---- max ----
    float m = -INFINITY;
    for (int i = threadIdx.x; i < a.V; i += TOPKP_THREADS) m = fmaxf(m, row[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_val[warp] = m;
    __syncthreads();
    m = s_val[0];
    for (int w = 1; w < TOPKP_THREADS / 32; ++w) m = fmaxf(m, s_val[w]);
    __syncthreads();

    // ---- top-k threshold ----
    unsigned floor_key = 0u;
    if (a.top_k > 0 && a.top_k < a.V)
        floor_key = topkp_select<false>(row, a.V, m, a.inv_temp, 0u, (unsigned long long)a.top_k, hist, s_bcast);

    // ---- top-p threshold inside the top-k set ----
    if (a.top_p > 0.f && a.top_p < 1.f) {
        unsigned long long mass = 0ull;
        for (int i = threadIdx.x; i < a.V; i += TOPKP_THREADS) {
            const float l = row[i];
            if (topkp_key(l) >= floor_key) mass += (unsigned long long)(expf((l - m) * a.inv_temp) * 1099511627776.0f);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mass += __shfl_xor_sync(0xffffffffu, mass, o);
        if (lane == 0) s_sum[warp] = mass;
        __syncthreads();
        mass = 0ull;
        for (int w = 0; w < TOPKP_THREADS / 32; ++w) mass += s_sum[w];
        __syncthreads();
        unsigned long long target = (unsigned long long)((double)a.top_p * (double)mass);
        if (target < 1ull) target = 1ull;
        floor_key = topkp_select<true>(row, a.V, m, a.inv_temp, floor_key, target, hist, s_bcast);
    }

    // ---- Gumbel arg-max over the kept set (lowest index wins ties) ----
    float best = -INFINITY;
    int bi = 0x7fffffff, kept = 0;
    for (int i = threadIdx.x; i < a.V; i += TOPKP_THREADS) {
        const float l = row[i];
        if (topkp_key(l) < floor_key) continue;
        ++kept;
        const float v = l * a.inv_temp + gumbel_noise(a.seed, step, t, i);
        if (v > best) { best = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float v2 = __shfl_xor_sync(0xffffffffu, best, o);
        const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
        kept += __shfl_xor_sync(0xffffffffu, kept, o);
        if (v2 > best || (v2 == best && i2 < bi)) { best = v2; bi = i2; }
    }
    if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; s_sum[warp] = (unsigned long long)kept; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < TOPKP_THREADS / 32; ++w) {
            if (s_val[w] > best || (s_val[w] == best && s_idx[w] < bi)) { best = s_val[w]; bi = s_idx[w]; }
            kept += (int)s_sum[w];
        }
        if (bi == 0x7fffffff) bi = 0;
        a.next_tok[t] = bi;
        if (a.tok) a.tok[t] = bi;
        if (a.pos) a.pos[t] += 1;
        if (a.out_tokens && step < a.max_out) a.out_tokens[(size_t)t * a.max_out + step] = bi;
        if (a.kept_out) a.kept_out[t] = kept;
        if (a.thresh_out) a.thresh_out[t] = floor_key ? topkp_unkey(floor_key) : -INFINITY;
        if (a.done_counter) {
            __threadfence();
            if (atomicAdd(a.done_counter, 1) == (int)gridDim.x - 1) {       // every CTA has read `step` by now
                if (a.step_w) *a.step_w = step + 1;
                *a.done_counter = 0;
            }
        }
    }
}
