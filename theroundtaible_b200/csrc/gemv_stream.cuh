// Decode projection kernel (v3): persistent, TMA-fed, stream-K.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an external
// server); SURVEY.md 8(a-K) rows "decode projections" and "lm_head".
//
// The CM weight layout (gemv.cuh) makes a matrix one long run of 1 KB units (16 rows x 32 k), tile-major.  The grid is
// one CTA per SM; CTA c owns the CONTIGUOUS unit range [c*U/G, (c+1)*U/G) -- equal bytes for every SM whatever the
// matrix shape (no wave quantisation).  Inside a CTA
//   * warp 8 is the producer: one lane streams the CTA's byte range through an 8-stage, 16 KB-per-stage shared-memory
//     ring with cp.async.bulk (TMA, 1-D) + mbarriers.  It starts BEFORE griddepcontrol.wait: weights do not depend on
//     the previous kernel, so with PDL the ring (128 KB per SM, 19 MB per chip) is already full when the activations
//     arrive, and the HBM stream keeps running while consumers reduce, run epilogues or wait for a neighbour;
//   * warps 0..7 consume: per 16-row tile they split the tile's units, LDS.128 the mma.sync fragments (conflict-free:
//     a half-unit is 32 lanes x 16 B), accumulate, add the 8 partial sums in fixed warp order;
//   * a tile cut by a CTA boundary is finished by the CTA holding its FIRST unit: the later CTAs publish the partial of
//     their first tile (16 x slots fp32) to scratch as soon as they have it, the owner adds them in CTA order at the
//     end of its own range -- deterministic, and nobody waits on a CTA that still has most of its work ahead.
//     All CTAs are co-resident (grid <= SMs, 1 CTA/SM), so the owner's short spin cannot deadlock.
// Epilogues are those of gemv.cuh (QKV+RoPE+KV pages, residual + RMSNorm tail, SwiGLU, logits + sampler tail).
#pragma once
#include "gemv.cuh"

constexpr int GS_CONSUMERS = 8;
constexpr int GS_THREADS = (GS_CONSUMERS + 1) * 32;
constexpr int GS_CHUNK_UNITS = 16;                 // 16 KB per stage
constexpr int GS_STAGES = 8;                       // 128 KB ring
constexpr int GS_RING_BYTES = GS_STAGES * GS_CHUNK_UNITS * 1024;

struct GemvStreamFix {
    float* scratch;     // [grid][16][32]  partial tile of a CTA that does not own its last tile
    int* flags;         // [grid]  1 while a published partial waits for its owner; the owner clears it (so the same
                        //         buffers serve every launch, also when a CUDA graph replays identical arguments)
};

__device__ __forceinline__ void gs_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void gs_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gs_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void gs_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "GS_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra GS_DONE;\n\t"
        "bra GS_WAIT;\n\t"
        "GS_DONE:\n\t"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void gs_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void gs_csync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }   // the 8 consumer warps
__device__ __forceinline__ int gs_ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void gs_st_release(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

__device__ __forceinline__ long long gs_u0(int c, long long U, int G) { return (long long)c * U / G; }

template <int NT, int EPI>
__global__ void __launch_bounds__(GS_THREADS, 1) gemv_stream_kernel(const GemvArgs a, const GemvStreamFix fx) {
    extern __shared__ __align__(1024) unsigned char gs_smem[];
    constexpr int NTHR = GS_CONSUMERS * 32;
    constexpr int NW = GS_CONSUMERS;
    __shared__ float4 red[NW][NT][32];
    __shared__ float res[NT][16][8];
    __shared__ float s_rn[8];
    __shared__ int s_last;
    __shared__ __align__(8) unsigned long long bars[2 * GS_STAGES];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const int nkb = a.K >> 5;
    const int n_tiles = a.N >> 4;
    const long long U = (long long)n_tiles * nkb;
    const int G = gridDim.x, cta = blockIdx.x;
    const long long u0 = gs_u0(cta, U, G), u1 = gs_u0(cta + 1, U, G);
    const int n_chunks = (int)((u1 - u0 + GS_CHUNK_UNITS - 1) / GS_CHUNK_UNITS);
    const uint32_t ring = smem_u32(gs_smem);
    const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[GS_STAGES]);

    if (threadIdx.x == 0) {
        for (int s = 0; s < GS_STAGES; ++s) {
            gs_mbar_init(bar_full + 8 * s, 1);
            gs_mbar_init(bar_empty + 8 * s, NW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    griddep_launch();

    if (warp == NW) {
        // ---------------- producer: stream this CTA's byte range, never waits for the previous kernel ----------------
        if (lane == 0) {
            const unsigned char* src = reinterpret_cast<const unsigned char*>(a.W) + u0 * 1024;
            for (int j = 0; j < n_chunks; ++j) {
                const int s = j % GS_STAGES;
                const uint32_t ph = (j / GS_STAGES) & 1;
                gs_mbar_wait(bar_empty + 8 * s, ph ^ 1);
                const long long left = u1 - u0 - (long long)j * GS_CHUNK_UNITS;
                const uint32_t bytes = (uint32_t)(left < GS_CHUNK_UNITS ? left : GS_CHUNK_UNITS) * 1024u;
                gs_mbar_expect_tx(bar_full + 8 * s, bytes);
                gs_bulk_load(ring + s * (GS_CHUNK_UNITS * 1024), src + (size_t)j * GS_CHUNK_UNITS * 1024, bytes, bar_full + 8 * s);
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    griddep_wait();     // activations come from the previous kernel
    int j_cur = -1;     // chunk this warp currently holds (waited full, not yet released)
    auto acquire_chunk = [&](int j) {
        while (j_cur < j) {
            if (j_cur >= 0 && lane == 0) gs_mbar_arrive(bar_empty + 8 * (j_cur % GS_STAGES));
            ++j_cur;
            gs_mbar_wait(bar_full + 8 * (j_cur % GS_STAGES), (j_cur / GS_STAGES) & 1);
        }
    };

    const int t_first = (int)(u0 / nkb), t_last = u1 > u0 ? (int)((u1 - 1) / nkb) : -1;
    for (int t = t_first; t <= t_last; ++t) {
        const long long tb = (long long)t * nkb;
        const long long ua = u0 > tb ? u0 : tb, ub = u1 < tb + nkb ? u1 : tb + nkb;
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[nt][i] = 0.f;

        constexpr int UB = 4;    // units per batch: the activation loads of a batch are issued together
        for (long long ubase_u = ua + warp; ubase_u < ub; ubase_u += NW * UB) {
            uint4 xv[UB][NT];
#pragma unroll
            for (int b = 0; b < UB; ++b) {
                const long long u = ubase_u + (long long)b * NW;
                const int ko = ((int)(u - tb) << 5) + q * 8;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
                    xv[b][nt] = u < ub ? *reinterpret_cast<const uint4*>(a.X + (size_t)(nt * 8 + g) * a.ldx + ko) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int b = 0; b < UB; ++b) {
                const long long u = ubase_u + (long long)b * NW;
                if (u >= ub) break;
                const int rel = (int)(u - u0);
                acquire_chunk(rel / GS_CHUNK_UNITS);
                const uint32_t ubase = ring + (j_cur % GS_STAGES) * (GS_CHUNK_UNITS * 1024) + (rel % GS_CHUNK_UNITS) * 1024 + (q * 8 + g) * 16;
                uint4 a0, a1;
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a0.x), "=r"(a0.y), "=r"(a0.z), "=r"(a0.w) : "r"(ubase));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a1.x), "=r"(a1.y), "=r"(a1.z), "=r"(a1.w) : "r"(ubase + 512));
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    mma_bf16_16816(acc[nt], a0.x, a1.x, a0.y, a1.y, xv[b][nt].x, xv[b][nt].y);
                    mma_bf16_16816(acc[nt], a0.z, a1.z, a0.w, a1.w, xv[b][nt].z, xv[b][nt].w);
                }
            }
        }

        // ---- CTA-level sum of the 8 warps (fixed order) ----
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) red[warp][nt][lane] = make_float4(acc[nt][0], acc[nt][1], acc[nt][2], acc[nt][3]);
        gs_csync();
        for (int e = threadIdx.x; e < NT * 32; e += NTHR) {
            const int tile = e >> 5, l2 = e & 31;
            float4 s = red[0][tile][l2];
#pragma unroll
            for (int w = 1; w < NW; ++w) {
                const float4 v = red[w][tile][l2];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            const int rr = l2 >> 2, cc = (l2 & 3) * 2;
            res[tile][rr][cc] = s.x;
            res[tile][rr][cc + 1] = s.y;
            res[tile][rr + 8][cc] = s.z;
            res[tile][rr + 8][cc + 1] = s.w;
        }
        gs_csync();

        const bool owner = ua == tb;                 // this CTA holds the tile's FIRST unit
        if (!owner) {
            // only a CTA's first tile can start in an earlier CTA: publish the partial right away (the owner needs it
            // at the END of its own range, so nobody waits on a CTA that has most of its work still ahead)
            float* dst = fx.scratch + (size_t)cta * (16 * 32);
            for (int e = threadIdx.x; e < 16 * NT * 8; e += NTHR) dst[e] = res[(e >> 3) / 16][(e >> 3) % 16][e & 7];
            __threadfence();
            gs_csync();
            if (threadIdx.x == 0) gs_st_release(fx.flags + cta, 1);
            continue;
        }
        if (ub < tb + nkb) {
            // the tile continues in later CTAs: add their partials in CTA order
            const long long last_u = tb + nkb - 1;
            int c_hi = (int)((last_u * G) / U);
            while (c_hi + 1 < G && gs_u0(c_hi + 1, U, G) <= last_u) ++c_hi;
            while (gs_u0(c_hi, U, G) > last_u) --c_hi;
            for (int c2 = cta + 1; c2 <= c_hi; ++c2) {
                if (threadIdx.x == 0)
                    while (gs_ld_acquire(fx.flags + c2) != 1) {}
                gs_csync();
                const float* src2 = fx.scratch + (size_t)c2 * (16 * 32);
                for (int e = threadIdx.x; e < 16 * NT * 8; e += NTHR) res[(e >> 3) / 16][(e >> 3) % 16][e & 7] += __ldcg(src2 + e);
                gs_csync();
                if (threadIdx.x == 0) gs_st_release(fx.flags + c2, 0);
            }
        }

        // ---- epilogue of a finished tile ----
        const int row0 = t * 16;
        if (EPI == EPI_STORE) {
            for (int e = threadIdx.x; e < 16 * NT * 8; e += NTHR) {
                const int r = e % 16, s = e / 16;
                if (s < a.T) {
                    float v = res[s >> 3][r][s & 7];
                    if (a.bias) v += a.bias[row0 + r];
                    a.out[(size_t)s * a.ldo + row0 + r] = v;
                }
            }
        } else if (EPI == EPI_RESID) {
            for (int e = threadIdx.x; e < 16 * NT * 8; e += NTHR) {
                const int r = e % 16, s = e / 16;
                if (s < a.T) {
                    float* o = a.out + (size_t)s * a.ldo + row0 + r;
                    *o = *o + res[s >> 3][r][s & 7];
                }
            }
        } else if (EPI == EPI_SWIGLU) {
            for (int e = threadIdx.x; e < 8 * NT * 8; e += NTHR) {
                const int rr = e % 8, s = e / 8;
                if (s < a.T) {
                    const float gt = res[s >> 3][rr][s & 7], up = res[s >> 3][rr + 8][s & 7];
                    a.out_bf[(size_t)s * a.ldob + (row0 >> 1) + rr] = __float2bfloat16_rn(gt / (1.0f + expf(-gt)) * up);
                }
            }
        } else if (EPI == EPI_QKV) {
            const int half = a.hd >> 1;
            for (int e = threadIdx.x; e < 8 * NT * 8; e += NTHR) {
                const int rr = e % 8, s = e / 8;
                if (s < a.T) {
                    float x1 = res[s >> 3][rr][s & 7], x2 = res[s >> 3][rr + 8][s & 7];
                    if (a.bias) {
                        x1 += a.bias[row0 + rr];
                        x2 += a.bias[row0 + rr + 8];
                    }
                    const int head = row0 / a.hd;
                    const int d = ((row0 % a.hd) >> 4) * 8 + rr;
                    const int pos = a.slot_pos[s];
                    float y1 = x1, y2 = x2;
                    if (head < a.n_q + a.n_kv) {
                        const float cs = a.cos_t[(size_t)pos * half + d], sn = a.sin_t[(size_t)pos * half + d];
                        y1 = x1 * cs - x2 * sn;
                        y2 = x2 * cs + x1 * sn;
                    }
                    bf16* dst;
                    if (head < a.n_q) {
                        dst = a.Q + (size_t)s * (a.n_q * a.hd) + (size_t)head * a.hd;
                    } else {
                        const bool is_k = head < a.n_q + a.n_kv;
                        const int kvh = head - a.n_q - (is_k ? 0 : a.n_kv);
                        const int page = a.page_table[(size_t)a.slot_seq[s] * a.pt_stride + pos / KV_PAGE];
                        dst = (is_k ? a.Kp : a.Vp) + (((size_t)page * a.n_kv + kvh) * KV_PAGE + (pos % KV_PAGE)) * a.hd;
                    }
                    dst[d] = __float2bfloat16_rn(y1);
                    dst[d + half] = __float2bfloat16_rn(y2);
                }
            }
        } else {  // EPI_LOGITS: candidates per TILE
            if (a.logits) {
                for (int e = threadIdx.x; e < 16 * NT * 8; e += NTHR) {
                    const int r = e % 16, s = e / 16;
                    if (s < a.T) a.logits[(size_t)s * a.ld_logits + row0 + r] = res[s >> 3][r][s & 7];
                }
            }
            if (threadIdx.x < NT * 8) {
                const int s = threadIdx.x;
                float best = -INFINITY;
                int bi = 0x7fffffff;
                if (s < a.T) {
                    const int step = (a.inv_temp > 0.f && a.step_ptr) ? *a.step_ptr : 0;
                    for (int r = 0; r < 16; ++r) {
                        float v = res[s >> 3][r][s & 7];
                        if (a.inv_temp > 0.f) v = v * a.inv_temp + gumbel_noise(a.seed, step, s, a.row_offset + row0 + r);
                        if (v > best) { best = v; bi = a.row_offset + row0 + r; }
                    }
                }
                a.cta_max[t * 32 + s] = best;
                a.cta_idx[t * 32 + s] = bi;
            }
        }
        gs_csync();     // res / red are reused by the next tile
    }
    // release whatever this warp still holds, so that the ring bookkeeping is balanced
    if (j_cur >= 0 && lane == 0) gs_mbar_arrive(bar_empty + 8 * (j_cur % GS_STAGES));
    for (int j = j_cur + 1; j < n_chunks; ++j) {
        gs_mbar_wait(bar_full + 8 * (j % GS_STAGES), (j / GS_STAGES) & 1);
        if (lane == 0) gs_mbar_arrive(bar_empty + 8 * (j % GS_STAGES));
    }

    // ---- tails: the LAST CTA to finish normalises the residual rows (RESID) / picks the winners (LOGITS) ----
    if ((EPI == EPI_RESID && a.norm_w) || (EPI == EPI_LOGITS && a.next_tok)) {
        __threadfence();
        gs_csync();
        if (threadIdx.x == 0) s_last = atomicAdd(a.done_counter, 1) == (int)gridDim.x - 1;
        gs_csync();
        if (!s_last) return;
        __threadfence();
        float* sred = &res[0][0][0];
        if (EPI == EPI_RESID) {
            const int n4 = a.N >> 2;
            constexpr int PER = 4, SG = 4;
            if (n4 <= PER * NTHR) {
                for (int s0 = 0; s0 < a.T; s0 += SG) {
                    float4 v[SG][PER];
                    float ss[SG];
#pragma unroll
                    for (int sl = 0; sl < SG; ++sl) {
                        ss[sl] = 0.f;
#pragma unroll
                        for (int k = 0; k < PER; ++k) {
                            const int c4 = threadIdx.x + k * NTHR;
                            if (s0 + sl < a.T && c4 < n4) {
                                v[sl][k] = __ldcg(reinterpret_cast<const float4*>(a.out + (size_t)(s0 + sl) * a.ldo) + c4);
                                ss[sl] += v[sl][k].x * v[sl][k].x + v[sl][k].y * v[sl][k].y + v[sl][k].z * v[sl][k].z + v[sl][k].w * v[sl][k].w;
                            } else {
                                v[sl][k] = make_float4(0.f, 0.f, 0.f, 0.f);
                            }
                        }
                    }
#pragma unroll
                    for (int sl = 0; sl < SG; ++sl) {
                        ss[sl] = warp_sum(ss[sl]);
                        if (lane == 0) sred[warp * SG + sl] = ss[sl];
                    }
                    gs_csync();
#pragma unroll
                    for (int sl = 0; sl < SG; ++sl) {
                        float tot = 0.f;
                        for (int w = 0; w < NW; ++w) tot += sred[w * SG + sl];
                        const float r = 1.0f / sqrtf(tot / (float)a.N + a.eps);
#pragma unroll
                        for (int k = 0; k < PER; ++k) {
                            const int c4 = threadIdx.x + k * NTHR;
                            if (s0 + sl < a.T && c4 < n4) {
                                const uint2 wv = *reinterpret_cast<const uint2*>(a.norm_w + c4 * 4);
                                uint2 o;
                                o.x = pack_bf16(v[sl][k].x * r * bf16lo(wv.x), v[sl][k].y * r * bf16hi(wv.x));
                                o.y = pack_bf16(v[sl][k].z * r * bf16lo(wv.y), v[sl][k].w * r * bf16hi(wv.y));
                                *reinterpret_cast<uint2*>(a.xn_out + (size_t)(s0 + sl) * a.N + c4 * 4) = o;
                            }
                        }
                    }
                    gs_csync();
                }
            } else {
                for (int s = 0; s < a.T; ++s) {
                    const float* xr = a.out + (size_t)s * a.ldo;
                    float ss = 0.f;
                    for (int c = threadIdx.x * 4; c < a.N; c += NTHR * 4) {
                        const float4 v = __ldcg(reinterpret_cast<const float4*>(xr + c));
                        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                    }
                    ss = warp_sum(ss);
                    if (lane == 0) sred[warp] = ss;
                    gs_csync();
                    float tot = 0.f;
                    for (int w = 0; w < NW; ++w) tot += sred[w];
                    const float r = 1.0f / sqrtf(tot / (float)a.N + a.eps);
                    for (int c = threadIdx.x * 4; c < a.N; c += NTHR * 4) {
                        const float4 v = __ldcg(reinterpret_cast<const float4*>(xr + c));
                        const uint2 wv = *reinterpret_cast<const uint2*>(a.norm_w + c);
                        uint2 o;
                        o.x = pack_bf16(v.x * r * bf16lo(wv.x), v.y * r * bf16hi(wv.x));
                        o.y = pack_bf16(v.z * r * bf16lo(wv.y), v.w * r * bf16hi(wv.y));
                        *reinterpret_cast<uint2*>(a.xn_out + (size_t)s * a.N + c) = o;
                    }
                    gs_csync();
                }
            }
        } else {
            float* sval = sred;
            int* sidx = reinterpret_cast<int*>(sval + 32);
            const int step = a.step_w ? *a.step_w : 0;
            for (int t = 0; t < a.T; ++t) {
                float best = -INFINITY;
                int bi = 0x7fffffff;
                for (int c0 = threadIdx.x; c0 < n_tiles; c0 += NTHR * 4) {
                    float v[4];
                    int ix[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int c = c0 + j * NTHR;
                        v[j] = c < n_tiles ? __ldcg(a.cta_max + c * 32 + t) : -INFINITY;
                        ix[j] = c < n_tiles ? __ldcg(a.cta_idx + c * 32 + t) : 0x7fffffff;
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (v[j] > best || (v[j] == best && ix[j] < bi)) { best = v[j]; bi = ix[j]; }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float v2 = __shfl_xor_sync(0xffffffffu, best, o);
                    const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (v2 > best || (v2 == best && i2 < bi)) { best = v2; bi = i2; }
                }
                if (lane == 0) { sval[warp] = best; sidx[warp] = bi; }
                gs_csync();
                if (threadIdx.x == 0) {
                    for (int w = 1; w < NW; ++w)
                        if (sval[w] > best || (sval[w] == best && sidx[w] < bi)) { best = sval[w]; bi = sidx[w]; }
                    if (bi == 0x7fffffff) bi = 0;
                    a.next_tok[t] = bi;
                    if (a.next_val) a.next_val[t] = best;
                    if (a.tok) a.tok[t] = bi;
                    if (a.pos) a.pos[t] += 1;
                    if (a.out_tokens && step < a.max_out) a.out_tokens[(size_t)t * a.max_out + step] = bi;
                }
                gs_csync();
            }
            if (threadIdx.x == 0 && a.step_w) *a.step_w = step + 1;
        }
        if (threadIdx.x == 0) *a.done_counter = 0;
    }
}
