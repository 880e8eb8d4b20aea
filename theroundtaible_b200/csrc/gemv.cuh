// Decode projection kernel (v2): the weight stream of the generation backend.
//
// No reference counterpart (/root/reference/src/adapters/local-llm.ts:159 delegates generation to an
// external server); SURVEY.md 8(a-K) row "decode QKV / O / gate-up / down projections" and "lm_head".
//
//   acc[t][n] = sum_k W[n][k] * X[t][k]        W bf16 [N][K] (HBM stream, read once), T <= 32 token slots
//
// Weight layout in HBM ("core-matrix-major", CM): the matrix is cut into 16-row tiles and 32-wide k-blocks;
// one (tile, k-block) is 1 KB = [half h: rows 0-7 / 8-15][k-chunk q: 4 x 8 elements][row r: 8][8 elements].
// Every 8-row x 16-byte "core matrix" is 128 contiguous bytes, so
//   * a warp's mma.sync A-fragment load (lane (g,q) reads piece q*8+g of a half) is 512 CONTIGUOUS bytes,
//     a whole tile is K/32 contiguous KB and a warp's k-range is a contiguous span of it;
//   * the same bytes are a valid no-swizzle K-major canonical operand for tcgen05.mma (prefill path).
// The 16 k's of one MMA are a permutation of memory order (a dot product does not care), chosen so that a
// lane's fragment is one 16-byte piece: MMA #0 uses elements 0..3 of the piece, MMA #1 elements 4..7.
//
// A CTA owns 16*MT rows over the full K; its NW warps split K into contiguous spans; per-warp partial sums are
// added in fixed warp order through shared memory (bit-reproducible, no atomics).  With PDL the first weight
// loads are issued BEFORE griddepcontrol.wait, so the HBM stream keeps running across kernel boundaries.
//
// Activations are fp16 VALUES (common.cuh) that their producer has split into bf16 hi + lo and stored in the XF layout
// (common.cuh xf_index): the lanes of a warp fetch the B fragments of a k-block as 512 contiguous bytes per part, and both
// parts go through the MMAs, so the product is exact in the fp16 input.  Measured alternatives (profiles/r2_summary.md): the
// split done in registers by the consumer costs ~36 ALU instructions per k-block and lane (+6-8 % per step); hi / lo as two
// ROW-MAJOR planes double the half-line requests to the few L2 lines every CTA reads (+18 % per kernel).
// With T <= 4 slots (template XM = 1) hi and lo share ONE MMA: columns 0..3 carry hi of slots 0..3, columns 4..7 their lo,
// and the epilogue adds the two columns of a slot.  With 5..8 slots (XM = 2) the kernel streams the fp16 twin of the matrix
// against the fp16 activations themselves: see the template's comment.
//
// Fusions: EPI_QKV    -> + bias, RoPE, q -> bf16 Q buffer, k/v -> bf16 into their KV page;
//          EPI_RESID  -> x += acc, xw = fp16(x * w_next) (XF layout) and per-tile partial sums of x^2: the next projection applies
//                        the RMSNorm scale to its accumulators (no RMSNorm launch, no tail);
//          EPI_SWIGLU -> fp16(silu(gate) * up);   EPI_LOGITS -> logits and per-CTA arg-max candidates;
//          EPI_PUSH   -> tensor parallel: the partial sums go straight into EVERY rank's receive buffer (peer memory over
//                        NVLink) as self-validating 16-byte packets (tp_fused.cuh tp_ll_reduce_norm_kernel).
#pragma once
#include "common.cuh"

enum { EPI_STORE = 0, EPI_RESID = 1, EPI_SWIGLU = 2, EPI_LOGITS = 3, EPI_QKV = 4, EPI_PUSH = 5 };

constexpr int KV_PAGE = 16;

struct GemvArgs {
    const bf16* W;            // CM layout (mode 2: the fp16 twin, same layout)
    int N, K, T;
    // B operand
    const bf16* X;            // activations of the K input columns in the XF layout (common.cuh): bf16 hi + lo of fp16 values
    int f16;                  // host side only: 1 = W is the fp16 twin and X / the XF outputs are in the XF16 form (mode 2)
    // PUSH (tensor parallel, <= 8 slots): every rank's receive buffer of this call site [tp][32][N] (peer-mapped; [rank] = local),
    // this rank's index, and the per-slot packet epochs of the site [32] (tp_fused.cuh)
    float* push_dst[8];
    int push_tp, push_rank;
    const int* push_epoch;
    int ldx;                  // (unused: the XF layout is fixed by K)
    // RMSNorm is split around the projections ("scale after the matmul"): the PRODUCER of the residual stream (RESID
    // epilogue, embed kernel) writes xw = bf16(x * norm_w) per element plus the per-16-row-tile partial sums of x^2; the
    // CONSUMER projection multiplies its accumulators by r = rsqrt(mean x^2 + eps) in the epilogue.  No kernel ever
    // waits for a whole row, so there is no RMSNorm launch and no last-CTA tail on the critical path.
    const bf16* norm_w;       // RESID: weight of the NEXT norm [N]; nullable => no xw / partial sums
    bf16* xn_out;             // RESID: xw out, XF layout over N columns
    float* ss_out;            // RESID: [N/16][32] partial sums of x^2 of the updated rows
    const float* ss_in;       // consumer: partial sums [n_parts_in][32] of the rows behind X (nullable)
    int n_parts_in;
    const float* inv_rms;     // consumer alternative: r per slot given directly [T] (tensor-parallel / prefill paths)
    int* done_counter;        // LOGITS sampler tail: arrival counter (zero before the first launch; reset by the last CTA)
    float eps;
    // outputs
    float* out;               // STORE: out[t*ldo+n] = acc (+bias) ; RESID: out[t*ldo+n] += acc
    int ldo;
    const float* bias;        // STORE / QKV (physical row order), nullable
    bf16* out_bf;             // SWIGLU: XF layout over ldob columns
    int ldob;
    // LOGITS
    float* logits;
    int ld_logits;
    float* cta_max;
    int* cta_idx;
    int row_offset;
    int late_launch;          // QKV only: griddepcontrol.launch_dependents after the wait instead of at the start
    L2Prefetch pf;            // a later kernel's weights: each CTA asks for its slice once its own stream is issued
    float inv_temp;
    unsigned long long seed;
    const int* step_ptr;
    // LOGITS "sampler tail" (next_tok nullable => candidates only)
    int* next_tok;            // [T]
    float* next_val;          // nullable: winning score per slot (tensor parallel: ranks compare their local winners)
    int* tok;                 // nullable: tok[t] = next (decode feeds the next step)
    int* pos;                 // nullable: pos[t] += 1
    int* out_tokens;          // nullable: out_tokens[t*max_out + *step_w] = next
    int max_out;
    int* step_w;              // nullable: incremented once per launch
    // QKV + RoPE + KV append
    const float* cos_t;
    const float* sin_t;
    const int* slot_seq;
    const int* slot_pos;
    const int* page_table;
    int pt_stride;
    bf16* Q;                  // [T][n_q*hd]
    bf16* Kp;
    bf16* Vp;
    int n_q, n_kv, hd;
};

__device__ __forceinline__ float gumbel_noise(unsigned long long seed, int step, int slot, int idx) {
    unsigned long long z = seed + 0x9E3779B97F4A7C15ull * ((unsigned long long)(unsigned)step * 64ull + (unsigned)slot) +
                           ((unsigned long long)(unsigned)idx << 20);
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u = ((float)(z >> 40) + 0.5f) * (1.0f / 16777216.0f);  // (0,1)
    return -logf(-logf(u));
}

// programmatic dependent launch (PDL): no-ops when the kernel was launched without the attribute
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;"); }

// CL = 2: a CLUSTER PAIR shares a row group and splits K between its two CTAs; the second CTA adds its reduced tile into the
// first one's shared memory (distributed shared memory) and the first runs the epilogue.  For matrices with fewer 16-row
// tiles than half the SMs -- the QKV projection of a tensor-parallel shard: 70B / TP=8 has 80 tiles, 21 MB on 80 of 148 SMs,
// 10 us in the in-kernel timeline -- it doubles the CTAs that stream the weights.
//
// XM (activation mode): 0 = bf16 weights, activations as bf16 hi + lo, two MMAs per weight fragment;
//                       1 = the same with <= 4 slots: hi and lo share one MMA (see the header);
//                       2 = the fp16 TWINS of the weights (engine.cu: the exact fp16 image of the bf16 matrices, which the prefill
//                           GEMMs read anyway) x the fp16 activations themselves (XF16 form, common.cuh xf_store*): one f16 MMA
//                           and one activation load per fragment, the same products.  5..8 slots cost 5-8 % per projection
//                           in mode 0 (profiles/r2_summary.md); mode 2 takes that back.
// OC = 3: three CTAs per SM (<= 85 registers: half the weight batch in flight per warp) for matrices with more 16-row tiles
// than two CTAs per SM hold but no more than three do -- every tile is then resident at once and the busiest SM carries 3
// tiles instead of 4 (Llama-3.1-8B QKV: 384 tiles on 148 SMs).
template <int NT, int MT, int EPI, int NW, int XM = 0, int CL = 1, int OC = 1>
__global__ void __launch_bounds__(NW * 32, OC == 3 ? 3 : (NW == 8 ? 2 : 1)) gemv_kernel(const GemvArgs a) {   // <= 128 / 85 registers
    constexpr int HM = XM == 1, F16 = XM == 2;
    constexpr int TRK = EPI == EPI_PUSH ? 9 : EPI;    // trace kind (5 is the mma.sync attention kernel's)
    static_assert(HM == 0 || NT == 1, "the shared-MMA mode serves <= 4 slots");
    static_assert(CL == 1 || CL == 2, "cluster pair or none");
    static_assert(NT * MT <= 4, "shared-memory reduction sized for NT*MT <= 4");
    // NT = groups of 8 B-operand columns (slots).  HM: one group whose columns 0..3 are the hi part of slots 0..3 and
    // columns 4..7 their lo part.
    constexpr int SL = HM ? 4 : NT * 8;               // slots this instantiation serves
    // k-blocks of weights a warp keeps in flight (registers): W takes 8*U*MT registers, X 4*U*NT
    constexpr int XP = (HM || F16) ? 1 : 2;           // activation loads per (k-block, slot group): HM lanes pick their part
    constexpr int U = OC == 3 ? 4 : (HM || (F16 && NT == 1)) ? 8 / MT : (NT == 1 ? (MT == 1 ? 6 : (MT == 2 ? 4 : 2)) : (NT == 2 && MT == 1 ? 4 : 2));
    constexpr int NTHR = NW * 32;
    __shared__ float4 red[NW][MT * NT][32];
    __shared__ float res[MT * NT][16][8];
    __shared__ float res_peer[CL > 1 ? MT * NT : 1][16][8];       // CL = 2: written by the pair's second CTA
    __shared__ float s_rn[SL];
    __shared__ int s_last;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    // A CTA owns the row groups tile0, tile0 + gridDim.x*MT, ... (one group when the grid covers the matrix: the usual
    // single-wave case; several for the matrices with more tiles than two CTAs per SM can hold at once -- lm_head, the
    // 8B+ gate/up -- where the host launches a persistent grid and the first weights of the NEXT group are requested
    // before this group's reduction and epilogue, so the weight stream never pauses at a group boundary).
    const int crank = CL > 1 ? (int)cl_rank() : 0;
    const int n_groups_grid = (int)gridDim.x / CL;        // row groups the grid covers at once
    int tile0 = ((int)blockIdx.x / CL) * MT;
    int row0 = tile0 * 16;
    const int n_tiles = a.N >> 4;
    const int nkb = a.K >> 5;
    const int nkb_r = nkb / CL;                            // this CTA's share of K (host: nkb % CL == 0)
    const int kb0 = crank * nkb_r + (warp * nkb_r) / NW, kb1 = crank * nkb_r + ((warp + 1) * nkb_r) / NW;

    // lane's byte offset inside a 512-byte half: piece (q*8 + g)
    const bf16* wbase[MT];
    uint4 av[U][MT][2];
    auto first_batch = [&](int t0) {
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
            int tile = t0 + mt;
            tile = tile < n_tiles ? tile : (n_tiles - 1);
            wbase[mt] = a.W + (size_t)tile * nkb * 512 + (q * 8 + g) * 8;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int kb = kb0 + u;
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
                if (kb < kb1) {
                    av[u][mt][0] = ld_stream16(wbase[mt] + (size_t)kb * 512);
                    av[u][mt][1] = ld_stream16(wbase[mt] + (size_t)kb * 512 + 256);
                } else {
                    av[u][mt][0] = make_uint4(0, 0, 0, 0);
                    av[u][mt][1] = make_uint4(0, 0, 0, 0);
                }
            }
        }
    };

    float acc[MT][NT][4];

    RT_TR(TRK * 10 + 0);
    // ---- first batch of weight loads: independent of the previous kernel, issued before the PDL wait ----
    first_batch(tile0);
    // The QKV projection releases its dependent (the attention kernel) only AFTER its own wait: whatever that kernel
    // does before ITS wait (index setup, KV pages of older tokens) then runs with every kernel before this one complete.
    if (!(EPI == EPI_QKV && a.late_launch)) griddep_launch();
    griddep_wait();
    if (EPI == EPI_QKV && a.late_launch) griddep_launch();
    RT_TR(TRK * 10 + 1);

    // RMSNorm scale of every slot (used in the epilogue only: these loads overlap the weight stream)
    if (a.ss_in || a.inv_rms) {
        for (int s = warp; s < SL; s += NW) {
            float v = 0.f;
            if (s < a.T) {
                if (a.inv_rms) {
                    v = lane == 0 ? a.inv_rms[s] : 0.f;
                } else {
                    for (int p = lane; p < a.n_parts_in; p += 32) v += a.ss_in[p * 32 + s];
                }
            }
            v = warp_sum(v);
            if (lane == 0) s_rn[s] = s >= a.T ? 0.f : (a.inv_rms ? v : 1.0f / sqrtf(v / (float)a.K + a.eps));
        }
    } else if (threadIdx.x < SL) {
        s_rn[threadIdx.x] = 1.0f;
    }
    // value of (row tile mt, row r, slot s) after the reduction: hi column + lo column
    auto val = [&](int mt, int r, int s) -> float {
        if (HM) return res[mt][r][s] + res[mt][r][s + 4];
        return res[mt * NT + (s >> 3)][r][s & 7];
    };

    // LOGITS: arg-max candidate of this CTA over all its row groups (threads < NT*8)
    float cand_best = -INFINITY;
    int cand_idx = 0x7fffffff;

    for (;;) {
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0.f;

    // ---- inputs of this thread's FIRST epilogue element, fetched now so that their L2 round trips (two dependent ones
    // for the QKV epilogue: position -> rope table / page table) hide behind the weight stream ----
    float pre_f0 = 0.f, pre_f1 = 0.f, pre_f2 = 0.f, pre_f3 = 0.f;
    bf16* pre_dst = nullptr;
    if (EPI == EPI_RESID) {
        const int e = threadIdx.x;
        const int r = e % (MT * 16), sl = e / (MT * 16), n = row0 + r;
        if (e < MT * 16 * SL && sl < a.T && n < a.N) {
            pre_f0 = a.out[(size_t)sl * a.ldo + n];
            if (a.norm_w) pre_f1 = __bfloat162float(a.norm_w[n]);
        }
    } else if (EPI == EPI_QKV) {
        const int e = threadIdx.x;
        const int r = e % (MT * 8), sl = e / (MT * 8);
        const int mt = r >> 3, rr = r & 7, n = row0 + mt * 16;
        if (e < MT * 8 * SL && sl < a.T && n < a.N) {
            const int half = a.hd >> 1;
            if (a.bias) { pre_f2 = a.bias[n + rr]; pre_f3 = a.bias[n + rr + 8]; }
            const int head = n / a.hd;
            const int d = ((n % a.hd) >> 4) * 8 + rr;
            const int pos = a.slot_pos[sl];
            pre_f0 = 1.f;
            if (head < a.n_q + a.n_kv) { pre_f0 = a.cos_t[(size_t)pos * half + d]; pre_f1 = a.sin_t[(size_t)pos * half + d]; }
            if (head < a.n_q) {
                pre_dst = a.Q + (size_t)sl * (a.n_q * a.hd) + (size_t)head * a.hd + d;
            } else {
                const bool is_k = head < a.n_q + a.n_kv;
                const int kvh = head - a.n_q - (is_k ? 0 : a.n_kv);
                const int page = a.page_table[(size_t)a.slot_seq[sl] * a.pt_stride + pos / KV_PAGE];
                pre_dst = (is_k ? a.Kp : a.Vp) + (((size_t)page * a.n_kv + kvh) * KV_PAGE + (pos % KV_PAGE)) * a.hd + d;
            }
        }
    }

    for (int kbb = kb0; kbb < kb1; kbb += U) {
        // B operand for this batch
        uint4 xv[U][NT][XP];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int kb = kbb + u;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int slot = HM ? (g & 3) : nt * 8 + g;
                // XF layout: [slot group][k-block][hi | lo][8 slots][32 k]; HM lanes 4..7 fetch the lo part of slots 0..3.
                // Rows of unused slots are never fetched: every CTA reads the SAME activation lines, which makes the few
                // L2 slices holding them the hot spot of a short kernel (measured: scripts/probes/gridbar.cu mode 2)
                const bf16* xp = a.X + (size_t)nt * ((size_t)a.K * 16) + (size_t)kb * 512 + (HM ? (g >> 2) * 256 + (g & 3) * 32 : g * 32) + q * 8;
#pragma unroll
                for (int p = 0; p < XP; ++p) {
                    if (kb < kb1 && slot < a.T) xv[u][nt][p] = *reinterpret_cast<const uint4*>(xp + p * 256);
                    else xv[u][nt][p] = make_uint4(0, 0, 0, 0);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int p = 0; p < XP; ++p) {       // hi part, then lo part, into the same accumulator
                        if (F16) {
                            mma_f16_16816(acc[mt][nt], av[u][mt][0].x, av[u][mt][1].x, av[u][mt][0].y, av[u][mt][1].y,
                                          xv[u][nt][p].x, xv[u][nt][p].y);
                            mma_f16_16816(acc[mt][nt], av[u][mt][0].z, av[u][mt][1].z, av[u][mt][0].w, av[u][mt][1].w,
                                          xv[u][nt][p].z, xv[u][nt][p].w);
                        } else {
                            mma_bf16_16816(acc[mt][nt], av[u][mt][0].x, av[u][mt][1].x, av[u][mt][0].y, av[u][mt][1].y,
                                           xv[u][nt][p].x, xv[u][nt][p].y);
                            mma_bf16_16816(acc[mt][nt], av[u][mt][0].z, av[u][mt][1].z, av[u][mt][0].w, av[u][mt][1].w,
                                           xv[u][nt][p].z, xv[u][nt][p].w);
                        }
                    }
        // next batch of weights
        if (kbb + U < kb1) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int kb = kbb + U + u;
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) {
                    if (kb < kb1) {
                        av[u][mt][0] = ld_stream16(wbase[mt] + (size_t)kb * 512);
                        av[u][mt][1] = ld_stream16(wbase[mt] + (size_t)kb * 512 + 256);
                    } else {
                        av[u][mt][0] = make_uint4(0, 0, 0, 0);
                        av[u][mt][1] = make_uint4(0, 0, 0, 0);
                    }
                }
            }
        }
    }

    // this CTA's own weights are in: ask for its slice of a later kernel's (the reduction, the epilogue and the gap
    // before the next launch then overlap with HBM traffic instead of leaving it idle)
    const int next_tile0 = tile0 + n_groups_grid * MT;
    const bool more = next_tile0 < n_tiles;          // block-uniform
    if (more) first_batch(next_tile0);               // in flight during this group's reduction and epilogue
    else if (threadIdx.x == 0) l2_prefetch_slice(a.pf, blockIdx.x, gridDim.x);
    RT_TR(TRK * 10 + 2);

#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
            red[warp][mt * NT + nt][lane] = make_float4(acc[mt][nt][0], acc[mt][nt][1], acc[mt][nt][2], acc[mt][nt][3]);
    __syncthreads();

    // fixed-order sum over the warps; element (tile, lane', i) -> row = lane'/4 + 8*(i>>1), slot = (lane'%4)*2 + (i&1)
    for (int e = threadIdx.x; e < MT * NT * 32; e += NTHR) {
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
    __syncthreads();
    RT_TR(TRK * 10 + 3);
    if (CL > 1) {
        // the second CTA of the pair hands its half-K sums to the first through distributed shared memory
        if (crank == 1) {
            const uint32_t dst = cl_mapa(smem_u32(&res_peer[0][0][0]), 0);
            for (int e = threadIdx.x; e < MT * NT * 128; e += NTHR) cl_st_f32(dst + 4 * e, (&res[0][0][0])[e]);
        }
        cl_sync();
        if (crank == 0) {
            for (int e = threadIdx.x; e < MT * NT * 128; e += NTHR) (&res[0][0][0])[e] += (&res_peer[0][0][0])[e];
        }
        __syncthreads();
    }

    if (CL > 1 && crank != 0) {
        // nothing to write: only the pair's first CTA runs the epilogue
    } else if (EPI == EPI_STORE) {
        for (int e = threadIdx.x; e < MT * 16 * SL; e += NTHR) {
            const int r = e % (MT * 16), s = e / (MT * 16);
            const int n = row0 + r;
            if (s < a.T && n < a.N) {
                float v = val(r >> 4, r & 15, s) * s_rn[s];
                if (a.bias) v += a.bias[n];
                a.out[(size_t)s * a.ldo + n] = v;
            }
        }
    } else if (EPI == EPI_PUSH) {
        // four consecutive rows of one slot per thread = one 16-byte packet; the lowest mantissa bit of each of its words
        // carries the parity of the slot's epoch, so a receiver can tell a packet of this exchange from the one that lay there
        // before without a flag, a fence or a second round trip (every exchange rewrites every word of the rows it uses)
        for (int e = threadIdx.x; e < MT * 4 * SL; e += NTHR) {
            const int r4 = e % (MT * 4), s = e / (MT * 4);
            const int r = r4 * 4, n = row0 + r;
            if (s < a.T && n < a.N) {
                const uint32_t par = (uint32_t)a.push_epoch[s] & 1u;
                uint32_t w[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) w[i] = (__float_as_uint(val(r >> 4, (r & 15) + i, s) * s_rn[s]) & ~1u) | par;
                const size_t off = ((size_t)(a.push_rank * 32 + s)) * a.N + n;
#pragma unroll
                for (int p = 0; p < 8; ++p)
                    if (p < a.push_tp)
                        asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(a.push_dst[p] + off), "r"(w[0]), "r"(w[1]),
                                     "r"(w[2]), "r"(w[3]) : "memory");
            }
        }
    } else if (EPI == EPI_RESID) {
        for (int e = threadIdx.x; e < MT * 16 * SL; e += NTHR) {
            const int r = e % (MT * 16), s = e / (MT * 16);
            const int n = row0 + r;
            if (s < a.T && n < a.N) {
                float* o = a.out + (size_t)s * a.ldo + n;
                const bool first = e == (int)threadIdx.x;
                const float v = (first ? pre_f0 : *o) + val(r >> 4, r & 15, s);
                *o = v;
                if (a.norm_w) {
                    xf_store1(a.xn_out, a.N, s, n, v * (first ? pre_f1 : __bfloat162float(a.norm_w[n])), F16);
                    // the updated residual goes back (HM: into the hi column, lo column zero) for the sums of squares below
                    if (HM) { res[r >> 4][r & 15][s] = v; res[r >> 4][r & 15][s + 4] = 0.f; }
                    else res[(r >> 4) * NT + (s >> 3)][r & 15][s & 7] = v;
                }
            }
        }
        if (a.norm_w) {
            __syncthreads();
            for (int e = threadIdx.x; e < MT * SL; e += NTHR) {
                const int mt = e / SL, s = e % SL;
                if (s < a.T && (tile0 + mt) * 16 < a.N) {
                    float ss = 0.f;
#pragma unroll
                    for (int r = 0; r < 16; ++r) {
                        const float v = val(mt, r, s);
                        ss += v * v;
                    }
                    a.ss_out[(size_t)(tile0 + mt) * 32 + s] = ss;
                }
            }
        }
    } else if (EPI == EPI_SWIGLU) {
        // rows 0..7 of every 16-row tile are gate_{8j+r}, rows 8..15 are up_{8j+r}
        for (int e = threadIdx.x; e < MT * 8 * SL; e += NTHR) {
            const int r = e % (MT * 8), s = e / (MT * 8);
            const int mt = r >> 3, rr = r & 7;
            const int n = row0 + mt * 16;
            if (s < a.T && n < a.N) {
                const float gt = val(mt, rr, s) * s_rn[s];
                const float up = val(mt, rr + 8, s) * s_rn[s];
                const float act = gt / (1.0f + expf(-gt)) * up;
                xf_store1(a.out_bf, a.ldob, s, (n >> 1) + rr, act, F16);
            }
        }
    } else if (EPI == EPI_QKV) {
        // tile j of a head holds dims [8j, 8j+8) in rows 0..7 and dims [hd/2+8j, hd/2+8j+8) in rows 8..15:
        // a rotate_half pair (d, d+hd/2) sits in rows (r, r+8) of one tile.
        const int half = a.hd >> 1;
        for (int e = threadIdx.x; e < MT * 8 * SL; e += NTHR) {
            const int r = e % (MT * 8), s = e / (MT * 8);
            const int mt = r >> 3, rr = r & 7;
            const int n = row0 + mt * 16;
            if (s < a.T && n < a.N) {
                float x1 = val(mt, rr, s) * s_rn[s];
                float x2 = val(mt, rr + 8, s) * s_rn[s];
                if (e == (int)threadIdx.x) {          // inputs fetched before the weight stream (see above)
                    x1 += pre_f2;
                    x2 += pre_f3;
                    pre_dst[0] = __float2bfloat16_rn(x1 * pre_f0 - x2 * pre_f1);
                    pre_dst[half] = __float2bfloat16_rn(x2 * pre_f0 + x1 * pre_f1);
                    continue;
                }
                if (a.bias) {
                    x1 += a.bias[n + rr];
                    x2 += a.bias[n + rr + 8];
                }
                const int head = n / a.hd;
                const int d = ((n % a.hd) >> 4) * 8 + rr;
                const int pos = a.slot_pos[s];
                float y1 = x1, y2 = x2;
                if (head < a.n_q + a.n_kv) {
                    const float c = a.cos_t[(size_t)pos * half + d], sn = a.sin_t[(size_t)pos * half + d];
                    y1 = x1 * c - x2 * sn;
                    y2 = x2 * c + x1 * sn;
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
    } else {  // EPI_LOGITS
        if (a.logits) {
            for (int e = threadIdx.x; e < MT * 16 * SL; e += NTHR) {
                const int r = e % (MT * 16), s = e / (MT * 16);
                const int n = row0 + r;
                if (s < a.T && n < a.N) a.logits[(size_t)s * a.ld_logits + n] = val(r >> 4, r & 15, s) * s_rn[s];
            }
        }
        if (threadIdx.x < SL) {
            const int s = threadIdx.x;
            if (s < a.T) {
                const int step = (a.inv_temp > 0.f && a.step_ptr) ? *a.step_ptr : 0;
                for (int r = 0; r < MT * 16; ++r) {
                    const int n = row0 + r;
                    if (n >= a.N) break;
                    float v = val(r >> 4, r & 15, s) * s_rn[s];
                    if (a.inv_temp > 0.f) v = v * a.inv_temp + gumbel_noise(a.seed, step, s, a.row_offset + n);
                    if (v > cand_best) { cand_best = v; cand_idx = a.row_offset + n; }      // rows ascend: lowest index wins ties
                }
            }
        }
    }
    if (!more) break;
    tile0 = next_tile0;
    row0 = tile0 * 16;
    if (CL > 1) cl_sync();                           // the first CTA has read res_peer before the second refills it
    __syncthreads();                                 // red / res are reused by the next group
    }

    if (EPI == EPI_LOGITS) {
        if (threadIdx.x < SL) {
            a.cta_max[blockIdx.x * 32 + threadIdx.x] = cand_best;
            a.cta_idx[blockIdx.x * 32 + threadIdx.x] = cand_idx;
        }
        if (a.next_tok) {
            // the LAST CTA to publish its candidates picks the winners (lowest index wins ties, like torch.argmax)
            // and advances the decode state on the device, so a decode step needs no separate sampler launch
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) s_last = atomicAdd(a.done_counter, 1) == (int)gridDim.x - 1;
            __syncthreads();
            if (s_last) {
                __threadfence();
                const int n_cta = gridDim.x;
                float* sval = &res[0][0][0];                        // [NW] values, then [NW] indices
                int* sidx = reinterpret_cast<int*>(sval + 32);
                const int step = a.step_w ? *a.step_w : 0;
                for (int t = 0; t < a.T; ++t) {
                    float best = -INFINITY;
                    int bi = 0x7fffffff;
                    for (int c0 = threadIdx.x; c0 < n_cta; c0 += NTHR * 4) {
                        float v[4];
                        int ix[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int c = c0 + j * NTHR;
                            v[j] = c < n_cta ? __ldcg(a.cta_max + c * 32 + t) : -INFINITY;
                            ix[j] = c < n_cta ? __ldcg(a.cta_idx + c * 32 + t) : 0x7fffffff;
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
                    __syncthreads();
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
                    __syncthreads();
                }
                if (threadIdx.x == 0) {
                    if (a.step_w) *a.step_w = step + 1;
                    *a.done_counter = 0;
                }
            }
        }
    }
    RT_TR(TRK * 10 + 4);
}
