// Interface of the tensor-core prefill path (prefill.cu): TMA tensor maps + tcgen05 GEMM launches.
#pragma once
#include "../../include/rt_engine.h"
#include "common.cuh"

constexpr int PF_BN = 256;   // weight rows per CTA tile (TMEM columns of the accumulator)

// opaque mirror of CUtensorMap (128 bytes, 64-byte aligned) so that engine code need not include <cuda.h>
struct alignas(64) PfTensorMap {
    unsigned long long opaque[16];
};

enum { G5_EPI_QKV = 0, G5_EPI_RESID = 1, G5_EPI_SWIGLU = 2, G5_EPI_STORE = 3, G5_EPI_STORE16 = 4 };

struct Gemm5Args {
    int S, N, K;
    int ab_f16;               // 1: BOTH operands hold fp16 values (engine: fp16 activations x the fp16 twin of the weights); 0: both bf16
    const float* row_scale;   // nullable: acc *= row_scale[token] first (split RMSNorm: r = rsqrt(mean x^2 + eps))
    // RESID: out[m*ldo + n] += acc ; STORE: out[m*ldo + n] = acc
    float* out;
    int ldo;
    // SWIGLU: out_bf[m*ldob + n/2] ; STORE16 (tensor parallel: the partial sums that cross NVLink): out_bf[m*ldob + n] = fp16(acc)
    f16* out_bf;              // fp16
    int ldob;
    // QKV
    const float* bias;
    const float* cos_t;
    const float* sin_t;
    const int* tok_seq;
    const int* tok_pos;
    const int* page_table;
    int pt_stride;
    bf16* Q;
    bf16* Kp;
    bf16* Vp;
    int n_q, n_kv, hd;
};

struct PfAttnArgs {
    const bf16* Q;          // [S][n_q*hd] (RoPE applied, bf16)
    const bf16* Kp;
    const bf16* Vp;
    const int* page_table;
    int pt_stride;
    const int* tok_pos;     // [S]
    const int4* work;       // {first token (flat), n tokens, sequence, kv head}
    int n_q, n_kv;
    float scale_log2;
    f16* out;               // [S][n_q*hd] fp16 (the O-projection's input)
};

void pf_attention(cudaStream_t st, bool pdl, int hd, int n_items, const PfAttnArgs& a);

// tcgen05 variant (attn_prefill_tc5.cuh, head_dim 128): Q as a 3-D map (dims, heads, tokens), K/V pools as 2-D maps
struct Pf5Args;
void pf_make_map_q3(PfTensorMap* out, const void* base, int tokens, int n_q, int G, int BT, int hd);
void pf_make_map_kv(PfTensorMap* out, const void* base, long long rows, int hd);
void pf_attention_tc5(cudaStream_t st, bool pdl, int n_items, const PfTensorMap& mq, const PfTensorMap& mk, const PfTensorMap& mv,
                      const int* page_table, int pt_stride, const int* tok_pos, const int4* work, int n_q, int n_kv, int G, int BT,
                      int n_pages_pool, int layer, float scale_log2, f16* out, int hd);
// tcgen05 decode attention for long contexts (attn_decode_tc5.cuh): query rows = (slot, q-head) pairs, work items = key ranges
struct Dec5Args;
struct Dec5Item;
void pf_attention_dec5(cudaStream_t st, bool pdl, int n_items, const PfTensorMap& mq, const PfTensorMap& mk, const PfTensorMap& mv,
                       const Dec5Args& a, const Dec5Item* host_items, int hd);
void pf_make_map_a(PfTensorMap* out, const void* base, int rows, int K);
void pf_make_map_w(PfTensorMap* out, const void* base, int N, int K, int BN);
void pf_gemm(cudaStream_t st, bool pdl, int epi, const PfTensorMap& ma, const PfTensorMap& mw, const Gemm5Args& a);
