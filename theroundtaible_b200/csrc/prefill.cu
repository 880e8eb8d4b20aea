// Host side of the tensor-core prefill path: TMA tensor maps and launches of gemm_tc5_kernel.
#include "prefill.cuh"

#include <cstring>

#include "attn_prefill.cuh"
#include "attn_decode_tc5.cuh"
#include "attn_prefill_tc5b.cuh"
#include "gemm_tc5.cuh"

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        RT_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !p) throw RtError(RT_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

template <int BN, int EPI>
void launch_gemm5(cudaStream_t st, bool pdl, const CUtensorMap& ma, const CUtensorMap& mw, const Gemm5Args& a) {
    const size_t smem = gemm5_smem_bytes(BN);
    RT_CUDA(cudaFuncSetAttribute(gemm_tc5_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    // token blocks fastest: the CTAs in flight (one per SM, dispatched in x-then-y order) then cover ALL token blocks of a
    // few weight-row blocks, so a weight tile comes from HBM once and from L2 for the other token blocks, and the
    // activations (<= 2,048 x K, L2-resident) are what every CTA shares.  Weight blocks fastest re-streamed an 8B
    // gate/up matrix (235 MB > L2) from HBM once per token block.
    // persistent: one CTA per SM walks the tiles (token blocks fastest, see above) with two accumulators in TMEM
    static int sm_count = 0;
    if (!sm_count) { int dev = 0; RT_CUDA(cudaGetDevice(&dev)); RT_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev)); }
    cfg.gridDim = dim3(std::min(cdiv(a.S, G5_BM) * cdiv(a.N, BN), sm_count));
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    RT_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc5_kernel<BN, EPI>, ma, mw, a));
}

}  // namespace

// A operand: bf16 [rows][K] row-major; box = 128 rows x 64 k, 128-byte swizzle
void pf_make_map_a(PfTensorMap* out, const void* base, int rows, int K) {
    static_assert(sizeof(PfTensorMap) == sizeof(CUtensorMap), "PfTensorMap must mirror CUtensorMap");
    RT_REQUIRE(K % G5_BK == 0, "prefill GEMM needs K to be a multiple of 64");
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {G5_BK, G5_BM};
    const cuuint32_t es[2] = {1, 1};
    const CUresult r = encode_fn()(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                                   strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw RtError(RT_ERR_CUDA, "cuTensorMapEncodeTiled failed for an activation matrix (code " + std::to_string((int)r) + ")");
}

// W operand: bf16 CM layout [N/16 tiles][K/32 k-blocks][2 halves][256 elems]; viewed as a 4-D tensor
// (256 elems, k-block, half, tile) so that one box lands as [tile][half][k-block][256] = the no-swizzle K-major
// canonical layout with 8-row groups 1024 B apart (BK = 64)
void pf_make_map_w(PfTensorMap* out, const void* base, int N, int K, int BN) {
    RT_REQUIRE(K % G5_BK == 0 && N % 16 == 0, "prefill GEMM needs K % 64 == 0 and N % 16 == 0");
    const cuuint64_t nkb = K / 32;
    const cuuint64_t dims[4] = {256, nkb, 2, (cuuint64_t)(N / 16)};
    const cuuint64_t strides[3] = {1024, 512, nkb * 1024};
    const cuuint32_t box[4] = {256, 2, 2, (cuuint32_t)(BN / 16)};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    const CUresult r = encode_fn()(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims,
                                   strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw RtError(RT_ERR_CUDA, "cuTensorMapEncodeTiled failed for a weight matrix (code " + std::to_string((int)r) + ")");
}

void pf_gemm(cudaStream_t st, bool pdl, int epi, const PfTensorMap& ma, const PfTensorMap& mw, const Gemm5Args& a) {
    const CUtensorMap& A = reinterpret_cast<const CUtensorMap&>(ma);
    const CUtensorMap& W = reinterpret_cast<const CUtensorMap&>(mw);
    RT_REQUIRE(a.K % G5_BK == 0 && a.N % 32 == 0, "prefill GEMM needs K % 64 == 0 and N % 32 == 0");
    switch (epi) {
        case G5_EPI_QKV: launch_gemm5<PF_BN, G5_EPI_QKV>(st, pdl, A, W, a); break;
        case G5_EPI_RESID: launch_gemm5<PF_BN, G5_EPI_RESID>(st, pdl, A, W, a); break;
        case G5_EPI_SWIGLU: launch_gemm5<PF_BN, G5_EPI_SWIGLU>(st, pdl, A, W, a); break;
        case G5_EPI_STORE16: launch_gemm5<PF_BN, G5_EPI_STORE16>(st, pdl, A, W, a); break;
        default: launch_gemm5<PF_BN, G5_EPI_STORE>(st, pdl, A, W, a); break;
    }
}

template <int HD>
static void launch_pfa(cudaStream_t st, bool pdl, int n_items, const PfAttnArgs& a) {
    const size_t smem = pfa_smem_bytes(HD);
    RT_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_items);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    RT_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_kernel<HD>, a));
}

void pf_attention(cudaStream_t st, bool pdl, int hd, int n_items, const PfAttnArgs& a) {
    if (n_items <= 0) return;
    if (hd == 128) launch_pfa<128>(st, pdl, n_items, a);
    else if (hd == 64) launch_pfa<64>(st, pdl, n_items, a);
    else if (hd == 32) launch_pfa<32>(st, pdl, n_items, a);
    else throw RtError(RT_ERR_INVALID, "head_dim must be 32, 64 or 128");
}

// ---- tcgen05 prefill attention (head_dim 128) ------------------------------------------------------------------
// Q: bf16 [tokens][n_q][128] viewed as (128 dims, n_q heads, tokens); box = 64 dims x G heads x BT tokens, 128-byte swizzle:
// lands as rows (token, head-in-group) of 128 bytes = the K-major operand of the QK^T product
void pf_make_map_q3(PfTensorMap* out, const void* base, int tokens, int n_q, int G, int BT, int hd) {
    const cuuint64_t dims[3] = {(cuuint64_t)hd, (cuuint64_t)n_q, (cuuint64_t)tokens};
    const cuuint64_t strides[2] = {(cuuint64_t)hd * 2, (cuuint64_t)n_q * hd * 2};
    const cuuint32_t box[3] = {64, (cuuint32_t)G, (cuuint32_t)BT};
    const cuuint32_t es[3] = {1, 1, 1};
    const CUresult r = encode_fn()(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims,
                                   strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw RtError(RT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the prefill queries (code " + std::to_string((int)r) + ")");
}
// K or V pool of all layers: bf16 [rows][128], rows = ((layer * pages + page) * kv heads + head) * 16 + key; box = one page of
// one head, half the head dimension
void pf_make_map_kv(PfTensorMap* out, const void* base, long long rows, int hd) {
    const cuuint64_t dims[2] = {(cuuint64_t)hd, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)hd * 2};
    const cuuint32_t box[2] = {64, KV_PAGE};
    const cuuint32_t es[2] = {1, 1};
    const CUresult r = encode_fn()(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                                   strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw RtError(RT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the KV pool (code " + std::to_string((int)r) + ")");
}

void pf_attention_tc5(cudaStream_t st, bool pdl, int n_items, const PfTensorMap& mq, const PfTensorMap& mk, const PfTensorMap& mv,
                      const int* page_table, int pt_stride, const int* tok_pos, const int4* work, int n_q, int n_kv, int G, int BT,
                      int n_pages_pool, int layer, float scale_log2, f16* out, int hd) {
    if (n_items <= 0) return;
    Pf5Args a{};
    a.page_table = page_table; a.pt_stride = pt_stride; a.tok_pos = tok_pos; a.work = work; a.n_q = n_q; a.n_kv = n_kv; a.G = G; a.BT = BT;
    a.n_pages_pool = n_pages_pool; a.layer = layer; a.scale_log2 = scale_log2; a.out = out;
    const size_t smem = p6_smem_bytes(hd);
    auto kern = hd == 128 ? attn_prefill_tc5b_kernel<128> : attn_prefill_tc5b_kernel<64>;
    RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_items);
    cfg.blockDim = dim3(P5_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    RT_CUDA(cudaLaunchKernelEx(&cfg, kern, reinterpret_cast<const CUtensorMap&>(mq), reinterpret_cast<const CUtensorMap&>(mk),
                               reinterpret_cast<const CUtensorMap&>(mv), a));
}

void pf_attention_dec5(cudaStream_t st, bool pdl, int n_items, const PfTensorMap& mq, const PfTensorMap& mk, const PfTensorMap& mv,
                       const Dec5Args& a, const Dec5Item* host_items, int hd) {
    if (n_items <= 0) return;
    RT_REQUIRE(n_items <= D5_MAX_ITEMS, "too many work items for the decode attention kernel");
    static thread_local Dec5ItemList list;            // kernel parameter: copied at launch (or into the graph node at capture)
    memcpy(list.it, host_items, (size_t)n_items * sizeof(Dec5Item));
    const size_t smem = p6_smem_bytes(hd) + (D5_PARTS - 2) * 512;   // the exchange area holds D5_PARTS values per row
    auto kern = hd == 128 ? attn_decode_tc5_kernel<128> : attn_decode_tc5_kernel<64>;
    RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_items);
    cfg.blockDim = dim3(D5_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    RT_CUDA(cudaLaunchKernelEx(&cfg, kern, reinterpret_cast<const CUtensorMap&>(mq), reinterpret_cast<const CUtensorMap&>(mk),
                               reinterpret_cast<const CUtensorMap&>(mv), a, list));
}

#ifdef RT_TRACE
#include <cstdio>
#include <vector>
void pf_trace_begin() {
    RtTraceRec* buf = nullptr;
    cudaMalloc(&buf, (size_t)RT_TRACE_CAP * sizeof(RtTraceRec));
    unsigned zero = 0;
    cudaMemcpyToSymbol(g_rt_trace_n, &zero, sizeof zero);
    cudaMemcpyToSymbol(g_rt_trace, &buf, sizeof buf);
}
void pf_trace_end(FILE* f) {
    RtTraceRec* buf = nullptr;
    unsigned n = 0;
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&buf, g_rt_trace, sizeof buf);
    cudaMemcpyFromSymbol(&n, g_rt_trace_n, sizeof n);
    n = n < RT_TRACE_CAP ? n : RT_TRACE_CAP;
    std::vector<RtTraceRec> h(n);
    cudaMemcpy(h.data(), buf, (size_t)n * sizeof(RtTraceRec), cudaMemcpyDeviceToHost);
    fwrite(h.data(), sizeof(RtTraceRec), n, f);
    RtTraceRec* null = nullptr;
    cudaMemcpyToSymbol(g_rt_trace, &null, sizeof null);
    cudaFree(buf);
}
#endif
