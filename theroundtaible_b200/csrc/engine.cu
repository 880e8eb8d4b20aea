// Host side of the generation backend: engine construction (weights, rope tables, paged KV pool),
// the forward step (kernel sequence of one decoder pass over <= 32 token slots), batch generation
// with a shared prefix, teacher forcing.  Replaces the server behind
// /root/reference/src/adapters/local-llm.ts:150-199 (POST /v1/chat/completions).
#include "engine.cuh"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <cuda_profiler_api.h>
#include <memory>

#include "attn_decode3.cuh"
#include "attn_decode_tc5.cuh"
#include "kernels.cuh"
#include "safetensors.h"
#include "sampler.cuh"
#include "tp_fused.cuh"
#include "weights.cuh"

#ifdef RT_TRACE
// the trace symbols are per translation unit: prefill.cu (tcgen05 kernels) keeps its own buffer
void pf_trace_begin();
void pf_trace_end(FILE* f);
#endif

namespace {

template <typename T>
T* dalloc(rt_engine* e, size_t n) {
    void* p = nullptr;
    cudaError_t er = cudaMalloc(&p, n * sizeof(T));
    if (er != cudaSuccess) {
        cudaGetLastError();
        throw RtError(RT_ERR_OOM, std::string("out of HBM allocating ") + std::to_string(n * sizeof(T)) + " bytes: " +
                                      cudaGetErrorString(er));
    }
    e->allocs.push_back(p);
    return reinterpret_cast<T*>(p);
}

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

void gen_plain(rt_engine* e, void* out, int rows, int cols, int tid, long long lrow0, long long lcols, long long lcol0,
               float std_, float offset, bool cm, bool as_f32 = false) {
    wgen::GenDesc d{};
    d.map = 0;
    d.nseg = 1;
    d.seg[0] = {0, rows, wgen::stream_key(e->cfg.seed, tid), lrow0, lcols, lcol0};
    d.rows = rows;
    d.cols = cols;
    d.cm = cm ? 1 : 0;
    d.scale = (float)((double)std_ / wgen::IH4_STD);
    d.offset = offset;
    d.as_f32 = as_f32 ? 1 : 0;
    wgen::launch_gen(d, out, e->st);
}

void build_rope_tables(rt_engine* e) {
    const rt_config& c = e->cfg;
    const int half = c.head_dim / 2;
    std::vector<float> inv(half);
    for (int i = 0; i < half; ++i) {
        double f = 1.0 / std::pow((double)c.rope_theta, (double)(2 * i) / (double)c.head_dim);
        if (c.rope_factor > 0.f) {
            const double low_wl = (double)c.rope_orig_ctx / (double)c.rope_low_freq;
            const double high_wl = (double)c.rope_orig_ctx / (double)c.rope_high_freq;
            const double wl = 2.0 * M_PI / f;
            double fl = wl > low_wl ? f / (double)c.rope_factor : f;
            const double smooth = ((double)c.rope_orig_ctx / wl - (double)c.rope_low_freq) /
                                  ((double)c.rope_high_freq - (double)c.rope_low_freq);
            const double smoothed = (1.0 - smooth) * fl / (double)c.rope_factor + smooth * fl;
            const bool medium = !(wl < high_wl) && !(wl > low_wl);
            f = medium ? smoothed : fl;
        }
        inv[i] = (float)f;
    }
    const size_t n = (size_t)c.max_ctx * half;
    std::vector<float> hc(n), hs(n);
    for (int p = 0; p < c.max_ctx; ++p)
        for (int i = 0; i < half; ++i) {
            volatile float ang = (float)p * inv[i];
            hc[(size_t)p * half + i] = (float)std::cos((double)ang);
            hs[(size_t)p * half + i] = (float)std::sin((double)ang);
        }
    e->cos_t = dalloc<float>(e, n);
    e->sin_t = dalloc<float>(e, n);
    RT_CUDA(cudaMemcpy(e->cos_t, hc.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    RT_CUDA(cudaMemcpy(e->sin_t, hs.data(), n * sizeof(float), cudaMemcpyHostToDevice));
}

// All projection matrices live in the CM layout of gemv.cuh (one copy serves decode and prefill).
void build_weights(rt_engine* e) {
    using namespace wgen;
    const int H = e->H, hd = e->hd, rank = e->cfg.tp_rank;
    const unsigned long long seed = e->cfg.seed;
    e->embed = dalloc<bf16>(e, (size_t)e->vocab * H);
    gen_plain(e, e->embed, e->vocab, H, tensor_id(EMBED, -1), 0, H, 0, 0.02f, 0.f, true);
    if (e->cfg.tied && e->cfg.tp_size == 1) {
        e->lm_head = e->embed;
    } else {
        e->lm_head = dalloc<bf16>(e, (size_t)e->vocab_l * H);
        gen_plain(e, e->lm_head, e->vocab_l, H, tensor_id(e->cfg.tied ? EMBED : LM_HEAD, -1), (long long)rank * e->vocab_l, H, 0,
                  0.02f, 0.f, true);
    }
    e->final_norm = dalloc<bf16>(e, H);
    gen_plain(e, e->final_norm, 1, H, tensor_id(FINAL_NORM, -1), 0, H, 0, 0.02f, 1.0f, false);

    e->layers.resize(e->L);
    const float sc = (float)(0.02 / IH4_STD);
    for (int l = 0; l < e->L; ++l) {
        LayerW& w = e->layers[l];
        const int qr = e->nq_l * hd, kr = e->nkv_l * hd;
        w.wqkv = dalloc<bf16>(e, (size_t)e->qkv_rows * H);
        {
            GenDesc d{};
            d.map = 2;
            d.hd = hd;
            d.cm = 1;
            d.nseg = 3;
            d.seg[0] = {0, qr, stream_key(seed, tensor_id(Q, l)), (long long)rank * qr, H, 0};
            d.seg[1] = {qr, kr, stream_key(seed, tensor_id(K, l)), (long long)rank * kr, H, 0};
            d.seg[2] = {qr + kr, kr, stream_key(seed, tensor_id(V, l)), (long long)rank * kr, H, 0};
            d.rows = e->qkv_rows;
            d.cols = H;
            d.scale = sc;
            launch_gen(d, w.wqkv, e->st);
        }
        if (e->cfg.qkv_bias) {
            // bias as a [rows][8] fp32 matrix whose row map equals wqkv's; column 0 is the value (lcols = 1 makes the
            // logical index the row itself), compacted below
            w.bqkv = dalloc<float>(e, e->qkv_rows);
            float* tmp = dalloc<float>(e, (size_t)e->qkv_rows * 8);
            GenDesc d{};
            d.map = 2;
            d.hd = hd;
            d.cm = 0;
            d.nseg = 3;
            d.seg[0] = {0, qr, stream_key(seed, tensor_id(Q_BIAS, l)), (long long)rank * qr, 1, 0};
            d.seg[1] = {qr, kr, stream_key(seed, tensor_id(K_BIAS, l)), (long long)rank * kr, 1, 0};
            d.seg[2] = {qr + kr, kr, stream_key(seed, tensor_id(V_BIAS, l)), (long long)rank * kr, 1, 0};
            d.rows = e->qkv_rows;
            d.cols = 8;
            d.scale = sc;
            d.as_f32 = 1;
            launch_gen(d, tmp, e->st);
            RT_CUDA(cudaMemcpy2DAsync(w.bqkv, sizeof(float), tmp, 8 * sizeof(float), sizeof(float), e->qkv_rows,
                                      cudaMemcpyDeviceToDevice, e->st));
        }
        w.wo = dalloc<bf16>(e, (size_t)H * qr);
        gen_plain(e, w.wo, H, qr, tensor_id(O, l), 0, (long long)e->nq * hd, (long long)rank * qr, 0.02f, 0.f, true);
        w.wgu = dalloc<bf16>(e, (size_t)2 * e->F_l * H);
        {
            GenDesc d{};
            d.map = 1;
            d.cm = 1;
            d.nseg = 2;
            d.seg[0] = {0, 0, stream_key(seed, tensor_id(GATE, l)), (long long)rank * e->F_l, H, 0};
            d.seg[1] = {0, 0, stream_key(seed, tensor_id(UP, l)), (long long)rank * e->F_l, H, 0};
            d.rows = 2 * e->F_l;
            d.cols = H;
            d.scale = sc;
            launch_gen(d, w.wgu, e->st);
        }
        w.wdown = dalloc<bf16>(e, (size_t)H * e->F_l);
        gen_plain(e, w.wdown, H, e->F_l, tensor_id(DOWN, l), 0, e->F, (long long)rank * e->F_l, 0.02f, 0.f, true);
        w.norm1 = dalloc<bf16>(e, H);
        w.norm2 = dalloc<bf16>(e, H);
        gen_plain(e, w.norm1, 1, H, tensor_id(IN_NORM, l), 0, H, 0, 0.02f, 1.0f, false);
        gen_plain(e, w.norm2, 1, H, tensor_id(POST_NORM, l), 0, H, 0, 0.02f, 1.0f, false);
    }
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaStreamSynchronize(e->st));
}

// ---- checkpoint weights (rt_config.weights_path) ------------------------------------------------------
// The Hugging Face tensors of a Llama / Qwen2 checkpoint go through the SAME layout kernel as the seeded generator
// (weights.cuh: row maps + CM layout); only the value source differs.  Each tensor -- for tensor parallel only this
// rank's rows (column-parallel matrices, lm_head) or columns (row-parallel matrices) -- is copied from the file
// mapping into a staging buffer in HBM and laid out from there.
namespace {
struct Stager {
    rt_engine* e;
    unsigned char* buf = nullptr;
    size_t cap = 0, used = 0;
    void reserve(size_t bytes) {
        RT_CUDA(cudaMalloc(&buf, bytes));
        cap = bytes;
    }
    void reset() {
        RT_CUDA(cudaStreamSynchronize(e->st));     // the layout kernel of the previous matrix has read its staging
        used = 0;
    }
    // rows [r0, r0+nr) x columns [c0, c0+nc) of a row-major [R][C] tensor -> contiguous [nr][nc] in the staging buffer
    const void* put(const st::Tensor& t, long long r0, long long nr, long long c0, long long nc) {
        const long long C = t.shape.size() == 2 ? t.shape[1] : t.shape[0];
        const size_t es = (size_t)t.elem_size();
        const size_t bytes = ((size_t)nr * nc * es + 255) & ~(size_t)255;
        if (used + bytes > cap) throw RtError(RT_ERR_IO, "checkpoint staging buffer too small");
        unsigned char* dst = buf + used;
        used += bytes;
        const unsigned char* src = t.data + ((size_t)r0 * C + c0) * es;
        // on the engine's stream: a legacy-stream cudaMemcpy from pageable memory may return before its DMA has landed
        // and is not ordered against the (non-blocking) stream the layout kernel runs on
        if (nc == C) RT_CUDA(cudaMemcpyAsync(dst, src, (size_t)nr * nc * es, cudaMemcpyHostToDevice, e->st));
        else RT_CUDA(cudaMemcpy2DAsync(dst, (size_t)nc * es, src, (size_t)C * es, (size_t)nc * es, (size_t)nr, cudaMemcpyHostToDevice, e->st));
        return dst;
    }
    ~Stager() { if (buf) cudaFree(buf); }
};
int st_dtype(const st::Tensor& t) { return t.dtype == "BF16" ? 0 : t.dtype == "F16" ? 1 : 2; }

// one logical tensor slice -> a matrix in row-major or CM layout
void load_plain(rt_engine* e, Stager& sg, const st::Tensor& t, void* out, long long r0, int rows, long long c0, int cols, bool cm) {
    sg.reset();
    wgen::GenDesc d{};
    d.map = 0;
    d.nseg = 1;
    d.seg[0] = {0, rows, 0ull, 0, cols, 0, sg.put(t, r0, rows, c0, cols)};
    d.rows = rows;
    d.cols = cols;
    d.cm = cm ? 1 : 0;
    d.src_dtype = st_dtype(t);
    wgen::launch_gen(d, out, e->st);
}
float st_scalar(const st::Tensor& t, long long i) {
    if (t.dtype == "F32") { float v; memcpy(&v, t.data + 4 * i, 4); return v; }
    uint16_t h; memcpy(&h, t.data + 2 * i, 2);
    if (t.dtype == "BF16") { const uint32_t u = (uint32_t)h << 16; float v; memcpy(&v, &u, 4); return v; }
    return __half2float(__ushort_as_half(h));
}
}  // namespace

void load_weights(rt_engine* e) {
    using namespace wgen;
    const rt_config& c = e->cfg;
    const int H = e->H, hd = e->hd, rank = c.tp_rank;
    const long long V = e->vocab, F = e->F, QR = (long long)e->nq * hd, KR = (long long)e->nkv * hd;
    std::unique_ptr<st::Files> files;
    try {
        files.reset(new st::Files(c.weights_path));
    } catch (const std::exception& ex) {
        throw RtError(RT_ERR_INVALID, ex.what());
    }
    auto get = [&](const std::string& name, std::initializer_list<long long> shape) -> const st::Tensor& {
        try {
            return files->get(name, shape);
        } catch (const std::exception& ex) {
            throw RtError(RT_ERR_IO, ex.what());
        }
    };
    Stager sg{e};
    {
        const size_t qkv = (size_t)e->qkv_rows * H, gu = (size_t)2 * e->F_l * H, emb = (size_t)V * H;
        sg.reserve(std::max({qkv, gu, emb}) * 4 + 4096);      // fp32 sources at worst
    }
    const st::Tensor& emb = get("model.embed_tokens.weight", {V, H});
    e->embed = dalloc<bf16>(e, (size_t)V * H);
    load_plain(e, sg, emb, e->embed, 0, (int)V, 0, H, true);
    const st::Tensor* head = files->find("lm_head.weight");
    if (c.tied || !head) {
        RT_REQUIRE(c.tied, "the checkpoint has no lm_head.weight but the configuration is not tied");
        if (c.tp_size == 1) {
            e->lm_head = e->embed;
        } else {
            e->lm_head = dalloc<bf16>(e, (size_t)e->vocab_l * H);
            load_plain(e, sg, emb, e->lm_head, (long long)rank * e->vocab_l, e->vocab_l, 0, H, true);
        }
    } else {
        e->lm_head = dalloc<bf16>(e, (size_t)e->vocab_l * H);
        load_plain(e, sg, get("lm_head.weight", {V, H}), e->lm_head, (long long)rank * e->vocab_l, e->vocab_l, 0, H, true);
    }
    e->final_norm = dalloc<bf16>(e, H);
    load_plain(e, sg, get("model.norm.weight", {H}), e->final_norm, 0, 1, 0, H, false);

    e->layers.resize(e->L);
    const int qr = e->nq_l * hd, kr = e->nkv_l * hd;
    for (int l = 0; l < e->L; ++l) {
        LayerW& w = e->layers[l];
        const std::string p = "model.layers." + std::to_string(l) + ".";
        {
            const st::Tensor& tq = get(p + "self_attn.q_proj.weight", {QR, H});
            const st::Tensor& tk = get(p + "self_attn.k_proj.weight", {KR, H});
            const st::Tensor& tv = get(p + "self_attn.v_proj.weight", {KR, H});
            RT_REQUIRE(tq.dtype == tk.dtype && tq.dtype == tv.dtype, "q/k/v projections of one layer must share a dtype");
            sg.reset();
            w.wqkv = dalloc<bf16>(e, (size_t)e->qkv_rows * H);
            GenDesc d{};
            d.map = 2; d.hd = hd; d.cm = 1; d.nseg = 3;
            d.seg[0] = {0, qr, 0ull, 0, H, 0, sg.put(tq, (long long)rank * qr, qr, 0, H)};
            d.seg[1] = {qr, kr, 0ull, 0, H, 0, sg.put(tk, (long long)rank * kr, kr, 0, H)};
            d.seg[2] = {qr + kr, kr, 0ull, 0, H, 0, sg.put(tv, (long long)rank * kr, kr, 0, H)};
            d.rows = e->qkv_rows; d.cols = H; d.src_dtype = st_dtype(tq);
            launch_gen(d, w.wqkv, e->st);
        }
        if (c.qkv_bias) {
            const st::Tensor* tb[3] = {&get(p + "self_attn.q_proj.bias", {QR}), &get(p + "self_attn.k_proj.bias", {KR}),
                                       &get(p + "self_attn.v_proj.bias", {KR})};
            GenDesc d{};                                            // the row map only (host)
            d.map = 2; d.hd = hd; d.nseg = 3;
            d.seg[0] = {0, qr, 0ull, (long long)rank * qr, 1, 0, nullptr};
            d.seg[1] = {qr, kr, 1ull, (long long)rank * kr, 1, 0, nullptr};
            d.seg[2] = {qr + kr, kr, 2ull, (long long)rank * kr, 1, 0, nullptr};
            std::vector<float> hb(e->qkv_rows);
            for (int r = 0; r < e->qkv_rows; ++r) {
                unsigned long long which; long long lrow, lcols, lcol0; const void* src;
                map_row(d, r, which, lrow, lcols, lcol0, src);
                hb[r] = __bfloat162float(__float2bfloat16_rn(st_scalar(*tb[which], lrow)));
            }
            w.bqkv = dalloc<float>(e, e->qkv_rows);
            RT_CUDA(cudaMemcpyAsync(w.bqkv, hb.data(), hb.size() * sizeof(float), cudaMemcpyHostToDevice, e->st));
            RT_CUDA(cudaStreamSynchronize(e->st));                   // hb goes out of scope
        }
        w.wo = dalloc<bf16>(e, (size_t)H * qr);
        load_plain(e, sg, get(p + "self_attn.o_proj.weight", {H, QR}), w.wo, 0, H, (long long)rank * qr, qr, true);
        {
            const st::Tensor& tg = get(p + "mlp.gate_proj.weight", {F, H});
            const st::Tensor& tu = get(p + "mlp.up_proj.weight", {F, H});
            RT_REQUIRE(tg.dtype == tu.dtype, "gate/up projections of one layer must share a dtype");
            sg.reset();
            w.wgu = dalloc<bf16>(e, (size_t)2 * e->F_l * H);
            GenDesc d{};
            d.map = 1; d.cm = 1; d.nseg = 2;
            d.seg[0] = {0, 0, 0ull, 0, H, 0, sg.put(tg, (long long)rank * e->F_l, e->F_l, 0, H)};
            d.seg[1] = {0, 0, 0ull, 0, H, 0, sg.put(tu, (long long)rank * e->F_l, e->F_l, 0, H)};
            d.rows = 2 * e->F_l; d.cols = H; d.src_dtype = st_dtype(tg);
            launch_gen(d, w.wgu, e->st);
        }
        w.wdown = dalloc<bf16>(e, (size_t)H * e->F_l);
        load_plain(e, sg, get(p + "mlp.down_proj.weight", {H, F}), w.wdown, 0, H, (long long)rank * e->F_l, e->F_l, true);
        w.norm1 = dalloc<bf16>(e, H);
        w.norm2 = dalloc<bf16>(e, H);
        load_plain(e, sg, get(p + "input_layernorm.weight", {H}), w.norm1, 0, 1, 0, H, false);
        load_plain(e, sg, get(p + "post_attention_layernorm.weight", {H}), w.norm2, 0, 1, 0, H, false);
    }
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaStreamSynchronize(e->st));
}

#ifdef RT_TRACE
// timeline dump (RT_TRACE_FILE): records of {globaltimer ns, tag, cta} written by thread 0 of every CTA
static void trace_begin() {
    pf_trace_begin();
    RtTraceRec* buf = nullptr;
    cudaMalloc(&buf, (size_t)RT_TRACE_CAP * sizeof(RtTraceRec));
    unsigned zero = 0;
    cudaMemcpyToSymbol(g_rt_trace_n, &zero, sizeof zero);
    cudaMemcpyToSymbol(g_rt_trace, &buf, sizeof buf);
}
static void trace_end(const char* path) {
    RtTraceRec* buf = nullptr;
    unsigned n = 0;
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&buf, g_rt_trace, sizeof buf);
    cudaMemcpyFromSymbol(&n, g_rt_trace_n, sizeof n);
    n = std::min(n, RT_TRACE_CAP);
    std::vector<RtTraceRec> h(n);
    cudaMemcpy(h.data(), buf, (size_t)n * sizeof(RtTraceRec), cudaMemcpyDeviceToHost);
    if (FILE* f = fopen(path, "wb")) { fwrite(h.data(), sizeof(RtTraceRec), n, f); pf_trace_end(f); fclose(f); }
    RtTraceRec* null = nullptr;
    cudaMemcpyToSymbol(g_rt_trace, &null, sizeof null);
    cudaFree(buf);
}
#endif

// ---- launch helpers -----------------------------------------------------------------------------------
// Every kernel of the step is launched with the programmatic-stream-serialization attribute (PDL): its CTAs may
// become resident while the previous kernel drains, issue their first weight loads, and block in
// griddepcontrol.wait until the previous kernel's writes are visible.
template <typename... KArgs, typename... Args>
void launch_pdl(rt_engine* e, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = e->st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = e->use_pdl ? 1 : 0;
    RT_CUDA(cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...));
    e->launches++;
}

// cluster pair that splits K (gemv.cuh CL = 2): cluster dimension + PDL on the same launch
template <int EPI, int NW, int XM>
void gemv_launch_pair(rt_engine* e, const GemvArgs& a, int groups) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * groups);
    cfg.blockDim = dim3(NW * 32);
    cfg.stream = e->st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = e->use_pdl ? 2 : 1;
    RT_CUDA(cudaLaunchKernelEx(&cfg, gemv_kernel<1, 1, EPI, NW, XM, 2>, a));
    e->launches++;
}

template <int EPI, int NT, int MT, int NW>
void gemv_launch(rt_engine* e, const GemvArgs& a, int grid) {
    // fp16 twins (gemv.cuh mode 2); otherwise, with <= 4 slots, the hi and lo parts of the activations share one MMA (mode 1)
    if (NT == 1 && a.f16) launch_pdl(e, gemv_kernel<1, MT, EPI, NW, 2>, dim3(grid), dim3(NW * 32), 0, a);
    else if (NT == 1 && a.T <= 4) launch_pdl(e, gemv_kernel<1, MT, EPI, NW, 1>, dim3(grid), dim3(NW * 32), 0, a);
    else launch_pdl(e, gemv_kernel<NT, MT, EPI, NW, 0>, dim3(grid), dim3(NW * 32), 0, a);
}

// 16 warps only for very long rows (down-projection, K = FFN >= 8192): measured on B200 (scripts/gemv_tune*.py)
// 16 warps split a long K (each warp then issues enough batches to hide the latency of the others) -- and also a
// matrix with fewer row tiles than SMs (tensor-parallel QKV / O shards): with one CTA on most SMs, twice the warps means
// twice the loads in flight per SM
inline int gemv_warps(int K, int tiles = 1 << 30) {
    static const int force = env_int("RT_GEMV_NW", 0), few = env_int("RT_GEMV_NW16_TILES", 0);
    return force ? force : ((K >= 8192 || (tiles <= few && K >= 2048)) ? 16 : 8);
}

template <int EPI, int NT, int MT>
void gemv_pick_warps(rt_engine* e, const GemvArgs& a, int grid) {
    if (gemv_warps(a.K, a.N / 16) == 16) gemv_launch<EPI, NT, MT, 16>(e, a, grid);
    else gemv_launch<EPI, NT, MT, 8>(e, a, grid);
}

template <int EPI>
int gemv_dispatch(rt_engine* e, const GemvArgs& a) {
    // few row tiles (the QKV projection of a tensor-parallel shard): a cluster pair of 8-warp CTAs per 16-row tile splits K,
    // as long as all pairs are resident at once (two such CTAs per SM)
    static const int pair_on = env_int("RT_GEMV_PAIR", 1);
    // measured on one GPU with the shards of tensor-parallel models (profiles/r2_summary.md): 32B/TP4 5.42 -> 5.06 ms/step,
    // 8B/TP4 1.79 -> 1.73, 8B/TP8 1.41 -> 1.38; NOT where a tile already has 16 warps (K >= 8192: 70B/TP8 5.47 -> 5.71)
    if (EPI == EPI_QKV && pair_on && a.T <= 8 && a.N / 16 <= e->sm_count && (a.K / 32) % 2 == 0 && a.K / 64 >= 8 &&
        gemv_warps(a.K, a.N / 16) == 8) {
        if (a.f16) gemv_launch_pair<EPI_QKV, 8, 2>(e, a, a.N / 16);
        else if (a.T <= 4) gemv_launch_pair<EPI_QKV, 8, 1>(e, a, a.N / 16);
        else gemv_launch_pair<EPI_QKV, 8, 0>(e, a, a.N / 16);
        return a.N / 16;
    }
    // more tiles than two 8-warp CTAs per SM hold, no more than three do (Llama-3.1-8B QKV: 384 tiles; Qwen2.5-32B/TP4 down: 320):
    // 16-row CTAs at three per SM, all resident at once, instead of 32-row CTAs that put 4 tiles on some SMs and 2 on others
    // (gemv.cuh OC; config 3 -1.0..-1.8 % per step, 3 knights on the 8B model -4.8 %)
    static const int oc3 = env_int("RT_GEMV_OC3", 1);
    if constexpr (EPI == EPI_QKV || EPI == EPI_RESID || EPI == EPI_PUSH) {
        if (oc3 && a.T <= 8 && (a.f16 || a.T <= 4) && gemv_warps(a.K, a.N / 16) == 8 && a.N / 16 > 2 * e->sm_count &&
            a.N / 16 <= 3 * e->sm_count) {
            if (a.f16) launch_pdl(e, gemv_kernel<1, 1, EPI, 8, 2, 1, 3>, dim3(a.N / 16), dim3(256), 0, a);
            else launch_pdl(e, gemv_kernel<1, 1, EPI, 8, 1, 1, 3>, dim3(a.N / 16), dim3(256), 0, a);
            return a.N / 16;
        }
    }
    static const int force_mt = env_int("RT_GEMV_MT", 0);
    RT_REQUIRE(!a.f16 || a.T <= 8, "the fp16-twin projection mode serves <= 8 slots");
    const int NT = a.T <= 8 ? 1 : (a.T <= 16 ? 2 : 4);
    // rows per CTA: the smallest 16*MT that fits the whole matrix into ONE wave of resident CTAs (2 per SM with 8 warps,
    // 1 per SM with 16 warps); matrices that need several waves anyway take 32-row CTAs, the lm_head 64-row CTAs
    const int tiles = a.N / 16;
    static const int cap16 = env_int("RT_GEMV_CAP16", 1);
    const int capacity = e->sm_count * (gemv_warps(a.K, a.N / 16) == 16 ? cap16 : 2);
    const int mt_max = 4 / NT;
    int mt = 1;
    while (mt < mt_max && (tiles + mt - 1) / mt > capacity) mt <<= 1;
    if ((tiles + mt - 1) / mt > capacity) mt = std::min(mt_max, tiles >= 4096 ? 4 : 2);
    while (mt > 1 && a.N % (16 * mt) != 0) mt >>= 1;
    // balance: the rule above may leave SMs with very different numbers of 16-row tiles (864 tiles as 216 CTAs of 64 rows
    // on 148 SMs: 68 SMs stream twice what the other 80 do -- Qwen2.5-32B/TP4 gate/up 28.4 us against 23.7 us as a
    // persistent grid of 16-row groups, scripts/gemv_shard_tune.py).  Take the rows per CTA whose busiest SM has the fewest
    // tiles, when that is at least a fifth fewer than under the rule above (32B/TP4 shard 4.75 -> 4.58 ms per step, 32B/TP2
    // 7.56 -> 7.22; a smaller edge did not survive inside the step: Llama-3.1-8B gate/up, 16 -> 14 tiles on the busiest SM, is
    // 5 % faster alone and 1 % slower per step).
    static const int balance = env_int("RT_GEMV_BALANCE", 1);
    if (balance && tiles < 4096) {
        const int cpsm = capacity / e->sm_count;
        auto busiest = [&](int m) {
            const int groups = (tiles + m - 1) / m;
            return groups <= capacity ? cdiv(groups, e->sm_count) * m : cpsm * cdiv(groups, capacity) * m;
        };
        int best = mt;
        for (int m = mt_max; m >= 1; m >>= 1)
            if (a.N % (16 * m) == 0 && busiest(m) < busiest(best) && 5 * busiest(m) <= 4 * busiest(mt)) best = m;
        mt = best;
    }
    // experiment: instead of growing the CTA until one wave covers the matrix, keep small row groups and walk them
    static const int pmode = env_int("RT_GEMV_PMODE", 0);
    if (pmode && tiles > capacity && pmode * NT <= 4 && a.N % (16 * pmode) == 0) mt = pmode;
    if (force_mt && force_mt * NT <= 4 && a.N % (16 * force_mt) == 0) mt = force_mt;
    int grid = cdiv(a.N, 16 * mt);
    // more row groups than resident CTAs: a persistent grid of one full wave walks them (gemv.cuh), requesting the next
    // group's first weights before the current group's reduction and epilogue
    static const int persist = env_int("RT_GEMV_PERSIST", 1), persist_mt = env_int("RT_GEMV_PMT", 0);
    if (persist && grid > capacity) {
        if (persist_mt && persist_mt * NT <= 4 && a.N % (16 * persist_mt) == 0) mt = persist_mt;
        grid = std::min(cdiv(a.N, 16 * mt), capacity);
    }
    if (NT == 1) {
        if (mt == 4) gemv_pick_warps<EPI, 1, 4>(e, a, grid);
        else if (mt == 2) gemv_pick_warps<EPI, 1, 2>(e, a, grid);
        else gemv_pick_warps<EPI, 1, 1>(e, a, grid);
    } else if (NT == 2) {
        if (mt == 2) gemv_pick_warps<EPI, 2, 2>(e, a, grid);
        else gemv_pick_warps<EPI, 2, 1>(e, a, grid);
    } else {
        gemv_pick_warps<EPI, 4, 1>(e, a, grid);
    }
    return grid;
}

size_t attn_smem_bytes(int hd) { return (size_t)ATTN_WARPS * ATTN_STAGES * 2 * KV_PAGE * hd * 2; }

}  // namespace

int rt_gemv_host_dispatch(rt_engine* e, const GemvArgs& a, int epi) {
    switch (epi) {
        case EPI_STORE: return gemv_dispatch<EPI_STORE>(e, a);
        case EPI_RESID: return gemv_dispatch<EPI_RESID>(e, a);
        case EPI_SWIGLU: return gemv_dispatch<EPI_SWIGLU>(e, a);
        default: throw RtError(RT_ERR_INVALID, "unsupported epilogue for the test hook");
    }
}

void rt_attn_host_setup() {
    // opt in to > 48 KB of dynamic shared memory (per device, idempotent)
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(128)));
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(64)));
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(32)));
    RT_CUDA(cudaFuncSetAttribute(attn_decode3_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, a3_smem_bytes<128>()));
    RT_CUDA(cudaFuncSetAttribute(attn_decode3_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, a3_smem_bytes<64>()));
    RT_CUDA(cudaFuncSetAttribute(attn_decode3_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, a3_smem_bytes<32>()));
}

// version 3 (attn_decode3.cuh): one cluster of `cs` CTAs per (slot, kv head), partials merged through distributed shared
// memory; PDL attribute and cluster dimension on the same launch
template <int HD>
static void attn3_launch_hd(rt_engine* e, const Attn3Args& a, int cs) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cs, a.n_kv, a.T);
    cfg.blockDim = dim3(A3_WARPS * 32);
    cfg.dynamicSmemBytes = a3_smem_bytes<HD>();
    cfg.stream = e->st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = e->use_pdl ? 2 : 1;
    RT_CUDA(cudaLaunchKernelEx(&cfg, attn_decode3_kernel<HD>, a));
    e->launches++;
}
void rt_attn3_host_launch(rt_engine* e, const Attn3Args& a, int hd, int cs) {
    if (hd == 128) attn3_launch_hd<128>(e, a, cs);
    else if (hd == 64) attn3_launch_hd<64>(e, a, cs);
    else attn3_launch_hd<32>(e, a, cs);
}

void rt_attn_host_launch(rt_engine* e, const AttnArgs& a, int hd, int n_chunk_blocks) {
    const dim3 grid(n_chunk_blocks, a.n_kv, a.T + a.n_shared_tiles);
    const size_t smem = attn_smem_bytes(hd);
    if (hd == 128) launch_pdl(e, attn_decode_kernel<128>, grid, dim3(ATTN_WARPS * 32), smem, a);
    else if (hd == 64) launch_pdl(e, attn_decode_kernel<64>, grid, dim3(ATTN_WARPS * 32), smem, a);
    else if (hd == 32) launch_pdl(e, attn_decode_kernel<32>, grid, dim3(ATTN_WARPS * 32), smem, a);
    else throw RtError(RT_ERR_INVALID, "head_dim must be 32, 64 or 128");
}

// ===================================================================================================
void engine_init(rt_engine* e) {
    const rt_config& c = e->cfg;
    RT_REQUIRE(c.struct_size == sizeof(rt_config), "rt_config.struct_size mismatch");
    RT_REQUIRE(c.layers > 0 && c.hidden > 0 && c.n_q > 0 && c.n_kv > 0 && c.ffn > 0 && c.vocab > 0, "architecture fields must be positive");
    RT_REQUIRE(c.head_dim == 32 || c.head_dim == 64 || c.head_dim == 128, "head_dim must be 32, 64 or 128");
    RT_REQUIRE(c.n_q % c.n_kv == 0 && c.n_q / c.n_kv <= 16, "n_q must be a multiple of n_kv with ratio <= 16");
    RT_REQUIRE(c.hidden % 32 == 0 && c.ffn % 32 == 0 && (c.n_q * c.head_dim) % 32 == 0, "hidden, ffn and n_q*head_dim must be multiples of 32");
    RT_REQUIRE(c.vocab % 16 == 0 && c.ffn % 8 == 0, "vocab must be a multiple of 16");
    RT_REQUIRE(c.max_batch >= 1 && c.max_batch <= 8, "max_batch must be 1..8");
    RT_REQUIRE(c.max_ctx >= 32, "max_ctx must be >= 32");
    RT_REQUIRE(c.tp_size >= 1 && c.tp_rank >= 0 && c.tp_rank < c.tp_size, "bad tp_rank/tp_size");
    RT_REQUIRE(c.n_q % c.tp_size == 0 && c.n_kv % c.tp_size == 0 && c.ffn % (8 * c.tp_size) == 0 && c.vocab % (16 * c.tp_size) == 0,
               "n_q, n_kv, ffn/8 and vocab/16 must be divisible by tp_size");

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        throw RtError(RT_ERR_CUDA, "no CUDA device available: this backend has no CPU fallback");
    }
    RT_REQUIRE(c.device >= 0 && c.device < ndev, "device ordinal out of range");
    e->dev = c.device;
    RT_CUDA(cudaSetDevice(e->dev));
    cudaDeviceProp prop{};
    RT_CUDA(cudaGetDeviceProperties(&prop, e->dev));
    if (prop.major != 10) throw RtError(RT_ERR_CUDA, std::string("device ") + prop.name + " is not sm_100: this library is built for B200 only");
    e->sm_count = prop.multiProcessorCount;
    RT_CUDA(cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking));
    for (auto& ev : e->ev) RT_CUDA(cudaEventCreate(&ev));
    rt_attn_host_setup();
    if (c.tp_size > 1) e->tp = tp_init(c.tp_rank, c.tp_size, c.nccl_id);

    e->L = c.layers; e->H = c.hidden; e->hd = c.head_dim; e->nq = c.n_q; e->nkv = c.n_kv; e->F = c.ffn; e->vocab = c.vocab;
    e->nq_l = e->nq / c.tp_size; e->nkv_l = e->nkv / c.tp_size; e->F_l = e->F / c.tp_size; e->vocab_l = e->vocab / c.tp_size;
    e->G = e->nq / e->nkv;
    e->qkv_rows = (e->nq_l + 2 * e->nkv_l) * e->hd;
    RT_REQUIRE(e->qkv_rows % 16 == 0 && (2 * e->F_l) % 16 == 0 && e->H % 16 == 0, "projection row counts must be multiples of 16");

    e->from_checkpoint = c.weights_path && c.weights_path[0];
    RT_REQUIRE(c.n_eos >= 0 && c.n_eos <= 3, "n_eos must be 0..3");
    if (c.n_eos > 0) {
        e->n_eos = c.n_eos;
        for (int i = 0; i < 3; ++i) e->eos_ids[i] = i < c.n_eos ? c.eos_ids[i] : -1;
    } else if (e->from_checkpoint) {
        e->n_eos = 0;                  // a checkpoint without ids in its configuration never stops early
    }
    if (e->from_checkpoint) load_weights(e);
    else build_weights(e);
    e->cfg.weights_path = nullptr;     // caller-owned strings: not kept beyond creation
    e->cfg.nccl_id = nullptr;
    build_rope_tables(e);

    // KV pool
    e->n_seq = c.max_batch;
    e->max_pages_per_seq = cdiv(c.max_ctx, KV_PAGE);
    const size_t page_elems = (size_t)e->nkv_l * KV_PAGE * e->hd;
    if (c.kv_pool_bytes > 0) {
        e->n_pages = (int)(c.kv_pool_bytes / ((size_t)2 * e->L * page_elems * sizeof(bf16)));
    } else {
        e->n_pages = e->n_seq * e->max_pages_per_seq;
    }
    RT_REQUIRE(e->n_pages >= e->max_pages_per_seq, "kv_pool_bytes too small for one sequence of max_ctx tokens");
    e->layer_pool_elems = (size_t)e->n_pages * page_elems;
    e->kpool = dalloc<bf16>(e, e->layer_pool_elems * e->L);
    e->vpool = dalloc<bf16>(e, e->layer_pool_elems * e->L);
    RT_CUDA(cudaMemsetAsync(e->kpool, 0, e->layer_pool_elems * e->L * sizeof(bf16), e->st));
    RT_CUDA(cudaMemsetAsync(e->vpool, 0, e->layer_pool_elems * e->L * sizeof(bf16), e->st));
    e->page_ref.assign(e->n_pages, 0);
    e->free_pages.clear();
    for (int p = e->n_pages - 1; p >= 0; --p) e->free_pages.push_back(p);
    e->seq_pages.assign(e->n_seq, {});
    e->seq_len.assign(e->n_seq, 0);
    RT_CUDA(cudaMallocHost(&e->h_page_table, (size_t)e->n_seq * e->max_pages_per_seq * sizeof(int)));
    memset(e->h_page_table, 0, (size_t)e->n_seq * e->max_pages_per_seq * sizeof(int));
    e->d_page_table = dalloc<int>(e, (size_t)e->n_seq * e->max_pages_per_seq);
    RT_CUDA(cudaMallocHost(&e->h_stage, 8 * MAX_SLOTS * sizeof(int)));
    engine_reset_sequences(e);

    // step state
    e->d_tok = dalloc<int>(e, MAX_SLOTS);
    e->d_seq = dalloc<int>(e, MAX_SLOTS);
    e->d_pos = dalloc<int>(e, MAX_SLOTS);
    e->d_next_tok = dalloc<int>(e, MAX_SLOTS);
    e->d_shared_len = dalloc<int>(e, 1);
    e->d_step = dalloc<int>(e, 1);
    e->max_out = c.max_ctx;
    e->d_out_tokens = dalloc<int>(e, (size_t)e->n_seq * e->max_out);
    RT_CUDA(cudaMemsetAsync(e->d_shared_len, 0, sizeof(int), e->st));
    RT_CUDA(cudaMemsetAsync(e->d_step, 0, sizeof(int), e->st));

    // activations: zero-filled so that the padded slots of a GEMV hold finite values
    auto zalloc_f = [&](size_t n) { float* p = dalloc<float>(e, n); RT_CUDA(cudaMemsetAsync(p, 0, n * sizeof(float), e->st)); return p; };
    auto zalloc_b = [&](size_t n) { bf16* p = dalloc<bf16>(e, n); RT_CUDA(cudaMemsetAsync(p, 0, n * sizeof(bf16), e->st)); return p; };
    auto zalloc_h = [&](size_t n) { f16* p = dalloc<f16>(e, n); RT_CUDA(cudaMemsetAsync(p, 0, n * sizeof(f16), e->st)); return p; };
    e->x = zalloc_f((size_t)MAX_SLOTS * e->H);
    e->delta = zalloc_f((size_t)MAX_SLOTS * e->H);
    e->qkv = zalloc_f((size_t)MAX_SLOTS * e->qkv_rows);
    e->xn = zalloc_b(xf_elems(e->H));
    e->qb = zalloc_b((size_t)MAX_SLOTS * e->nq_l * e->hd);
    e->attn_bf = zalloc_b(xf_elems(e->nq_l * e->hd));
    e->act_bf = zalloc_b(xf_elems(e->F_l));
    e->part_o = zalloc_f((size_t)MAX_SLOTS * e->nq_l * MAX_CHUNKS * e->hd);
    e->part_ml = zalloc_f((size_t)MAX_SLOTS * e->nq_l * MAX_CHUNKS * 2);
    e->ss_part = zalloc_f((size_t)(e->H / 16) * 32);
    e->inv_rms = zalloc_f(MAX_SLOTS);
    e->done_counter = dalloc<int>(e, 1);
    RT_CUDA(cudaMemsetAsync(e->done_counter, 0, sizeof(int), e->st));
    e->use_pdl = getenv("RT_NO_PDL") == nullptr;
    e->pf_attn_bytes = (long long)env_int("RT_PF_ATTN_MB", 0) << 20;
    e->pf_tail_bytes = (long long)env_int("RT_PF_TAIL_MB", 0) << 20;
    e->lm_ctas = e->vocab_l / 16;
    e->cta_max = zalloc_f((size_t)e->lm_ctas * 32);
    e->cta_idx = dalloc<int>(e, (size_t)e->lm_ctas * 32);
    e->logits = zalloc_f((size_t)MAX_SLOTS * e->vocab_l);

    if (c.tp_size > 1) {
        e->loc_best = zalloc_f(64);
        e->all_best = zalloc_f((size_t)64 * c.tp_size);
        e->logits_full = zalloc_f((size_t)MAX_SLOTS * e->vocab);
        e->tp_stop = zalloc_f(1);
        e->tpf_on = env_int("RT_TP_FUSED", 1) != 0 && c.tp_size <= TPF_MAX && e->H <= 4 * TPF_THREADS * TPF_CMAX && e->H % 4 == 0;
        if (e->tpf_on) {
            // separate allocations: CUDA IPC shares whole allocations
            for (int j = 0; j < 2; ++j) e->tpf_delta[j] = zalloc_f((size_t)MAX_SLOTS * e->H);
            e->tpf_flags = dalloc<int>(e, MAX_SLOTS * TPF_CMAX * TPF_MAX);
            e->tpf_epoch = dalloc<int>(e, MAX_SLOTS * TPF_CMAX);
            RT_CUDA(cudaMemsetAsync(e->tpf_flags, 0, MAX_SLOTS * TPF_CMAX * TPF_MAX * sizeof(int), e->st));
            RT_CUDA(cudaMemsetAsync(e->tpf_epoch, 0, MAX_SLOTS * TPF_CMAX * sizeof(int), e->st));
            RT_CUDA(cudaStreamSynchronize(e->st));
            for (int j = 0; j < 2; ++j) tp_ipc_exchange(e->tp, e->tpf_delta[j], e->tpf_peer_delta[j], e->st);
            tp_ipc_exchange(e->tp, e->tpf_flags, e->tpf_peer_flags, e->st);
            e->tpl_on = env_int("RT_TP_PUSH", 1) != 0;
            if (e->tpl_on) {
                for (int j = 0; j < 2; ++j) {
                    e->tpl_recv[j] = zalloc_f((size_t)c.tp_size * MAX_SLOTS * e->H);          // zero = parity 0 = "no packet yet"
                    e->tpl_cell_epoch[j] = dalloc<int>(e, MAX_SLOTS * TPF_CMAX);
                    e->tpl_send_epoch[j] = dalloc<int>(e, MAX_SLOTS);
                    RT_CUDA(cudaMemsetAsync(e->tpl_cell_epoch[j], 0, MAX_SLOTS * TPF_CMAX * sizeof(int), e->st));
                    int init[MAX_SLOTS];
                    for (int t = 0; t < MAX_SLOTS; ++t) init[t] = j == 0 ? 1 : 0;      // site 1 is advanced to 1 before its first use
                    RT_CUDA(cudaMemcpyAsync(e->tpl_send_epoch[j], init, sizeof(init), cudaMemcpyHostToDevice, e->st));
                    RT_CUDA(cudaStreamSynchronize(e->st));
                    tp_ipc_exchange(e->tp, e->tpl_recv[j], e->tpl_peer_recv[j], e->st);
                }
            }
            // nobody may raise a flag before every rank has cleared its own: one more collective as a barrier
            tp_allreduce_sum_f32(e->tp, e->loc_best, 64, e->st);
            RT_CUDA(cudaStreamSynchronize(e->st));
        }
    }

    // tensor-core prefill path
    e->pf_cap = std::min(2048, cdiv(std::min(c.max_ctx * c.max_batch, 1 << 20), 128) * 128);
    const int bt = std::max(1, 64 / e->G);
    e->pf_work_cap = (e->pf_cap / bt + 2 * e->n_seq + 2) * e->nkv_l;
    e->pf_x = zalloc_f((size_t)e->pf_cap * e->H);
    e->pf_xn = zalloc_h((size_t)e->pf_cap * e->H);
    e->pf_q = zalloc_b((size_t)e->pf_cap * e->nq_l * e->hd);
    e->pf_attn = zalloc_h((size_t)e->pf_cap * e->nq_l * e->hd);
    e->pf_act = zalloc_h((size_t)e->pf_cap * e->F_l);
    if (c.tp_size > 1) e->pf_delta = zalloc_h((size_t)e->pf_cap * e->H);
    e->pf_inv_rms = zalloc_f(e->pf_cap);
    e->pf_tok = dalloc<int>(e, e->pf_cap);
    e->pf_seq = dalloc<int>(e, e->pf_cap);
    e->pf_pos = dalloc<int>(e, e->pf_cap);
    e->pf_work = dalloc<int4>(e, e->pf_work_cap);
    RT_CUDA(cudaMallocHost(&e->pf_host, ((size_t)3 * e->pf_cap + (size_t)4 * e->pf_work_cap) * sizeof(int)));
    e->pf_ok = e->H % 64 == 0 && (e->nq_l * e->hd) % 64 == 0 && e->F_l % 64 == 0;
    if (e->pf_ok) {
        pf_make_map_a(&e->map_xn, e->pf_xn, e->pf_cap, e->H);
        pf_make_map_a(&e->map_attn, e->pf_attn, e->pf_cap, e->nq_l * e->hd);
        pf_make_map_a(&e->map_act, e->pf_act, e->pf_cap, e->F_l);
        e->map_w.resize((size_t)4 * e->L);
        for (int l = 0; l < e->L; ++l) {
            // fp16 twins of the layer matrices (kernels.cuh bf16_to_f16_kernel): what the prefill GEMMs read
            LayerW& w = e->layers[l];
            auto twin = [&](const bf16* src, size_t n) {
                f16* t = dalloc<f16>(e, n);
                bf16_to_f16_kernel<<<e->sm_count * 4, 256, 0, e->st>>>(src, t, n / 8);
                return t;
            };
            w.wqkv16 = twin(w.wqkv, (size_t)e->qkv_rows * e->H);
            w.wo16 = twin(w.wo, (size_t)e->H * e->nq_l * e->hd);
            w.wgu16 = twin(w.wgu, (size_t)2 * e->F_l * e->H);
            w.wdown16 = twin(w.wdown, (size_t)e->H * e->F_l);
            pf_make_map_w(&e->map_w[4 * l + 0], w.wqkv16, e->qkv_rows, e->H, PF_BN);
            pf_make_map_w(&e->map_w[4 * l + 1], w.wo16, e->H, e->nq_l * e->hd, PF_BN);
            pf_make_map_w(&e->map_w[4 * l + 2], w.wgu16, 2 * e->F_l, e->H, PF_BN);
            pf_make_map_w(&e->map_w[4 * l + 3], w.wdown16, e->H, e->F_l, PF_BN);
        }
        // the decode projections read the twins too when a step has 5..8 slots (gemv.cuh mode 2): the lm_head then needs one
        e->dec_f16 = env_int("RT_DEC_F16", 1) != 0;
        if (e->dec_f16) {
            e->lm_head16 = dalloc<f16>(e, (size_t)e->vocab_l * e->H);
            bf16_to_f16_kernel<<<e->sm_count * 4, 256, 0, e->st>>>(e->lm_head, e->lm_head16, (size_t)e->vocab_l * e->H / 8);
        }
        RT_CUDA(cudaGetLastError());
        // tcgen05 prefill attention (attn_prefill_tc5.cuh): head_dim 128, a q-head group that fits 128 rows
        // head_dim 64 also runs (RT_PFA_TC5=2) but gains nothing measurable: the kernel is softmax-paced and a 64-dim head
        // only halves its MMA work (30k-token 1B prompt: 378 vs 391 ms)
        const int tc5 = env_int("RT_PFA_TC5", 1);
        e->pfa_tc5 = tc5 != 0 && (e->hd == 128 || (e->hd == 64 && tc5 == 2)) && e->G <= 16;   // RT_PFA_TC5=0: the mma.sync kernel
        // tcgen05 decode attention for long contexts (attn_decode_tc5.cuh): head_dim 64 / 128, rows = (slot, head) pairs
        e->dec5_ok = env_int("RT_DEC5", 1) != 0 && (e->hd == 128 || e->hd == 64) && e->G <= 16;
        if (e->pfa_tc5 || e->dec5_ok) {
            const long long rows = (long long)e->L * e->n_pages * e->nkv_l * KV_PAGE;
            pf_make_map_kv(&e->map_kpool, e->kpool, rows, e->hd);
            pf_make_map_kv(&e->map_vpool, e->vpool, rows, e->hd);
        }
        if (e->pfa_tc5) pf_make_map_q3(&e->map_q3, e->pf_q, e->pf_cap, e->nq_l, e->G, 128 / e->G, e->hd);
        if (e->dec5_ok) {
            e->dec5_bt = std::min(128 / e->G, MAX_SLOTS);
            pf_make_map_q3(&e->map_qdec, e->qb, MAX_SLOTS, e->nq_l, e->G, e->dec5_bt, e->hd);
            e->dec5_cap = std::min(2 * e->sm_count, D5_MAX_ITEMS);
            RT_CUDA(cudaMallocHost(&e->dec5_items_host, (size_t)e->dec5_cap * sizeof(Dec5Item)));
        }
    }
    RT_CUDA(cudaStreamSynchronize(e->st));
}

void engine_free(rt_engine* e) {
    if (e->dev >= 0) cudaSetDevice(e->dev);
    if (e->st) cudaStreamSynchronize(e->st);
    if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
    if (e->graph_exec_multi) cudaGraphExecDestroy(e->graph_exec_multi);
    if (e->tp && e->tpf_on) {
        for (int j = 0; j < 2; ++j) tp_ipc_close(e->tp, e->tpf_peer_delta[j]);
        if (e->tpl_on) for (int j = 0; j < 2; ++j) tp_ipc_close(e->tp, e->tpl_peer_recv[j]);
        tp_ipc_close(e->tp, e->tpf_peer_flags);
    }
    tp_destroy(e->tp);
    e->tp = nullptr;
    for (void* p : e->allocs) cudaFree(p);
    if (e->h_page_table) cudaFreeHost(e->h_page_table);
    if (e->h_stage) cudaFreeHost(e->h_stage);
    if (e->pf_host) cudaFreeHost(e->pf_host);
    if (e->dec5_items_host) cudaFreeHost(e->dec5_items_host);
    for (auto& ev : e->ev)
        if (ev) cudaEventDestroy(ev);
    if (e->st) cudaStreamDestroy(e->st);
}

// ---- page management ---------------------------------------------------------------------------------
static void page_release(rt_engine* e, int p) {
    if (--e->page_ref[p] == 0) e->free_pages.push_back(p);
}

// Drop every sequence's pages; pages still referenced by the prefix cache stay alive.
void engine_reset_sequences(rt_engine* e) {
    for (auto& v : e->seq_pages) {
        for (int p : v) page_release(e, p);
        v.clear();
    }
    std::fill(e->seq_len.begin(), e->seq_len.end(), 0);
    e->scratch_active = false;
}

// Evict the least recently used prefix-cache entry; false if the cache is empty.
static bool cache_evict_one(rt_engine* e) {
    if (e->cache.empty()) return false;
    size_t lru = 0;
    for (size_t i = 1; i < e->cache.size(); ++i)
        if (e->cache[i].stamp < e->cache[lru].stamp) lru = i;
    for (int p : e->cache[lru].pages) page_release(e, p);
    e->cache.erase(e->cache.begin() + lru);
    return true;
}

void engine_alloc_pages(rt_engine* e, int seq, int n_tokens_total) {
    const int need = cdiv(n_tokens_total, KV_PAGE);
    if (need > e->max_pages_per_seq)
        throw RtError(RT_ERR_OOM, "context length exceeded: sequence needs " + std::to_string(n_tokens_total) +
                                      " tokens but max_ctx is " + std::to_string(e->cfg.max_ctx));
    auto& pages = e->seq_pages[seq];
    while ((int)pages.size() < need) {
        while (e->free_pages.empty())
            if (!cache_evict_one(e)) throw RtError(RT_ERR_OOM, "context length exceeded: KV page pool exhausted");
        const int p = e->free_pages.back();
        e->free_pages.pop_back();
        e->page_ref[p] = 1;
        e->h_page_table[(size_t)seq * e->max_pages_per_seq + pages.size()] = p;
        pages.push_back(p);
    }
}

// ---- cross-call prefix cache -------------------------------------------------------
// The reference rebuilds every knight's prompt each turn; its head is constant and the transcript is append-only
// (/root/reference/src/utils/prompt.ts:60-77), so a knight's round r+1 prompt starts with its round r prompt up to
// the end of the old transcript.  Entries keep the token ids of a past prompt and the pages that hold its KV (whole
// pages only); a new prompt reuses the longest page-aligned common token prefix and prefills only the rest.
static int cache_lookup(rt_engine* e, const int32_t* prompt, int len, int* entry_out) {
    int best = 0, best_i = -1;
    for (size_t i = 0; i < e->cache.size(); ++i) {
        const auto& t = e->cache[i].tokens;
        const int lim = std::min((int)e->cache[i].pages.size() * KV_PAGE, std::min((int)t.size(), len - 1));
        int l = 0;
        while (l < lim && t[l] == prompt[l]) ++l;
        l = (l / KV_PAGE) * KV_PAGE;
        if (l > best) { best = l; best_i = (int)i; }
    }
    *entry_out = best_i;
    return best;
}

// seq (empty) takes the first n_tokens/16 pages of a cache entry
static void cache_attach(rt_engine* e, int entry, int seq, int n_tokens) {
    auto& dst = e->seq_pages[seq];
    RT_REQUIRE(dst.empty(), "cache_attach target must be empty");
    auto& ce = e->cache[entry];
    ce.stamp = ++e->cache_clock;
    for (int i = 0; i < n_tokens / KV_PAGE; ++i) {
        const int p = ce.pages[i];
        e->page_ref[p]++;
        e->h_page_table[(size_t)seq * e->max_pages_per_seq + i] = p;
        dst.push_back(p);
    }
}

// remember the prompt of seq (its KV is complete for positions < len)
static void cache_insert(rt_engine* e, int seq, const int32_t* prompt, int len) {
    const int n_pages = len / KV_PAGE;
    if (n_pages == 0) return;
    // an existing entry that already covers this prompt is only refreshed; entries this one extends are replaced
    for (size_t i = 0; i < e->cache.size();) {
        auto& ce = e->cache[i];
        const int common = std::min((int)ce.pages.size(), n_pages) * KV_PAGE;
        bool same = (int)ce.tokens.size() >= common;
        for (int l = 0; same && l < common; ++l) same = ce.tokens[l] == prompt[l];
        if (same && (int)ce.pages.size() >= n_pages) { ce.stamp = ++e->cache_clock; return; }
        if (same) {
            for (int p : ce.pages) page_release(e, p);
            e->cache.erase(e->cache.begin() + i);
        } else {
            ++i;
        }
    }
    while (e->cache.size() >= 16) cache_evict_one(e);
    rt_engine::CacheEntry ne;
    ne.tokens.assign(prompt, prompt + n_pages * KV_PAGE);
    ne.pages.assign(e->seq_pages[seq].begin(), e->seq_pages[seq].begin() + n_pages);
    for (int p : ne.pages) e->page_ref[p]++;
    ne.stamp = ++e->cache_clock;
    e->cache.push_back(std::move(ne));
}

void engine_share_prefix(rt_engine* e, int from_seq, int to_seq, int n_pages) {
    auto& dst = e->seq_pages[to_seq];
    RT_REQUIRE(dst.empty(), "share_prefix target must be empty");
    for (int i = 0; i < n_pages; ++i) {
        const int p = e->seq_pages[from_seq][i];
        e->page_ref[p]++;
        e->h_page_table[(size_t)to_seq * e->max_pages_per_seq + i] = p;
        dst.push_back(p);
    }
}

void engine_upload_page_table(rt_engine* e) {
    RT_CUDA(cudaMemcpyAsync(e->d_page_table, e->h_page_table, (size_t)e->n_seq * e->max_pages_per_seq * sizeof(int),
                            cudaMemcpyHostToDevice, e->st));
}

void engine_set_slots(rt_engine* e, const int* tok, const int* seq, const int* pos, int T) {
    // the pinned staging buffer is reused: wait for the previous upload to be consumed
    RT_CUDA(cudaStreamSynchronize(e->st));
    memcpy(e->h_stage, tok, T * sizeof(int));
    memcpy(e->h_stage + MAX_SLOTS, seq, T * sizeof(int));
    memcpy(e->h_stage + 2 * MAX_SLOTS, pos, T * sizeof(int));
    RT_CUDA(cudaMemcpyAsync(e->d_tok, e->h_stage, T * sizeof(int), cudaMemcpyHostToDevice, e->st));
    RT_CUDA(cudaMemcpyAsync(e->d_seq, e->h_stage + MAX_SLOTS, T * sizeof(int), cudaMemcpyHostToDevice, e->st));
    RT_CUDA(cudaMemcpyAsync(e->d_pos, e->h_stage + 2 * MAX_SLOTS, T * sizeof(int), cudaMemcpyHostToDevice, e->st));
}

// 5..8 token slots: the decode projections read the fp16 twins of the weights and the activations travel in the XF16 form
// (gemv.cuh mode 2); <= 4 slots keep the bf16 weights (hi and lo share one MMA there, nothing to gain)
static inline int xf_mode(const rt_engine* e, int T) { return (e->dec_f16 && T > 4 && T <= 8) ? 2 : 1; }

// row-parallel projection under tensor parallelism: partial sums -> all-reduce -> residual add -> RMSNorm.
// Fused path: ONE kernel pulls every rank's partial over NVLink peer memory (tp_fused.cuh); otherwise NCCL + rmsnorm.
static void tp_reduce_residual_norm(rt_engine* e, GemvArgs g, int site, int T, const bf16* norm_w) {
    const int H = e->H;
    if (e->tpl_on && T <= 8 && !g.bias) {
        // push form: the projection delivers its partial sums into every rank's receive buffer, the receiver does not wait for it
        g.push_tp = e->cfg.tp_size; g.push_rank = e->cfg.tp_rank; g.push_epoch = e->tpl_send_epoch[site];
        for (int p = 0; p < e->cfg.tp_size; ++p) g.push_dst[p] = reinterpret_cast<float*>(e->tpl_peer_recv[site][p]);
        gemv_dispatch<EPI_PUSH>(e, g);
        TpLLArgs a{};
        a.recv = e->tpl_recv[site]; a.cell_epoch = e->tpl_cell_epoch[site]; a.next_site_epoch = e->tpl_send_epoch[1 - site];
        a.rank = e->cfg.tp_rank; a.tp = e->cfg.tp_size; a.H = H;
        a.x = e->x; a.norm_w = norm_w; a.xn = e->xn; a.ss_part = e->ss_part; a.raw = g.f16;
        launch_pdl(e, tp_ll_reduce_norm_kernel, dim3(cdiv(H / 4, TPF_THREADS), T), dim3(TPF_THREADS), 0, a);
    } else if (e->tpf_on) {
        g.out = e->tpf_delta[site]; g.ldo = H;
        gemv_dispatch<EPI_STORE>(e, g);
        TpFusedArgs a{};
        for (int p = 0; p < e->cfg.tp_size; ++p) {
            a.partial[p] = reinterpret_cast<const float*>(e->tpf_peer_delta[site][p]);
            a.flags[p] = reinterpret_cast<int*>(e->tpf_peer_flags[p]);
        }
        a.epoch = e->tpf_epoch; a.rank = e->cfg.tp_rank; a.tp = e->cfg.tp_size; a.H = H;
        a.x = e->x; a.norm_w = norm_w; a.xn = e->xn; a.ss_part = e->ss_part; a.raw = g.f16;
        launch_pdl(e, tp_fused_allreduce_norm_kernel, dim3(cdiv(H / 4, TPF_THREADS), T), dim3(TPF_THREADS), 0, a);
    } else {
        g.out = e->delta; g.ldo = H;
        gemv_dispatch<EPI_STORE>(e, g);
        tp_allreduce_sum_f32(e->tp, e->delta, (size_t)T * H, e->st);
        rms_xw_kernel<float><<<T, 256, 0, e->st>>>(e->x, e->delta, norm_w, e->xn, g.f16 ? 2 : 1, e->inv_rms, H, e->cfg.rms_eps);
        e->launches += 2;
    }
}

// where the consuming projection finds the RMSNorm scale of e->xn's rows: per-tile partial sums (single-GPU decode:
// the producing projection left them) or r itself (tensor parallel / rows copied from a prefill chunk)
static void set_norm_scale(rt_engine* e, GemvArgs& g, bool direct) {
    g.eps = e->cfg.rms_eps;
    if (direct || (e->tp && !e->tpf_on)) g.inv_rms = e->inv_rms;
    else if (e->tp) { g.ss_in = e->ss_part; g.n_parts_in = cdiv(e->H / 4, TPF_THREADS); }   // the fused exchange's column blocks
    else { g.ss_in = e->ss_part; g.n_parts_in = e->H / 16; }
}

// lm_head over the T rows of e->xn + sampling.  Single GPU: the projection's last CTA picks the winners and advances
// the decode state.  Tensor parallel: vocab-parallel local winners -> all-gather -> tp_pick_kernel on every rank.
static void lm_head_and_sample(rt_engine* e, int T, bool sample, bool advance, bool store_logits, float inv_temp,
                               unsigned long long seed, bool direct_scale = false, int top_k = 0, float top_p = 0.f) {
    // top-k / top-p: the projection only stores the logits; sampler.cuh filters and draws (one CTA per slot)
    const bool filtered = sample && inv_temp > 0.f && ((top_k > 0 && top_k < e->vocab) || (top_p > 0.f && top_p < 1.f));
    if (filtered) {
        lm_head_and_sample(e, T, false, false, true, 0.f, seed, direct_scale);
        TopkpArgs k{};
        k.logits = e->logits; k.ld = e->vocab_l; k.V = e->vocab_l; k.T = T;
        if (e->tp) {
            // vocab-parallel shards -> full rows on every rank (one grouped all-gather); every rank then filters and draws
            // from identical inputs with the same noise stream, so all ranks pick the same token
            tp_allgather_rows(e->tp, e->logits, (size_t)e->vocab_l * sizeof(float), e->logits_full, (size_t)e->vocab * sizeof(float),
                              (size_t)e->vocab_l * sizeof(float), T, e->st);
            e->launches++;
            k.logits = e->logits_full; k.ld = e->vocab; k.V = e->vocab;
        }
        k.inv_temp = inv_temp; k.top_k = top_k; k.top_p = top_p; k.seed = seed; k.step_ptr = e->d_step;
        k.next_tok = e->d_next_tok; k.done_counter = e->done_counter;
        if (advance) { k.tok = e->d_tok; k.pos = e->d_pos; k.out_tokens = e->d_out_tokens; k.max_out = e->max_out; k.step_w = e->d_step; }
        launch_pdl(e, topkp_sample_kernel, dim3(T), dim3(TOPKP_THREADS), 0, k);
        return;
    }
    GemvArgs gl{};
    gl.f16 = xf_mode(e, T) == 2;
    gl.W = gl.f16 ? reinterpret_cast<const bf16*>(e->lm_head16) : e->lm_head; gl.N = e->vocab_l; gl.K = e->H; gl.T = T; gl.X = e->xn; gl.ldx = e->H;
    set_norm_scale(e, gl, direct_scale);
    gl.logits = store_logits ? e->logits : nullptr; gl.ld_logits = e->vocab_l;
    gl.cta_max = e->cta_max; gl.cta_idx = e->cta_idx; gl.row_offset = e->cfg.tp_rank * e->vocab_l;
    gl.inv_temp = inv_temp; gl.seed = seed; gl.step_ptr = e->d_step;
    const bool tp = e->tp != nullptr;
    if (sample) {
        gl.done_counter = e->done_counter;
        if (tp) {
            gl.next_val = e->loc_best;
            gl.next_tok = reinterpret_cast<int*>(e->loc_best) + 32;
        } else {
            gl.next_tok = e->d_next_tok;
            if (advance) { gl.tok = e->d_tok; gl.pos = e->d_pos; gl.out_tokens = e->d_out_tokens; gl.max_out = e->max_out; gl.step_w = e->d_step; }
        }
    }
    gemv_dispatch<EPI_LOGITS>(e, gl);
    if (sample && tp) {
        tp_allgather_bytes(e->tp, e->loc_best, e->all_best, 64 * sizeof(float), e->st);
        TpPickArgs p{};
        p.all_val = e->all_best; p.tp = e->cfg.tp_size; p.T = T; p.next_tok = e->d_next_tok;
        if (advance) { p.tok = e->d_tok; p.pos = e->d_pos; p.out_tokens = e->d_out_tokens; p.max_out = e->max_out; p.step_w = e->d_step; }
        launch_pdl(e, tp_pick_kernel, dim3(1), dim3(32), 0, p);
        e->launches++;   // the NCCL kernel
    }
}

// ===================================================================================================
// One decoder pass over T token slots (slot = (token, sequence, position) held on the device).
// 5 kernels per layer:  [norm + QKV + bias + RoPE + KV append] -> [attention + merge] -> [O-proj + residual]
//                       -> [norm + gate/up + SwiGLU] -> [down-proj + residual]
// ===================================================================================================
void engine_forward(rt_engine* e, const StepOpts& o) {
    const int T = o.T, H = e->H, hd = e->hd;
    RT_REQUIRE(T >= 1 && T <= MAX_SLOTS, "1..32 token slots per step");
    // RT_SKIP_MASK (profiling only, results are then wrong): leave kernel classes out of the step to measure their
    // marginal cost in the dependent chain: 2 qkv, 4 attention, 8 o-proj, 16 gate/up, 32 down, 64 lm_head, 128 sampler
    static const int skip = env_int("RT_SKIP_MASK", 0);

    launch_pdl(e, embed_norm_kernel, dim3(T), dim3(256), 0, (const bf16*)e->embed, (const int*)e->d_tok, e->x,
               (const bf16*)e->layers[0].norm1, (void*)e->xn, xf_mode(e, T), e->ss_part, e->inv_rms, H, e->vocab, e->cfg.rms_eps);
    const int raw = xf_mode(e, T) == 2;
    // the matrix a projection streams (and the one a kernel before it pulls into L2): the bf16 original or its fp16 twin
    auto wsel = [raw](const bf16* w, const f16* w16) { return raw ? reinterpret_cast<const bf16*>(w16) : w; };

    for (int l = 0; l < e->L; ++l) {
        const LayerW& w = e->layers[l];
        const bf16 *Wqkv = wsel(w.wqkv, w.wqkv16), *Wo = wsel(w.wo, w.wo16), *Wgu = wsel(w.wgu, w.wgu16), *Wdown = wsel(w.wdown, w.wdown16);
        bf16* kp = e->kpool + (size_t)l * e->layer_pool_elems;
        bf16* vp = e->vpool + (size_t)l * e->layer_pool_elems;

        static const int late = env_int("RT_QKV_LATE_LAUNCH", 0);
        GemvArgs g{};
        set_norm_scale(e, g, /*direct=*/l == 0 && e->tp != nullptr);     // tensor parallel: layer 0 takes r from the embed kernel
        g.W = Wqkv; g.f16 = raw; g.N = e->qkv_rows; g.K = H; g.T = T; g.X = e->xn; g.ldx = H;
        g.bias = w.bqkv; g.cos_t = e->cos_t; g.sin_t = e->sin_t; g.slot_seq = e->d_seq; g.slot_pos = e->d_pos;
        g.page_table = e->d_page_table; g.pt_stride = e->max_pages_per_seq; g.Q = e->qb; g.Kp = kp; g.Vp = vp;
        g.n_q = e->nq_l; g.n_kv = e->nkv_l; g.hd = hd;
        // weight stream pipelining through L2 (common.cuh): what each kernel of the layer asks for ahead of its consumer
        const size_t wo_bytes = (size_t)H * e->nq_l * hd * sizeof(bf16), wgu_bytes = (size_t)2 * e->F_l * H * sizeof(bf16);
        const size_t wd_bytes = (size_t)H * e->F_l * sizeof(bf16), wqkv_bytes = (size_t)e->qkv_rows * H * sizeof(bf16);
        const size_t attn_budget = (size_t)e->pf_attn_bytes, tail = (size_t)e->pf_tail_bytes;
        const size_t pf_wo = std::min(wo_bytes, attn_budget), pf_wgu = std::min(wgu_bytes, attn_budget - pf_wo);
        g.late_launch = o.qkv_late || late;
        if (!(skip & 2)) gemv_dispatch<EPI_QKV>(e, g);

        AttnArgs a{};
        a.Q = e->qb; a.ldq = e->nq_l * hd; a.Kp = kp; a.Vp = vp; a.page_table = e->d_page_table; a.pt_stride = e->max_pages_per_seq;
        a.slot_seq = e->d_seq; a.slot_pos = e->d_pos; a.shared_len_ptr = e->d_shared_len;
        a.T = T; a.n_q = e->nq_l; a.n_kv = e->nkv_l; a.chunk_pages = o.chunk_pages; a.n_shared_tiles = o.n_shared_tiles;
        a.part_o = e->part_o; a.part_ml = e->part_ml; a.max_chunks = MAX_CHUNKS;
        a.scale_log2 = (float)(1.4426950408889634 / std::sqrt((double)hd));
        // Attention may set up its indices and fetch KV pages of older tokens before its PDL wait when nothing it reads
        // there can still be in flight: inside the decode graph (a graph launch starts after the previous one, whose
        // sampler advanced slot_pos, has completed), or after a QKV projection that releases it late (gemv.cuh).
        static const int attn_early = env_int("RT_ATTN_EARLY", 1);
        a.early = attn_early && !(skip & 2) && (o.in_graph || late);
        a.pf[0] = {Wo, pf_wo};
        a.pf[1] = {Wgu, pf_wgu};
        if (!(skip & 4)) {
            if (o.attn3_cs > 0 && T <= 8) {
                Attn3Args a3{};
                a3.Q = a.Q; a3.ldq = a.ldq; a3.Kp = kp; a3.Vp = vp; a3.page_table = a.page_table; a3.pt_stride = a.pt_stride;
                a3.slot_seq = a.slot_seq; a3.slot_pos = a.slot_pos; a3.T = T; a3.n_q = a.n_q; a3.n_kv = a.n_kv;
                a3.scale_log2 = a.scale_log2; a3.out = e->attn_bf; a3.ldo = e->nq_l * hd; a3.early = a.early; a3.raw = raw;
                rt_attn3_host_launch(e, a3, hd, o.attn3_cs);
            } else {
            MergeArgs m{};
            m.n_sh_parts = -1;
            m.chunk_keys = o.chunk_pages * KV_PAGE;
            m.cascade = o.n_shared_tiles > 0;
            if (o.dec5_items > 0) {
                Dec5Args d{};
                d.page_table = e->d_page_table; d.pt_stride = e->max_pages_per_seq; d.slot_seq = e->d_seq; d.slot_pos = e->d_pos;
                d.shared_len_ptr = e->d_shared_len;
                d.T = T; d.n_q = e->nq_l; d.n_kv = e->nkv_l; d.G = e->G; d.BT = e->dec5_bt; d.n_pages_pool = e->n_pages; d.layer = l;
                d.scale_log2 = a.scale_log2; d.part_o = e->part_o; d.part_ml = e->part_ml; d.max_chunks = MAX_CHUNKS;
                static const int dec5_early = env_int("RT_DEC5_EARLY", 1);
                d.early = a.early && dec5_early;
                pf_attention_dec5(e->st, e->use_pdl, o.dec5_items, e->map_qdec, e->map_kpool, e->map_vpool, d,
                                  reinterpret_cast<const Dec5Item*>(e->dec5_items_host), hd);
                e->launches++;
                m.n_sh_parts = o.dec5_nsh; m.chunk_keys = o.dec5_own_span; m.cascade = 1;
            } else {
                rt_attn_host_launch(e, a, hd, o.n_chunk_blocks);
            }
            m.part_o = e->part_o; m.part_ml = e->part_ml; m.max_chunks = MAX_CHUNKS; m.slot_pos = e->d_pos;
            m.shared_len_ptr = e->d_shared_len;
            m.n_q = e->nq_l; m.hd = hd; m.out_bf = e->attn_bf; m.out_f32 = nullptr; m.ldo = e->nq_l * hd; m.raw = raw;
            const dim3 mg(e->nq_l, T);
            m.early = a.early;
            // many partials per row (long contexts): 16 warps hold all of them in flight at once
            static const int wide_from = env_int("RT_MERGE_WIDE", 24);
            const int n_parts = o.dec5_items > 0 ? o.dec5_nsh + o.dec5_nown : o.n_partials_max;
            // ... when the merge has few CTAs: a wide merge on most SMs keeps the NEXT projection's CTAs from becoming resident
            // and requesting their first weights early (config 3, 160 CTAs: 3.19 -> 3.30 ms per step with 8 or 16 warps)
            static const int wide_ctas = env_int("RT_MERGE_WIDE_CTAS", 64);
            if (n_parts >= wide_from && e->nq_l * T <= wide_ctas) {
                if (hd == 128) launch_pdl(e, attn_merge2_kernel<128, 16>, mg, dim3(512), 0, m);
                else if (hd == 64) launch_pdl(e, attn_merge2_kernel<64, 16>, mg, dim3(512), 0, m);
                else launch_pdl(e, attn_merge2_kernel<32, 16>, mg, dim3(512), 0, m);
            } else {
                if (hd == 128) launch_pdl(e, attn_merge2_kernel<128, 4>, mg, dim3(128), 0, m);
                else if (hd == 64) launch_pdl(e, attn_merge2_kernel<64, 4>, mg, dim3(128), 0, m);
                else launch_pdl(e, attn_merge2_kernel<32, 4>, mg, dim3(128), 0, m);
            }
            }
        }

        GemvArgs go{};
        go.W = Wo; go.f16 = raw; go.X = e->attn_bf; go.N = H; go.K = e->nq_l * hd; go.ldx = e->nq_l * hd; go.T = T;
        if (pf_wgu < wgu_bytes) go.pf = {reinterpret_cast<const char*>(Wgu) + pf_wgu, std::min(tail, wgu_bytes - pf_wgu)};
        if (!e->tp) {
            go.out = e->x; go.ldo = H;
            go.norm_w = w.norm2; go.xn_out = e->xn; go.ss_out = e->ss_part;
            if (!(skip & 8)) gemv_dispatch<EPI_RESID>(e, go);
        } else {
            tp_reduce_residual_norm(e, go, 0, T, w.norm2);
        }

        GemvArgs gu{};
        gu.W = Wgu; gu.f16 = raw; gu.N = 2 * e->F_l; gu.K = H; gu.T = T; gu.X = e->xn; gu.ldx = H;
        set_norm_scale(e, gu, false);
        gu.out_bf = e->act_bf; gu.ldob = e->F_l;
        gu.pf = {Wdown, std::min(tail, wd_bytes)};
        if (!(skip & 16)) gemv_dispatch<EPI_SWIGLU>(e, gu);

        GemvArgs gd{};
        gd.W = Wdown; gd.f16 = raw; gd.X = e->act_bf; gd.N = H; gd.K = e->F_l; gd.ldx = e->F_l; gd.T = T;
        const bf16* next_norm = (l + 1 < e->L) ? e->layers[l + 1].norm1 : e->final_norm;   // feeds the next layer's QKV, or the lm_head
        if (l + 1 < e->L) gd.pf = {wsel(e->layers[l + 1].wqkv, e->layers[l + 1].wqkv16), std::min(tail, wqkv_bytes)};
        else gd.pf = {wsel(e->lm_head, e->lm_head16), std::min(tail, (size_t)e->vocab_l * H * sizeof(bf16))};
        if (!e->tp) {
            gd.out = e->x; gd.ldo = H;
            gd.norm_w = next_norm; gd.xn_out = e->xn; gd.ss_out = e->ss_part;
            if (!(skip & 32)) gemv_dispatch<EPI_RESID>(e, gd);
        } else {
            tp_reduce_residual_norm(e, gd, 1, T, next_norm);
        }
    }

    if ((o.sample || o.store_logits) && !(skip & 64)) lm_head_and_sample(e, T, o.sample, o.advance, o.store_logits, o.inv_temp, o.seed, false, o.top_k, o.top_p);
    RT_CUDA(cudaGetLastError());
}

// ---- attention split geometry for a call whose longest sequence reaches max_len tokens ----------------
static void plan_chunks(int max_len, int shared_len, StepOpts& o, int hd = 64) {
    const int total_pages = cdiv(max_len, KV_PAGE);
    // measured (scripts/attn_bench.py): short chunks for head_dim 64 (latency-bound, tiny KV), 128-key chunks for 128
    o.chunk_pages = std::max(env_int("RT_ATTN_MIN_CHUNK_PAGES", hd >= 128 ? 8 : 2), cdiv(total_pages, 64));
    const int C = o.chunk_pages * KV_PAGE;
    const int n_sh = cdiv(shared_len, C);
    const int n_own = cdiv(std::max(max_len - shared_len, 1), C);
    o.n_chunk_blocks = cdiv(std::max(n_sh, n_own), ATTN_WARPS);
    o.n_chunks = std::max(n_sh, n_own);
    o.n_sh = n_sh;
    o.n_own = n_own;
    if (n_sh + n_own + 1 > MAX_CHUNKS) throw RtError(RT_ERR_INVALID, "attention split exceeds MAX_CHUNKS");
    o.n_partials_max = n_sh + n_own;
    // short contexts: the cluster-merged kernel (no merge launch, no cascade); <= RT_ATTN3_PPW pages per warp
    static const int a3_on = env_int("RT_ATTN_V3", 1), a3_ppw = env_int("RT_ATTN3_PPW", 4);
    // ... and while what the knights would re-read of a shared prefix (version 3 has no cascade phase) stays small:
    // (T - 1) x shared pages x page bytes per kv head <= RT_ATTN3_REDUNDANT_KB (measured: config 2, 3 knights x 1k shared
    // at head_dim 64 = 0.5 MB: version 3 is 9 % faster per step; config 4, 8 knights x 2k at head_dim 128 = 7 MB: 3 % slower)
    static const int a3_red_kb = env_int("RT_ATTN3_REDUNDANT_KB", 2048);
    const long long redundant = (long long)std::max(o.T - 1, 0) * (shared_len / KV_PAGE) * (2ll * KV_PAGE * hd * 2);
    o.attn3_cs = 0;
    if (a3_on && total_pages <= 8 * A3_WARPS * a3_ppw && redundant <= (long long)a3_red_kb * 1024) {
        int cs = 1;
        while (cs * A3_WARPS * a3_ppw < total_pages) cs <<= 1;
        o.attn3_cs = cs;
    }
}

// Under tensor parallel the decision is COLLECTIVE: every rank contributes "my clock says stop" and all of them stop
// if any does, at the same point of the call -- otherwise one rank would throw while its peers launch kernels that wait
// for its partial sums (tp_fused.cuh).  Hence deadline_ms must be the same on every rank.
// Work items of the tcgen05 decode attention (attn_decode_tc5.cuh) for a call whose sequences share `shared` keys and reach
// at most max_len: one CTA per item, all items within ONE wave, the key ranges sized so that every item walks about the
// same number of 64-key tiles.  Uploaded once per call; the kernels read the positions themselves.
static void plan_dec5(rt_engine* e, int max_len, int shared, StepOpts& o) {
    o.dec5_items = 0;
    if (!e->dec5_ok || o.attn3_cs > 0 || o.T > e->dec5_bt || o.T > 8) return;
    // measured (profiles/r2_summary.md): with >= 8 kv heads on the GPU and <= 32 query rows per kv head on a shared prefix
    // the mma.sync split-KV kernel has enough warps in flight and is 0-3 % faster (config 3: 3.52 vs 3.63 ms per step); with
    // fewer kv heads per GPU (tensor-parallel shards: 70B/TP8 -7 %, 32B/TP4 -5 %), more rows, or no shared prefix (SEQ mode,
    // -3 %) the tensor-memory kernel wins.  RT_DEC5=2 forces it.
    const int dec5_mode = env_int("RT_DEC5", 1);       // read per call (tests switch it between engines)
    if (dec5_mode != 2 && shared > 0 && o.T * e->G <= 32 && e->nkv_l >= 8) return;
    const int T = o.T, nkv = e->nkv_l;
    const int sh_tiles = cdiv(shared, P6_KEYS), own_tiles = cdiv(std::max(max_len - shared, 1), P6_KEYS);
    int L = std::max(1, env_int("RT_DEC5_MINL", 1));        // tiles per item: the smallest that fits one wave
    auto count = [&](int l) { return nkv * cdiv(sh_tiles, l) + T * nkv * cdiv(own_tiles, l); };
    while (count(L) > e->sm_count && L < 4096) ++L;
    const int n_sh = sh_tiles ? cdiv(sh_tiles, L) : 0, n_own = cdiv(own_tiles, L);
    if (count(L) > e->dec5_cap || n_sh + n_own + 1 > MAX_CHUNKS) return;
    Dec5Item* it = reinterpret_cast<Dec5Item*>(e->dec5_items_host);
    int n = 0;
    // own items first: they are the ones that may end early, the shared ranges are always full
    for (int t = 0; t < T; ++t)
        for (int h = 0; h < nkv; ++h)
            for (int c = 0; c < n_own; ++c) it[n++] = Dec5Item{t, h, c * L * P6_KEYS, L * P6_KEYS, n_sh + c, t, {0, 0}};
    for (int h = 0; h < nkv; ++h)
        for (int c = 0; c < n_sh; ++c) it[n++] = Dec5Item{-1, h, c * L * P6_KEYS, L * P6_KEYS, c, 0, {0, 0}};
    // (the items go to the kernel as a parameter: prefill.cu pf_attention_dec5 copies them from this host list at launch /
    // capture; a generate call's decode slot t IS sequence t, and slot 0's sequence holds the shared pages)
    o.dec5_items = n;
    o.dec5_nsh = n_sh;
    o.dec5_own_span = L * P6_KEYS;
    o.dec5_nown = n_own;
}

static void check_deadline(rt_engine* e, double t0, uint64_t deadline_ms) {
    if (deadline_ms == 0) return;
    const double el = now_ms() - t0;
    bool stop = el > (double)deadline_ms;
    if (e->tp) {
        const float mine = stop ? 1.f : 0.f;
        float any = 0.f;
        RT_CUDA(cudaMemcpyAsync(e->tp_stop, &mine, sizeof(float), cudaMemcpyHostToDevice, e->st));
        tp_allreduce_sum_f32(e->tp, e->tp_stop, 1, e->st);
        RT_CUDA(cudaMemcpyAsync(&any, e->tp_stop, sizeof(float), cudaMemcpyDeviceToHost, e->st));
        RT_CUDA(cudaStreamSynchronize(e->st));
        stop = any > 0.f;
    }
    if (stop)
        throw RtError(RT_ERR_TIMEOUT, "generation timed out after " + std::to_string((long long)el) + " ms (deadline " +
                                          std::to_string((unsigned long long)deadline_ms) + " ms)");
}

// Run a flat stream of (token, seq, pos) slots through the decode kernels in slabs of <= 32.
// `want_first[i]` >= 0 marks the last prompt token of sequence want_first[i]: its arg-max becomes that
// sequence's first generated token (copied device-to-device into d_out_tokens[seq][0]).
static void prefill_via_slabs(rt_engine* e, const std::vector<int>& tok, const std::vector<int>& seq, const std::vector<int>& pos,
                              const std::vector<int>& want_first, StepOpts base, double t0, uint64_t deadline_ms) {
    const int n = (int)tok.size();
    for (int s0 = 0; s0 < n; s0 += MAX_SLOTS) {
        const int T = std::min(MAX_SLOTS, n - s0);
        engine_set_slots(e, tok.data() + s0, seq.data() + s0, pos.data() + s0, T);
        StepOpts o = base;
        o.T = T;
        o.n_shared_tiles = 0;
        o.advance = false;
        bool any = false;
        for (int i = 0; i < T; ++i) any = any || want_first[s0 + i] >= 0;
        o.sample = any;
        engine_forward(e, o);
        if (any)
            for (int i = 0; i < T; ++i)
                if (want_first[s0 + i] >= 0)
                    RT_CUDA(cudaMemcpyAsync(e->d_out_tokens + (size_t)want_first[s0 + i] * e->max_out, e->d_next_tok + i,
                                            sizeof(int), cudaMemcpyDeviceToDevice, e->st));
        check_deadline(e, t0, deadline_ms);
    }
}

// ===================================================================================================
// Tensor-core prefill: a flat stream of prompt tokens (token, seq, pos) in chunks of <= pf_cap.  Per layer:
//   tcgen05 GEMM [QKV + bias + RoPE + KV pages] -> causal flash attention over the pages -> tcgen05 GEMM [O + residual]
//   -> RMSNorm -> tcgen05 GEMM [gate/up + SwiGLU] -> tcgen05 GEMM [down + residual] -> RMSNorm.
// Rows whose `want_first` is a sequence id get their next token from the decode lm_head kernel.
// With logits_out != null (teacher forcing) the logits of every token are produced instead.
// ===================================================================================================
static void prefill_tensor(rt_engine* e, const std::vector<int>& tok, const std::vector<int>& seq, const std::vector<int>& pos,
                           const std::vector<int>& want_first, const StepOpts& base, double t0, uint64_t deadline_ms,
                           float* logits_out = nullptr) {
    const int n = (int)tok.size(), H = e->H, hd = e->hd;
    const int BT = e->pfa_tc5 ? 128 / e->G : std::max(1, 64 / e->G);
    const bool pdl = e->use_pdl;
    for (int c0 = 0; c0 < n; c0 += e->pf_cap) {
        const int S = std::min(e->pf_cap, n - c0);
        RT_CUDA(cudaStreamSynchronize(e->st));     // staging buffers are free again
        int* ht = e->pf_host;
        int* hs = ht + e->pf_cap;
        int* hp = hs + e->pf_cap;
        int4* hw = reinterpret_cast<int4*>(hp + e->pf_cap);
        memcpy(ht, tok.data() + c0, S * sizeof(int));
        memcpy(hs, seq.data() + c0, S * sizeof(int));
        memcpy(hp, pos.data() + c0, S * sizeof(int));
        int n_items = 0;
        for (int s0 = 0; s0 < S;) {
            int s1 = s0 + 1;
            while (s1 < S && hs[s1] == hs[s0] && hp[s1] == hp[s1 - 1] + 1) ++s1;
            for (int b = s0; b < s1; b += BT)
                for (int kh = 0; kh < e->nkv_l; ++kh) {
                    if (n_items >= e->pf_work_cap) throw RtError(RT_ERR_INVALID, "prefill work list overflow");
                    hw[n_items++] = make_int4(b, std::min(BT, s1 - b), hs[s0], kh);
                }
            s0 = s1;
        }
        // longest first: an item walks the keys 0 .. position of its last token, so the late tokens of a long prompt are
        // the heavy items; issuing them first keeps the last wave of CTAs short
        std::stable_sort(hw, hw + n_items, [&](const int4& x, const int4& y) { return hp[x.x] + x.y > hp[y.x] + y.y; });
        RT_CUDA(cudaMemcpyAsync(e->pf_tok, ht, S * sizeof(int), cudaMemcpyHostToDevice, e->st));
        RT_CUDA(cudaMemcpyAsync(e->pf_seq, hs, S * sizeof(int), cudaMemcpyHostToDevice, e->st));
        RT_CUDA(cudaMemcpyAsync(e->pf_pos, hp, S * sizeof(int), cudaMemcpyHostToDevice, e->st));
        RT_CUDA(cudaMemcpyAsync(e->pf_work, hw, (size_t)n_items * sizeof(int4), cudaMemcpyHostToDevice, e->st));

        launch_pdl(e, embed_norm_kernel, dim3(S), dim3(256), 0, (const bf16*)e->embed, (const int*)e->pf_tok, e->pf_x,
                   (const bf16*)e->layers[0].norm1, (void*)e->pf_xn, 0, (float*)nullptr, e->pf_inv_rms, H, e->vocab, e->cfg.rms_eps);
        for (int l = 0; l < e->L; ++l) {
            const LayerW& w = e->layers[l];
            bf16* kp = e->kpool + (size_t)l * e->layer_pool_elems;
            bf16* vp = e->vpool + (size_t)l * e->layer_pool_elems;
            Gemm5Args g{};
            g.ab_f16 = 1; g.S = S; g.N = e->qkv_rows; g.K = H; g.row_scale = e->pf_inv_rms; g.bias = w.bqkv; g.cos_t = e->cos_t; g.sin_t = e->sin_t;
            g.tok_seq = e->pf_seq; g.tok_pos = e->pf_pos; g.page_table = e->d_page_table; g.pt_stride = e->max_pages_per_seq;
            g.Q = e->pf_q; g.Kp = kp; g.Vp = vp; g.n_q = e->nq_l; g.n_kv = e->nkv_l; g.hd = hd;
            pf_gemm(e->st, pdl, G5_EPI_QKV, e->map_xn, e->map_w[4 * l + 0], g);
            e->launches++;

            PfAttnArgs a{};
            a.Q = e->pf_q; a.Kp = kp; a.Vp = vp; a.page_table = e->d_page_table; a.pt_stride = e->max_pages_per_seq;
            a.tok_pos = e->pf_pos; a.work = e->pf_work; a.n_q = e->nq_l; a.n_kv = e->nkv_l;
            a.scale_log2 = (float)(1.4426950408889634 / std::sqrt((double)hd)); a.out = e->pf_attn;
            if (e->pfa_tc5)
                pf_attention_tc5(e->st, pdl, n_items, e->map_q3, e->map_kpool, e->map_vpool, e->d_page_table, e->max_pages_per_seq, e->pf_pos,
                                 e->pf_work, e->nq_l, e->nkv_l, e->G, BT, e->n_pages, l, a.scale_log2, e->pf_attn, hd);
            else
                pf_attention(e->st, pdl, hd, n_items, a);
            e->launches++;

            Gemm5Args go{};
            go.ab_f16 = 1; go.S = S; go.N = H; go.K = e->nq_l * hd; go.ldo = H;
            // tensor parallel: the partial sums cross NVLink as fp16 (half the bytes of the [S][H] all-reduce; the rounding
            // is that of every other projection input) and are added to the fp32 residual stream by the norm kernel
            go.out = e->pf_x; go.out_bf = e->pf_delta; go.ldob = H;
            pf_gemm(e->st, pdl, e->tp ? G5_EPI_STORE16 : G5_EPI_RESID, e->map_attn, e->map_w[4 * l + 1], go);
            if (e->tp) tp_allreduce_sum_f16(e->tp, e->pf_delta, (size_t)S * H, e->st);
            rms_xw_kernel<f16><<<S, 256, 0, e->st>>>(e->pf_x, e->tp ? e->pf_delta : nullptr, w.norm2, e->pf_xn, 0, e->pf_inv_rms, H, e->cfg.rms_eps);
            e->launches += 2;

            Gemm5Args gu{};
            gu.ab_f16 = 1; gu.S = S; gu.N = 2 * e->F_l; gu.K = H; gu.row_scale = e->pf_inv_rms; gu.out_bf = e->pf_act; gu.ldob = e->F_l;
            pf_gemm(e->st, pdl, G5_EPI_SWIGLU, e->map_xn, e->map_w[4 * l + 2], gu);
            Gemm5Args gd{};
            gd.ab_f16 = 1; gd.S = S; gd.N = H; gd.K = e->F_l; gd.ldo = H;
            gd.out = e->pf_x; gd.out_bf = e->pf_delta; gd.ldob = H;
            pf_gemm(e->st, pdl, e->tp ? G5_EPI_STORE16 : G5_EPI_RESID, e->map_act, e->map_w[4 * l + 3], gd);
            if (e->tp) tp_allreduce_sum_f16(e->tp, e->pf_delta, (size_t)S * H, e->st);
            const bf16* nw = (l + 1 < e->L) ? e->layers[l + 1].norm1 : e->final_norm;
            rms_xw_kernel<f16><<<S, 256, 0, e->st>>>(e->pf_x, e->tp ? e->pf_delta : nullptr, nw, e->pf_xn, 0, e->pf_inv_rms, H, e->cfg.rms_eps);
            e->launches += 3;
        }
        RT_CUDA(cudaGetLastError());

        // next tokens (or all logits) through the decode lm_head kernel, <= 32 rows at a time
        auto lm_head_rows = [&](const std::vector<int>& rows, bool sample, bool store) {
            const int T = (int)rows.size();
            GatherRowsArgs ga{};
            ga.src = e->pf_xn; ga.src_r = e->pf_inv_rms; ga.dst = e->xn; ga.dst_r = e->inv_rms; ga.H = H; ga.raw = xf_mode(e, T) == 2;
            for (int j = 0; j < T; ++j) ga.rows[j] = rows[j];
            gather_rows_kernel<<<T, 256, 0, e->st>>>(ga);
            e->launches++;
            lm_head_and_sample(e, T, sample, false, store, base.inv_temp, base.seed, true, base.top_k, base.top_p);
        };
        if (logits_out) {
            for (int r0 = 0; r0 < S; r0 += MAX_SLOTS) {
                std::vector<int> rows;
                for (int r = r0; r < std::min(S, r0 + MAX_SLOTS); ++r) rows.push_back(r);
                lm_head_rows(rows, false, true);
                RT_CUDA(cudaMemcpyAsync(logits_out + (size_t)(c0 + r0) * e->vocab_l, e->logits, rows.size() * (size_t)e->vocab_l * sizeof(float),
                                        cudaMemcpyDeviceToHost, e->st));
                RT_CUDA(cudaStreamSynchronize(e->st));
            }
        } else {
            std::vector<int> rows, seqs;
            for (int i = 0; i < S; ++i)
                if (want_first[c0 + i] >= 0) { rows.push_back(i); seqs.push_back(want_first[c0 + i]); }
            if (!rows.empty()) {
                lm_head_rows(rows, true, false);
                for (size_t j = 0; j < rows.size(); ++j)
                    RT_CUDA(cudaMemcpyAsync(e->d_out_tokens + (size_t)seqs[j] * e->max_out, e->d_next_tok + j, sizeof(int),
                                            cudaMemcpyDeviceToDevice, e->st));
            }
        }
        check_deadline(e, t0, deadline_ms);
    }
}

static cudaGraphExec_t capture_steps(rt_engine* e, const StepOpts& o, int steps, int* nodes) {
    const long long before = e->launches;
    cudaGraph_t graph = nullptr;
    RT_CUDA(cudaStreamBeginCapture(e->st, cudaStreamCaptureModeThreadLocal));
    try {
        StepOpts og = o;
        og.in_graph = true;
        // Several steps in one graph chain through PDL too: step k+1's attention may then start while step k's sampler
        // (which advances slot_pos) is still running, unless its QKV projection releases it late (gemv.cuh).
        og.qkv_late = steps > 1;
        for (int k = 0; k < steps; ++k) engine_forward(e, og);
    } catch (...) {
        cudaStreamEndCapture(e->st, &graph);
        if (graph) cudaGraphDestroy(graph);
        throw;
    }
    RT_CUDA(cudaStreamEndCapture(e->st, &graph));
    if (nodes) *nodes = (int)(e->launches - before) / steps;
    e->launches = before;
    cudaGraphExec_t exec = nullptr;
    RT_CUDA(cudaGraphInstantiate(&exec, graph, 0));
    cudaGraphDestroy(graph);
    return exec;
}

static void ensure_graph(rt_engine* e, const StepOpts& o) {
    const StepOpts& g = e->graph_opts;
    if (e->graph_exec && g.dec5_items == o.dec5_items && g.dec5_nsh == o.dec5_nsh && g.dec5_own_span == o.dec5_own_span && g.dec5_nown == o.dec5_nown && g.T == o.T && g.chunk_pages == o.chunk_pages && g.n_chunk_blocks == o.n_chunk_blocks && g.n_sh == o.n_sh && g.n_own == o.n_own &&
        g.n_shared_tiles == o.n_shared_tiles && g.attn3_cs == o.attn3_cs && g.inv_temp == o.inv_temp && g.seed == o.seed && g.top_k == o.top_k && g.top_p == o.top_p)
        return;
    if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
    if (e->graph_exec_multi) { cudaGraphExecDestroy(e->graph_exec_multi); e->graph_exec_multi = nullptr; }
    e->graph_exec = capture_steps(e, o, 1, &e->graph_nodes);
    // several steps per graph launch: measured no faster (1B 635 vs 630 us/step, 8B 3,318 vs 3,299: the ~3 us launch gap
    // it removes is paid back by the late QKV release it needs), so one step per launch stays the default
    e->graph_steps = std::max(1, env_int("RT_GRAPH_STEPS", 1));
    if (e->graph_steps > 1) e->graph_exec_multi = capture_steps(e, o, e->graph_steps, nullptr);
    e->graph_opts = o;
}

// ===================================================================================================
// Batch generation: n token-id prompts -> n greedy (or Gumbel-sampled) continuations.
// ===================================================================================================
void engine_generate_ids(rt_engine* e, const int32_t* const* prompts, const int32_t* lens, int n, const rt_sampling* sp,
                         uint64_t deadline_ms, rt_result* outs) {
    RT_REQUIRE(n >= 1 && n <= e->n_seq, "batch size must be 1..max_batch");
    rt_sampling s{};
    if (sp) s = *sp;
    if (s.max_new_tokens <= 0) s.max_new_tokens = 512;
    RT_REQUIRE(s.top_k >= 0 && s.top_p >= 0.f && s.top_p <= 1.f, "top_k must be >= 0 and top_p within [0, 1]");
    RT_REQUIRE(s.temperature >= 0.f, "temperature must be >= 0");
    for (int i = 0; i < n; ++i) {
        RT_REQUIRE(lens[i] >= 1, "empty prompt");
        for (int j = 0; j < lens[i]; ++j) RT_REQUIRE(prompts[i][j] >= 0 && prompts[i][j] < e->vocab, "token id out of range");
    }
    RT_CUDA(cudaSetDevice(e->dev));
    const double t0 = now_ms();
    const long long launches0 = e->launches;
    engine_reset_sequences(e);

    int max_len = 0, min_len = 1 << 30;
    for (int i = 0; i < n; ++i) {
        if (lens[i] + s.max_new_tokens > e->cfg.max_ctx)
            throw RtError(RT_ERR_OOM, "context length exceeded: prompt of " + std::to_string(lens[i]) + " tokens + " +
                                          std::to_string(s.max_new_tokens) + " new tokens does not fit max_ctx " +
                                          std::to_string(e->cfg.max_ctx));
        max_len = std::max(max_len, lens[i] + s.max_new_tokens);
        min_len = std::min(min_len, lens[i]);
    }
    // longest common token prefix, page aligned, leaving >= 1 own token per sequence
    int shared = 0;
    if (n > 1 && !(e->cfg.flags & RT_FLAG_NO_PREFIX_SHARE)) {
        int lc = 0;
        for (; lc < min_len - 1; ++lc) {
            bool same = true;
            for (int i = 1; i < n && same; ++i) same = prompts[i][lc] == prompts[0][lc];
            if (!same) break;
        }
        shared = (lc / KV_PAGE) * KV_PAGE;
    }
    // pages: cached prefix (if any) -> batch-shared prefix -> fresh pages
    const bool use_cache = (e->cfg.flags & RT_FLAG_PREFIX_CACHE) != 0;
    std::vector<int> start(n, 0);          // first prompt position each sequence still has to prefill itself
    std::vector<int> cached(n, 0);
    for (int i = 0; i < n; ++i) {
        int entry = -1;
        const int hit = use_cache ? cache_lookup(e, prompts[i], lens[i], &entry) : 0;
        if (i == 0) {
            if (hit > 0) cache_attach(e, entry, 0, hit);
            cached[0] = hit;
            start[0] = std::max(hit, shared);       // [hit, shared) is prefilled once, as the batch-shared part
        } else if (hit > shared) {
            cache_attach(e, entry, i, hit);
            cached[i] = hit;
            start[i] = hit;
        } else {
            engine_share_prefix(e, 0, i, shared / KV_PAGE);
            start[i] = shared;
        }
        engine_alloc_pages(e, i, lens[i] + s.max_new_tokens);
    }
    engine_upload_page_table(e);
    RT_CUDA(cudaMemsetAsync(e->d_step, 0, sizeof(int), e->st));   // sampling noise is keyed on the step counter

    StepOpts base{};
    plan_chunks(max_len, 0, base, e->hd);   // prefill slabs: no cascade phase
    base.inv_temp = s.temperature > 0.f ? 1.0f / s.temperature : 0.f;
    base.seed = s.seed;
    base.top_k = s.top_k;
    base.top_p = s.top_p;

    // ---- prefill ----
    const char* prof_phase = getenv("RT_PROFILE_PHASE");     // ncu --profile-from-start off: "prefill" | "decode"
    if (prof_phase && !strcmp(prof_phase, "prefill")) cudaProfilerStart();
    RT_CUDA(cudaEventRecord(e->ev[0], e->st));
    {
        std::vector<int> tok, seq, pos, want;
        for (int p = cached[0]; p < shared; ++p) { tok.push_back(prompts[0][p]); seq.push_back(0); pos.push_back(p); want.push_back(-1); }
        for (int i = 0; i < n; ++i)
            for (int p = start[i]; p < lens[i]; ++p) {
                tok.push_back(prompts[i][p]); seq.push_back(i); pos.push_back(p);
                want.push_back(p == lens[i] - 1 ? i : -1);
            }
        if (e->pf_ok && !(e->cfg.flags & RT_FLAG_PREFILL_VIA_DECODE) && (int)tok.size() >= 32)
            prefill_tensor(e, tok, seq, pos, want, base, t0, deadline_ms);
        else
            prefill_via_slabs(e, tok, seq, pos, want, base, t0, deadline_ms);
    }
    RT_CUDA(cudaEventRecord(e->ev[1], e->st));
    if (prof_phase && !strcmp(prof_phase, "prefill")) {
        RT_CUDA(cudaStreamSynchronize(e->st));
        cudaProfilerStop();
    }

    // ---- decode state: slot b = sequence b, token = its first generated token ----
    std::vector<int> first(n);
    RT_CUDA(cudaStreamSynchronize(e->st));
    for (int i = 0; i < n; ++i)
        RT_CUDA(cudaMemcpy(&first[i], e->d_out_tokens + (size_t)i * e->max_out, sizeof(int), cudaMemcpyDeviceToHost));
    {
        std::vector<int> seq(n), pos(n);
        for (int i = 0; i < n; ++i) { seq[i] = i; pos[i] = lens[i]; }
        engine_set_slots(e, first.data(), seq.data(), pos.data(), n);
        const int one = 1;
        e->h_stage[3 * MAX_SLOTS] = one;
        e->h_stage[3 * MAX_SLOTS + 1] = shared;
        RT_CUDA(cudaMemcpyAsync(e->d_step, e->h_stage + 3 * MAX_SLOTS, sizeof(int), cudaMemcpyHostToDevice, e->st));
        RT_CUDA(cudaMemcpyAsync(e->d_shared_len, e->h_stage + 3 * MAX_SLOTS + 1, sizeof(int), cudaMemcpyHostToDevice, e->st));
    }
    StepOpts dec = base;
    dec.T = n;
    dec.sample = true;
    dec.advance = true;
    dec.n_shared_tiles = shared > 0 ? cdiv(n * e->G, 16) : 0;
    plan_chunks(max_len, shared, dec, e->hd);
    plan_dec5(e, max_len, shared, dec);

    const int n_steps = s.max_new_tokens - 1;
    auto is_eos = [&](int t) { for (int j = 0; j < e->n_eos; ++j) if (t == e->eos_ids[j]) return true; return false; };
    const bool use_graph = !(e->cfg.flags & RT_FLAG_NO_GRAPH);
    if (use_graph && n_steps > 0) ensure_graph(e, dec);
    std::vector<int> host_out((size_t)n * s.max_new_tokens);
    std::vector<int> n_out(n, s.max_new_tokens);
    std::vector<int> finish(n, 0);
    int steps_done = 0;

    const bool prof = getenv("RT_PROFILE_DECODE") != nullptr || (prof_phase && !strcmp(prof_phase, "decode"));
    if (prof) cudaProfilerStart();
#ifdef RT_TRACE
    if (getenv("RT_TRACE_FILE")) { RT_CUDA(cudaStreamSynchronize(e->st)); trace_begin(); }
#endif
    RT_CUDA(cudaEventRecord(e->ev[2], e->st));
    while (steps_done < n_steps) {
        const int burst = std::min(16, n_steps - steps_done);
        for (int k = 0; k < burst;) {
            if (use_graph && e->graph_exec_multi && burst - k >= e->graph_steps) {
                RT_CUDA(cudaGraphLaunch(e->graph_exec_multi, e->st));
                e->launches += (long long)e->graph_nodes * e->graph_steps;
                k += e->graph_steps;
            } else if (use_graph) {
                RT_CUDA(cudaGraphLaunch(e->graph_exec, e->st));
                e->launches += e->graph_nodes;
                ++k;
            } else {
                engine_forward(e, dec);
                ++k;
            }
        }
        steps_done += burst;
        if (deadline_ms || !s.ignore_eos) {
            RT_CUDA(cudaStreamSynchronize(e->st));
            check_deadline(e, t0, deadline_ms);
            if (!s.ignore_eos) {
                bool all_done = true;
                for (int i = 0; i < n; ++i) {
                    RT_CUDA(cudaMemcpy(host_out.data() + (size_t)i * s.max_new_tokens, e->d_out_tokens + (size_t)i * e->max_out,
                                       (size_t)(steps_done + 1) * sizeof(int), cudaMemcpyDeviceToHost));
                    bool found = false;
                    for (int k = 0; k <= steps_done && !found; ++k) found = is_eos(host_out[(size_t)i * s.max_new_tokens + k]);
                    all_done = all_done && found;
                }
                if (all_done) break;
            }
        }
    }
    RT_CUDA(cudaEventRecord(e->ev[3], e->st));
    RT_CUDA(cudaStreamSynchronize(e->st));
    if (prof) cudaProfilerStop();
#ifdef RT_TRACE
    if (const char* tf = getenv("RT_TRACE_FILE")) trace_end(tf);
#endif
    float pre_ms = 0.f, dec_ms = 0.f;
    RT_CUDA(cudaEventElapsedTime(&pre_ms, e->ev[0], e->ev[1]));
    RT_CUDA(cudaEventElapsedTime(&dec_ms, e->ev[2], e->ev[3]));

    if (use_cache)
        for (int i = 0; i < n; ++i) cache_insert(e, i, prompts[i], lens[i]);
    const int produced = steps_done + 1;
    for (int i = 0; i < n; ++i) {
        RT_CUDA(cudaMemcpy(host_out.data() + (size_t)i * s.max_new_tokens, e->d_out_tokens + (size_t)i * e->max_out,
                           (size_t)produced * sizeof(int), cudaMemcpyDeviceToHost));
        n_out[i] = produced;
        if (!s.ignore_eos)
            for (int k = 0; k < produced; ++k)
                if (is_eos(host_out[(size_t)i * s.max_new_tokens + k])) { n_out[i] = k; finish[i] = 1; break; }
    }
    for (int i = 0; i < n; ++i) {
        rt_result& r = outs[i];
        memset(&r, 0, sizeof(r));
        r.n_tokens = n_out[i];
        r.tokens = (int32_t*)malloc(sizeof(int32_t) * std::max(1, n_out[i]));
        memcpy(r.tokens, host_out.data() + (size_t)i * s.max_new_tokens, sizeof(int32_t) * n_out[i]);
        std::string text = rt_tok_decode(r.tokens, r.n_tokens);
        if (text.empty()) text = "<eos>";   // the adapter treats empty content as an error (local-llm.ts:189-191)
        r.text = (char*)malloc(text.size() + 1);
        memcpy(r.text, text.c_str(), text.size() + 1);
        r.text_len = text.size();
        r.n_prompt_tokens = lens[i];
        r.prefix_shared_tokens = (i == 0 || cached[i] > shared) ? 0 : shared;
        r.prefix_cached_tokens = cached[i];
        r.finish_reason = finish[i];
        r.prefill_ms = pre_ms;
        r.decode_ms = dec_ms;
        r.decode_steps = steps_done;
        r.gpu_launches = (int32_t)(e->launches - launches0);
    }
}

// ===================================================================================================
// Teacher forcing on the scratch sequence (seq 0): append ids, return logits of every appended position.
// ===================================================================================================
void engine_step_logits(rt_engine* e, const int32_t* ids, int n, int mode, float* logits_out) {
    RT_REQUIRE(n >= 1, "n must be >= 1");
    RT_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (decode kernels) or 1 (tensor-core prefill kernels)");
    RT_REQUIRE(mode == 0 || e->pf_ok, "the tensor-core prefill path needs hidden, n_q*head_dim and ffn to be multiples of 64");
    RT_CUDA(cudaSetDevice(e->dev));
    if (!e->scratch_active) {
        engine_reset_sequences(e);
        e->scratch_active = true;
    }
    const int p0 = e->seq_len[0];
    if (p0 + n > e->cfg.max_ctx) throw RtError(RT_ERR_OOM, "context length exceeded in rt_step_logits");
    for (int j = 0; j < n; ++j) RT_REQUIRE(ids[j] >= 0 && ids[j] < e->vocab, "token id out of range");
    engine_alloc_pages(e, 0, p0 + n);
    engine_upload_page_table(e);
    const int zero = 0;
    RT_CUDA(cudaMemcpyAsync(e->d_shared_len, &zero, sizeof(int), cudaMemcpyHostToDevice, e->st));
    StepOpts o{};
    plan_chunks(p0 + n, 0, o, e->hd);
    o.store_logits = true;
    if (mode == 1) {
        std::vector<int> tok(ids, ids + n), seq(n, 0), pos(n), want(n, -1);
        for (int i = 0; i < n; ++i) pos[i] = p0 + i;
        prefill_tensor(e, tok, seq, pos, want, o, now_ms(), 0, logits_out);
        e->seq_len[0] = p0 + n;
        return;
    }
    for (int s0 = 0; s0 < n; s0 += MAX_SLOTS) {
        const int T = std::min(MAX_SLOTS, n - s0);
        int tok[MAX_SLOTS], seq[MAX_SLOTS], pos[MAX_SLOTS];
        for (int i = 0; i < T; ++i) { tok[i] = ids[s0 + i]; seq[i] = 0; pos[i] = p0 + s0 + i; }
        engine_set_slots(e, tok, seq, pos, T);
        o.T = T;
        engine_forward(e, o);
        RT_CUDA(cudaMemcpyAsync(logits_out + (size_t)s0 * e->vocab_l, e->logits, (size_t)T * e->vocab_l * sizeof(float),
                                cudaMemcpyDeviceToHost, e->st));
        RT_CUDA(cudaStreamSynchronize(e->st));
    }
    e->seq_len[0] = p0 + n;
}

// ===================================================================================================
void engine_export_weight(rt_engine* e, int tid, uint16_t* out, long long n_elems) {
    RT_REQUIRE(e->cfg.tp_size == 1, "weight export needs tp_size == 1");
    RT_CUDA(cudaSetDevice(e->dev));
    const int H = e->H, hd = e->hd;
    // fetch a CM matrix and return it row-major (physical row order)
    auto fetch_cm = [&](const bf16* src, int rows, int cols) {
        std::vector<uint16_t> cm((size_t)rows * cols), rm((size_t)rows * cols);
        RT_CUDA(cudaMemcpy(cm.data(), src, cm.size() * 2, cudaMemcpyDeviceToHost));
        wgen::cm_to_rowmajor(cm.data(), rm.data(), rows, cols);
        return rm;
    };
    auto copy_vec = [&](const bf16* src, long long n) {
        RT_REQUIRE(n_elems == n, "n_elems does not match tensor shape");
        RT_CUDA(cudaMemcpy(out, src, (size_t)n * 2, cudaMemcpyDeviceToHost));
    };
    auto copy_all = [&](const std::vector<uint16_t>& rm) {
        RT_REQUIRE(n_elems == (long long)rm.size(), "n_elems does not match tensor shape");
        memcpy(out, rm.data(), rm.size() * 2);
    };
    // physical row of logical dim d inside a head (inverse of the RoPE pair permutation)
    auto rope_row = [&](int d) { const int half = hd / 2; return d < half ? (d / 8) * 16 + d % 8 : ((d - half) / 8) * 16 + 8 + (d - half) % 8; };
    if (tid == 0) return copy_all(fetch_cm(e->embed, e->vocab, H));
    if (tid == 1) return copy_all(fetch_cm(e->lm_head, e->vocab, H));
    if (tid == 2) return copy_vec(e->final_norm, H);
    const int l = tid / 16 - 1, k = tid % 16;
    RT_REQUIRE(l >= 0 && l < e->L && k <= 11, "bad tensor id");
    const LayerW& w = e->layers[l];
    const int qr = e->nq * hd, kr = e->nkv * hd;
    switch (k) {
        case 0: case 1: case 2: {
            const int off = k == 0 ? 0 : (k == 1 ? qr : qr + kr), cnt = k == 0 ? qr : kr;
            RT_REQUIRE(n_elems == (long long)cnt * H, "n_elems does not match tensor shape");
            const std::vector<uint16_t> rm = fetch_cm(w.wqkv, e->qkv_rows, H);
            for (int r = 0; r < cnt; ++r) {
                const int prow = off + (r / hd) * hd + rope_row(r % hd);
                memcpy(out + (size_t)r * H, rm.data() + (size_t)prow * H, (size_t)H * 2);
            }
            return;
        }
        case 3: return copy_all(fetch_cm(w.wo, H, qr));
        case 6: return copy_all(fetch_cm(w.wdown, H, e->F));
        case 7: return copy_vec(w.norm1, H);
        case 8: return copy_vec(w.norm2, H);
        case 4: case 5: {
            RT_REQUIRE(n_elems == (long long)e->F * H, "n_elems does not match tensor shape");
            const std::vector<uint16_t> rm = fetch_cm(w.wgu, 2 * e->F, H);
            for (int f = 0; f < e->F; ++f) {
                const size_t prow = (size_t)(f / 8) * 16 + (k == 5 ? 8 : 0) + (f % 8);
                memcpy(out + (size_t)f * H, rm.data() + prow * H, (size_t)H * 2);
            }
            return;
        }
        default: {
            RT_REQUIRE(w.bqkv != nullptr, "model has no qkv bias");
            const int off = k == 9 ? 0 : (k == 10 ? qr : qr + kr), cnt = k == 9 ? qr : kr;
            RT_REQUIRE(n_elems == cnt, "n_elems does not match tensor shape");
            std::vector<float> tmp(e->qkv_rows);
            RT_CUDA(cudaMemcpy(tmp.data(), w.bqkv, tmp.size() * 4, cudaMemcpyDeviceToHost));
            for (int i = 0; i < cnt; ++i) {
                const float f = tmp[off + (i / hd) * hd + rope_row(i % hd)];
                uint32_t u;
                memcpy(&u, &f, 4);
                out[i] = (uint16_t)(u >> 16);
            }
            return;
        }
    }
}
