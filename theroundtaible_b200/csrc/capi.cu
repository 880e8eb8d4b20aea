// extern "C" boundary of the generation backend (include/rt_engine.h).  No exception crosses it.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "engine.cuh"
#include "kernels.cuh"
#include "sampler.cuh"
#include "weights.cuh"
#include "prefill.cuh"

int rt_gemv_host_dispatch(rt_engine* e, const GemvArgs& a, int epi);
void rt_attn_host_setup();
void rt_attn_host_launch(rt_engine* e, const AttnArgs& a, int hd, int n_chunk_blocks);

namespace {
thread_local std::string g_err;

template <typename F>
int guard(rt_engine* e, F&& f) {
    try {
        f();
        return RT_OK;
    } catch (const RtError& x) {
        (e ? e->err : g_err) = x.what();
        if (e) g_err = x.what();
        return x.code;
    } catch (const std::exception& x) {
        (e ? e->err : g_err) = std::string("internal error: ") + x.what();
        if (e) g_err = e->err;
        return RT_ERR_INVALID;
    } catch (...) {
        (e ? e->err : g_err) = "internal error";
        return RT_ERR_INVALID;
    }
}

struct ArchRow {
    const char* name;
    int family, layers, hidden, n_q, n_kv, head_dim, ffn, vocab, tied, qkv_bias;
    float rope_theta, rms_eps, rope_factor, rope_low, rope_high;
    int rope_orig;
};
// public architecture constants; tiny-* are test shapes shared with oracle/archs.py
const ArchRow kArchs[] = {
    {"qwen2.5-0.5b", 1, 24, 896, 14, 2, 64, 4864, 151936, 1, 1, 1e6f, 1e-6f, 0.f, 1.f, 4.f, 8192},
    {"llama-3.2-1b", 0, 16, 2048, 32, 8, 64, 8192, 128256, 1, 0, 5e5f, 1e-5f, 32.f, 1.f, 4.f, 8192},
    {"llama-3.1-8b", 0, 32, 4096, 32, 8, 128, 14336, 128256, 0, 0, 5e5f, 1e-5f, 8.f, 1.f, 4.f, 8192},
    {"qwen2.5-32b", 1, 64, 5120, 40, 8, 128, 27648, 152064, 0, 1, 1e6f, 1e-5f, 0.f, 1.f, 4.f, 8192},
    {"llama-3.1-70b", 0, 80, 8192, 64, 8, 128, 28672, 128256, 0, 0, 5e5f, 1e-5f, 8.f, 1.f, 4.f, 8192},
    {"tiny-llama", 0, 2, 128, 4, 2, 32, 256, 512, 1, 0, 5e5f, 1e-5f, 32.f, 1.f, 4.f, 64},
    {"tiny-llama-hd128", 0, 2, 256, 4, 1, 128, 512, 640, 0, 0, 5e5f, 1e-5f, 8.f, 1.f, 4.f, 64},
    {"tiny-qwen", 1, 3, 448, 14, 2, 32, 384, 768, 1, 1, 1e6f, 1e-6f, 0.f, 1.f, 4.f, 8192},
    {"tiny-qwen-hd128", 1, 2, 640, 5, 1, 128, 768, 1024, 0, 1, 1e6f, 1e-5f, 0.f, 1.f, 4.f, 8192},
};
}  // namespace

void rt_internal_set_error(const std::string& msg) { g_err = msg; }

extern "C" {

int rt_arch_by_name(const char* name, rt_config* cfg) {
    return guard(nullptr, [&] {
        RT_REQUIRE(name && cfg, "null argument");
        std::string n(name);
        int trunc = -1;
        const size_t at = n.find("@L");
        if (at != std::string::npos) {
            trunc = atoi(n.c_str() + at + 2);
            n = n.substr(0, at);
            RT_REQUIRE(trunc > 0, "bad @L<n> suffix");
        }
        for (char& ch : n) ch = (char)tolower((unsigned char)ch);
        for (const ArchRow& r : kArchs) {
            if (n == r.name) {
                cfg->struct_size = sizeof(rt_config);
                cfg->family = r.family; cfg->layers = trunc > 0 ? std::min(trunc, r.layers) : r.layers;
                cfg->hidden = r.hidden; cfg->n_q = r.n_q; cfg->n_kv = r.n_kv; cfg->head_dim = r.head_dim;
                cfg->ffn = r.ffn; cfg->vocab = r.vocab; cfg->tied = r.tied; cfg->qkv_bias = r.qkv_bias;
                cfg->rope_theta = r.rope_theta; cfg->rms_eps = r.rms_eps; cfg->rope_factor = r.rope_factor;
                cfg->rope_low_freq = r.rope_low; cfg->rope_high_freq = r.rope_high; cfg->rope_orig_ctx = r.rope_orig;
                cfg->n_eos = 0;     // the byte tokenizer's <eos>
                return;
            }
        }
        throw RtError(RT_ERR_INVALID, "model '" + std::string(name) + "' not found in the architecture table");
    });
}

int rt_engine_create(const rt_config* cfg, rt_engine** out) {
    if (!cfg || !out) {
        g_err = "invalid argument: null pointer";
        return RT_ERR_INVALID;
    }
    *out = nullptr;
    rt_engine* e = new rt_engine();
    e->cfg = *cfg;
    e->dev = -1;
    const int rc = guard(nullptr, [&] { engine_init(e); });
    if (rc != RT_OK) {
        engine_free(e);
        delete e;
        return rc;
    }
    *out = e;
    return RT_OK;
}

void rt_engine_destroy(rt_engine* e) {
    if (!e) return;
    engine_free(e);
    delete e;
}

const char* rt_last_error(const rt_engine* e) { return e ? e->err.c_str() : g_err.c_str(); }

int rt_model_info_get(const rt_engine* ce, rt_model_info* out) {
    rt_engine* e = const_cast<rt_engine*>(ce);
    return guard(e, [&] {
        RT_REQUIRE(e && out, "null argument");
        memset(out, 0, sizeof(*out));
        out->context_length = e->cfg.max_ctx;
        out->vocab = e->vocab;
        const long long per_layer = (long long)e->H * (e->nq + 2 * e->nkv) * e->hd + (long long)e->nq * e->hd * e->H +
                                    3LL * e->H * e->F + 2LL * e->H + (e->cfg.qkv_bias ? (e->nq + 2 * e->nkv) * e->hd : 0);
        out->n_params = e->L * per_layer + (long long)e->vocab * e->H * (e->cfg.tied ? 1 : 2) + e->H;
        const long long per_layer_l = (long long)e->qkv_rows * e->H + (long long)e->nq_l * e->hd * e->H + 3LL * e->H * e->F_l + 2LL * e->H;
        out->weight_bytes_per_step = 2 * (e->L * per_layer_l + (long long)e->vocab_l * e->H + e->H);
        out->kv_bytes_per_token = 2LL * e->L * e->nkv_l * e->hd * 2;
        out->max_batch = e->cfg.max_batch;
        out->page_size = KV_PAGE;
        out->n_pages = e->n_pages;
        out->sm_count = e->sm_count;
    });
}

int rt_generate_ids(rt_engine* e, const int32_t* const* prompts, const int32_t* lens, int32_t n, const rt_sampling* s,
                    uint64_t deadline_ms, rt_result* outs) {
    if (!e) { g_err = "invalid argument: null engine"; return RT_ERR_INVALID; }
    return guard(e, [&] {
        RT_REQUIRE(prompts && lens && outs, "null argument");
        engine_generate_ids(e, prompts, lens, n, s, deadline_ms, outs);
    });
}

int rt_generate_batch(rt_engine* e, const char* const* prompts, const size_t* lens, int32_t n, const rt_sampling* s,
                      uint64_t deadline_ms, rt_result* outs) {
    if (!e) { g_err = "invalid argument: null engine"; return RT_ERR_INVALID; }
    return guard(e, [&] {
        RT_REQUIRE(prompts && lens && outs && n >= 1, "null argument");
        // the byte-level tokenizer only matches the seeded weights: a checkpoint's vocabulary is its own tokenizer's
        if (e->from_checkpoint)
            throw RtError(RT_ERR_INVALID, "text generation is unavailable for an engine created from a checkpoint: the built-in "
                                          "tokenizer is byte-level; use rt_generate_ids with the checkpoint's tokenizer");
        std::vector<std::vector<int>> ids(n);
        std::vector<const int32_t*> ptrs(n);
        std::vector<int32_t> ls(n);
        for (int i = 0; i < n; ++i) {
            RT_REQUIRE(prompts[i] != nullptr, "null prompt");
            rt_tok_encode_chat(prompts[i], lens[i], ids[i]);
            ptrs[i] = ids[i].data();
            ls[i] = (int32_t)ids[i].size();
        }
        engine_generate_ids(e, ptrs.data(), ls.data(), n, s, deadline_ms, outs);
    });
}

int rt_generate(rt_engine* e, const char* prompt_utf8, size_t len, const rt_sampling* s, uint64_t deadline_ms, rt_result* out) {
    const char* ps[1] = {prompt_utf8};
    const size_t ls[1] = {len};
    return rt_generate_batch(e, ps, ls, 1, s, deadline_ms, out);
}

int rt_seq_reset(rt_engine* e) {
    if (!e) { g_err = "invalid argument: null engine"; return RT_ERR_INVALID; }
    return guard(e, [&] {
        engine_reset_sequences(e);
        e->scratch_active = true;
    });
}

int rt_step_logits(rt_engine* e, const int32_t* ids, int32_t n, int32_t mode, float* logits_out) {
    if (!e) { g_err = "invalid argument: null engine"; return RT_ERR_INVALID; }
    return guard(e, [&] {
        RT_REQUIRE(ids && logits_out, "null argument");
        engine_step_logits(e, ids, n, mode, logits_out);
    });
}

void rt_result_free(rt_result* r) {
    if (!r) return;
    free(r->text);
    free(r->tokens);
    r->text = nullptr;
    r->tokens = nullptr;
    r->text_len = 0;
    r->n_tokens = 0;
}

int32_t rt_tokenize_chat(const char* utf8, size_t len, int32_t* ids, int32_t cap) {
    std::vector<int> v;
    rt_tok_encode_chat(utf8 ? utf8 : "", utf8 ? len : 0, v);
    if (ids)
        for (int i = 0; i < (int)v.size() && i < cap; ++i) ids[i] = v[i];
    return (int32_t)v.size();
}

int64_t rt_detokenize(const int32_t* ids, int32_t n, char* out, int64_t cap) {
    const std::string s = rt_tok_decode(ids, n);
    if (out && cap > 0) {
        const size_t c = std::min((size_t)(cap - 1), s.size());
        memcpy(out, s.data(), c);
        out[c] = 0;
    }
    return (int64_t)s.size();
}

int rt_nccl_unique_id(void* out128) {
    return guard(nullptr, [&] {
        RT_REQUIRE(out128 != nullptr, "null argument");
        tp_unique_id(out128);
    });
}

int rt_weights_export(rt_engine* e, int32_t tensor_id, uint16_t* out, int64_t n_elems) {
    if (!e) { g_err = "invalid argument: null engine"; return RT_ERR_INVALID; }
    return guard(e, [&] {
        RT_REQUIRE(out != nullptr, "null argument");
        engine_export_weight(e, tensor_id, out, n_elems);
    });
}

}  // extern "C"

// ---- kernel-level hooks -------------------------------------------------------------------------------
namespace {
struct Scratch {
    rt_engine e;
    std::vector<void*> bufs;
    explicit Scratch(int device) {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            throw RtError(RT_ERR_CUDA, "no CUDA device available: this backend has no CPU fallback");
        }
        RT_REQUIRE(device >= 0 && device < ndev, "device ordinal out of range");
        RT_CUDA(cudaSetDevice(device));
        cudaDeviceProp prop{};
        RT_CUDA(cudaGetDeviceProperties(&prop, device));
        e.dev = device;
        e.sm_count = prop.multiProcessorCount;
        RT_CUDA(cudaStreamCreateWithFlags(&e.st, cudaStreamNonBlocking));
    }
    template <typename T>
    T* alloc(size_t n, bool zero = true) {
        void* p = nullptr;
        RT_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
        bufs.push_back(p);
        if (zero) RT_CUDA(cudaMemsetAsync(p, 0, std::max<size_t>(n, 1) * sizeof(T), e.st));
        return reinterpret_cast<T*>(p);
    }
    ~Scratch() {
        if (e.st) cudaStreamSynchronize(e.st);
        for (void* p : bufs) cudaFree(p);
        for (void* p : e.allocs) cudaFree(p);
        if (e.st) cudaStreamDestroy(e.st);
    }
};
}  // namespace

extern "C" {

// the projection hooks follow the engine's rule for which form of the kernel a step of T slots runs (engine.cu xf_mode)
static bool gemv_hook_twin(int T) {
    const char* v = getenv("RT_DEC_F16");
    return T > 4 && T <= 8 && !(v && atoi(v) == 0);
}

int rt_k_gemv(const uint16_t* w, const uint16_t* x, float* y, int32_t N, int32_t K, int32_t T, int32_t device) {
    return guard(nullptr, [&] {
        RT_REQUIRE(w && x && y, "null argument");
        RT_REQUIRE(N > 0 && N % 16 == 0 && K > 0 && K % 32 == 0 && T >= 1 && T <= MAX_SLOTS, "need N%16==0, K%32==0, 1<=T<=32");
        Scratch s(device);
        bf16* dw = s.alloc<bf16>((size_t)N * K, false);
        bf16* dx = s.alloc<bf16>(xf_elems(K));
        float* dy = s.alloc<float>((size_t)MAX_SLOTS * N);
        std::vector<uint16_t> cm((size_t)N * K);
        wgen::rowmajor_to_cm(w, cm.data(), N, K);     // the kernel streams the core-matrix-major layout (gemv.cuh)
        // the engine's own rule (engine.cu xf_mode): 5..8 slots stream the fp16 twin of the matrix against the fp16 inputs
        // themselves (gemv.cuh mode 2), otherwise the bf16 matrix against their bf16 hi + lo parts; XF layout either way
        const bool twin = gemv_hook_twin(T);
        if (twin)
            for (auto& v : cm) v = __half_as_ushort(__float2half_rn(__bfloat162float(__ushort_as_bfloat16(v))));
        std::vector<bf16> xf(xf_elems(K), __float2bfloat16_rn(0.f));
        for (int t = 0; t < T; ++t)
            for (int k = 0; k < K; ++k) {
                if (twin) {
                    xf[xf_index(t, k, K)] = __ushort_as_bfloat16(x[(size_t)t * K + k]);      // the fp16 bits, as they are
                    continue;
                }
                const float h = __half2float(__ushort_as_half(x[(size_t)t * K + k]));
                const bf16 hi = __float2bfloat16_rn(h);
                xf[xf_index(t, k, K)] = hi;
                xf[xf_index(t, k, K) + 256] = __float2bfloat16_rn(h - __bfloat162float(hi));
            }
        RT_CUDA(cudaMemcpyAsync(dw, cm.data(), (size_t)N * K * 2, cudaMemcpyHostToDevice, s.e.st));
        RT_CUDA(cudaMemcpyAsync(dx, xf.data(), xf.size() * 2, cudaMemcpyHostToDevice, s.e.st));
        GemvArgs a{};
        a.W = dw; a.X = dx; a.N = N; a.K = K; a.ldx = K; a.T = T; a.out = dy; a.ldo = N; a.f16 = twin;
        rt_gemv_host_dispatch(&s.e, a, EPI_STORE);
        RT_CUDA(cudaGetLastError());
        RT_CUDA(cudaMemcpyAsync(y, dy, (size_t)T * N * 4, cudaMemcpyDeviceToHost, s.e.st));
        RT_CUDA(cudaStreamSynchronize(s.e.st));
    });
}

int rt_k_rmsnorm(const float* x, const uint16_t* w, uint16_t* y, int32_t T, int32_t H, float eps, int32_t device) {
    return guard(nullptr, [&] {
        RT_REQUIRE(x && w && y, "null argument");
        RT_REQUIRE(T >= 1 && H > 0 && H % 8 == 0, "need H%8==0");
        Scratch s(device);
        float* dx = s.alloc<float>((size_t)T * H, false);
        bf16* dw = s.alloc<bf16>(H, false);
        bf16* dy = s.alloc<bf16>((size_t)T * H);
        RT_CUDA(cudaMemcpyAsync(dx, x, (size_t)T * H * 4, cudaMemcpyHostToDevice, s.e.st));
        RT_CUDA(cudaMemcpyAsync(dw, w, (size_t)H * 2, cudaMemcpyHostToDevice, s.e.st));
        rmsnorm_kernel<<<T, 256, 0, s.e.st>>>(dx, nullptr, dw, dy, H, H, eps);
        RT_CUDA(cudaGetLastError());
        RT_CUDA(cudaMemcpyAsync(y, dy, (size_t)T * H * 2, cudaMemcpyDeviceToHost, s.e.st));
        RT_CUDA(cudaStreamSynchronize(s.e.st));
    });
}

int rt_k_sample_topkp(const float* logits, int32_t T, int32_t V, float temperature, int32_t top_k, float top_p, uint64_t seed,
                      int32_t step, int32_t* tok_out, int32_t* kept_out, float* thresh_out, int32_t device) {
    return guard(nullptr, [&] {
        RT_REQUIRE(logits && tok_out, "null argument");
        RT_REQUIRE(T >= 1 && T <= MAX_SLOTS && V >= 1, "bad shape");
        RT_REQUIRE(temperature > 0.f && top_k >= 0 && top_p >= 0.f && top_p <= 1.f, "need temperature > 0, top_k >= 0, 0 <= top_p <= 1");
        Scratch s(device);
        float* dl = s.alloc<float>((size_t)T * V, false);
        int* dt = s.alloc<int>(T);
        int* dk = s.alloc<int>(T);
        float* dth = s.alloc<float>(T);
        RT_CUDA(cudaMemcpyAsync(dl, logits, (size_t)T * V * 4, cudaMemcpyHostToDevice, s.e.st));
        TopkpArgs k{};
        k.logits = dl; k.ld = V; k.V = V; k.T = T; k.inv_temp = 1.0f / temperature; k.top_k = top_k; k.top_p = top_p;
        k.seed = seed; k.step_const = step; k.next_tok = dt; k.kept_out = dk; k.thresh_out = dth;
        topkp_sample_kernel<<<T, TOPKP_THREADS, 0, s.e.st>>>(k);
        RT_CUDA(cudaGetLastError());
        RT_CUDA(cudaMemcpyAsync(tok_out, dt, T * 4, cudaMemcpyDeviceToHost, s.e.st));
        if (kept_out) RT_CUDA(cudaMemcpyAsync(kept_out, dk, T * 4, cudaMemcpyDeviceToHost, s.e.st));
        if (thresh_out) RT_CUDA(cudaMemcpyAsync(thresh_out, dth, T * 4, cudaMemcpyDeviceToHost, s.e.st));
        RT_CUDA(cudaStreamSynchronize(s.e.st));
    });
}

int rt_k_attention(const uint16_t* q, const uint16_t* k, const uint16_t* v, const int32_t* lens, float* out, int32_t T,
                   int32_t ctx, int32_t n_q, int32_t n_kv, int32_t hd, int32_t shared, int32_t chunk, int32_t device) {
    return guard(nullptr, [&] {
        RT_REQUIRE(q && k && v && lens && out, "null argument");
        RT_REQUIRE(T >= 1 && T <= MAX_SLOTS && ctx >= 1 && n_kv >= 1 && n_q % n_kv == 0 && n_q / n_kv <= 16, "bad shape");
        RT_REQUIRE(hd == 32 || hd == 64 || hd == 128, "head_dim must be 32, 64 or 128");
        RT_REQUIRE(shared >= 0 && shared % KV_PAGE == 0, "shared must be a multiple of the page size (16)");
        RT_REQUIRE(chunk >= KV_PAGE && chunk % KV_PAGE == 0, "chunk must be a multiple of 16");
        int max_len = 0;
        for (int t = 0; t < T; ++t) {
            RT_REQUIRE(lens[t] >= 1 && lens[t] <= ctx && lens[t] > shared, "need shared < lens[t] <= ctx");
            max_len = std::max(max_len, lens[t]);
        }
        const int G = n_q / n_kv;
        const int n_sh = cdiv(shared, chunk), n_own = cdiv(max_len - shared, chunk);
        RT_REQUIRE(n_sh + n_own + 1 <= MAX_CHUNKS, "too many chunks");
        Scratch s(device);
        rt_attn_host_setup();
        // paginate on the host: sequence t owns pages for [shared, len); sequence 0 owns the shared pages
        const int pps = cdiv(ctx, KV_PAGE);
        std::vector<int> pt((size_t)T * pps, 0);
        int next_page = 0;
        const int sh_pages = shared / KV_PAGE;
        for (int p = 0; p < sh_pages; ++p) {
            for (int t = 0; t < T; ++t) pt[(size_t)t * pps + p] = next_page;
            next_page++;
        }
        for (int t = 0; t < T; ++t)
            for (int p = sh_pages; p < cdiv(lens[t], KV_PAGE); ++p) pt[(size_t)t * pps + p] = next_page++;
        const size_t page_elems = (size_t)n_kv * KV_PAGE * hd;
        std::vector<uint16_t> hk((size_t)next_page * page_elems, 0), hv((size_t)next_page * page_elems, 0);
        for (int t = 0; t < T; ++t)
            for (int pos = (t == 0 ? 0 : shared); pos < lens[t]; ++pos) {
                const int page = pt[(size_t)t * pps + pos / KV_PAGE];
                for (int h = 0; h < n_kv; ++h) {
                    const size_t src = (((size_t)t * ctx + pos) * n_kv + h) * hd;
                    const size_t dst = ((size_t)page * n_kv + h) * KV_PAGE * hd + (size_t)(pos % KV_PAGE) * hd;
                    memcpy(&hk[dst], k + src, (size_t)hd * 2);
                    memcpy(&hv[dst], v + src, (size_t)hd * 2);
                }
            }
        bf16* dk = s.alloc<bf16>(hk.size(), false);
        bf16* dv = s.alloc<bf16>(hv.size(), false);
        bf16* dq = s.alloc<bf16>((size_t)T * n_q * hd, false);
        int* dpt = s.alloc<int>(pt.size(), false);
        std::vector<int> hseq(T), hpos(T);
        for (int t = 0; t < T; ++t) { hseq[t] = t; hpos[t] = lens[t] - 1; }
        int* dseq = s.alloc<int>(T, false);
        int* dpos = s.alloc<int>(T, false);
        int* dsh = s.alloc<int>(1, false);
        float* po = s.alloc<float>((size_t)T * n_q * MAX_CHUNKS * hd);
        float* pml = s.alloc<float>((size_t)T * n_q * MAX_CHUNKS * 2);
        float* dout = s.alloc<float>((size_t)T * n_q * hd);
        cudaStream_t st = s.e.st;
        RT_CUDA(cudaMemcpyAsync(dk, hk.data(), hk.size() * 2, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dv, hv.data(), hv.size() * 2, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dq, q, (size_t)T * n_q * hd * 2, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dpt, pt.data(), pt.size() * 4, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dseq, hseq.data(), T * 4, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dpos, hpos.data(), T * 4, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(dsh, &shared, 4, cudaMemcpyHostToDevice, st));
        AttnArgs a{};
        a.Q = dq; a.ldq = n_q * hd; a.Kp = dk; a.Vp = dv; a.page_table = dpt; a.pt_stride = pps; a.slot_seq = dseq; a.slot_pos = dpos;
        a.shared_len_ptr = dsh; a.T = T; a.n_q = n_q; a.n_kv = n_kv; a.chunk_pages = chunk / KV_PAGE;
        a.n_shared_tiles = shared > 0 ? cdiv(T * G, 16) : 0;
        a.part_o = po; a.part_ml = pml; a.max_chunks = MAX_CHUNKS;
        a.scale_log2 = (float)(1.4426950408889634 / std::sqrt((double)hd));
        rt_attn_host_launch(&s.e, a, hd, cdiv(std::max(n_sh, n_own), ATTN_WARPS));
        if (const char* it = getenv("RT_ATTN_BENCH_ITERS")) {   // profiling aid: time the partials kernel alone
            const int iters = atoi(it);
            cudaEvent_t e0, e1;
            RT_CUDA(cudaEventCreate(&e0));
            RT_CUDA(cudaEventCreate(&e1));
            RT_CUDA(cudaEventRecord(e0, st));
            for (int i = 0; i < iters; ++i) rt_attn_host_launch(&s.e, a, hd, cdiv(std::max(n_sh, n_own), ATTN_WARPS));
            RT_CUDA(cudaEventRecord(e1, st));
            RT_CUDA(cudaStreamSynchronize(st));
            float ms = 0.f;
            RT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            fprintf(stderr, "[attn bench] T=%d n_q=%d n_kv=%d hd=%d max_len=%d shared=%d chunk=%d: %.2f us per launch\n", T, n_q, n_kv, hd,
                    max_len, shared, chunk, 1e3 * ms / iters);
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        }
        MergeArgs m{};
        m.part_o = po; m.part_ml = pml; m.max_chunks = MAX_CHUNKS; m.slot_pos = dpos; m.shared_len_ptr = dsh;
        m.cascade = shared > 0; m.chunk_keys = chunk; m.n_q = n_q; m.hd = hd; m.out_f32 = dout; m.ldo = n_q * hd;
        attn_merge_kernel<<<dim3(n_q, T), hd, 0, st>>>(m);
        RT_CUDA(cudaGetLastError());
        RT_CUDA(cudaMemcpyAsync(out, dout, (size_t)T * n_q * hd * 4, cudaMemcpyDeviceToHost, st));
        RT_CUDA(cudaStreamSynchronize(st));
    });
}

int rt_k_gemm_tc5(const uint16_t* a_bits, const uint16_t* w_bits, float* c, int32_t S, int32_t N, int32_t K, int32_t device) {
    return guard(nullptr, [&] {
        RT_REQUIRE(a_bits && w_bits && c, "null argument");
        RT_REQUIRE(S >= 1 && N > 0 && N % 32 == 0 && K > 0 && K % 64 == 0, "need N % 32 == 0 and K % 64 == 0");
        Scratch s(device);
        const int S_pad = cdiv(S, 128) * 128;
        f16* da = s.alloc<f16>((size_t)S_pad * K);
        bf16* dw = s.alloc<bf16>((size_t)N * K, false);
        f16* dw16 = s.alloc<f16>((size_t)N * K, false);
        float* dc = s.alloc<float>((size_t)S * N);
        std::vector<uint16_t> cm((size_t)N * K);
        wgen::rowmajor_to_cm(w_bits, cm.data(), N, K);
        RT_CUDA(cudaMemcpyAsync(da, a_bits, (size_t)S * K * 2, cudaMemcpyHostToDevice, s.e.st));
        RT_CUDA(cudaMemcpyAsync(dw, cm.data(), (size_t)N * K * 2, cudaMemcpyHostToDevice, s.e.st));
        PfTensorMap ma, mw;
        pf_make_map_a(&ma, da, S_pad, K);
        bf16_to_f16_kernel<<<1024, 256, 0, s.e.st>>>(dw, dw16, (size_t)N * K / 8);     // the fp16 twin, as the engine makes it
        pf_make_map_w(&mw, dw16, N, K, PF_BN);
        Gemm5Args g{};
        g.ab_f16 = 1; g.S = S; g.N = N; g.K = K; g.out = dc; g.ldo = N;
        pf_gemm(s.e.st, false, G5_EPI_STORE, ma, mw, g);
        RT_CUDA(cudaGetLastError());
        if (const char* it = getenv("RT_GEMM_BENCH_ITERS")) {
            const int iters = atoi(it);
            cudaEvent_t e0, e1;
            RT_CUDA(cudaEventCreate(&e0));
            RT_CUDA(cudaEventCreate(&e1));
            RT_CUDA(cudaEventRecord(e0, s.e.st));
            for (int i = 0; i < iters; ++i) pf_gemm(s.e.st, false, G5_EPI_STORE, ma, mw, g);
            RT_CUDA(cudaEventRecord(e1, s.e.st));
            RT_CUDA(cudaStreamSynchronize(s.e.st));
            float ms = 0.f;
            RT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            fprintf(stderr, "[gemm bench] S=%d N=%d K=%d: %.2f us per launch, %.1f TFLOP/s\n", S, N, K, 1e3 * ms / iters,
                    2.0 * S * N * K / (ms / iters * 1e-3) / 1e12);
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        }
        RT_CUDA(cudaMemcpyAsync(c, dc, (size_t)S * N * 4, cudaMemcpyDeviceToHost, s.e.st));
        RT_CUDA(cudaStreamSynchronize(s.e.st));
    });
}

int rt_k_gemv_bench(int32_t N, int32_t K, int32_t T, int32_t iters, int32_t epi, int32_t device, double* ms_out) {
    return guard(nullptr, [&] {
        RT_REQUIRE(ms_out != nullptr, "null argument");
        RT_REQUIRE(N > 0 && N % 16 == 0 && K > 0 && K % 32 == 0 && T >= 1 && T <= MAX_SLOTS && iters >= 1, "bad shape");
        RT_REQUIRE(epi == EPI_STORE || epi == EPI_SWIGLU, "epi must be 0 (store) or 2 (SwiGLU)");
        Scratch s(device);
        const size_t mat = (size_t)N * K;
        const int copies = (int)std::max<size_t>(2, (size_t)400e6 / (mat * 2) + 1);   // > 3x L2 in rotation
        bf16* dw = s.alloc<bf16>(mat * copies);
        bf16* dx = s.alloc<bf16>(xf_elems(K));
        float* dy = s.alloc<float>((size_t)MAX_SLOTS * N);
        GemvArgs a{};
        a.X = dx; a.N = N; a.K = K; a.ldx = K; a.T = T; a.out = dy; a.ldo = N; a.f16 = gemv_hook_twin(T);
        a.out_bf = s.alloc<bf16>(xf_elems(N / 2)); a.ldob = N / 2;               // SwiGLU: gate/up row pairs -> N/2 outputs (XF layout)
        cudaEvent_t e0, e1;
        RT_CUDA(cudaEventCreate(&e0));
        RT_CUDA(cudaEventCreate(&e1));
        for (int i = 0; i < 3; ++i) { a.W = dw + mat * (i % copies); rt_gemv_host_dispatch(&s.e, a, epi); }
        RT_CUDA(cudaEventRecord(e0, s.e.st));
        for (int i = 0; i < iters; ++i) { a.W = dw + mat * (i % copies); rt_gemv_host_dispatch(&s.e, a, epi); }
        RT_CUDA(cudaEventRecord(e1, s.e.st));
        RT_CUDA(cudaStreamSynchronize(s.e.st));
        RT_CUDA(cudaGetLastError());
        float ms = 0.f;
        RT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *ms_out = (double)ms / iters;
    });
}

}  // extern "C"
