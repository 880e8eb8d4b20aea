// Engine state of the generation backend: weights, paged KV pool, per-step device state, workspaces.
// Host-side C++; replaces the external server process behind local-llm.ts:159 (see include/rt_engine.h).
#pragma once
#include <string>
#include <vector>
#include "../../include/rt_engine.h"
#include "common.cuh"
#include "prefill.cuh"
#include "tp.cuh"

constexpr int MAX_SLOTS = 32;      // token slots per forward step (decode: knights; slab prefill: tokens)
constexpr int MAX_CHUNKS = 136;    // attention partials per (slot, head): <= 64 shared + <= 64 own (+ rounding)

struct LayerW {
    bf16* wqkv = nullptr;   // [(nq_l + 2*nkv_l)*hd][H]  rows: q heads, k heads, v heads
    float* bqkv = nullptr;  // [(nq_l + 2*nkv_l)*hd] or null
    bf16* wo = nullptr;     // [H][nq_l*hd]
    bf16* wgu = nullptr;    // [2*F_l][H]  per 16-row tile: 8 gate rows then the 8 matching up rows
    bf16* wdown = nullptr;  // [H][F_l]
    // fp16 twins of the four matrices (same CM layout), read by the tcgen05 prefill GEMMs: kind::f16 wants both operands
    // in ONE 16-bit format and the activations are fp16 (common.cuh).  Every bf16 weight with |w| >= 2^-14 is exact in fp16.
    f16 *wqkv16 = nullptr, *wo16 = nullptr, *wgu16 = nullptr, *wdown16 = nullptr;
    bf16* norm1 = nullptr;  // [H]
    bf16* norm2 = nullptr;  // [H]
};

struct StepOpts {
    int T = 0;                 // valid slots
    bool sample = false;       // run lm_head + arg-max
    bool advance = false;      // decode: feed the sampled token back, pos += 1, record it
    bool store_logits = false; // write logits[T][vocab]
    int chunk_pages = 4;
    int n_chunk_blocks = 1;    // grid.x of the attention kernel (version 1: 4 chunks per CTA)
    int n_chunks = 1;          // max(n_sh, n_own)
    int n_sh = 0, n_own = 1;   // chunks of the shared prefix / of the longest own region (version 2 grid)
    int n_shared_tiles = 0;    // > 0 => cascade phase on
    int n_partials_max = 1;    // most partials any (slot, head) will have (fused merge capacity check)
    float inv_temp = 0.f;
    unsigned long long seed = 0;
    int attn3_cs = 0;          // > 0: cluster-merged decode attention (attn_decode3.cuh) with this cluster size
    bool qkv_late = false;     // QKV projections release their dependent after their own wait (multi-step graphs)
    bool in_graph = false;     // the step is being captured into the decode graph (consecutive steps fully serialised)
    int top_k = 0;             // sampler.cuh; only with inv_temp > 0
    // tcgen05 decode attention for long contexts (attn_decode_tc5.cuh): work items of the call, 0 = not used
    int dec5_items = 0, dec5_nsh = 0, dec5_own_span = 0, dec5_nown = 0;
    float top_p = 0.f;
};

struct rt_engine {
    rt_config cfg{};
    int dev = 0, sm_count = 148;
    cudaStream_t st = nullptr;
    std::string err;
    long long launches = 0;
    bool from_checkpoint = false;   // weights came from rt_config.weights_path (its vocabulary is not the byte tokenizer's)
    int n_eos = 1, eos_ids[3] = {2, -1, -1};

    // global dims and this rank's shard
    int L = 0, H = 0, hd = 0, nq = 0, nkv = 0, F = 0, vocab = 0;
    int nq_l = 0, nkv_l = 0, F_l = 0, vocab_l = 0, G = 1, qkv_rows = 0;

    bf16 *embed = nullptr, *lm_head = nullptr, *final_norm = nullptr;
    f16* lm_head16 = nullptr;      // fp16 twin of the lm_head (decode steps with 5..8 slots, gemv.cuh mode 2)
    bool dec_f16 = false;          // twins exist and RT_DEC_F16 != 0
    std::vector<LayerW> layers;
    std::vector<void*> allocs;     // everything to cudaFree
    float *cos_t = nullptr, *sin_t = nullptr;

    // paged KV: per layer [n_pages][nkv_l][16][hd]
    bf16 *kpool = nullptr, *vpool = nullptr;
    size_t layer_pool_elems = 0;
    int n_pages = 0, max_pages_per_seq = 0, n_seq = 0;
    std::vector<int> free_pages;
    std::vector<std::vector<int>> seq_pages;  // pages owned (or shared) by each sequence, in order
    std::vector<int> page_ref;
    int* h_page_table = nullptr;  // pinned [n_seq][max_pages_per_seq]
    int* d_page_table = nullptr;
    std::vector<int> seq_len;
    // cross-call prefix cache: token ids of past prompts + the pages holding their KV (ref-counted)
    struct CacheEntry {
        std::vector<int> tokens;
        std::vector<int> pages;
        long long stamp = 0;
    };
    std::vector<CacheEntry> cache;
    long long cache_clock = 0;

    // step state (device)
    int *d_tok = nullptr, *d_seq = nullptr, *d_pos = nullptr, *d_shared_len = nullptr, *d_step = nullptr;
    int *d_next_tok = nullptr, *d_out_tokens = nullptr;
    int max_out = 0;
    int* h_stage = nullptr;       // pinned staging, 4 * MAX_SLOTS + misc ints

    // activations (MAX_SLOTS rows each)
    float *x = nullptr, *qkv = nullptr, *delta = nullptr, *logits = nullptr;
    // projection inputs of the decode kernels: fp16 values as bf16 hi + lo in the XF layout (common.cuh)
    bf16 *xn = nullptr, *attn_bf = nullptr, *act_bf = nullptr;
    bf16* qb = nullptr;
    float *part_o = nullptr, *part_ml = nullptr;
    long long pf_attn_bytes = 0;   // weights pulled into L2 while a layer's attention runs (RT_PF_ATTN_MB)
    long long pf_tail_bytes = 0;   // head of the NEXT projection's weights each projection asks for at its tail (RT_PF_TAIL_MB)
    float* ss_part = nullptr;      // [H/16][32] per-tile partial sums of x^2 (split RMSNorm, gemv.cuh)
    float* inv_rms = nullptr;      // [32] r per slot when a whole row was available (tensor parallel, prefill rows)
    float* pf_inv_rms = nullptr;   // [pf_cap]
    int* done_counter = nullptr;   // arrival counter of the sampler tails (lm_head, top-k/top-p)
    bool use_pdl = true;
    float* cta_max = nullptr;
    int* cta_idx = nullptr;
    int lm_ctas = 0;

    // tensor-core prefill path: activations of one chunk of <= pf_cap prompt tokens, TMA tensor maps
    int pf_cap = 0, pf_work_cap = 0;
    bool pf_ok = false;
    float* pf_x = nullptr;
    f16 *pf_xn = nullptr, *pf_attn = nullptr, *pf_act = nullptr;
    bf16* pf_q = nullptr;
    int *pf_tok = nullptr, *pf_seq = nullptr, *pf_pos = nullptr;
    int4* pf_work = nullptr;
    int* pf_host = nullptr;        // pinned staging: tok, seq, pos (pf_cap each) + work items
    PfTensorMap map_xn{}, map_attn{}, map_act{};
    PfTensorMap map_q3{}, map_kpool{}, map_vpool{};   // tcgen05 prefill attention (head_dim 128, opt-in)
    PfTensorMap map_qdec{};           // the decode queries qb as (dims, heads, slots) for attn_decode_tc5.cuh
    bool dec5_ok = false;
    int dec5_bt = 0;
    void* dec5_items_host = nullptr;  // Dec5Item[dec5_cap]: the work items of the current plan (a kernel parameter at launch)
    int dec5_cap = 0;
    bool pfa_tc5 = false;
    std::vector<PfTensorMap> map_w;   // per layer: qkv, o, gate/up, down

    // tensor parallel
    TpComm* tp = nullptr;
    float* loc_best = nullptr;     // [64 words]: 32 local winning scores, 32 local winning ids
    float* all_best = nullptr;     // [tp][64 words]
    float* logits_full = nullptr;  // [32][vocab]: all ranks' logits shards, gathered for top-k / top-p sampling
    float* tp_stop = nullptr;      // [1] the ranks' collective "deadline passed" flag
    f16* pf_delta = nullptr;       // [pf_cap][H] partial sums of the row-parallel prefill GEMMs, fp16 (what crosses NVLink)
    // fused peer-memory all-reduce (tp_fused.cuh): two partial buffers (O-proj / down-proj call sites), flags, epochs
    bool tpf_on = false;
    float* tpf_delta[2] = {nullptr, nullptr};
    int *tpf_flags = nullptr, *tpf_epoch = nullptr;
    void* tpf_peer_delta[2][8] = {};
    void* tpf_peer_flags[8] = {};
    // push form of the exchange (tp_fused.cuh tp_ll_reduce_norm_kernel, <= 8 slots): per call site a receive buffer
    // [tp][32][H], the receiver CTAs' private run counters, the senders' per-slot epochs
    bool tpl_on = false;
    float* tpl_recv[2] = {nullptr, nullptr};
    void* tpl_peer_recv[2][8] = {};
    int* tpl_cell_epoch[2] = {nullptr, nullptr};
    int* tpl_send_epoch[2] = {nullptr, nullptr};

    // cached decode graph
    cudaGraphExec_t graph_exec = nullptr;       // one decode step
    cudaGraphExec_t graph_exec_multi = nullptr; // graph_steps consecutive decode steps (no launch gap between them)
    StepOpts graph_opts{};
    int graph_nodes = 0, graph_steps = 1;

    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};

    // teacher-forcing scratch sequence (seq 0)
    bool scratch_active = false;
};

// engine.cu
void engine_init(rt_engine* e);
void engine_free(rt_engine* e);
void engine_reset_sequences(rt_engine* e);
void engine_alloc_pages(rt_engine* e, int seq, int n_tokens_total);
void engine_share_prefix(rt_engine* e, int from_seq, int to_seq, int n_pages);
void engine_upload_page_table(rt_engine* e);
void engine_forward(rt_engine* e, const StepOpts& o);
void engine_set_slots(rt_engine* e, const int* tok, const int* seq, const int* pos, int T);
void engine_generate_ids(rt_engine* e, const int32_t* const* prompts, const int32_t* lens, int n,
                         const rt_sampling* s, uint64_t deadline_ms, rt_result* outs);
void engine_step_logits(rt_engine* e, const int32_t* ids, int n, int mode, float* logits_out);
void engine_export_weight(rt_engine* e, int tensor_id, uint16_t* out, long long n_elems);

// tokenizer.cpp
int rt_tok_encode_chat(const char* s, size_t len, std::vector<int>& ids);
std::string rt_tok_decode(const int* ids, int n);
