/*
 * rt_engine.h -- C ABI of the B200-native generation backend behind TheRoundtAIble's
 * `local-llm` adapter.
 *
 * The reference adapter (src/adapters/local-llm.ts) is an HTTP client: `execute(prompt)`
 * POSTs `{model, messages:[{role:"user",content:prompt}]}` to an external server and returns
 * `choices[0].message.content`.  This library replaces what happens on the far side of that
 * request.  Every entry point cites the reference interface it stands in for; the bindings a
 * maintainer would add (N-API, HTTP shim, ctypes) are shown in INTEGRATION.md.
 *
 * Conventions: plain C types only; every function returns 0 on success and a negative
 * rt_status on failure; the text of the last failure is available through rt_last_error();
 * no exception crosses the boundary; all pointers are caller-owned unless stated; one engine
 * is used by one thread at a time (the reference never issues concurrent execute() calls,
 * src/orchestrator.ts:361-442).  There is NO CPU fallback: creating an engine without a
 * CUDA device fails with RT_ERR_CUDA.
 */
#ifndef RT_ENGINE_H
#define RT_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_API __attribute__((visibility("default")))

typedef struct rt_engine rt_engine;

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,   /* bad argument / unsupported shape */
    RT_ERR_CUDA = -2,      /* CUDA runtime / driver failure, or no device */
    RT_ERR_OOM = -3,       /* KV pool or HBM exhausted; text contains "context length exceeded" when a prompt does not fit */
    RT_ERR_TIMEOUT = -4,   /* deadline passed; text contains "timed out" (src/utils/errors.ts:102 -> kind "timeout") */
    RT_ERR_IO = -5,        /* weight file problems; text contains "not found" where a file is missing */
    RT_ERR_NCCL = -6
} rt_status;

enum { RT_FAMILY_LLAMA = 0, RT_FAMILY_QWEN2 = 1 };

/* rt_config.flags */
enum {
    RT_FLAG_NO_GRAPH = 1,        /* launch decode steps kernel by kernel instead of replaying a CUDA graph */
    RT_FLAG_NO_PREFIX_SHARE = 2, /* batch calls: do not share the common token prefix (every knight prefills and reads its own copy) */
    RT_FLAG_PREFILL_VIA_DECODE = 4, /* debug: run prompts through the decode kernels in 32-token slabs instead of the tcgen05 prefill path */
    RT_FLAG_PREFIX_CACHE = 8     /* keep the KV pages of past prompts and reuse the longest page-aligned common token prefix of a new
                                    prompt (a knight's next-round prompt extends its last one: src/utils/prompt.ts:60-77) */
};

/* Model + runtime description.  Stands in for AdapterLocalConfig {endpoint, model, name?, source?}
 * (src/types.ts:29-34): `model` there is only a name; the dimensions it implies live in the
 * external server.  rt_arch_by_name() fills the architecture half from such a name. */
typedef struct rt_config {
    uint32_t struct_size;        /* = sizeof(rt_config) */
    /* architecture */
    int32_t family;              /* RT_FAMILY_* */
    int32_t layers, hidden, n_q, n_kv, head_dim, ffn, vocab;
    int32_t tied;                /* lm_head shares the embedding table */
    int32_t qkv_bias;
    float rope_theta, rms_eps;
    float rope_factor;           /* llama3 rope scaling; 0 = plain rope */
    float rope_low_freq, rope_high_freq;
    int32_t rope_orig_ctx;
    /* weights */
    uint64_t seed;               /* seeded random-init weights (see DESIGN.md "weights") */
    const char* weights_path;    /* optional: a .safetensors file or a directory of them (Hugging Face Llama / Qwen2 tensor
                                    names, BF16 / F16 / F32); NULL or "" = seeded random weights.  Under tensor parallel
                                    every rank reads its own rows / columns from the same path */
    /* runtime */
    int32_t device;              /* CUDA ordinal */
    int32_t tp_rank, tp_size;    /* tensor parallel: one engine per GPU/process */
    const void* nccl_id;         /* tp_size > 1: the 128-byte ncclUniqueId from rt_nccl_unique_id() on rank 0 */
    int32_t max_batch;           /* knights in flight, 1..8 */
    int32_t max_ctx;             /* tokens per sequence (prompt + generated) */
    int64_t kv_pool_bytes;       /* 0 = size for max_batch * max_ctx */
    int32_t flags;               /* RT_FLAG_* */
    int32_t n_eos;               /* end-of-sequence ids in eos_ids (0..3); 0 = the byte tokenizer's <eos> (id 2), which is
                                    what the seeded weights are used with.  A checkpoint's ids come from its config.json /
                                    generation_config.json (Llama-3: 128001, 128009; Qwen2.5: 151645, 151643) */
    int32_t eos_ids[3];
    int32_t reserved;
} rt_config;

/* Sampling controls.  The reference request carries none (local-llm.ts:162-165: no max_tokens,
 * no temperature); the server picks.  Defaults here: greedy, 512 new tokens (templates/system-prompt.md:17
 * asks for <= 500 words), stop at EOS. */
typedef struct rt_sampling {
    int32_t max_new_tokens;      /* <= 0 -> 512 */
    float temperature;           /* 0 = greedy (argmax, lowest index wins ties) */
    int32_t top_k;               /* 0 = off; with temperature > 0: keep the k highest logits (ties at the k-th kept) */
    float top_p;                 /* 0 or 1 = off; with temperature > 0: keep the smallest set of highest-probability tokens
                                    (after top_k) whose mass reaches top_p (csrc/sampler.cuh; under tensor parallel the
                                    vocab-parallel logits are all-gathered and every rank draws the same token) */
    uint64_t seed;
    int32_t ignore_eos;
    int32_t reserved;
} rt_sampling;

/* One generation result (engine-owned buffers; release with rt_result_free).  `text` is what
 * the adapter returns from execute() (local-llm.ts:184-193); it is never empty on success. */
typedef struct rt_result {
    char* text;                  /* UTF-8, NUL-terminated */
    size_t text_len;
    int32_t* tokens;             /* generated token ids */
    int32_t n_tokens;
    int32_t n_prompt_tokens;
    int32_t prefix_shared_tokens;/* prompt tokens served from the shared prefix instead of being prefilled again */
    int32_t finish_reason;       /* 0 length, 1 eos */
    double prefill_ms;           /* CUDA-event time of the prefill phase of the CALL */
    double decode_ms;            /* CUDA-event time of the decode phase of the CALL */
    int32_t decode_steps;        /* decode steps timed in decode_ms */
    int32_t gpu_launches;        /* kernels this call launched (graph replays count their nodes) */
    int32_t prefix_cached_tokens;/* prompt tokens whose KV came from the cross-call prefix cache (RT_FLAG_PREFIX_CACHE) */
    int32_t reserved;
} rt_result;

typedef struct rt_model_info {
    int32_t context_length;      /* what /api/show's "*.context_length" would say (local-llm.ts:205-235) */
    int32_t vocab;
    int64_t n_params;
    int64_t weight_bytes_per_step;  /* bf16 weight bytes one decode step reads on this rank */
    int64_t kv_bytes_per_token;     /* on this rank */
    int32_t max_batch;
    int32_t page_size;
    int32_t n_pages;
    int32_t sm_count;
} rt_model_info;

/* Fill the architecture fields of *cfg from a model name ("llama-3.2-1b", "llama-3.1-8b",
 * "llama-3.1-70b", "qwen2.5-0.5b", "qwen2.5-32b", "tiny-*"); optional "@L<n>" suffix truncates
 * the layer count (large-config parity).  Replaces: the `model` string of
 * AdapterLocalConfig (src/types.ts:31) being resolved by the external server. */
RT_API int rt_arch_by_name(const char* name, rt_config* cfg);

/* Replaces: the external server process the adapter's `endpoint` points at
 * (src/utils/adapters.ts:47-51 constructs the adapter; local-llm.ts:15-22 stores endpoint+model). */
RT_API int rt_engine_create(const rt_config* cfg, rt_engine** out);
RT_API void rt_engine_destroy(rt_engine* e);

/* Text of the last failure on this engine (or of the last failed call without an engine when
 * e == NULL).  Replaces: the HTTP error body the adapter wraps into
 * `Local LLM error (status): body` (local-llm.ts:181) before classifyError (errors.ts:86-126). */
RT_API const char* rt_last_error(const rt_engine* e);

/* Replaces: GET {endpoint}/v1/models as availability probe (local-llm.ts:24-36) and POST /api/show
 * for the context window (local-llm.ts:205-235). */
RT_API int rt_model_info_get(const rt_engine* e, rt_model_info* out);

/* SEQ path -- one prompt, what `roundtable discuss` really issues (one execute() per knight per
 * round, src/orchestrator.ts:442).  Replaces: POST /v1/chat/completions (local-llm.ts:150-199) /
 * POST /api/chat (local-llm.ts:95-144).  `deadline_ms` = timeoutMs of execute() (0 = none); under tensor parallel
 * every rank must pass the same value (the ranks agree on "stop" collectively, so none is left spinning).
 * The two text entry points use the built-in byte-level tokenizer and therefore refuse (RT_ERR_INVALID) an engine
 * created from a checkpoint (rt_config.weights_path): its vocabulary is not the byte tokenizer's; drive such an
 * engine with rt_generate_ids and the checkpoint's own tokenizer. */
RT_API int rt_generate(rt_engine* e, const char* prompt_utf8, size_t len, const rt_sampling* s,
                       uint64_t deadline_ms, rt_result* out);

/* BATCH path -- the N knight prompts of one round as one batched generation; the longest common
 * token prefix is prefilled once and shared.  No counterpart call in the reference today (knights
 * are strictly serial, orchestrator.ts:361); it is the north-star capability and SURVEY.md 8f row 4. */
RT_API int rt_generate_batch(rt_engine* e, const char* const* prompts, const size_t* lens, int32_t n,
                             const rt_sampling* s, uint64_t deadline_ms, rt_result* outs);

/* Same as rt_generate_batch but on token ids (no chat wrapper): the parity / bench harness entry. */
RT_API int rt_generate_ids(rt_engine* e, const int32_t* const* prompts, const int32_t* lens, int32_t n,
                           const rt_sampling* s, uint64_t deadline_ms, rt_result* outs);

/* Teacher forcing for parity: rt_seq_reset() clears the scratch sequence; rt_step_logits() appends
 * `n` tokens to it and writes the logits of each appended position to logits_out[n][vocab] (host,
 * fp32).  mode 0 = decode kernels, 1 = prefill kernels. */
RT_API int rt_seq_reset(rt_engine* e);
RT_API int rt_step_logits(rt_engine* e, const int32_t* ids, int32_t n, int32_t mode, float* logits_out);

RT_API void rt_result_free(rt_result* r);

/* Byte-level tokenizer shared with the oracle (DESIGN.md "tokenizer").  rt_tokenize_chat returns the
 * number of ids (writes at most cap); rt_detokenize returns the byte length (writes at most cap-1 + NUL). */
RT_API int32_t rt_tokenize_chat(const char* utf8, size_t len, int32_t* ids, int32_t cap);
RT_API int64_t rt_detokenize(const int32_t* ids, int32_t n, char* out, int64_t cap);

/* tp_size > 1: rank 0 calls this and hands the 128 bytes to every rank's rt_config.nccl_id. */
RT_API int rt_nccl_unique_id(void* out128);

/* ---- parity / measurement hooks (host pointers in, host pointers out) ------------------------ */

/* Copy a weight tensor of the engine back to the host as bf16 bits in HF layout [rows][cols]
 * (kind: 0 embed, 1 lm_head, 2 final_norm; per layer: 16*(layer+1) + {0 q,1 k,2 v,3 o,4 gate,5 up,
 * 6 down,7 in_norm,8 post_norm,9 q_bias,10 k_bias,11 v_bias}).  Proves the GPU generator equals the oracle's. */
RT_API int rt_weights_export(rt_engine* e, int32_t tensor_id, uint16_t* out, int64_t n_elems);

/* y[t][n] = sum_k W[n][k] * x[t][k]; W bf16 bits [N][K], x FP16 bits [T][K] (projection inputs are fp16 in this engine:
 * DESIGN.md 3 "numerics"), y fp32 [T][N]; T <= 32.  The product is exact in both operands, the sum is fp32.
 * The hook runs the form of the kernel the engine runs for a step of T slots: T = 5..8 streams the fp16 twin of W against
 * the fp16 inputs themselves, every other T streams the bf16 matrix against the inputs' bf16 hi + lo parts (RT_DEC_F16=0
 * forces the latter, as in the engine). */
RT_API int rt_k_gemv(const uint16_t* w, const uint16_t* x, float* y, int32_t N, int32_t K, int32_t T, int32_t device);
/* y = rmsnorm(x) * w, rounded to bf16: x fp32 [T][H], w bf16 bits [H], y bf16 bits [T][H]. */
RT_API int rt_k_rmsnorm(const float* x, const uint16_t* w, uint16_t* y, int32_t T, int32_t H, float eps, int32_t device);
/* top-k / top-p sampling (csrc/sampler.cuh) on host logits [T][V] at the given decode step: tok_out[T] the drawn
 * tokens, kept_out[T] (nullable) the size of the kept set, thresh_out[T] (nullable) the lowest kept logit (-inf when no
 * filter is active).  Parity hook for oracle/sampling.py; no reference counterpart (local-llm.ts:162-165 sends none). */
RT_API int rt_k_sample_topkp(const float* logits, int32_t T, int32_t V, float temperature, int32_t top_k, float top_p,
                             uint64_t seed, int32_t step, int32_t* tok_out, int32_t* kept_out, float* thresh_out,
                             int32_t device);
/* Paged GQA decode attention on contiguous inputs: q bf16 [T][n_q][hd] (T query tokens of T
 * sequences), k/v bf16 [T][ctx][n_kv][hd], lens[T]; `shared` leading tokens are identical in all
 * sequences and are read once (cascade) when shared > 0.  out fp32 [T][n_q][hd]. */
RT_API int rt_k_attention(const uint16_t* q, const uint16_t* k, const uint16_t* v, const int32_t* lens,
                          float* out, int32_t T, int32_t ctx, int32_t n_q, int32_t n_kv, int32_t hd,
                          int32_t shared, int32_t chunk, int32_t device);
/* Prefill projection on the tcgen05 tensor cores: c[s][n] = sum_k a[s][k] * w[n][k]; a FP16 bits [S][K], w bf16 bits
 * [N][K] (row-major; converted internally to the engine's core-matrix-major layout and to its fp16 twin), c fp32 [S][N]. */
RT_API int rt_k_gemm_tc5(const uint16_t* a, const uint16_t* w, float* c, int32_t S, int32_t N, int32_t K, int32_t device);
/* Time `iters` back-to-back launches of the decode projection kernel on an [N][K] bf16 matrix that is re-streamed
 * from HBM (rotating over enough copies to exceed L2); epi 0 = plain store, 2 = the fused SwiGLU epilogue of the
 * gate/up projection (the dominant kernel of a decode step); returns mean ms per launch.  Same choice of kernel form by
 * T as rt_k_gemv. */
RT_API int rt_k_gemv_bench(int32_t N, int32_t K, int32_t T, int32_t iters, int32_t epi, int32_t device, double* ms_out);

/* Paged GQA decode attention for callers that must not allocate.  It has no counterpart in the
 * reference.  It exists for two kinds of caller: a process that has already given (almost) all of
 * HBM to its own pool allocator -- a serving framework embedding this library -- where a cudaMalloc
 * made inside the library would fail or fragment that pool; and a latency-critical token loop that
 * cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply.  Everything is passed as
 * DEVICE pointers the caller owns:
 *   q_dev          bf16 [T][n_q][hd]           one query token per sequence, RoPE already applied
 *   kpool_dev/vpool_dev  bf16 [pages][n_kv][16][hd]   the paged KV pool (page = 16 tokens)
 *   page_table_dev int32 [T][pt_stride]        page ids of sequence t, in order
 *   query_pos_dev  int32 [T]                   position of query t = number of earlier tokens; it attends keys 0..pos
 *   out_dev        fp32 [T][n_q][hd]
 * max_len is an upper bound of query_pos+1 over the batch (a position beyond it is clamped on the
 * device: that query then attends only its first max_len keys); chunk (multiple of 16) is the split-KV
 * granularity.  The launch is asynchronous on `stream` (a cudaStream_t; NULL = default stream), and
 * `device` must be the calling thread's current device.
 * The split-KV partial results go to a 4.25 MiB scratch array built into the library (every GPU that
 * loads the library keeps it until exit).  Limit: T * n_q * ceil(max_len/chunk) * hd <= 1,048,576
 * floats, else RT_ERR_INVALID with an explanatory rt_last_error(NULL).  No cudaMalloc, cudaFree or
 * other allocating call is made.  The first call on each device resolves the scratch address
 * (cudaGetSymbolAddress) and sets kernel attributes: make that first call before starting any stream
 * capture.  All calls share the one scratch array, and it keeps the previous call's contents: nothing
 * in it needs to start at zero (each call overwrites every partial it later reads), but two calls
 * whose kernels overlap on the GPU (different streams, or different host threads) race on it -- keep
 * calls on one device ordered. */
RT_API int rt_k_attention_noalloc(const void* q_dev, const void* kpool_dev, const void* vpool_dev,
                                  const int32_t* page_table_dev, int32_t pt_stride, const int32_t* query_pos_dev,
                                  float* out_dev, int32_t T, int32_t max_len, int32_t n_q, int32_t n_kv, int32_t hd,
                                  int32_t chunk, int32_t device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RT_ENGINE_H */
