// rt_k_attention_noalloc: the paged decode-attention kernel for callers that may not allocate.
//
// Who cannot allocate: a host process that has handed (nearly) all of HBM to its own pool allocator --
// a serving framework embedding this library -- so that a cudaMalloc from inside the library would fail
// or fragment the pool; and a latency-critical token loop, where cudaMalloc / cudaFree are off limits
// because they synchronise the whole device.  Such callers pass every input and output as DEVICE
// pointers they already own; the only thing the split-KV attention needs beyond them is scratch for
// the per-chunk partial results, and that scratch lives here in a file-scope device array.
//
// Cost: every GPU that loads this library holds RT_NOALLOC_SCRATCH_BYTES (4.25 MiB) until the process
// exits, used or not.
#include <cmath>
#include <mutex>
#include <string>

#include "engine.cuh"
#include "kernels.cuh"

void rt_internal_set_error(const std::string& msg);   // capi.cu: feeds rt_last_error(NULL)

namespace {

constexpr size_t NOALLOC_O_FLOATS = 1u << 20;                  // partial outputs: T*n_q*n_chunks*hd <= 1 Mi floats
constexpr size_t NOALLOC_ML_FLOATS = 1u << 16;                 // partial (max, sum) pairs: T*n_q*n_chunks*2 <= 64 Ki floats
constexpr int NOALLOC_MAX_DEVICES = 64;

// The scratch itself.  In HOST code the name `g_noalloc_scratch` is only a handle for the runtime -- it
// is not a device address and must never be passed to a kernel; the address comes from
// cudaGetSymbolAddress (below).  Device code could name the array directly.
static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_O_FLOATS + NOALLOC_ML_FLOATS];
// slot -> sequence map of the kernel: query t belongs to sequence t
static __device__ __align__(256) int g_noalloc_identity[MAX_SLOTS] = {0,  1,  2,  3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15,
                                                                      16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31};

struct DevSetup {
    std::once_flag once;
    float* scratch = nullptr;
    int* identity = nullptr;
};
DevSetup g_setup[NOALLOC_MAX_DEVICES];

// once per device, with that device current: resolve the device addresses of the two arrays and opt the
// kernels of THIS translation unit in to their dynamic shared memory size.  Nothing here allocates.
const DevSetup& setup_device(int device) {
    DevSetup& s = g_setup[device];
    std::call_once(s.once, [&] {       // two host threads may make their first call on a device at the same time
    RT_CUDA(cudaGetSymbolAddress((void**)&s.scratch, g_noalloc_scratch));
    RT_CUDA(cudaGetSymbolAddress((void**)&s.identity, g_noalloc_identity));
    const size_t per_hd = (size_t)ATTN_WARPS * ATTN_STAGES * 2 * KV_PAGE * 2;
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(per_hd * 128)));
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(per_hd * 64)));
    RT_CUDA(cudaFuncSetAttribute(attn_decode_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(per_hd * 32)));
    });
    return s;
}

}  // namespace

extern "C" {

RT_API int rt_k_attention_noalloc(const void* q_dev, const void* kpool_dev, const void* vpool_dev, const int32_t* page_table_dev,
                                  int32_t pt_stride, const int32_t* query_pos_dev, float* out_dev, int32_t T, int32_t max_len,
                                  int32_t n_q, int32_t n_kv, int32_t hd, int32_t chunk, int32_t device, void* stream) {
    try {
        RT_REQUIRE(q_dev && kpool_dev && vpool_dev && page_table_dev && query_pos_dev && out_dev, "null device pointer");
        RT_REQUIRE(device >= 0 && device < NOALLOC_MAX_DEVICES, "device ordinal out of range");
        RT_REQUIRE(T >= 1 && T <= MAX_SLOTS && max_len >= 1 && n_kv >= 1 && n_q % n_kv == 0 && n_q / n_kv <= 16, "bad shape");
        RT_REQUIRE(hd == 32 || hd == 64 || hd == 128, "head_dim must be 32, 64 or 128");
        RT_REQUIRE(chunk >= KV_PAGE && chunk % KV_PAGE == 0, "chunk must be a multiple of 16");
        RT_REQUIRE(pt_stride >= cdiv(max_len, KV_PAGE), "pt_stride smaller than the pages max_len needs");
        const int n_chunks = cdiv(max_len, chunk);
        const size_t need_o = (size_t)T * n_q * n_chunks * hd, need_ml = (size_t)T * n_q * n_chunks * 2;
        if (need_o > NOALLOC_O_FLOATS || need_ml > NOALLOC_ML_FLOATS)
            throw RtError(RT_ERR_INVALID, "rt_k_attention_noalloc: T*n_q*ceil(max_len/chunk)*head_dim = " + std::to_string(need_o) +
                                              " floats of split-KV scratch exceed the built-in maximum of " +
                                              std::to_string(NOALLOC_O_FLOATS) + " (4 MiB); use a larger chunk or rt_k_attention");
        int cur = -1;
        RT_CUDA(cudaGetDevice(&cur));
        RT_REQUIRE(cur == device, "the given device must be the current CUDA device of the calling thread");
        const DevSetup& s = setup_device(device);
        cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

        AttnArgs a{};
        a.Q = reinterpret_cast<const bf16*>(q_dev);
        a.ldq = n_q * hd;
        a.Kp = reinterpret_cast<const bf16*>(kpool_dev);
        a.Vp = reinterpret_cast<const bf16*>(vpool_dev);
        a.page_table = page_table_dev;
        a.pt_stride = pt_stride;
        a.slot_seq = s.identity;
        a.slot_pos = query_pos_dev;
        a.shared_len_ptr = nullptr;
        a.T = T; a.n_q = n_q; a.n_kv = n_kv;
        a.chunk_pages = chunk / KV_PAGE;
        a.n_shared_tiles = 0;
        a.part_o = s.scratch;
        a.part_ml = s.scratch + NOALLOC_O_FLOATS;
        a.max_chunks = n_chunks;
        a.scale_log2 = (float)(1.4426950408889634 / std::sqrt((double)hd));
        // the positions live on the device: the kernels clamp them to max_len, so that a position beyond what the caller
        // declared truncates that query's context instead of writing partials past the shared scratch array
        a.max_len = max_len;
        const dim3 grid(cdiv(n_chunks, ATTN_WARPS), n_kv, T);
        const size_t smem = (size_t)ATTN_WARPS * ATTN_STAGES * 2 * KV_PAGE * hd * 2;
        if (hd == 128) attn_decode_kernel<128><<<grid, ATTN_WARPS * 32, smem, st>>>(a);
        else if (hd == 64) attn_decode_kernel<64><<<grid, ATTN_WARPS * 32, smem, st>>>(a);
        else attn_decode_kernel<32><<<grid, ATTN_WARPS * 32, smem, st>>>(a);
        MergeArgs m{};
        m.part_o = a.part_o; m.part_ml = a.part_ml; m.max_chunks = n_chunks; m.slot_pos = query_pos_dev;
        m.shared_len_ptr = nullptr; m.cascade = 0; m.max_len = max_len; m.chunk_keys = chunk; m.n_q = n_q; m.hd = hd;
        m.out_bf = nullptr; m.out_f32 = out_dev; m.ldo = n_q * hd;
        attn_merge_kernel<<<dim3(n_q, T), hd, 0, st>>>(m);
        RT_CUDA(cudaGetLastError());
        return RT_OK;
    } catch (const RtError& x) {
        rt_internal_set_error(x.what());
        return x.code;
    } catch (const std::exception& x) {
        rt_internal_set_error(std::string("internal error: ") + x.what());
        return RT_ERR_INVALID;
    }
}

}  // extern "C"
