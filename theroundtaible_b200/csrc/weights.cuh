// Seeded, counter-based weight generator evaluated directly into the engine's fused HBM layouts.
// Bit-exact restatement of oracle/weights.py (integer hash + one fp32 multiply (+ one fp32 add) +
// RNE to bf16), so the CPU oracle and the GPU engine hold identical bytes without a 139 GB file.
// No counterpart in the reference (it holds no weights; local-llm.ts:162 only names a model).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace wgen {

constexpr double IH4_STD = 37837.22722;
enum Kind { EMBED = 0, LM_HEAD = 1, FINAL_NORM = 2 };
enum LKind { Q = 0, K = 1, V = 2, O = 3, GATE = 4, UP = 5, DOWN = 6, IN_NORM = 7, POST_NORM = 8, Q_BIAS = 9, K_BIAS = 10, V_BIAS = 11 };

__host__ __device__ inline unsigned long long splitmix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline int tensor_id(int kind, int layer) { return layer < 0 ? kind : (layer + 1) * 16 + kind; }
inline unsigned long long stream_key(unsigned long long seed, int tid) {
    return splitmix64(seed + 0x100000001B3ull * (unsigned long long)tid);
}

__host__ __device__ inline float gen_value(unsigned long long key, unsigned long long idx, float scale, float offset) {
    const unsigned long long h = splitmix64(key + idx);
    const int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) + (int)(h >> 48);
    const float u = (float)(s - 131070);
#ifdef __CUDA_ARCH__
    float v = __fmul_rn(u, scale);
    if (offset != 0.f) v = __fadd_rn(v, offset);
#else
    volatile float vm = u * scale;
    float v = vm;
    if (offset != 0.f) { volatile float va = v + offset; v = va; }
#endif
    return v;
}

// A fused matrix [rows][cols] is described by a ROW MAP (which logical HF tensor row a physical row holds) and
// a physical LAYOUT (row-major, or the core-matrix-major "CM" layout of gemv.cuh).
//   map 0: up to 3 row segments, each taken from a logical tensor [lrows][lcols] starting at (lrow0, lcol0)
//   map 1: gate/up 8-row interleave (seg[0] = gate, seg[1] = up)
//   map 2: segments (q, k, v heads) with the RoPE pair permutation inside every head of `hd` rows:
//          physical row p of a head -> dim (p%16 < 8 ? 8*(p/16) + p%16 : hd/2 + 8*(p/16) + p%16 - 8)
struct Seg {
    int prow0, nrows;             // physical rows [prow0, prow0+nrows)
    unsigned long long key;
    long long lrow0, lcols, lcol0;
    const void* src;              // non-null: copy from this row-major [*][lcols] device tensor (a staged checkpoint
                                  // tensor, dtype GenDesc.src_dtype) instead of generating
};
struct GenDesc {
    int map;
    int nseg;
    Seg seg[3];
    int rows, cols;
    int hd;                       // map 2
    int cm;                       // 1 = CM layout, 0 = row-major
    float scale, offset;
    int as_f32;                   // write fp32 (biases; row-major only) instead of bf16
    int src_dtype;                // Seg.src element type: 0 bf16, 1 fp16, 2 fp32 (converted RNE to bf16)
};

__host__ __device__ inline void map_row(const GenDesc& d, int pr, unsigned long long& key, long long& lrow, long long& lcols,
                                        long long& lcol0, const void*& src) {
    if (d.map == 1) {
        const int tile = pr >> 4, w = pr & 15;
        const Seg& s = d.seg[w < 8 ? 0 : 1];
        key = s.key; lrow = s.lrow0 + tile * 8 + (w & 7); lcols = s.lcols; lcol0 = s.lcol0; src = s.src;
        return;
    }
    int si = 0;
    for (int j = 1; j < d.nseg; ++j)
        if (pr >= d.seg[j].prow0) si = j;
    const Seg& s = d.seg[si];
    int local = pr - s.prow0;
    if (d.map == 2) {
        const int head = local / d.hd, p = local % d.hd, j = p >> 4, w = p & 15;
        local = head * d.hd + (w < 8 ? 8 * j + w : d.hd / 2 + 8 * j + (w - 8));
    }
    key = s.key; lrow = s.lrow0 + local; lcols = s.lcols; lcol0 = s.lcol0; src = s.src;
}

// physical 16-byte piece index -> (row, first column) for the CM layout
__host__ __device__ inline void cm_piece(long long piece, int cols, int& row, int& col) {
    const int nkb = cols >> 5;
    const int r = (int)(piece & 7);
    const int q = (int)((piece >> 3) & 3);
    const int h = (int)((piece >> 5) & 1);
    const long long tk = piece >> 6;          // tile * nkb + kb
    const int kb = (int)(tk % nkb);
    const int tile = (int)(tk / nkb);
    row = tile * 16 + h * 8 + r;
    col = kb * 32 + q * 8;
}

static __global__ void gen_kernel(const GenDesc d, void* out) {
    const long long n8 = (long long)d.rows * d.cols / 8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        int pr, pc;
        if (d.cm) {
            cm_piece(i, d.cols, pr, pc);
        } else {
            pr = (int)((i * 8) / d.cols);
            pc = (int)((i * 8) % d.cols);
        }
        unsigned long long key;
        long long lrow, lcols, lcol0;
        const void* src;
        map_row(d, pr, key, lrow, lcols, lcol0, src);
        const long long lidx = lrow * lcols + lcol0 + pc;
        float v[8];
        if (src) {
            // checkpoint tensor: the same row map and layout, values copied (bf16) or rounded RNE (fp16 / fp32)
            if (d.src_dtype == 0) {
                const bf16* sp = reinterpret_cast<const bf16*>(src) + lidx;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = __bfloat162float(sp[j]);
            } else if (d.src_dtype == 1) {
                const __half* sp = reinterpret_cast<const __half*>(src) + lidx;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = __half2float(sp[j]);
            } else {
                const float* sp = reinterpret_cast<const float*>(src) + lidx;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sp[j];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = gen_value(key, (unsigned long long)(lidx + j), d.scale, d.offset);
        }
        if (d.as_f32) {
            float* o = reinterpret_cast<float*>(out) + i * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
        } else {
            uint4 p;
            p.x = pack_bf16(v[0], v[1]);
            p.y = pack_bf16(v[2], v[3]);
            p.z = pack_bf16(v[4], v[5]);
            p.w = pack_bf16(v[6], v[7]);
            *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(out) + i * 8) = p;
        }
    }
}

inline void launch_gen(const GenDesc& d, void* out, cudaStream_t st) {
    const long long n8 = (long long)d.rows * d.cols / 8;
    int blocks = (int)((n8 + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    gen_kernel<<<blocks, 256, 0, st>>>(d, out);
}

// host: CM buffer [rows][cols] (bf16 bits) -> row-major
inline void cm_to_rowmajor(const uint16_t* cm, uint16_t* rm, int rows, int cols) {
    const long long n8 = (long long)rows * cols / 8;
    for (long long p = 0; p < n8; ++p) {
        int r, c;
        cm_piece(p, cols, r, c);
        for (int e = 0; e < 8; ++e) rm[(size_t)r * cols + c + e] = cm[p * 8 + e];
    }
}
inline void rowmajor_to_cm(const uint16_t* rm, uint16_t* cm, int rows, int cols) {
    const long long n8 = (long long)rows * cols / 8;
    for (long long p = 0; p < n8; ++p) {
        int r, c;
        cm_piece(p, cols, r, c);
        for (int e = 0; e < 8; ++e) cm[p * 8 + e] = rm[(size_t)r * cols + c + e];
    }
}

}  // namespace wgen
