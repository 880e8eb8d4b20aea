// Seeded, counter-based weight generator evaluated directly into the engine's fused HBM layouts.
// Bit-exact restatement of oracle/weights.py (integer hash + one fp32 multiply (+ one fp32 add) +
// RNE to bf16), so the CPU oracle and the GPU engine hold identical bytes without a 139 GB file.
// No counterpart in the reference (it holds no weights; local-llm.ts:162 only names a model).
#pragma once
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

// A fused matrix [rows][cols] is described by up to 3 row segments, each taken from a logical HF
// tensor [lrows][lcols] starting at (lrow0, lcol0); or by the gate/up 8-row interleave.
struct Seg {
    int prow0, nrows;             // physical rows [prow0, prow0+nrows)
    unsigned long long key;
    long long lrow0, lcols, lcol0;
};
struct GenDesc {
    int mode;                     // 0 segments, 1 gate/up interleave (seg[0]=gate, seg[1]=up; lrow0 = first logical row)
    int nseg;
    Seg seg[3];
    int rows, cols;
    float scale, offset;
    int as_f32;                   // write fp32 (biases) instead of bf16
};

static __global__ void gen_kernel(const GenDesc d, void* out) {
    const long long n8 = (long long)d.rows * d.cols / 8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        const long long e = i * 8;
        const int pr = (int)(e / d.cols), pc = (int)(e % d.cols);
        unsigned long long key;
        long long lidx;
        if (d.mode == 1) {
            const int tile = pr >> 4, w = pr & 15;
            const Seg& s = d.seg[w < 8 ? 0 : 1];
            key = s.key;
            lidx = (s.lrow0 + tile * 8 + (w & 7)) * s.lcols + s.lcol0 + pc;
        } else {
            int si = 0;
            for (int j = 1; j < d.nseg; ++j)
                if (pr >= d.seg[j].prow0) si = j;
            const Seg& s = d.seg[si];
            key = s.key;
            lidx = (s.lrow0 + (pr - s.prow0)) * s.lcols + s.lcol0 + pc;
        }
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = gen_value(key, (unsigned long long)(lidx + j), d.scale, d.offset);
        if (d.as_f32) {
            float* o = reinterpret_cast<float*>(out) + e;
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
        } else {
            uint4 p;
            p.x = pack_bf16(v[0], v[1]);
            p.y = pack_bf16(v[2], v[3]);
            p.z = pack_bf16(v[4], v[5]);
            p.w = pack_bf16(v[6], v[7]);
            *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(out) + e) = p;
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

}  // namespace wgen
