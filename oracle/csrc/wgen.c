/* Fast C restatement of oracle/weights.py::gen_bits (TEST INFRASTRUCTURE; see oracle/__init__.py).
 * Same integer hash, one fp32 multiply (+ one fp32 add), RNE to bf16.  Built by oracle/build.py with
 * -ffp-contract=off so that the multiply and the add are never fused.  No reference counterpart
 * (the reference holds no weights: /root/reference/src/adapters/local-llm.ts:162). */
#include <stdint.h>
#include <string.h>

static inline uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

/* out_f32 (nullable) receives the bf16 values widened to fp32; out_bits (nullable) the bit patterns */
void oracle_gen(uint64_t key, uint64_t start, uint64_t count, float scale, float offset, float* out_f32,
                uint16_t* out_bits) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)count; ++i) {
        const uint64_t h = splitmix64(key + start + (uint64_t)i);
        const int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) + (int)(h >> 48);
        volatile float v = (float)(s - 131070) * scale;
        float r = v;
        if (offset != 0.0f) {
            volatile float w = r + offset;
            r = w;
        }
        const uint16_t b = f32_to_bf16(r);
        if (out_bits) out_bits[i] = b;
        if (out_f32) {
            const uint32_t u = (uint32_t)b << 16;
            memcpy(&out_f32[i], &u, 4);
        }
    }
}
