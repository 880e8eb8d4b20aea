// Probe: cost of a software grid barrier + one dependent L2 round trip on B200, for sizing a persistent decode kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gridbar gridbar.cu && ./gridbar
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// monotonic counter barrier: generation g completes when counter == g * gridDim.x
__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        red_release(counter, 1u);
        long long spins = 0;
        while (ld_acquire(counter) < target) {
            if (++spins > (1ll << 26)) { printf("barrier timeout\n"); __trap(); }
        }
    }
    __syncthreads();
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(unsigned* counter, float* buf, int iters, float* sink) {
    float acc = 0.f;
    const int n = gridDim.x;
    for (int it = 0; it < iters; ++it) {
        if (MODE >= 1) {   // "epilogue": every thread writes one float; then after the barrier reads another CTA's
            buf[(size_t)(it & 1) * n * 256 + blockIdx.x * 256 + threadIdx.x] = acc + it;
        }
        grid_barrier(counter, (unsigned)(it + 1) * n);
        if (MODE >= 1) {
            const int other = (blockIdx.x * 7 + 3) % n;
            acc += __ldcg(buf + (size_t)(it & 1) * n * 256 + other * 256 + threadIdx.x);
        }
        if (MODE >= 2) {   // every CTA reads 32 KB produced by all the others (an activation vector)
            for (int j = threadIdx.x; j < 8192; j += 256) acc += __ldcg(buf + (size_t)(it & 1) * n * 256 + (j % (n * 256)));
        }
    }
    if (acc == 12345.f) *sink = acc;
}

// the same dependent stage as a chain of kernels launched with programmatic dependent launch inside a CUDA graph
__global__ void __launch_bounds__(256) stage(float* buf, int it, float* sink) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int n = gridDim.x;
    const int other = (blockIdx.x * 7 + 3) % n;
    float acc = __ldcg(buf + (size_t)((it + 1) & 1) * n * 256 + other * 256 + threadIdx.x);
    buf[(size_t)(it & 1) * n * 256 + blockIdx.x * 256 + threadIdx.x] = acc + it;
    if (acc == 12345.f) *sink = acc;
}

static void chain(int grid, bool pdl, float* buf, float* sink) {
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int it = 0; it < 100; ++it) {
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1; cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, stage, buf, it, sink);
    }
    cudaStreamEndCapture(st, &g);
    cudaGraphInstantiate(&ge, g, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0, st);
        for (int k = 0; k < 20; ++k) cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st);
        cudaError_t er = cudaStreamSynchronize(st);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) printf("kernel chain grid %d pdl %d: %.3f us per stage (%s)\n", grid, (int)pdl, 1e3 * ms / 2000, cudaGetErrorString(er));
    }
}

int main() {
    unsigned* counter; float *buf, *sink;
    cudaMalloc(&counter, 4); cudaMalloc(&buf, 2 * 1024 * 256 * 4); cudaMalloc(&sink, 4);
    cudaMemset(buf, 0, 2 * 1024 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000;
    for (int mode = 0; mode < 3; ++mode)
        for (int k = 1; k <= 2; ++k) {
            const int grid = 148 * k;
            for (int rep = 0; rep < 2; ++rep) {
                cudaMemset(counter, 0, 4);
                cudaEventRecord(e0);
                if (mode == 0) probe<0><<<grid, 256>>>(counter, buf, iters, sink);
                else if (mode == 1) probe<1><<<grid, 256>>>(counter, buf, iters, sink);
                else probe<2><<<grid, 256>>>(counter, buf, iters, sink);
                cudaEventRecord(e1);
                cudaError_t er = cudaDeviceSynchronize();
                float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
                if (rep) printf("mode %d grid %d: %.3f us per barrier stage (%s)\n", mode, grid, 1e3 * ms / iters, cudaGetErrorString(er));
            }
        }
    for (int k = 1; k <= 2; ++k) { chain(148 * k, false, buf, sink); chain(148 * k, true, buf, sink); }
    return 0;
}
