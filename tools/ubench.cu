// Pipe-rate microbenchmark for B200 (sm_100a): sustained FFMA / FFMA2 / FADD2 / DFMA / MUFU /
// SHFL / LDS rates.  SURVEY.md 8(d) asks for a measured non-tensor FP32 peak (the roofline
// denominator of the Griffin-Lim kernels) instead of the nominal 148*128*2*f figure.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu ; run: ./ubench
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int ILP = 8;

__global__ void k_ffma(float* out, float a, float b) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = fmaf(v[i], a, b);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fadd(float* out, float a) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = v[i] + a;
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b) {
    float2 v[ILP]; float2 aa = make_float2(a, a * 1.0001f), bb = make_float2(b, b * 0.999f);
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = make_float2(threadIdx.x + i, i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = __ffma2_rn(v[i], aa, bb);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fadd2(float* out, float a) {
    float2 v[ILP]; float2 aa = make_float2(a, a * 1.0001f);
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = make_float2(threadIdx.x + i, i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = __fadd2_rn(v[i], aa);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dfma(float* out, double a, double b) {
    double v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = fma(v[i], a, b);
    }
    double s = 0; for (int i = 0; i < ILP; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)s;
}
__global__ void k_mufu(float* out, float a) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i + a;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = __log2f(v[i]) + 3.0f;
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_shfl(float* out, int d) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) v[i] = __shfl_xor_sync(0xffffffffu, v[i], d);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_lds64(float* out, int stride) {
    __shared__ float2 sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = make_float2(i, i);
    __syncthreads();
    float2 acc = make_float2(0, 0);
    int idx = (threadIdx.x * stride) & 4095;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) { float2 t = sm[(idx + i * 32 * stride) & 4095]; acc.x += t.x; acc.y += t.y; }
        idx = (idx + (int)acc.x) & 4095;   // dependent index: defeats hoisting, adds ~0 (acc.x stays integral, masked)
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y;
}
// mixed: FADD2 + scalar FFMA interleaved (do they share the pipe?)
__global__ void k_mix(float* out, float a, float b) {
    float2 v[ILP]; float w[ILP]; float2 aa = make_float2(a, a * 1.0001f);
#pragma unroll
    for (int i = 0; i < ILP; ++i) { v[i] = make_float2(threadIdx.x + i, i); w[i] = i; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) { v[i] = __fadd2_rn(v[i], aa); w[i] = fmaf(w[i], a, b); }
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += v[i].x + v[i].y + w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F> static float time_ms(F launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount, clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d,\n", p.name, sms, clk);
    const int threads = 512, blocks = sms * 4;
    float* out; CK(cudaMalloc(&out, sizeof(float) * threads * blocks));
    double lanes = (double)threads * blocks, n = (double)ITERS * ILP;
    float ms;
    ms = time_ms([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    printf(" \"ffma_tflops\": %.2f,\n", 2 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_fadd<<<blocks, threads>>>(out, 0.5f); });
    printf(" \"fadd_tops\": %.2f,\n", lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    printf(" \"ffma2_tflops\": %.2f,\n", 4 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_fadd2<<<blocks, threads>>>(out, 0.5f); });
    printf(" \"fadd2_tops\": %.2f,\n", 2 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_mix<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    printf(" \"mix_fadd2_ffma_tops\": %.2f,\n", 4 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_dfma<<<blocks, threads>>>(out, 1.0001, 0.5); });
    printf(" \"dfma_tflops\": %.2f,\n", 2 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_mufu<<<blocks, threads>>>(out, 0.5f); });
    printf(" \"mufu_lg2_tops\": %.2f,\n", lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_shfl<<<blocks, threads>>>(out, 5); });
    printf(" \"shfl_tops\": %.2f,\n", lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_lds64<<<blocks, threads>>>(out, 1); });
    printf(" \"lds64_TBps\": %.2f,\n", 8 * lanes * n / ms * 1e-9);
    ms = time_ms([&] { k_lds64<<<blocks, threads>>>(out, 33); });
    printf(" \"lds64_stride33_TBps\": %.2f\n}\n", 8 * lanes * n / ms * 1e-9);
    CK(cudaGetLastError());
    return 0;
}
