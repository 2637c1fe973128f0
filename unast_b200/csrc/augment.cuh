// K6: the two masking ops the training step applies to the padded mel batch right after the loader (SURVEY.md 8(f)
// rank 2): `specaugment` (reference src/utils.py:51-75, called at src/train.py:236) and `noise_fn`
// (src/utils.py:40-49, called at src/network.py:205,307).
//
// What the reference computes, per utterance i of a [B, T, D] batch:
//   specaugment: res = mel.clone(); mean_ = res[i].mean() over the whole padded [T, D] slab; then
//       res[i][:][f_zero:f_zero+f] = mean_   -- `[:]` is a full view, so this slices the TIME axis, like
//       res[i][t_zero:t_zero+t][:] = mean_   -- the second one: two row spans, every channel (0 if replace_with_zero)
//   noise_fn:    to_noise * bernoulli(1 - mask_p)[B, T, 1]   -- whole rows dropped
// The random draws (torch.randint, random.randrange, torch.bernoulli) stay on the host side of the boundary, in the
// reference's order (unast_b200/augment.py), so a seeded run selects the same rows; the kernels take the spans and
// the keep mask as device arrays.
//
// HBM-bound element work: one float4 per thread and step, rows resolved per vector.  The mean is a fixed-order
// two-stage float64 sum (deterministic; the reference's float32 torch.mean differs from it by rounding only).
#pragma once
#include "common.cuh"

namespace unast {

constexpr int AUG_PARTS = 32;       // partial sums per utterance
constexpr int AUG_THREADS = 256;

// stage 1: partial sums of utterance blockIdx.y over a strided share of its T*D floats (n4 float4 + tail)
__global__ void __launch_bounds__(AUG_THREADS) aug_sum_kernel(const float* __restrict__ x, long long slab, double* __restrict__ parts) {
    const int b = blockIdx.y, p = blockIdx.x;
    const float* xb = x + (size_t)b * slab;
    const long long n4 = slab >> 2;
    const bool vec = (reinterpret_cast<uintptr_t>(xb) & 15) == 0;
    double acc = 0.0;
    if (vec) {
        const float4* x4 = reinterpret_cast<const float4*>(xb);
        for (long long i = (long long)p * AUG_THREADS + threadIdx.x; i < n4; i += (long long)AUG_PARTS * AUG_THREADS) {
            const float4 q = x4[i];
            acc += ((double)q.x + (double)q.y) + ((double)q.z + (double)q.w);
        }
        if (p == 0) for (long long i = 4 * n4 + threadIdx.x; i < slab; i += AUG_THREADS) acc += (double)xb[i];
    } else {
        for (long long i = (long long)p * AUG_THREADS + threadIdx.x; i < slab; i += (long long)AUG_PARTS * AUG_THREADS) acc += (double)xb[i];
    }
    __shared__ double red[AUG_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < AUG_THREADS / 32; ++w) s += red[w];
        parts[b * AUG_PARTS + p] = s;
    }
}

// stage 2: out = (row in a span ? fill : x) * keep[row]
//   spans: int32 [B][4] = {f_zero, f, t_zero, t} or nullptr; keep: float [B][T] or nullptr
__global__ void __launch_bounds__(AUG_THREADS) aug_apply_kernel(const float* __restrict__ x, float* __restrict__ out, int T, int D,
                                                                 const int* __restrict__ spans, const float* __restrict__ keep,
                                                                 const double* __restrict__ parts, int replace_with_zero) {
    const int b = blockIdx.y;
    const long long slab = (long long)T * D;
    const float* xb = x + (size_t)b * slab;
    float* ob = out + (size_t)b * slab;
    int a0 = 0, a1 = 0, c0 = 0, c1 = 0;
    float fill = 0.0f;
    if (spans) {
        const int4 s = reinterpret_cast<const int4*>(spans)[b];
        a0 = s.x; a1 = s.x + s.y; c0 = s.z; c1 = s.z + s.w;
        if (!replace_with_zero) {
            double sum = 0.0;
            for (int p = 0; p < AUG_PARTS; ++p) sum += parts[b * AUG_PARTS + p];     // same order in every thread
            fill = (float)(sum / (double)slab);
        }
    }
    const float* kb = keep ? keep + (size_t)b * T : nullptr;
    auto value = [&](float v, int t) {
        if ((t >= a0 && t < a1) || (t >= c0 && t < c1)) v = fill;
        return kb ? v * kb[t] : v;
    };
    const bool vec = (D & 3) == 0 && ((reinterpret_cast<uintptr_t>(xb) | reinterpret_cast<uintptr_t>(ob)) & 15) == 0;
    const long long stride = (long long)gridDim.x * AUG_THREADS;
    if (vec) {
        const int d4 = D >> 2;
        const long long n4 = slab >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(xb);
        float4* o4 = reinterpret_cast<float4*>(ob);
        for (long long i = (long long)blockIdx.x * AUG_THREADS + threadIdx.x; i < n4; i += stride) {
            const int t = (int)(i / d4);                                             // a float4 never straddles rows
            float4 q = x4[i];
            q.x = value(q.x, t); q.y = value(q.y, t); q.z = value(q.z, t); q.w = value(q.w, t);
            o4[i] = q;
        }
    } else {
        for (long long i = (long long)blockIdx.x * AUG_THREADS + threadIdx.x; i < slab; i += stride)
            ob[i] = value(xb[i], (int)(i / D));
    }
}

}  // namespace unast
