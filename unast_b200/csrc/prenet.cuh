// K7: the first layer of SpeechPrenet fused behind the masking ops (last piece): what the supervised
// step does to the padded mel batch between the loader and the encoder, in one pass over HBM --
//     specaugment (src/utils.py:51-75, src/train.py:236)  ->  noise_fn (src/utils.py:40-49, src/network.py:205,307)
//     ->  fc1 = Linear(num_mels = 80 -> s_pre_hid) + bias  ->  ReLU          (src/module.py:93-95, :109)
//   y[b][t][:] = relu(W ((t in a span ? fill_b : x[b][t][:]) * keep[b][t]) + bias)
//
// This one IS GEMM-shaped ([B T, 80] x [80, N], N = 256 or 128), so it runs on the 5th-generation tensor cores:
//   * one thread per row builds the masked row and writes it straight into TENSOR MEMORY as the A operand
//     (tcgen05.st 32x32b: thread i of warp w <-> TMEM lane 32 w + i, K along the columns) -- A never touches shared memory;
//   * W sits in shared memory for the whole kernel in the canonical K-major, no-swizzle core-matrix layout (packed once per
//     weight update by prenet_pack_kernel, brought in by one TMA bulk copy);
//   * one elected thread issues tcgen05.mma (M = 128, N, K = 8 per instruction) into a TMEM accumulator, signals an
//     mbarrier through tcgen05.commit, and the same 128 threads read the accumulator back with tcgen05.ld, add the bias,
//     apply the ReLU and store the row.
// float32 fidelity (the reference's nn.Linear is float32, TF32 off by default) comes from an operand split:
//     x = xh + xl,  w = wh + wl  (xh, wh: TF32 = top 19 bits; xl, wl: exact remainders)
//     y = xh wh + xl wh + xh wl      (three kind::tf32 products per K step; the missing xl wl term is 2^-22 |x||w|)
// -- measured against float64: 2e-6 of |x||w|, ~5x the error of cuBLAS's float32 GEMM (the tensor cores' float32
// accumulation truncates; a single TF32 pass would be ~1e-3).  Dropout (src/module.py:96) is a training-time draw and stays with torch.
#pragma once
#include "augment.cuh"

namespace unast {

constexpr int PRE_K = NMELS;                       // 80 input features
constexpr int PRE_MAXN = 256;
constexpr int PRE_TILE_M = 128;
constexpr int PRE_THREADS = 128;
// TMEM columns: accumulator [0, 256), A hi [256, 336), A lo [336, 416)
constexpr int PRE_COL_D = 0, PRE_COL_AH = 256, PRE_COL_AL = 336;

__host__ __device__ constexpr int pre_wh_bytes(int N) { return (PRE_K / 4) * (N / 8) * 128; }    // tf32: 4 per 16-byte chunk
__host__ __device__ constexpr int pre_wl_bytes(int N) { return pre_wh_bytes(N); }                 // the remainders, same layout
__host__ __device__ constexpr int pre_packed_bytes(int N) { return pre_wh_bytes(N) + pre_wl_bytes(N) + N * 4; }

// W [N][80] float32 (torch nn.Linear.weight) + bias [N] -> the shared-memory image the kernel copies in:
//   wh: core matrices of 8 rows x 16 bytes, K-chunk-major: byte offset ((k / 4) (N / 8) + n / 8) 128 + (n % 8) 16 + (k % 4) 4
//   wl: the same for the remainders w - wh;  then the bias.
__global__ void prenet_pack_kernel(const float* __restrict__ W, const float* __restrict__ bias, int N, unsigned char* __restrict__ img) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N * PRE_K) {
        const int n = i / PRE_K, k = i % PRE_K;
        const float w = W[i];
        const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
        const float lo = w - hi;                                             // exact
        *reinterpret_cast<float*>(img + ((k / 4) * (N / 8) + n / 8) * 128 + (n % 8) * 16 + (k % 4) * 4) = hi;
        *reinterpret_cast<float*>(img + pre_wh_bytes(N) + ((k / 4) * (N / 8) + n / 8) * 128 + (n % 8) * 16 + (k % 4) * 4) = lo;
    }
    if (i < N) reinterpret_cast<float*>(img + pre_wh_bytes(N) + pre_wl_bytes(N))[i] = bias ? bias[i] : 0.0f;
}

struct PrenetParams {
    const float* x;              // [B][T][80]
    float* y;                    // [B][T][N]
    int B, T, N;
    const int* spans;            // [B][4] or nullptr
    const float* keep;           // [B][T] or nullptr
    const double* parts;         // [B][AUG_PARTS] slab sums (fill = mean) or nullptr
    int replace_with_zero;
    int relu;
    const unsigned char* packed; // prenet_pack_kernel's image (16-byte aligned)
};

// shared-memory matrix descriptor, K-major, no swizzle: start address, leading (K-chunk) and stride (8-row group) byte
// offsets in 16-byte units, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t pre_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) |
           (1ull << 46);
}
// instruction descriptor: D = F32, A / B format (0 F16, 1 BF16, 2 TF32), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t pre_idesc(int fmt, int N) {
    return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(PRE_TILE_M >> 4) << 24);
}
__device__ __forceinline__ void pre_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void pre_tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void pre_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    // load and wait in ONE statement: the registers may only be read after tcgen05.wait::ld
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
                 "tcgen05.wait::ld.sync.aligned;"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr)
                 : "memory");
}

__global__ void __launch_bounds__(PRE_THREADS, 1) prenet_fc1_kernel(const PrenetParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5;
    const int N = P.N;
    const int wh_bytes = pre_wh_bytes(N), wl_bytes = pre_wl_bytes(N);
    unsigned char* s_wh = smem;
    unsigned char* s_wl = smem + wh_bytes;
    const float* s_bias = reinterpret_cast<const float*>(smem + wh_bytes + wl_bytes);
    uint64_t* bar_w = reinterpret_cast<uint64_t*>(smem + ((pre_packed_bytes(N) + 15) & ~15));     // weights landed
    uint64_t* bar_d = bar_w + 1;                                                                     // accumulator complete
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_w + 2);

    const long long M = (long long)P.B * P.T;
    const int tiles = (int)((M + PRE_TILE_M - 1) / PRE_TILE_M);
    if ((int)blockIdx.x >= tiles) return;

    if (tid == 0) {
        mbar_init(bar_w, 1);
        mbar_init(bar_d, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tma_load_1d(smem, P.packed, (uint32_t)pre_packed_bytes(N), bar_w);   // W hi, W lo, bias: one bulk copy
    }
    if (warp == 0) {                                                          // one warp allocates all 512 columns (one CTA per SM)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *s_tmem;
    const uint32_t t_lane = tbase + ((uint32_t)(32 * warp) << 16);           // this warp's 32 lanes of tensor memory
    mbar_wait(bar_w, 0);

    const uint32_t idesc_tf32 = pre_idesc(2, N);
    const uint32_t chunk_stride = (uint32_t)(N / 8) * 128u;                  // bytes between consecutive 16-byte K chunks
    uint32_t ph_d = 0;

    // A row in flight: the NEXT tile's row is fetched (20 x 128-bit loads per thread) right after this tile's MMAs have been
    // issued, so its HBM latency hides behind the tensor cores and the epilogue.
    struct RowIn { float4 q[PRE_K / 4]; float fill, kp; bool masked, valid; long long row; };
    auto fetch_row = [&](int tile) {
        RowIn r;
        r.row = (long long)tile * PRE_TILE_M + tid;                          // this thread's row == its TMEM lane
        r.valid = tile < tiles && r.row < M;
        r.fill = 0.0f; r.kp = 1.0f; r.masked = false;
        if (r.valid) {
            const int b = (int)(r.row / P.T), t = (int)(r.row % P.T);
            const float4* xr = reinterpret_cast<const float4*>(P.x + (size_t)r.row * PRE_K);
#pragma unroll
            for (int i = 0; i < PRE_K / 4; ++i) r.q[i] = __ldcs(xr + i);     // streamed once
            if (P.spans) {
                const int4 sp = reinterpret_cast<const int4*>(P.spans)[b];
                r.masked = (t >= sp.x && t < sp.x + sp.y) || (t >= sp.z && t < sp.z + sp.w);
                if (r.masked && !P.replace_with_zero) {
                    double sum = 0.0;
                    for (int q = 0; q < AUG_PARTS; ++q) sum += P.parts[b * AUG_PARTS + q];     // the order aug_apply_kernel uses
                    r.fill = (float)(sum / ((double)P.T * PRE_K));
                }
            }
            if (P.keep) r.kp = P.keep[r.row];
        } else {
#pragma unroll
            for (int i = 0; i < PRE_K / 4; ++i) r.q[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return r;
    };
    RowIn nxt = fetch_row(blockIdx.x);

    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const RowIn cur = nxt;
        const long long row = cur.row;
        const bool valid = cur.valid, masked = cur.masked;
        const float fill = cur.fill, kp = cur.kp;
        // ---- the masked input row -> A operand in tensor memory (hi, lo) ----
#pragma unroll
        for (int c = 0; c < PRE_K / 8; ++c) {                                // 8 features = one K step of the tf32 MMA
            const float4 q0 = cur.q[2 * c], q1 = cur.q[2 * c + 1];
            const float v[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float a = (masked ? fill : v[i]) * kp;                 // aug_apply_kernel's arithmetic: one float32 product
                hi[i] = __float_as_uint(a) & 0xffffe000u;
                lo[i] = __float_as_uint(a - __uint_as_float(hi[i]));
            }
            pre_tmem_st8(t_lane + PRE_COL_AH + 8 * c, hi);
            pre_tmem_st8(t_lane + PRE_COL_AL + 8 * c, lo);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        // ---- one thread issues the 30 MMAs of the tile; tcgen05.commit arrives on bar_d when they are done ----
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t wh = smem_u32(s_wh), wl = smem_u32(s_wl);
#pragma unroll 1
            for (int s = 0; s < PRE_K / 8; ++s)                              // xh wh
                pre_mma_tf32_ts(tbase + PRE_COL_D, tbase + PRE_COL_AH + 8 * s, pre_smem_desc(wh + 2 * s * chunk_stride, chunk_stride, 128),
                                idesc_tf32, s > 0);
#pragma unroll 1
            for (int s = 0; s < PRE_K / 8; ++s)                              // xl wh
                pre_mma_tf32_ts(tbase + PRE_COL_D, tbase + PRE_COL_AL + 8 * s, pre_smem_desc(wh + 2 * s * chunk_stride, chunk_stride, 128),
                                idesc_tf32, 1);
#pragma unroll 1
            for (int s = 0; s < PRE_K / 8; ++s)                              // xh wl
                pre_mma_tf32_ts(tbase + PRE_COL_D, tbase + PRE_COL_AH + 8 * s, pre_smem_desc(wl + 2 * s * chunk_stride, chunk_stride, 128),
                                idesc_tf32, 1);
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar_d)) : "memory");
        }
        nxt = fetch_row(tile + gridDim.x);                                   // in flight during the MMAs and the epilogue
        mbar_wait(bar_d, ph_d);
        ph_d ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // ---- epilogue: accumulator row -> + bias -> ReLU -> y[row][:] ----
        float* yr = P.y + (size_t)(valid ? row : 0) * N;
        for (int c0 = 0; c0 < N; c0 += 16) {
            uint32_t acc[16];
            pre_tmem_ld16(t_lane + PRE_COL_D + c0, acc);
            if (valid) {
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    float4 o;
                    o.x = __uint_as_float(acc[i]) + s_bias[c0 + i];
                    o.y = __uint_as_float(acc[i + 1]) + s_bias[c0 + i + 1];
                    o.z = __uint_as_float(acc[i + 2]) + s_bias[c0 + i + 2];
                    o.w = __uint_as_float(acc[i + 3]) + s_bias[c0 + i + 3];
                    if (P.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
                    __stcs(reinterpret_cast<float4*>(yr + c0 + i), o);      // written once, not re-read here
                }
            }
        }
        // the next tile overwrites the A columns and the accumulator: every thread must be done reading
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

}  // namespace unast
