// libunast_spectral.so -- host side of the C ABI declared in include/unast_spectral.h:
// constant tables (built in float64 on the host), launch logic, diagnostics.  No torch types here.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/unast_spectral.h"
#include "frontend_au.cuh"
#include "griffinlim_pk.cuh"
#include "trim.cuh"
#include "augment.cuh"
#include "prenet.cuh"

using namespace unast;

// warps per CTA of the fused front end (frontend_au.cuh): 12 x 168 registers in float64, 16 x 128 in float32
constexpr int FE64_WARPS = 12;
constexpr int FE32_WARPS = 16;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

#define CU(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(UNAST_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));         \
    } while (0)

constexpr double kPi = 3.14159265358979323846264338327950288;

size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// Every launching entry point runs on the context's device whatever the caller's current device is, and leaves the
// caller's current device as it found it.
struct DeviceGuard {
    int prev = -1;
    bool good = true;
    explicit DeviceGuard(int dev) {
        int cur = -1;
        good = cudaGetDevice(&cur) == cudaSuccess;
        if (good && cur != dev) { good = cudaSetDevice(dev) == cudaSuccess; if (good) prev = cur; }
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    bool ok() const { return good; }
};

// ---- librosa.filters.mel(22050, 2048, 80): Slaney scale, area-normalised, float32 ----
double hz_to_mel(double f) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return f >= min_log_hz ? min_log_mel + std::log(f / min_log_hz) / logstep : f / f_sp;
}
double mel_to_hz(double m) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return m >= min_log_mel ? min_log_hz * std::exp(logstep * (m - min_log_mel)) : f_sp * m;
}
std::vector<float> build_mel_basis() {
    const double fmax = SR / 2.0;
    std::vector<double> fftfreqs(NBINS), mel_f(NMELS + 2);
    for (int i = 0; i < NBINS; ++i) fftfreqs[i] = i * (fmax / (NBINS - 1));
    fftfreqs[NBINS - 1] = fmax;
    const double m0 = hz_to_mel(0.0), m1 = hz_to_mel(fmax), step = (m1 - m0) / (NMELS + 1);
    for (int i = 0; i < NMELS + 2; ++i) mel_f[i] = mel_to_hz(i == NMELS + 1 ? m1 : m0 + step * i);
    std::vector<float> w((size_t)NMELS * NBINS);
    for (int i = 0; i < NMELS; ++i) {
        const double fd0 = mel_f[i + 1] - mel_f[i], fd1 = mel_f[i + 2] - mel_f[i + 1];
        const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
        for (int k = 0; k < NBINS; ++k) {
            const double lower = -(mel_f[i] - fftfreqs[k]) / fd0, upper = (mel_f[i + 2] - fftfreqs[k]) / fd1;
            const float tri = (float)std::fmax(0.0, std::fmin(lower, upper));
            w[(size_t)i * NBINS + k] = (float)((double)tri * enorm);
        }
    }
    return w;
}

template <class T>
cudaError_t upload(T** dptr, const std::vector<T>& h) {
    cudaError_t e = cudaMalloc((void**)dptr, h.size() * sizeof(T));
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

}  // namespace

struct unast_ctx {
    int device = 0;
    int num_sms = 0;
    // analysis windows carry the 1/2 of the real-FFT post-processing (and 2^-15 for PCM16 input)
    float2 *win2_f = nullptr, *win2i_f = nullptr, *wini2_f = nullptr, *tw_f = nullptr, *wp_f = nullptr, *phi_f = nullptr;
    float4 *tw4a_f = nullptr, *wp4_f = nullptr;
    float4* tw4_f = nullptr;                                    // tw_f in the packed layout of fft_pk.cuh
    double2 *win2_d = nullptr, *win2i_d = nullptr, *tw_d = nullptr, *wp_d = nullptr;
    double* hann_d = nullptr;               // float64 periodic Hann (gl_init_f64_kernel)
    float *wss_per = nullptr, *rinv_l = nullptr, *rinv_r = nullptr;
    float2* wa2_f = nullptr;
    double* w2 = nullptr;
    MelItems items{};
    float4* it_w = nullptr; int* it_k0 = nullptr; int2* it_mel = nullptr;
    int* d_err = nullptr;
    int* d_abort = nullptr;
    int gl_block = 0;                       // canonical Griffin-Lim summation block in frames (0 = off), unast_griffinlim_set_block
    double* aug_static = nullptr;           // device address of g_aug_static_parts on this context's device (unast_create)
    cudaError_t aug_static_err = cudaSuccess;
    std::vector<float> mel_basis;
};

// ---- scratch of unast_mel_augment_static: file scope, UNAST_AUG_STATIC_MAX_B x AUG_PARTS partial sums = 256 KB, held by every GPU
// that loads the library until the process exits.  In HOST code this name is a stand-in, not a device address: the address
// each device gives it is taken once per context with cudaGetSymbolAddress (unast_create, that device current).
static __device__ __align__(256) double g_aug_static_parts[UNAST_AUG_STATIC_MAX_B * unast::AUG_PARTS];

// ---- diagnostics kernels ----------------------------------------------------------------------
namespace unast {

// The transform cores exactly as the product kernels run them, one frame per warp:
//   float64: warp_fft1024_split + rfft_post_pairs (frontend_au_kernel<double>)
//   float32: warp_fft1024_pk + pk_post_pairs / pk_pre_pairs + warp_fft1024_pk_b (gl_persist_kernel)
__global__ void __launch_bounds__(128) debug_rfft_f64_kernel(const float* __restrict__ frames, int F, const double2* tw,
                                                             const double2* wp, float2* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* xs = reinterpret_cast<double*>(smem + warp * xs_bytes<double>());
    const int f = blockIdx.x * 4 + warp;
    if (f >= F) return;
    const float2* src = reinterpret_cast<const float2*>(frames + (size_t)f * NFFT);
    double2 v[32];
    static_for<0, 18>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const int n = lane + 32 * j;
        if (j < 17 || n < NLIVE_C) { const float2 q = src[n]; v[j] = mk(0.5 * (double)q.x, 0.5 * (double)q.y); } else v[j] = mk(0.0, 0.0);
    });
    static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; v[j] = mk(0.0, 0.0); });
    warp_fft1024_split<double, 18>(v, xs, TwMem<double>{tw, lane}, lane);
    float2* dst = out + (size_t)f * NBINS;
    rfft_post_pairs<double>(v, reinterpret_cast<double2*>(xs), WpLane<double>{wp[lane]}, lane, [&](auto jj, double2 xa, double2 xb) {
        constexpr int j = decltype(jj)::value;
        dst[lane + 32 * j] = make_float2((float)xa.x, (float)xa.y);
        dst[1024 - lane - 32 * j] = make_float2((float)xb.x, (float)xb.y);
    }, [&](double2 x) { dst[512] = make_float2((float)x.x, (float)x.y); });
}

__global__ void __launch_bounds__(128) debug_rfft_pk_kernel(const float* __restrict__ frames, int F, const float4* tw4,
                                                            const float4* wp4, float2* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* xs = reinterpret_cast<float*>(smem + warp * GLP_XS_BYTES);
    const int f = blockIdx.x * 4 + warp;
    if (f >= F) return;
    const float2* src = reinterpret_cast<const float2*>(frames + (size_t)f * NFFT);
    Pk v;
    static_for<0, 18>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const int n = lane + 32 * j;
        if (j < 17 || n < NLIVE_C) { const float2 q = src[n]; reA<j>(v) = 0.5f * q.x; imA<j>(v) = 0.5f * q.y; } else { reA<j>(v) = 0.f; imA<j>(v) = 0.f; }
    });
    static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; reA<j>(v) = 0.f; imA<j>(v) = 0.f; });
    warp_fft1024_pk<18>(v, xs, TabMem{tw4, lane}, lane);
    pk_post_pairs(v, xs, TabMem{wp4, lane}, lane);
    float2* dst = out + (size_t)f * NBINS;
    static_for<0, 8>([&](auto mm) {
        constexpr int m = decltype(mm)::value;
        const int ka = lane + 64 * m;
        dst[ka] = make_float2(v.re[m].x, v.im[m].x);
        dst[ka + 32] = make_float2(v.re[m].y, v.im[m].y);
        dst[1024 - ka] = make_float2(v.re[m + 8].x, v.im[m + 8].x);
        dst[992 - ka] = make_float2(v.re[m + 8].y, v.im[m + 8].y);
    });
    if (lane == 0) dst[512] = make_float2(2.0f * xs[32 * PK_PARK_S + 16], -2.0f * xs[PK_PARK_PLANE + 32 * PK_PARK_S + 16]);
}

__global__ void __launch_bounds__(128) debug_irfft_kernel(const float2* __restrict__ spec, int F, const float4* tw4a,
                                                          const float4* wp4, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* xs = reinterpret_cast<float*>(smem + warp * GLP_XS_BYTES);
    const int f = blockIdx.x * 4 + warp;
    if (f >= F) return;
    const float2* src = spec + (size_t)f * NBINS;
    Pk v;
    static_for<0, 8>([&](auto mm) {
        constexpr int m = decltype(mm)::value;
        const int ka = lane + 64 * m;
        const float2 a = src[ka], b = src[ka + 32], c = src[1024 - ka], d = src[992 - ka];
        v.re[m] = make_float2(a.x, b.x); v.im[m] = make_float2(a.y, b.y);
        v.re[m + 8] = make_float2(c.x, d.x); v.im[m + 8] = make_float2(c.y, d.y);
    });
    float2 x512 = make_float2(0.f, 0.f);
    if (lane == 0) { v.im[0].x = 0.f; v.im[8].x = 0.f; x512 = src[512]; }      // irfft ignores Im of DC and Nyquist
    pk_pre_pairs(v, xs, TabMem{wp4, lane}, lane);
    if (lane == 0) { xs[0] = 2.0f * x512.x; xs[PK_PARK_PLANE] = 2.0f * x512.y; }
    __syncwarp();
    pk_pre_collect(v, xs, lane);
    __syncwarp();
    warp_fft1024_pk_b(v, xs, TabMem{tw4a, lane}, lane, []() {});
    float2* dst = reinterpret_cast<float2*>(out + (size_t)f * WIN);
    static_for<0, 18>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const int n = lane + 32 * j;
        if (j < 17 || n < NLIVE_C) dst[n] = make_float2(reB<j>(v) * (1.0f / 2048.0f), -imB<j>(v) * (1.0f / 2048.0f));
    });
}

// Sustained issue rate of one arithmetic pipe: every thread runs ITERS x 8 independent FMAs.
// kind 0: FFMA (float), 1: DFMA (double), 2: FFMA2 (packed float2, Blackwell fma.rn.f32x2).
constexpr int PEAK_ILP = 8;
template <int KIND>
__global__ void __launch_bounds__(256) pipe_peak_kernel(int iters, float a, float b, float* __restrict__ out) {
    float acc = 0.f;
    if constexpr (KIND == 0) {
        float v[PEAK_ILP];
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) v[i] = threadIdx.x + i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < PEAK_ILP; ++i) v[i] = fmaf(v[i], a, b);
        }
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) acc += v[i];
    } else if constexpr (KIND == 1) {
        double v[PEAK_ILP];
        const double da = a, db = b;
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) v[i] = threadIdx.x + i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < PEAK_ILP; ++i) v[i] = fma(v[i], da, db);
        }
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) acc += (float)v[i];
    } else {
        float2 v[PEAK_ILP];
        const float2 aa = make_float2(a, a * 1.0001f), bb = make_float2(b, b * 0.999f);
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) v[i] = make_float2(threadIdx.x + i, i);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < PEAK_ILP; ++i) v[i] = __ffma2_rn(v[i], aa, bb);
        }
#pragma unroll
        for (int i = 0; i < PEAK_ILP; ++i) acc += v[i].x + v[i].y;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

}  // namespace unast

// ---- C ABI -------------------------------------------------------------------------------------
extern "C" {

int unast_abi_version(void) { return UNAST_ABI_VERSION; }
const char* unast_last_error(void) { return g_err.c_str(); }

int unast_create(int device, unast_ctx** out) {
    if (!out) return fail(UNAST_E_ARG, "unast_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(UNAST_E_ARG, "unast_create: no such CUDA device (there is no CPU fallback)");
    DeviceGuard guard(device);                                      // the caller's current device is restored on return
    if (!guard.ok()) return fail(UNAST_E_CUDA, "unast_create: cannot select the device");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(UNAST_E_ARG, "unast_create: built for sm_100a (Blackwell B200) only");
    unast_ctx* c = new unast_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;

    std::vector<double> hann(WIN);
    for (int k = 0; k < WIN; ++k) hann[k] = 0.5 - 0.5 * std::cos(2.0 * kPi * k / WIN);
    std::vector<float2> win2_f(NLIVE_C), win2i_f(NLIVE_C), wini2_f(NLIVE_C);
    std::vector<double2> win2_d(NLIVE_C), win2i_d(NLIVE_C);
    for (int n = 0; n < NLIVE_C; ++n) {
        const double a = hann[2 * n], b = hann[2 * n + 1];
        win2_d[n] = make_double2(0.5 * a, 0.5 * b);                               // power-of-two scalings are exact
        win2i_d[n] = make_double2(0.5 * a / 32768.0, 0.5 * b / 32768.0);
        win2_f[n] = make_float2((float)(0.5 * a), (float)(0.5 * b));
        win2i_f[n] = make_float2((float)(0.5 * a / 32768.0), (float)(0.5 * b / 32768.0));
        wini2_f[n] = make_float2((float)(a / 2048.0), (float)(b / 2048.0));
    }
    std::vector<float2> tw_f(1024), wp_f(1024), phi_f(NBINS);
    std::vector<double2> tw_d(1024), wp_d(1024);
    for (int k1 = 0; k1 < 32; ++k1)
        for (int n1 = 0; n1 < 32; ++n1) {
            const double th = 2.0 * kPi * ((k1 * n1) % 1024) / 1024.0;
            tw_d[k1 * 32 + n1] = make_double2(std::cos(th), -std::sin(th));
            tw_f[k1 * 32 + n1] = make_float2((float)std::cos(th), (float)-std::sin(th));
        }
    std::vector<float4> tw4_f(512);
    for (int m = 0; m < 16; ++m)
        for (int n1 = 0; n1 < 32; ++n1) {
            const float2 a = tw_f[(2 * m) * 32 + n1], b = tw_f[(2 * m + 1) * 32 + n1];
            tw4_f[m * 32 + n1] = make_float4(a.x, b.x, a.y, b.y);
        }
    for (int k = 0; k < 1024; ++k) {
        const double th = 2.0 * kPi * k / 2048.0;
        wp_d[k] = make_double2(-std::sin(th), -std::cos(th));
        wp_f[k] = make_float2((float)(-std::sin(th)), (float)(-std::cos(th)));
    }
    std::vector<float4> tw4a_f(512), wp4_f(256);
    for (int j = 0; j < 16; ++j)
        for (int n1 = 0; n1 < 32; ++n1) {
            const float2 a = tw_f[j * 32 + n1], b = tw_f[(j + 16) * 32 + n1];
            tw4a_f[j * 32 + n1] = make_float4(a.x, b.x, a.y, b.y);
        }
    for (int m = 0; m < 8; ++m)
        for (int l = 0; l < 32; ++l) {
            const float2 a = wp_f[l + 64 * m], b = wp_f[l + 64 * m + 32];
            wp4_f[m * 32 + l] = make_float4(a.x, b.x, a.y, b.y);
        }
    for (int k = 0; k < NBINS; ++k) {
        const double th = 2.0 * kPi * ((WIN_LPAD * k) % NFFT) / NFFT;
        phi_f[k] = make_float2((float)std::cos(th), (float)std::sin(th));
    }
    std::vector<double> w2(WIN);
    for (int k = 0; k < WIN; ++k) w2[k] = hann[k] * hann[k];
    std::vector<float> wss_per(HOP);
    for (int r = 0; r < HOP; ++r) {
        float acc = 0.f;
        int mmax = (WIN - 1 - r) / HOP;
        for (int m = mmax; m >= 0; --m) acc = (float)((double)acc + w2[r + HOP * m]);   // ascending frame index
        wss_per[r] = acc;
    }
    // Griffin-Lim load tables: interior frames multiply by analysis window / envelope in one go; edge frames look the
    // reciprocal envelope up by padded position (left edge) or by offset from the last frame's first tap (right edge).
    auto wss_host = [&](int q, int T) -> float {                   // librosa.filters.window_sumsquare at padded position q
        const int c = q - WIN_LPAD;
        if (c < 0) return 0.f;
        const int t_hi = std::min(T - 1, c / HOP);
        const int t_lo = c < WIN ? 0 : (c - WIN) / HOP + 1;
        float acc = 0.f;
        for (int t = t_lo; t <= t_hi; ++t) acc = (float)((double)acc + w2[c - HOP * t]);
        return acc;
    };
    auto rinv_of = [](float w) -> float { return w > FLT_MIN ? (float)(1.0 / (double)w) : 1.0f; };
    std::vector<float2> wa2_f(NLIVE_C);
    for (int n = 0; n < NLIVE_C; ++n)
        wa2_f[n] = make_float2((float)(0.5 * hann[2 * n] / (double)wss_per[(2 * n) % HOP]),
                               (float)(0.5 * hann[2 * n + 1] / (double)wss_per[(2 * n + 1) % HOP]));
    const int T_big = 64;
    std::vector<float> rinv_l(GL_EDGE_L), rinv_r(GL_EDGE_R);
    for (int q = 0; q < GL_EDGE_L; ++q) rinv_l[q] = rinv_of(wss_host(q, T_big));
    for (int u = 0; u < GL_EDGE_R; ++u) rinv_r[u] = rinv_of(wss_host(u - 1100 + WIN_LPAD + HOP * (T_big - 1), T_big));

    c->mel_basis = build_mel_basis();
    for (int m = 0; m < NMELS; ++m)
        if (c->mel_basis[(size_t)m * NBINS + 1024] != 0.f) { delete c; return fail(UNAST_E_ARG, "filterbank has weight on the Nyquist bin"); }
    // filterbank items of the warp-autonomous kernel (frontend_au.cuh): every mel's run of non-zero bins cut into
    // pieces of 8 bins that start at a multiple of 4, dealt to (round, lane) slots in (mel, piece) order
    std::vector<float4> it_w;
    std::vector<int> it_k0;
    std::vector<int2> it_mel(NMELS);
    {
        std::vector<std::pair<int, int>> items;                          // (mel, k0)
        for (int m = 0; m < NMELS; ++m) {
            int lo = -1, hi = -1;
            for (int k = 0; k < NBINS; ++k) if (c->mel_basis[(size_t)m * NBINS + k] != 0.f) { if (lo < 0) lo = k; hi = k; }
            it_mel[m] = make_int2((int)items.size(), 0);
            if (lo < 0) { items.push_back({m, 0}); it_mel[m].y = 1; continue; }     // empty filter: one all-zero item
            for (int k0 = lo & ~3; k0 <= hi; k0 += 8) { items.push_back({m, k0}); ++it_mel[m].y; }
        }
        for (int m = 0; m < NMELS; ++m)
            if (it_mel[m].y > (m < 32 ? AU_MEL_CAP0 : m < 64 ? AU_MEL_CAP1 : AU_MEL_CAP2)) { delete c; return fail(UNAST_E_ARG, "unast_create: filterbank item plan exceeds AU_MEL_CAP"); }
        const int rounds = (((int)items.size() + 31) / 32 + AU_BATCH - 1) / AU_BATCH * AU_BATCH;
        if (rounds > AU_MAX_ROUNDS) { delete c; return fail(UNAST_E_ARG, "unast_create: filterbank item plan exceeds AU_MAX_ROUNDS"); }
        it_w.assign((size_t)rounds * 64, make_float4(0.f, 0.f, 0.f, 0.f));
        it_k0.assign((size_t)rounds * 32, 0);
        for (size_t i = 0; i < items.size(); ++i) {
            const int r = (int)(i / 32), l = (int)(i % 32), m = items[i].first, k0 = items[i].second;
            auto wt = [&](int k) { return k < NBINS ? c->mel_basis[(size_t)m * NBINS + k] : 0.f; };
            it_k0[i] = k0;
            it_w[(size_t)r * 64 + l] = make_float4(wt(k0), wt(k0 + 1), wt(k0 + 2), wt(k0 + 3));
            it_w[(size_t)r * 64 + 32 + l] = make_float4(wt(k0 + 4), wt(k0 + 5), wt(k0 + 6), wt(k0 + 7));
        }
        c->items.rounds = rounds;
    }

    cudaError_t e = cudaSuccess;
    auto up = [&](auto** d, const auto& h) { if (e == cudaSuccess) e = upload(d, h); };
    up(&c->win2_f, win2_f); up(&c->win2i_f, win2i_f); up(&c->wini2_f, wini2_f); up(&c->tw_f, tw_f); up(&c->wp_f, wp_f); up(&c->phi_f, phi_f);
    up(&c->win2_d, win2_d); up(&c->win2i_d, win2i_d); up(&c->tw_d, tw_d); up(&c->wp_d, wp_d);
    up(&c->hann_d, hann);
    up(&c->tw4_f, tw4_f); up(&c->tw4a_f, tw4a_f); up(&c->wp4_f, wp4_f);
    up(&c->wa2_f, wa2_f); up(&c->rinv_l, rinv_l); up(&c->rinv_r, rinv_r);
    up(&c->it_w, it_w); up(&c->it_k0, it_k0); up(&c->it_mel, it_mel);
    up(&c->wss_per, wss_per); up(&c->w2, w2);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_err, sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(c->d_err, 0, sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_abort, sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(c->d_abort, 0, sizeof(int));
    if (e == cudaSuccess) {                                             // a failure here only disables unast_mel_augment_static
        void* sym = nullptr;
        c->aug_static_err = cudaGetSymbolAddress(&sym, g_aug_static_parts);
        c->aug_static = c->aug_static_err == cudaSuccess ? (double*)sym : nullptr;
        if (c->aug_static_err != cudaSuccess) (void)cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaFuncSetAttribute(frontend_au_kernel<float, FE32_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeAuLayout<float, FE32_WARPS>::bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(frontend_au_kernel<double, FE64_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeAuLayout<double, FE64_WARPS>::bytes);
    c->items.w = c->it_w; c->items.k0 = c->it_k0; c->items.mel = c->it_mel;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(prenet_fc1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, pre_smem_bytes(PRE_MAXN));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gl_persist_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, glp_smem_bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gl_persist_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, glp_smem_bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gl_persist_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, glp_smem_bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gl_persist_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, glp_smem_bytes);
    if (e != cudaSuccess) { unast_destroy(c); return fail(UNAST_E_CUDA, std::string("unast_create: ") + cudaGetErrorString(e)); }
    *out = c;
    return UNAST_OK;
}

int unast_destroy(unast_ctx* c) {
    if (!c) return UNAST_OK;
    DeviceGuard guard(c->device);
    cudaFree(c->win2_f); cudaFree(c->win2i_f); cudaFree(c->win2i_d); cudaFree(c->wini2_f); cudaFree(c->tw_f); cudaFree(c->wp_f); cudaFree(c->phi_f);
    cudaFree(c->win2_d); cudaFree(c->tw_d); cudaFree(c->wp_d); cudaFree(c->hann_d);
    cudaFree(c->tw4_f); cudaFree(c->tw4a_f); cudaFree(c->wp4_f);
    cudaFree(c->wa2_f); cudaFree(c->rinv_l); cudaFree(c->rinv_r);
    cudaFree(c->it_w); cudaFree(c->it_k0); cudaFree(c->it_mel);
    cudaFree(c->wss_per); cudaFree(c->w2); cudaFree(c->d_err); cudaFree(c->d_abort);
    delete c;
    return UNAST_OK;
}

int unast_num_sms(const unast_ctx* c) { return c ? c->num_sms : 0; }

int unast_check(unast_ctx* c, void* stream) {
    if (!c) return fail(UNAST_E_ARG, "unast_check: ctx is NULL");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    int h = 0;
    CU(cudaMemcpyAsync(&h, c->d_err, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    if (h != ERR_NONE) {
        CU(cudaMemsetAsync(c->d_err, 0, sizeof(int), (cudaStream_t)stream));
        if (h == ERR_TOO_SHORT) return fail(UNAST_E_DATA, "an utterance is too short for reflect padding (needs > 1024 samples / >= 5 frames)");
        if (h == ERR_EMPTY) return fail(UNAST_E_DATA, "an utterance is empty after trimming");
        if (h == ERR_SYNC) return fail(UNAST_E_DATA, "Griffin-Lim: a wait for a neighbouring run / peer rank timed out");
        return fail(UNAST_E_DATA, "kernel reported data error " + std::to_string(h));
    }
    return UNAST_OK;
}

int unast_get_mel_basis(const unast_ctx* c, float* out_host) {
    if (!c || !out_host) return fail(UNAST_E_ARG, "unast_get_mel_basis: NULL argument");
    memcpy(out_host, c->mel_basis.data(), c->mel_basis.size() * sizeof(float));
    return UNAST_OK;
}

// ---- trim ----
static int nblk_of(int max_len) { return (max_len + TRIM_HOP - 1) / TRIM_HOP; }

size_t unast_trim_workspace_bytes(int B, int max_len) { return align256((size_t)B * nblk_of(max_len) * sizeof(double)); }

int unast_trim_bounds(unast_ctx* c, const void* wave, int dtype, const int64_t* wave_off, int B, int max_len,
                      int32_t* bounds_out, void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !wave || !wave_off || !bounds_out || B <= 0 || max_len <= 0) return fail(UNAST_E_ARG, "unast_trim_bounds: bad argument");
    if (dtype != UNAST_F32 && dtype != UNAST_I16) return fail(UNAST_E_ARG, "unast_trim_bounds: dtype must be UNAST_F32 or UNAST_I16");
    if (workspace_bytes < unast_trim_workspace_bytes(B, max_len) || !workspace) return fail(UNAST_E_WORKSPACE, "unast_trim_bounds: workspace too small");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    const int nblk = nblk_of(max_len);
    dim3 grid((nblk + 7) / 8, B);
    block_energy_kernel<<<grid, 256, 0, st>>>(wave, dtype == UNAST_I16, (const long long*)wave_off, B, nblk, (double*)workspace);
    trim_decide_kernel<<<B, 128, 0, st>>>(wave, dtype == UNAST_I16, (const long long*)wave_off, nblk, (const double*)workspace, bounds_out, c->d_err);
    CU(cudaGetLastError());
    return UNAST_OK;
}

int unast_frame_offsets(unast_ctx* c, const int32_t* bounds, int B, int32_t* frame_off_out, void* stream) {
    if (!c || !bounds || !frame_off_out || B <= 0) return fail(UNAST_E_ARG, "unast_frame_offsets: bad argument");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    frame_offsets_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(bounds, B, frame_off_out, c->d_err);
    CU(cudaGetLastError());
    return UNAST_OK;
}

// ---- front end ----
size_t unast_frontend_workspace_bytes(int B) { (void)B; return 256; }    // the fused kernel needs no scratch; kept for ABI stability

int unast_frontend(unast_ctx* c, const void* wave, int dtype, const int64_t* wave_off, const int32_t* bounds,
                   const int32_t* frame_off, int B, int precision, float* mel_out, int mel_layout, int T_max,
                   float* mag_out, void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !wave || !wave_off || !bounds || !frame_off || !mel_out || B <= 0) return fail(UNAST_E_ARG, "unast_frontend: bad argument");
    if (dtype != UNAST_F32 && dtype != UNAST_I16) return fail(UNAST_E_ARG, "unast_frontend: dtype must be UNAST_F32 or UNAST_I16");
    if (precision != UNAST_PREC_F32 && precision != UNAST_PREC_F64) return fail(UNAST_E_ARG, "unast_frontend: bad precision");
    if (mel_layout != UNAST_MEL_MAJOR_PER_UTT && mel_layout != UNAST_MEL_PADDED_BTM) return fail(UNAST_E_ARG, "unast_frontend: bad mel_layout");
    if (mel_layout == UNAST_MEL_PADDED_BTM && T_max <= 0) return fail(UNAST_E_ARG, "unast_frontend: T_max required for the padded layout");
    if (!workspace || workspace_bytes < unast_frontend_workspace_bytes(B)) return fail(UNAST_E_WORKSPACE, "unast_frontend: workspace too small");
    // the PCM16 path reads the wave as 32-bit words and prefetches it in 16-byte cp.async chunks relative to the base
    if (dtype == UNAST_I16 && ((uintptr_t)wave & 15)) return fail(UNAST_E_ARG, "unast_frontend: an int16 wave must be 16-byte aligned");
    if (dtype == UNAST_F32 && ((uintptr_t)wave & 3)) return fail(UNAST_E_ARG, "unast_frontend: a float32 wave must be 4-byte aligned");
    if (((uintptr_t)mel_out & 3) || ((uintptr_t)mag_out & 3)) return fail(UNAST_E_ARG, "unast_frontend: outputs must be 4-byte aligned");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    FrontendParams P;
    P.wave = wave; P.wave_is_i16 = dtype == UNAST_I16; P.wave_off = (const long long*)wave_off; P.bounds = bounds;
    P.frame_off = frame_off; P.B = B; P.mag_out = mag_out; P.mel_out = mel_out;
    P.mel_layout = mel_layout; P.T_max = T_max;
    P.err = c->d_err;
    const bool i16 = dtype == UNAST_I16;
    if (precision == UNAST_PREC_F32) {
        P.win2 = i16 ? c->win2i_f : c->win2_f; P.tw1024 = c->tw_f; P.wp2 = c->wp_f;
        frontend_au_kernel<float, FE32_WARPS><<<c->num_sms, FE32_WARPS * 32, FeAuLayout<float, FE32_WARPS>::bytes, st>>>(P, c->items);
    } else {
        P.win2 = i16 ? (const void*)c->win2i_d : (const void*)c->win2_d; P.tw1024 = c->tw_d; P.wp2 = c->wp_d;
        frontend_au_kernel<double, FE64_WARPS><<<c->num_sms, FE64_WARPS * 32, FeAuLayout<double, FE64_WARPS>::bytes, st>>>(P, c->items);
    }
    CU(cudaGetLastError());
    return UNAST_OK;
}

// ---- Griffin-Lim ----
namespace {
constexpr int GL_MAX_RUNS = 8192;                                  // >= SMs x GL_WARPS: one run of frames per warp
#ifndef GL_COST_EDGE
#define GL_COST_EDGE 6
#endif
constexpr long long GL_COST_I = 4, GL_COST_E = GL_COST_EDGE;     // relative cost of an interior / an edge frame (reflect padding, envelope tables)
constexpr int GL_MIN_RUN = 9;                                      // frames: reflect reads and seams then reach only the adjacent run

// A fragment is a contiguous range [t0, t0 + n) of the frames of an utterance of Tg frames.  Whole utterances are
// fragments with t0 = 0, n = Tg; the sharded long-form run hands every rank one fragment of the waveform.
struct GlPlan {
    std::vector<int> frame_off, t0, Tg;
    std::vector<long long> pad_off, cost_off;
    long long total_frames = 0, total_pad = 0;
    size_t o_runs = 0, o_meta_i = 0, o_meta_l = 0, o_sync = 0, o_mail = 0, o_S = 0, o_ola = 0, ola_stride = 0, o_tprev = 0, o_be = 0, bytes = 0;
    size_t o_f0 = 0;
};
// edge frames (reflect padding, envelope tables) cost 6, interior frames 4 (measured on C3: 4 / 5 / 6 / 7 / 8 / 10 -> 15.1 / 14.0 / 13.4 / 13.5 / 13.5 / 13.9 ms)
void frag_edges(int t0, int n, int Tg, int& head, int& tail) {
    head = std::max(0, std::min(t0 + n, 4) - t0);
    tail = std::max(0, t0 + n - std::max(t0, Tg - 4));
    if (head + tail > n) tail = n - head;
}
bool gl_plan(const int32_t* frames, const int32_t* t0, const int32_t* tglob, int B, float momentum, int max_wav, GlPlan& p) {
    p.frame_off.assign(B + 1, 0); p.pad_off.assign(B + 1, 0); p.cost_off.assign(B + 1, 0); p.t0.assign(B, 0); p.Tg.assign(B, 0);
    for (int b = 0; b < B; ++b) {
        const int n = frames[b];
        p.t0[b] = t0 ? t0[b] : 0;
        p.Tg[b] = tglob ? tglob[b] : n;
        if (p.Tg[b] < 5 || n < 1 || p.t0[b] < 0 || p.t0[b] + n > p.Tg[b]) return false;   // reflect padding needs 275 (T-1) > 1024
        if (n != p.Tg[b] && n < GL_MIN_RUN) return false;           // a fragment must hold a whole run
        int head, tail;
        frag_edges(p.t0[b], n, p.Tg[b], head, tail);
        p.frame_off[b + 1] = p.frame_off[b] + n;
        p.pad_off[b + 1] = p.pad_off[b] + NFFT + (long long)HOP * (n - 1);
        p.cost_off[b + 1] = p.cost_off[b] + GL_COST_I * n + (GL_COST_E - GL_COST_I) * (head + tail);
    }
    p.total_frames = p.frame_off[B]; p.total_pad = p.pad_off[B];
    size_t o = 0;
    p.o_runs = o; o += align256((size_t)GL_MAX_RUNS * sizeof(int4));
    p.o_meta_i = o; o += align256(3 * (size_t)(B + 1) * sizeof(int));
    p.o_meta_l = o; o += align256(2 * (size_t)(B + 1) * sizeof(long long));
    p.o_sync = o; o += align256((size_t)GL_MAX_RUNS * sizeof(int));
    p.o_mail = o; o += align256((size_t)GL_MAX_LINKS * sizeof(int));
    p.o_S = o; o += align256((size_t)p.total_frames * S_STRIDE * sizeof(float));
    p.ola_stride = align256((size_t)p.total_pad * sizeof(float));
    p.o_ola = o; o += 3 * p.ola_stride;
    p.o_tprev = o; if (momentum != 0.f) o += align256((size_t)p.total_frames * NBINS * sizeof(float2));
    p.o_be = o; o += align256((size_t)B * ((max_wav + TRIM_HOP - 1) / TRIM_HOP) * sizeof(double));
    p.o_f0 = o; o += align256((size_t)p.total_frames * GL_F0_STRIDE * sizeof(float));
    p.bytes = o;
    return true;
}
int max_wav_of(const int32_t* frames, int B) {
    int m = 0;
    for (int b = 0; b < B; ++b) m = std::max(m, HOP * (frames[b] - 1));
    return m;
}

// the frame at cost point cpt (costs as in gl_plan)
int frame_at(const GlPlan& p, int B, long long cpt) {
    if (cpt <= 0) return 0;
    if (cpt >= p.cost_off[B]) return (int)p.total_frames;
    int lo = 0, hi = B;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (p.cost_off[mid] <= cpt) lo = mid; else hi = mid; }
    const int n = p.frame_off[lo + 1] - p.frame_off[lo];
    int head, tail;
    frag_edges(p.t0[lo], n, p.Tg[lo], head, tail);
    const long long ci = cpt - p.cost_off[lo];
    const long long inner = GL_COST_I * (n - head - tail);
    long long j;
    if (ci < GL_COST_E * head) j = ci / GL_COST_E;
    else if (ci < GL_COST_E * head + inner) j = head + (ci - GL_COST_E * head) / GL_COST_I;
    else j = n - tail + (ci - GL_COST_E * head - inner) / GL_COST_E;
    return p.frame_off[lo] + (int)std::min<long long>(j, n - 1);
}
}  // namespace

size_t unast_griffinlim_workspace_bytes(const int32_t* frames_host, int B, float momentum) {
    GlPlan p;
    if (!frames_host || B <= 0 || !gl_plan(frames_host, nullptr, nullptr, B, momentum, max_wav_of(frames_host, B), p)) return 0;
    return p.bytes;
}

int unast_griffinlim_set_block(unast_ctx* c, int frames) {
    if (!c || (frames != 0 && (frames < 4 || frames > 64))) return fail(UNAST_E_ARG, "unast_griffinlim_set_block: 0 (off) or 4..64 frames");
    c->gl_block = frames;
    return UNAST_OK;
}

int unast_griffinlim_layout(const int32_t* frames_host, int B, float momentum, size_t* out5) {
    GlPlan p;
    if (!frames_host || !out5 || B <= 0) return fail(UNAST_E_ARG, "unast_griffinlim_layout: bad argument");
    std::vector<int32_t> big(B);
    for (int b = 0; b < B; ++b) big[b] = frames_host[b] + 2 * GL_MIN_RUN;     // any valid (t0, Tg): the layout depends on n only
    if (!gl_plan(frames_host, nullptr, big.data(), B, momentum, max_wav_of(frames_host, B), p)) return fail(UNAST_E_ARG, "unast_griffinlim_layout: bad frame counts");
    out5[0] = p.o_ola; out5[1] = p.o_ola + p.ola_stride; out5[2] = p.o_ola + 2 * p.ola_stride; out5[3] = p.o_mail; out5[4] = p.bytes;
    return UNAST_OK;
}

namespace {
struct GlDev {                      // device views of one call's workspace
    int* frame_off; int* t0; int* Tg; long long* pad_off; long long* aux_off; float* ola[3]; int final_buf; double* be;
};

// Everything up to and including the persistent launch; shared by the whole-utterance and the fragment entry points.
int gl_run(unast_ctx* c, const float* mag_norm, const GlPlan& p, int B, const unast_gl_link* links, int n_links, int epoch, int n_iter,
           float momentum, const float* angles0, float power, char* ws, cudaStream_t st, GlDev& d) {
    d.frame_off = (int*)(ws + p.o_meta_i);
    d.t0 = d.frame_off + (B + 1);
    d.Tg = d.t0 + (B + 1);
    d.pad_off = (long long*)(ws + p.o_meta_l);
    d.aux_off = d.pad_off + (B + 1);
    d.be = (double*)(ws + p.o_be);
    CU(cudaMemcpyAsync(d.frame_off, p.frame_off.data(), (B + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d.t0, p.t0.data(), B * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d.Tg, p.Tg.data(), B * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d.pad_off, p.pad_off.data(), (B + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    float* S = (float*)(ws + p.o_S);
    for (int k = 0; k < 3; ++k) d.ola[k] = (float*)(ws + p.o_ola + k * p.ola_stride);
    d.final_buf = (n_iter + 1) % 3;
    CU(cudaMemsetAsync(d.ola[0], 0, 3 * p.ola_stride, st));
    CU(cudaMemsetAsync(ws + p.o_sync, 0, (size_t)GL_MAX_RUNS * sizeof(int), st));
    CU(cudaMemsetAsync(c->d_abort, 0, sizeof(int), st));
    gl_prepare_kernel<<<c->num_sms * 8, 256, 0, st>>>(mag_norm, p.total_frames, power, S);

    float* frames0 = nullptr;
    if (!angles0) {                                                      // zero-phase start: the ill-conditioned first inverse in float64
        frames0 = (float*)(ws + p.o_f0);
        GlInitParams I;
        I.S = S; I.F = p.total_frames; I.tw = c->tw_d; I.wp = c->wp_d; I.hann = c->hann_d; I.frames0 = frames0;
        const int grid = (int)std::min<long long>((p.total_frames + GL_INIT_WARPS - 1) / GL_INIT_WARPS, (long long)c->num_sms * 3);
        gl_init_f64_kernel<<<grid, GL_INIT_WARPS * 32, GL_INIT_WARPS * xs_bytes<double>(), st>>>(I);
    }

    GlParams P;
    memset(&P, 0, sizeof(P));
    P.S = S; P.frame_off = d.frame_off; P.pad_off = d.pad_off; P.t0 = d.t0; P.Tg = d.Tg; P.B = B;
    for (int k = 0; k < 3; ++k) P.ola[k] = d.ola[k];
    const int blk = c->gl_block;
    P.block = blk;
    P.n_iter = n_iter; P.done = (int*)(ws + p.o_sync); P.epoch = epoch; P.abort = c->d_abort; P.err = c->d_err;
    P.tprev = momentum != 0.f ? (float2*)(ws + p.o_tprev) : nullptr;
    P.angles0 = (const float2*)angles0;
    P.frames0 = frames0;
    P.alpha = momentum / (1.0f + momentum);
    P.wa2 = c->wa2_f; P.win2 = c->win2_f; P.wini2 = c->wini2_f; P.tw4 = c->tw4_f; P.tw4a = c->tw4a_f; P.wp4 = c->wp4_f; P.phi = c->phi_f;
    P.rinv_l = c->rinv_l; P.rinv_r = c->rinv_r; P.wss_per = c->wss_per; P.w2 = c->w2;

    // ---- segments: linked seams cut the frame list; a run never spans one ----
    if (n_links < 0 || n_links > GL_MAX_LINKS) return fail(UNAST_E_ARG, "unast_griffinlim: too many links");
    std::vector<int> cut(1, 0);                                        // frame indices where a segment starts
    std::vector<int> link_at_start(B + 1, -1), link_at_end(B + 1, -1); // link index by fragment
    int* mail = (int*)(ws + p.o_mail);
    for (int i = 0; i < n_links; ++i) {
        const unast_gl_link& L = links[i];
        if (L.fragment < 0 || L.fragment >= B || (L.side != 0 && L.side != 1) || L.my_mail < 0 || L.my_mail >= GL_MAX_LINKS || !L.peer_mail)
            return fail(UNAST_E_ARG, "unast_griffinlim: bad link");
        const int b = L.fragment, n = p.frame_off[b + 1] - p.frame_off[b];
        if (blk > 0 && (L.side == 0 ? p.t0[b] : p.t0[b] + n) % blk)
            return fail(UNAST_E_ARG, "unast_griffinlim: with canonical blocks a linked seam must sit on a multiple of the block");
        long long seam0;                                               // local index of the seam's first sample
        if (L.side == 0) {
            if (p.t0[b] == 0 || link_at_start[b] >= 0) return fail(UNAST_E_ARG, "unast_griffinlim: left link on a fragment that starts its utterance");
            link_at_start[b] = i; cut.push_back(p.frame_off[b]); seam0 = p.pad_off[b] + WIN_LPAD;
        } else {
            if (p.t0[b] + n == p.Tg[b] || link_at_end[b] >= 0) return fail(UNAST_E_ARG, "unast_griffinlim: right link on a fragment that ends its utterance");
            link_at_end[b] = i; cut.push_back(p.frame_off[b + 1]); seam0 = p.pad_off[b] + (long long)HOP * n + WIN_LPAD;
        }
        for (int k = 0; k < 3; ++k) {
            if (!L.peer_ola[k] || ((uintptr_t)L.peer_ola[k] & 3)) return fail(UNAST_E_ARG, "unast_griffinlim: bad peer buffer");
            P.links[i].peer[k] = (float*)L.peer_ola[k] - seam0;
        }
        P.links[i].push = (int*)L.peer_mail;
        P.links[i].poll = mail + L.my_mail;
    }
    cut.push_back((int)p.total_frames);
    std::sort(cut.begin(), cut.end());
    cut.erase(std::unique(cut.begin(), cut.end()), cut.end());

    // ---- one run of frames per warp: equal shares of the cost inside every segment, runs of >= GL_MIN_RUN frames ----
    const int Wmax = std::min(GL_MAX_RUNS, c->num_sms * GL_WARPS);
    const int nseg = (int)cut.size() - 1;
    if (nseg > Wmax) return fail(UNAST_E_ARG, "unast_griffinlim: more segments than warps");
    auto cost_of_frame_index = [&](int f) -> long long {              // cost prefix at frame f
        if (f >= p.total_frames) return p.cost_off[B];
        int lo = 0, hi = B;
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (p.frame_off[mid] <= f) lo = mid; else hi = mid; }
        const int n = p.frame_off[lo + 1] - p.frame_off[lo], j = f - p.frame_off[lo];
        int head, tail;
        frag_edges(p.t0[lo], n, p.Tg[lo], head, tail);
        return p.cost_off[lo] + GL_COST_I * j + (GL_COST_E - GL_COST_I) * (std::min(j, head) + std::max(0, j - (n - tail)));
    };
    std::vector<int> wseg(nseg);
    int wsum = 0;
    for (int s = 0; s < nseg; ++s) {
        const long long cs = cost_of_frame_index(cut[s + 1]) - cost_of_frame_index(cut[s]);
        const int frames_s = cut[s + 1] - cut[s];
        long long w = (long long)Wmax * cs / std::max<long long>(1, p.cost_off[B]);
        w = std::min<long long>(w, frames_s / GL_MIN_RUN);
        wseg[s] = (int)std::max<long long>(1, w);
        wsum += wseg[s];
    }
    while (wsum > Wmax) { int big = 0; for (int s = 1; s < nseg; ++s) if (wseg[s] > wseg[big]) big = s; --wseg[big]; --wsum; }
    // canonical mode: run boundaries sit on block boundaries of their utterance (or on a fragment's edge)
    auto frag_of = [&](int f) {                                          // fragment holding frame f < total
        int lo = 0, hi = B;
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (p.frame_off[mid] <= f) lo = mid; else hi = mid; }
        return lo;
    };
    auto boundary_after = [&](int f) {                                   // smallest boundary > f
        const int bb = frag_of(f), t = f - p.frame_off[bb] + p.t0[bb];
        return std::min(f + ((t / blk + 1) * blk - t), p.frame_off[bb + 1]);
    };
    auto snap = [&](int f, int prev, int end) {                          // boundary nearest to f, above prev, at most end
        if (f >= end) return end;
        const int bb = frag_of(f), t = f - p.frame_off[bb] + p.t0[bb];
        int g = f;
        if (t % blk) {
            const int lo = std::max(f - t % blk, p.frame_off[bb]), hi = boundary_after(f);
            g = (f - lo <= hi - f) ? lo : hi;
        }
        while (g <= prev) g = boundary_after(g);
        return std::min(g, end);
    };
    std::vector<int4> runs;
    runs.reserve(wsum);
    int b = 0;
    for (int s = 0; s < nseg; ++s) {
        const long long c0 = cost_of_frame_index(cut[s]), c1 = cost_of_frame_index(cut[s + 1]);
        int prev = cut[s];
        for (int i = 0; i < wseg[s] && prev < cut[s + 1]; ++i) {
            int f1 = i + 1 == wseg[s] ? cut[s + 1] : frame_at(p, B, c0 + (c1 - c0) * (i + 1) / wseg[s]);
            if (blk > 0) {
                f1 = snap(f1, prev, cut[s + 1]);
                if (cut[s + 1] - f1 < 2 * blk) f1 = cut[s + 1];              // no sliver at the end of a segment
            } else {
                f1 = std::max(f1, prev + 1);
                f1 = std::min(f1, cut[s + 1] - (wseg[s] - 1 - i));
            }
            while (b + 1 < B && p.frame_off[b + 1] <= prev) ++b;
            int ll = -1, lr = -1;
            if (prev == cut[s] && prev == p.frame_off[b]) ll = link_at_start[b];
            if (f1 == cut[s + 1]) {                                       // the fragment this run ends in
                int be = b;
                while (be + 1 < B && p.frame_off[be + 1] < f1) ++be;
                if (f1 == p.frame_off[be + 1]) lr = link_at_end[be];
            }
            runs.push_back(make_int4(prev, f1, b, (ll + 1) | ((lr + 1) << 16)));
            prev = f1;
        }
    }
    P.n_runs = (int)runs.size();
    P.runs = (const int4*)(ws + p.o_runs);
    CU(cudaMemcpyAsync(ws + p.o_runs, runs.data(), runs.size() * sizeof(int4), cudaMemcpyHostToDevice, st));

    // one cooperative launch: every CTA must be resident, warps wait for one another
    const int grid = (P.n_runs + GL_WARPS - 1) / GL_WARPS;
    void* args[] = {(void*)&P};
    const void* fn = blk > 0 ? (momentum != 0.f ? (const void*)gl_persist_kernel<true, true> : (const void*)gl_persist_kernel<false, true>)
                             : (momentum != 0.f ? (const void*)gl_persist_kernel<true, false> : (const void*)gl_persist_kernel<false, false>);
    CU(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(GL_WARPS * 32), args, (size_t)glp_smem_bytes, st));
    return UNAST_OK;
}
}  // namespace

int unast_griffinlim(unast_ctx* c, const float* mag_norm, const int32_t* frames_host, int B, int n_iter, float momentum,
                     const float* angles0, float power, int deemphasis, float* wav_out, int32_t* bounds_out,
                     void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !mag_norm || !frames_host || !wav_out || B <= 0 || n_iter < 0) return fail(UNAST_E_ARG, "unast_griffinlim: bad argument");
    if (momentum < 0.f || momentum >= 1.f) return fail(UNAST_E_ARG, "unast_griffinlim: momentum must be in [0, 1)");
    GlPlan p;
    const int max_wav = max_wav_of(frames_host, B);
    if (!gl_plan(frames_host, nullptr, nullptr, B, momentum, max_wav, p)) return fail(UNAST_E_ARG, "unast_griffinlim: every utterance needs >= 5 frames (librosa reflect padding)");
    if (!workspace || workspace_bytes < p.bytes) return fail(UNAST_E_WORKSPACE, "unast_griffinlim: workspace too small");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    GlDev d;
    const int rc = gl_run(c, mag_norm, p, B, nullptr, 0, 1, n_iter, momentum, angles0, power, ws, st, d);
    if (rc != UNAST_OK) return rc;
    std::vector<long long> wav_off(B + 1, 0);
    for (int b = 0; b < B; ++b) wav_off[b + 1] = wav_off[b] + (long long)HOP * (frames_host[b] - 1);
    CU(cudaMemcpyAsync(d.aux_off, wav_off.data(), (B + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));

    DeemphParams D;
    D.ola = d.ola[d.final_buf]; D.x = nullptr; D.frame_off = d.frame_off; D.pad_off = d.pad_off; D.wav_off = d.aux_off;
    D.out = wav_out; D.wss_per = c->wss_per; D.w2 = c->w2; D.coef = deemphasis ? 0.97 : 0.0;
    const int chunks = (max_wav + DE_CHUNK - 1) / DE_CHUNK;
    deemph_kernel<<<dim3((chunks + 3) / 4, B), 128, 0, st>>>(D);
    if (bounds_out) {
        const int nblk = nblk_of(max_wav);
        block_energy_kernel<<<dim3((nblk + 7) / 8, B), 256, 0, st>>>(wav_out, 0, d.aux_off, B, nblk, d.be);
        trim_decide_kernel<<<B, 128, 0, st>>>(wav_out, 0, d.aux_off, nblk, d.be, bounds_out, c->d_err);
    }
    CU(cudaGetLastError());
    return UNAST_OK;
}

// samples of the whole utterance's waveform a fragment owns: [275 t0 - 551, 275 (t0 + n) - 551), clipped to [0, 275 (Tg - 1))
static void frag_samples(int t0, int n, int Tg, long long& a, long long& b) {
    a = t0 == 0 ? 0 : (long long)HOP * t0 - (HALF - WIN_LPAD);
    b = t0 + n == Tg ? (long long)HOP * (Tg - 1) : (long long)HOP * (t0 + n) - (HALF - WIN_LPAD);
}

int unast_griffinlim_fragment_samples(int32_t t0, int32_t n, int32_t tglob, int64_t* first_out, int64_t* end_out) {
    if (!first_out || !end_out || n < 1 || t0 < 0 || t0 + n > tglob) return fail(UNAST_E_ARG, "unast_griffinlim_fragment_samples: bad argument");
    long long a, b;
    frag_samples(t0, n, tglob, a, b);
    *first_out = a; *end_out = b;
    return UNAST_OK;
}

int unast_griffinlim_fragments(unast_ctx* c, const float* mag_norm, const int32_t* frames_host, const int32_t* t0_host,
                               const int32_t* tglob_host, int B, const unast_gl_link* links_host, int n_links, int epoch, int n_iter,
                               float momentum, const float* angles0, float power, float* raw_out, void* workspace,
                               size_t workspace_bytes, void* stream) {
    if (!c || !mag_norm || !frames_host || !t0_host || !tglob_host || !raw_out || B <= 0 || n_iter < 0 || epoch < 1)
        return fail(UNAST_E_ARG, "unast_griffinlim_fragments: bad argument");
    if (n_links > 0 && !links_host) return fail(UNAST_E_ARG, "unast_griffinlim_fragments: links is NULL");
    if (momentum < 0.f || momentum >= 1.f) return fail(UNAST_E_ARG, "unast_griffinlim_fragments: momentum must be in [0, 1)");
    GlPlan p;
    if (!gl_plan(frames_host, t0_host, tglob_host, B, momentum, max_wav_of(frames_host, B), p))
        return fail(UNAST_E_ARG, "unast_griffinlim_fragments: bad fragment (utterances need >= 5 frames, partial fragments >= 9)");
    if (!workspace || workspace_bytes < p.bytes) return fail(UNAST_E_WORKSPACE, "unast_griffinlim_fragments: workspace too small");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    GlDev d;
    const int rc = gl_run(c, mag_norm, p, B, links_host, n_links, epoch, n_iter, momentum, angles0, power, ws, st, d);
    if (rc != UNAST_OK) return rc;
    std::vector<long long> raw_off(B + 1, 0);
    long long longest = 0;
    for (int b = 0; b < B; ++b) {
        long long a, e;
        frag_samples(p.t0[b], frames_host[b], p.Tg[b], a, e);
        raw_off[b + 1] = raw_off[b] + (e - a);
        longest = std::max(longest, e - a);
    }
    CU(cudaMemcpyAsync(d.aux_off, raw_off.data(), (B + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    GlExtractParams X;
    X.ola = d.ola[d.final_buf]; X.frame_off = d.frame_off; X.pad_off = d.pad_off; X.t0 = d.t0; X.Tg = d.Tg; X.raw_off = d.aux_off;
    X.out = raw_out; X.wss_per = c->wss_per; X.w2 = c->w2;
    gl_extract_kernel<<<dim3((unsigned)std::min<long long>(4096, (longest + 1023) / 1024), B), 256, 0, st>>>(X);
    CU(cudaGetLastError());
    return UNAST_OK;
}

/* de-emphasis (float64 IIR, utils.py:301) + librosa.effects.trim bounds of plain, already normalised waveforms: the
 * epilogue of the sharded run once the fragments' samples have been gathered */
size_t unast_deemph_trim_workspace_bytes(int B, int max_len) { return unast_trim_workspace_bytes(B, max_len) + align256((size_t)(B + 1) * sizeof(long long)); }

int unast_deemph_trim(unast_ctx* c, const float* x, const int64_t* off, int B, int max_len, int deemphasis, float* wav_out,
                      int32_t* bounds_out, void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !x || !off || !wav_out || B <= 0 || max_len <= 0) return fail(UNAST_E_ARG, "unast_deemph_trim: bad argument");
    if (!workspace || workspace_bytes < unast_deemph_trim_workspace_bytes(B, max_len)) return fail(UNAST_E_WORKSPACE, "unast_deemph_trim: workspace too small");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    DeemphParams D;
    D.ola = nullptr; D.x = x; D.frame_off = nullptr; D.pad_off = nullptr; D.wav_off = (const long long*)off;
    D.out = wav_out; D.wss_per = c->wss_per; D.w2 = c->w2; D.coef = deemphasis ? 0.97 : 0.0;
    const int chunks = (max_len + DE_CHUNK - 1) / DE_CHUNK;
    deemph_kernel<<<dim3((chunks + 3) / 4, B), 128, 0, st>>>(D);
    if (bounds_out) {
        const int nblk = nblk_of(max_len);
        block_energy_kernel<<<dim3((nblk + 7) / 8, B), 256, 0, st>>>(wav_out, 0, (const long long*)off, B, nblk, (double*)workspace);
        trim_decide_kernel<<<B, 128, 0, st>>>(wav_out, 0, (const long long*)off, nblk, (const double*)workspace, bounds_out, c->d_err);
    }
    CU(cudaGetLastError());
    return UNAST_OK;
}

/* ---- peer-shareable device memory (CUDA IPC) for the linked overlap-add buffers ---- */
int unast_peer_alloc(unast_ctx* c, size_t bytes, void** dptr, unsigned char* handle64) {
    if (!c || !dptr || !handle64 || bytes == 0) return fail(UNAST_E_ARG, "unast_peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    void* p = nullptr;
    CU(cudaMalloc(&p, bytes));
    CU(cudaMemset(p, 0, bytes));                                         // mailboxes start at zero; epochs only grow
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, 64);
    *dptr = p;
    return UNAST_OK;
}
int unast_peer_open(unast_ctx* c, const unsigned char* handle64, void** dptr) {
    if (!c || !dptr || !handle64) return fail(UNAST_E_ARG, "unast_peer_open: bad argument");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return UNAST_OK;
}
int unast_peer_close(unast_ctx* c, void* dptr) {
    if (!c || !dptr) return fail(UNAST_E_ARG, "unast_peer_close: bad argument");
    DeviceGuard guard(c->device);
    CU(cudaIpcCloseMemHandle(dptr));
    return UNAST_OK;
}
int unast_peer_free(unast_ctx* c, void* dptr) {
    if (!c || !dptr) return fail(UNAST_E_ARG, "unast_peer_free: bad argument");
    DeviceGuard guard(c->device);
    CU(cudaFree(dptr));
    return UNAST_OK;
}

// ---- diagnostics ----
// ---- masking ops on the padded mel batch (8(f) rank 2) ----
size_t unast_mel_augment_workspace_bytes(int B) { return B > 0 ? align256((size_t)B * AUG_PARTS * sizeof(double)) : 0; }

int unast_mel_augment(unast_ctx* c, const float* x, int B, int T, int D, const int32_t* spans, const float* keep,
                      int replace_with_zero, float* out, void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !x || !out || B <= 0 || T <= 0 || D <= 0) return fail(UNAST_E_ARG, "unast_mel_augment: bad argument");
    if (B > 65535) return fail(UNAST_E_ARG, "unast_mel_augment: B must be <= 65535");
    const bool need_mean = spans && !replace_with_zero;
    if (need_mean && x == out) return fail(UNAST_E_ARG, "unast_mel_augment: in-place needs replace_with_zero or no spans");
    if (need_mean && (!workspace || workspace_bytes < unast_mel_augment_workspace_bytes(B)))
        return fail(UNAST_E_WORKSPACE, "unast_mel_augment: workspace too small");
    if (spans && ((uintptr_t)spans & 15)) return fail(UNAST_E_ARG, "unast_mel_augment: spans must be 16-byte aligned");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    const long long slab = (long long)T * D;
    double* parts = (double*)workspace;
    if (need_mean) aug_sum_kernel<<<dim3(AUG_PARTS, B), AUG_THREADS, 0, st>>>(x, slab, parts);
    // enough CTAs per utterance to fill the machine a few times over, at least 8 float4 per thread
    long long per = (slab / 4 + AUG_THREADS * 8 - 1) / (AUG_THREADS * 8);
    const long long want = ((long long)c->num_sms * 8 + B - 1) / B;
    if (per > want) per = want;
    if (per < 1) per = 1;
    aug_apply_kernel<<<dim3((unsigned)per, B), AUG_THREADS, 0, st>>>(x, out, T, D, spans, keep, need_mean ? parts : nullptr,
                                                                     replace_with_zero);
    CU(cudaGetLastError());
    return UNAST_OK;
}

// The same two launches as unast_mel_augment with the partial sums of the mean in g_aug_static_parts: no workspace argument,
// no cudaMalloc / cudaFree / allocating library call on this path.  aug_sum_kernel overwrites every partial sum of utterances
// 0..B-1 with a plain store before aug_apply_kernel reads it, so nothing has to start at zero; calls that overlap on the GPU
// (two streams, two contexts of one device) share the array and race.
int unast_mel_augment_static(unast_ctx* c, const float* x, int B, int T, int D, const int32_t* spans, const float* keep,
                             int replace_with_zero, float* out, void* stream) {
    if (!c || !x || !out || B <= 0 || T <= 0 || D <= 0) return fail(UNAST_E_ARG, "unast_mel_augment_static: bad argument");
    if (B > UNAST_AUG_STATIC_MAX_B)
        return fail(UNAST_E_ARG, "unast_mel_augment_static: B exceeds UNAST_AUG_STATIC_MAX_B (1024 utterances); use unast_mel_augment with a workspace");
    const bool need_mean = spans && !replace_with_zero;
    if (need_mean && x == out) return fail(UNAST_E_ARG, "unast_mel_augment_static: in-place needs replace_with_zero or no spans");
    if (spans && ((uintptr_t)spans & 15)) return fail(UNAST_E_ARG, "unast_mel_augment_static: spans must be 16-byte aligned");
    if (!c->aug_static)
        return fail(UNAST_E_CUDA, std::string("unast_mel_augment_static: cudaGetSymbolAddress failed at set-up: ") + cudaGetErrorString(c->aug_static_err));
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    const long long slab = (long long)T * D;
    double* parts = c->aug_static;                                       // the device address, never the symbol's name
    if (need_mean) aug_sum_kernel<<<dim3(AUG_PARTS, B), AUG_THREADS, 0, st>>>(x, slab, parts);
    long long per = (slab / 4 + AUG_THREADS * 8 - 1) / (AUG_THREADS * 8);
    const long long want = ((long long)c->num_sms * 8 + B - 1) / B;
    if (per > want) per = want;
    if (per < 1) per = 1;
    aug_apply_kernel<<<dim3((unsigned)per, B), AUG_THREADS, 0, st>>>(x, out, T, D, spans, keep, need_mean ? parts : nullptr,
                                                                     replace_with_zero);
    CU(cudaGetLastError());
    return UNAST_OK;
}

// ---- masking ops + first prenet layer on the tensor cores (8(f) rank 2) ----
size_t unast_prenet_packed_bytes(int n_out) {
    if (n_out < 32 || n_out > PRE_MAXN || (n_out & 31)) return 0;
    return (size_t)pre_packed_bytes(n_out);
}

int unast_prenet_pack(unast_ctx* c, const float* weight, const float* bias, int n_out, void* packed_out, void* stream) {
    if (!c || !weight || !packed_out || unast_prenet_packed_bytes(n_out) == 0) return fail(UNAST_E_ARG, "unast_prenet_pack: bad argument (n_out must be a multiple of 32, 32..256)");
    if ((uintptr_t)packed_out & 15) return fail(UNAST_E_ARG, "unast_prenet_pack: packed_out must be 16-byte aligned");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    const int n = n_out * PRE_K;
    prenet_pack_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(weight, bias, n_out, (unsigned char*)packed_out);
    CU(cudaGetLastError());
    return UNAST_OK;
}

int unast_mel_prenet_fc1(unast_ctx* c, const float* x, int B, int T, const int32_t* spans, const float* keep, int replace_with_zero,
                         const void* packed, int n_out, int relu, float* y, void* workspace, size_t workspace_bytes, void* stream) {
    if (!c || !x || !y || !packed || B <= 0 || T <= 0 || unast_prenet_packed_bytes(n_out) == 0) return fail(UNAST_E_ARG, "unast_mel_prenet_fc1: bad argument");
    if (B > 65535) return fail(UNAST_E_ARG, "unast_mel_prenet_fc1: B must be <= 65535");
    if (((uintptr_t)x | (uintptr_t)y | (uintptr_t)packed) & 15) return fail(UNAST_E_ARG, "unast_mel_prenet_fc1: x, y and packed must be 16-byte aligned");
    if (spans && ((uintptr_t)spans & 15)) return fail(UNAST_E_ARG, "unast_mel_prenet_fc1: spans must be 16-byte aligned");
    const bool need_mean = spans && !replace_with_zero;
    if (need_mean && (!workspace || workspace_bytes < unast_mel_augment_workspace_bytes(B))) return fail(UNAST_E_WORKSPACE, "unast_mel_prenet_fc1: workspace too small");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    double* parts = (double*)workspace;
    if (need_mean) aug_sum_kernel<<<dim3(AUG_PARTS, B), AUG_THREADS, 0, st>>>(x, (long long)T * PRE_K, parts);
    PrenetParams P;
    P.x = x; P.y = y; P.B = B; P.T = T; P.N = n_out; P.spans = spans; P.keep = keep; P.parts = need_mean ? parts : nullptr;
    P.replace_with_zero = replace_with_zero; P.relu = relu; P.packed = (const unsigned char*)packed;
    const long long tiles = ((long long)B * T + PRE_TILE_M - 1) / PRE_TILE_M;
    const int grid = (int)std::min<long long>(tiles, c->num_sms);
    const size_t smem = (size_t)pre_smem_bytes(n_out);
    prenet_fc1_kernel<<<grid, PRE_THREADS, smem, st>>>(P);
    CU(cudaGetLastError());
    return UNAST_OK;
}

int unast_debug_rfft(unast_ctx* c, const float* frames, int F, int precision, float* out, void* stream) {
    if (!c || !frames || !out || F <= 0) return fail(UNAST_E_ARG, "unast_debug_rfft: bad argument");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    cudaStream_t st = (cudaStream_t)stream;
    if (precision == UNAST_PREC_F64)
        debug_rfft_f64_kernel<<<(F + 3) / 4, 128, 4 * xs_bytes<double>(), st>>>(frames, F, c->tw_d, c->wp_d, (float2*)out);
    else
        debug_rfft_pk_kernel<<<(F + 3) / 4, 128, 4 * GLP_XS_BYTES, st>>>(frames, F, c->tw4_f, c->wp4_f, (float2*)out);
    CU(cudaGetLastError());
    return UNAST_OK;
}

int unast_debug_irfft(unast_ctx* c, const float* spec, int F, float* out, void* stream) {
    if (!c || !spec || !out || F <= 0) return fail(UNAST_E_ARG, "unast_debug_irfft: bad argument");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    debug_irfft_kernel<<<(F + 3) / 4, 128, 4 * GLP_XS_BYTES, (cudaStream_t)stream>>>((const float2*)spec, F, c->tw4a_f, c->wp4_f, out);
    CU(cudaGetLastError());
    return UNAST_OK;
}

int unast_debug_pipe_peak(unast_ctx* c, int kind, int iters, float* out, double* flop_out, void* stream) {
    if (!c || !out || iters <= 0 || kind < 0 || kind > 2) return fail(UNAST_E_ARG, "unast_debug_pipe_peak: bad argument");
    DeviceGuard guard(c->device);
    if (!guard.ok()) return fail(UNAST_E_CUDA, "cannot select the context's device");
    const int grid = c->num_sms * 8, block = 256;
    cudaStream_t st = (cudaStream_t)stream;
    if (kind == 0) pipe_peak_kernel<0><<<grid, block, 0, st>>>(iters, 0.999f, 0.001f, out);
    else if (kind == 1) pipe_peak_kernel<1><<<grid, block, 0, st>>>(iters, 0.999f, 0.001f, out);
    else pipe_peak_kernel<2><<<grid, block, 0, st>>>(iters, 0.999f, 0.001f, out);
    CU(cudaGetLastError());
    if (flop_out) *flop_out = 2.0 * (kind == 2 ? 2.0 : 1.0) * PEAK_ILP * (double)iters * grid * block;
    return UNAST_OK;
}

}  // extern "C"
