// Internal helpers shared by all kernels of libmedvae_b200.so (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace mvae {

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------- error plumbing
void set_error(const char* fmt, ...);
#define MVAE_CHECK(cond, ...)              \
    do {                                   \
        if (!(cond)) {                     \
            mvae::set_error(__VA_ARGS__);  \
            return 1;                      \
        }                                  \
    } while (0)
#define MVAE_CUDA(expr)                                                                   \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            mvae::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                            __FILE__, __LINE__);                                          \
            return 1;                                                                     \
        }                                                                                 \
    } while (0)
void count_launch();   // api.cu: every kernel launch of this library is counted (mvae_launch_count)
#define MVAE_LAUNCH_CHECK()          \
    do {                             \
        mvae::count_launch();        \
        MVAE_CUDA(cudaGetLastError()); \
    } while (0)

int debug_knob(int key);   // api.cu

// ---------------------------------------------------------------- dtype helpers
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// 8-element vector of T: 16 B for bf16, 32 B for float
template <typename T> struct Vec8;
template <> struct Vec8<bf16> {
    uint4 raw;
    __device__ __forceinline__ void load(const bf16* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(bf16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ void unpack(float* f) const {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 t = __bfloat1622float2(h[i]);
            f[2 * i] = t.x;
            f[2 * i + 1] = t.y;
        }
    }
    __device__ __forceinline__ void pack(const float* f) {
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    }
};
template <> struct Vec8<float> {
    float4 a, b;
    __device__ __forceinline__ void load(const float* p) {
        a = *reinterpret_cast<const float4*>(p);
        b = *reinterpret_cast<const float4*>(p + 4);
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = a;
        *reinterpret_cast<float4*>(p + 4) = b;
    }
    __device__ __forceinline__ void unpack(float* f) const {
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
        f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    }
    __device__ __forceinline__ void pack(const float* f) {
        a = make_float4(f[0], f[1], f[2], f[3]);
        b = make_float4(f[4], f[5], f[6], f[7]);
    }
};

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; every thread gets the result.  `sh` needs 32 floats.
__device__ __forceinline__ float block_sum(float v, float* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    float r = (lane < nw) ? sh[lane] : 0.f;
    r = warp_sum(r);
    return r;
}
__device__ __forceinline__ double block_sum_d(double v, double* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double r = (lane < nw) ? sh[lane] : 0.0;
    r = warp_sum_d(r);
    return r;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
// sigmoid for the bf16 path: 0.5*tanh(x/2)+0.5 on the MUFU tanh unit (3 instructions, ~5e-4 relative
// error, an eighth of a bf16 ulp); the fp32 verification mode keeps the exact form.  The GroupNorm
// kernels are issue-bound, not HBM-bound, with the IEEE division of the exact form.
__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <typename T> __device__ __forceinline__ float sigmoid_t(float x);
template <> __device__ __forceinline__ float sigmoid_t<float>(float x) { return sigmoidf_(x); }
template <> __device__ __forceinline__ float sigmoid_t<bf16>(float x) { return fmaf(tanh_approx(0.5f * x), 0.5f, 0.5f); }

// ---------------------------------------------------------------- conv "tap programs"
// A convolution (any of fprop / dgrad / wgrad, stride 1, strided or up-sampling) is lowered on
// the host to gathers over strided sub-grids ("views") of NHWC tensors plus a list of taps.
constexpr int MAX_TAPS = 16;
constexpr int MAX_VIEWS = 4;

struct View {
    long long off;          // element offset of (n=0, y=0, x=0, c=0)
    int Hs, Ws;             // extent of the sub-grid
    long long sn, sy, sx;   // element strides
};
struct Tap {
    int view;   // which input view
    int dx, dy; // offset added to the iteration coordinate
    int slab;   // which packed-weight slab
    int oview;  // wgrad only: which dy view
};
struct TapProg {
    int ntaps;
    Tap t[MAX_TAPS];
};

// out[oview](n,y,x,:) = bias + sum_t in[view_t](n, y+dy_t, x+dx_t, :) * W[slab_t]^T (+ residual)
struct GatherProblem {
    int B, OH, OW;      // iteration grid
    int K, N;           // channels reduced per tap, output channels
    int nviews;
    View in_views[MAX_VIEWS];
    View out_view;
    TapProg taps;
};
// dW[slab](co,ci) = sum_{n,y,x} dy[oview_t](n,y,x,co) * in[view_t](n, y+dy_t, x+dx_t, ci)
struct ReduceProblem {
    int B, GH, GW;      // iteration grid
    int Cout, Cin;
    int n_inviews, n_dyviews;
    View in_views[MAX_VIEWS];
    View dy_views[MAX_VIEWS];
    TapProg taps;       // one tap per slab
};

}  // namespace mvae
