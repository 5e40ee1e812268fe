// Fused GroupNorm (+SiLU) forward / backward on NHWC activations.
// Reference: Normalize = nn.GroupNorm(min(32,C), C, eps=1e-6, affine) followed by
// nonlinearity = x*sigmoid(x)  (encoder_decoder.py:13-15, 28-33).
//
// HBM-bound.  Layout [B][HW][C]; a CTA streams whole pixel rows (all channels, coalesced 16 B
// vectors), so group statistics are reduced from per-thread channel accumulators:
//   pass 1 (stats):  grid (splits, B): per-(sample, split, group) shifted sums -> scratch
//   pass 2 (apply):  grid (splits, B): folds the split partials (fp32, shifted by a per-group
//                    pivot so E[x^2]-E[x]^2 does not cancel), normalises, applies SiLU.
// Backward has the same two-pass shape and also fuses the residual-branch gradient add.
#include "../../include/medvae_b200.h"
#include "common.cuh"

namespace mvae {

constexpr int GN_THREADS = 256;
constexpr int MAX_G = 32;

// The two passes of each direction (statistics, then apply) run over CHUNKS of samples small enough
// that the slab(s) read by the first pass are still in the 126 MB L2 when the second pass reads
// them again: HBM then sees each tensor once.
int gn_chunk(int B, int HW, int C, int tensors) {
    // elements per chunk over all re-read tensors; debug knob 3 (in Mi elements) overrides; 0 = one chunk
    const long long budget = debug_knob(3) > 0 ? ((long long)debug_knob(3) << 20) : (1LL << 40);
    long long per = (long long)HW * C * tensors;
    long long nb = budget / (per > 0 ? per : 1);
    if (nb < 1) nb = 1;
    if (nb > B) nb = B;
    return (int)nb;
}
static int gn_splits_for(int nb, int HW, int C) {
    // ~12 CTAs per SM over the launch (4 waves at 3 resident CTAs); a CTA gets at least 16 K
    // elements (32 rows of 512 channels, 8 rows of 2048 ...) and at least 4 pixel rows
    long long want = (148LL * 12 + nb - 1) / nb;
    long long min_rows = 16384 / (C > 0 ? C : 1);
    if (min_rows < 4) min_rows = 4;
    if (min_rows > 32) min_rows = 32;
    long long cap = (HW + min_rows - 1) / min_rows;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    if (want > 64) want = 64;
    return (int)want;
}
// floats of scratch the forward (backward = 0) or backward (= 1) entry point needs
long long gn_scratch_floats(int B, int HW, int C, int G, int backward) {
    const int nb = gn_chunk(B, HW, C, backward ? 2 : 1);
    const int S = gn_splits_for(nb, HW, C);
    const long long chunks = (B + nb - 1) / nb;
    long long n = (long long)nb * S * G * 2;
    if (backward) n += chunks * nb * S * 2 * C;
    return n;
}
int gn_splits(int B, int HW, int C) {
    // legacy sizing helper: the larger of the two directions' split counts
    int a = gn_splits_for(gn_chunk(B, HW, C, 1), HW, C), b = gn_splits_for(gn_chunk(B, HW, C, 2), HW, C);
    return a > b ? a : b;
}
// thread mapping for a [rows][C] slab processed VEC channels at a time
struct RowMap {
    int nvec, tpr, rpi, v, r;   // vectors per row, threads per row, rows per iteration, my vector, my row lane
    bool active;
};
template <int VEC>
__device__ __forceinline__ RowMap row_map(int C) {
    RowMap m;
    m.nvec = C / VEC;
    m.tpr = m.nvec < GN_THREADS ? m.nvec : GN_THREADS;
    m.rpi = GN_THREADS / m.tpr;
    m.v = threadIdx.x % m.tpr;
    m.r = threadIdx.x / m.tpr;
    m.active = m.r < m.rpi;
    return m;
}

template <typename T, int VEC> struct Ld;
template <typename T> struct Ld<T, 8> {
    typedef Vec8<T> Raw;
    static __device__ __forceinline__ void load_raw(const T* p, Raw& r) { r.load(p); }
    static __device__ __forceinline__ void unpack(const Raw& r, float* f) { r.unpack(f); }
    static __device__ __forceinline__ void load(const T* p, float* f) { Vec8<T> v; v.load(p); v.unpack(f); }
    static __device__ __forceinline__ void store(T* p, const float* f) { Vec8<T> v; v.pack(f); v.store(p); }
};
template <typename T> struct Ld<T, 1> {
    typedef T Raw;
    static __device__ __forceinline__ void load_raw(const T* p, Raw& r) { r = *p; }
    static __device__ __forceinline__ void unpack(const Raw& r, float* f) { f[0] = to_f<T>(r); }
    static __device__ __forceinline__ void load(const T* p, float* f) { f[0] = to_f<T>(*p); }
    static __device__ __forceinline__ void store(T* p, const float* f) { *p = from_f<T>(f[0]); }
};

// ------------------------------------------------------------------ forward: statistics
template <typename T, int VEC>
__global__ void __launch_bounds__(GN_THREADS) gn_stats_kernel(const T* __restrict__ x, float* __restrict__ scratch,
                                                              int HW, int C, int G, int S) {
    __shared__ float sh_s[MAX_G], sh_q[MAX_G], pivot[MAX_G];
    extern __shared__ float part[];   // [rows per iteration][C][2]
    const int n = blockIdx.y, sp = blockIdx.x;
    const int cpg = C / G;
    const T* xn = x + (long long)n * HW * C;
    if (threadIdx.x < G) {
        sh_s[threadIdx.x] = 0.f; sh_q[threadIdx.x] = 0.f;
        pivot[threadIdx.x] = to_f<T>(xn[threadIdx.x * cpg]);
    }
    __syncthreads();
    const int p0 = (int)(((long long)HW * sp) / S), p1 = (int)(((long long)HW * (sp + 1)) / S);
    for (int vb = 0; vb < C / VEC; vb += GN_THREADS) {
        RowMap m = row_map<VEC>(C);
        int v = vb + m.v;
        if (m.active && v < m.nvec) {
            float s[VEC], q[VEC], pv[VEC];
#pragma unroll
            for (int j = 0; j < VEC; ++j) { s[j] = 0.f; q[j] = 0.f; pv[j] = pivot[(v * VEC + j) / cpg]; }
            int p = p0 + m.r;
            for (; p + 3 * m.rpi < p1; p += 4 * m.rpi) {
                float f[4][VEC];
#pragma unroll
                for (int u = 0; u < 4; ++u) Ld<T, VEC>::load(xn + (long long)(p + u * m.rpi) * C + v * VEC, f[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int j = 0; j < VEC; ++j) { float d = f[u][j] - pv[j]; s[j] += d; q[j] = fmaf(d, d, q[j]); }
            }
            for (; p < p1; p += m.rpi) {
                float f[VEC];
                Ld<T, VEC>::load(xn + (long long)p * C + v * VEC, f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) { float d = f[j] - pv[j]; s[j] += d; q[j] = fmaf(d, d, q[j]); }
            }
            // per-thread channel partials -> shared [row lane][channel]; folded below in a fixed order
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                part[(m.r * C + v * VEC + j) * 2] = s[j];
                part[(m.r * C + v * VEC + j) * 2 + 1] = q[j];
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < G) {
        RowMap m = row_map<VEC>(C);
        float as = 0.f, aq = 0.f;
        for (int r = 0; r < m.rpi; ++r)
            for (int c = threadIdx.x * cpg; c < (threadIdx.x + 1) * cpg; ++c) {
                as += part[(r * C + c) * 2]; aq += part[(r * C + c) * 2 + 1];
            }
        sh_s[threadIdx.x] = as; sh_q[threadIdx.x] = aq;
    }
    __syncthreads();
    if (threadIdx.x < G) {
        float* o = scratch + (((long long)n * S + sp) * G + threadIdx.x) * 2;
        o[0] = sh_s[threadIdx.x]; o[1] = sh_q[threadIdx.x];
    }
}

// ------------------------------------------------------------------ forward: normalise (+SiLU)
template <typename T, int VEC>
__global__ void __launch_bounds__(GN_THREADS) gn_apply_kernel(const T* __restrict__ x, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, T* __restrict__ y,
                                                              float* __restrict__ stats, const float* __restrict__ scratch,
                                                              int HW, int C, int G, int S, float eps, int silu) {
    __shared__ float mean[MAX_G], rstd[MAX_G];
    const int n = blockIdx.y, sp = blockIdx.x;
    const int cpg = C / G;
    const T* xn = x + (long long)n * HW * C;
    T* yn = y + (long long)n * HW * C;
    if (threadIdx.x < G) {
        float s = 0.f, q = 0.f;
        for (int k = 0; k < S; ++k) {
            const float* o = scratch + (((long long)n * S + k) * G + threadIdx.x) * 2;
            s += o[0]; q += o[1];
        }
        float cnt = (float)HW * (float)cpg;
        float m = s / cnt;
        float var = q / cnt - m * m;
        if (var < 0.f) var = 0.f;
        float mu = to_f<T>(xn[threadIdx.x * cpg]) + m;
        float rs = rsqrtf(var + eps);
        mean[threadIdx.x] = mu; rstd[threadIdx.x] = rs;
        if (sp == 0) { stats[((long long)n * G + threadIdx.x) * 2] = mu; stats[((long long)n * G + threadIdx.x) * 2 + 1] = rs; }
    }
    __syncthreads();
    const int p0 = (int)(((long long)HW * sp) / S), p1 = (int)(((long long)HW * (sp + 1)) / S);
    for (int vb = 0; vb < C / VEC; vb += GN_THREADS) {
        RowMap m = row_map<VEC>(C);
        int v = vb + m.v;
        if (m.active && v < m.nvec) {
            float a[VEC], b[VEC];   // y = x*a + b
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                int c = v * VEC + j, g = c / cpg;
                a[j] = rstd[g] * gamma[c];
                b[j] = beta[c] - mean[g] * a[j];
            }
            int p = p0 + m.r;
            for (; p + 3 * m.rpi < p1; p += 4 * m.rpi) {
                float f[4][VEC];
#pragma unroll
                for (int u = 0; u < 4; ++u) Ld<T, VEC>::load(xn + (long long)(p + u * m.rpi) * C + v * VEC, f[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        float t = fmaf(f[u][j], a[j], b[j]);
                        f[u][j] = silu ? t * sigmoid_t<T>(t) : t;
                    }
                    Ld<T, VEC>::store(yn + (long long)(p + u * m.rpi) * C + v * VEC, f[u]);
                }
            }
            for (; p < p1; p += m.rpi) {
                float f[VEC];
                Ld<T, VEC>::load(xn + (long long)p * C + v * VEC, f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    float u = fmaf(f[j], a[j], b[j]);
                    f[j] = silu ? u * sigmoid_t<T>(u) : u;
                }
                Ld<T, VEC>::store(yn + (long long)p * C + v * VEC, f);
            }
        }
    }
}

// ------------------------------------------------------------------ backward: reductions
// With u = x*P + Q the pre-activation (P = rstd*gamma, Q = beta - mean*rstd*gamma), xhat = x*R + M
// (R = rstd, M = -mean*rstd) and du = dy * silu'(u) (or dy):
//   per channel  dbeta = sum du,  dgamma = sum du*xhat
//   per group    s1 = sum du*gamma = sum_c gamma_c*dbeta_c,   s2 = sum du*gamma*xhat = sum_c gamma_c*dgamma_c
// so one pass over (x, dy) accumulates just two numbers per channel.  Everything is folded in a
// fixed order (bitwise run-to-run reproducible): thread partials -> shared -> per-CTA rows of
// `scratch` / `chan_scratch` -> gn_param_reduce_kernel.
template <typename T, int VEC>
__global__ void __launch_bounds__(GN_THREADS, 3) gn_bwd_stats_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                                     const float* __restrict__ gamma,
                                                                     const float* __restrict__ beta,
                                                                     const float* __restrict__ stats,
                                                                     float* __restrict__ scratch,
                                                                     float* __restrict__ chan_scratch, int HW, int C,
                                                                     int G, int S, int silu) {
    __shared__ float mean[MAX_G], rstd[MAX_G];
    extern __shared__ float part[];   // [rows per iteration][C][2] : dgamma, dbeta partials
    const int n = blockIdx.y, sp = blockIdx.x;
    const int cpg = C / G;
    if (threadIdx.x < G) {
        mean[threadIdx.x] = stats[((long long)n * G + threadIdx.x) * 2];
        rstd[threadIdx.x] = stats[((long long)n * G + threadIdx.x) * 2 + 1];
    }
    __syncthreads();
    const T* xn = x + (long long)n * HW * C;
    const T* dn = dy + (long long)n * HW * C;
    const int p0 = (int)(((long long)HW * sp) / S), p1 = (int)(((long long)HW * (sp + 1)) / S);
    const RowMap m = row_map<VEC>(C);
    for (int vb = 0; vb < C / VEC; vb += GN_THREADS) {
        const int v = vb + m.v;
        if (m.active && v < m.nvec) {
            // accumulate t = sum du*x and db = sum du; dgamma = R*t + M*db (xhat = x*R + M) afterwards
            float P[VEC], Q[VEC], dg[VEC], db[VEC];
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const int c = v * VEC + j, g = c / cpg;
                const float rs = rstd[g], ga = gamma[c];
                P[j] = rs * ga; Q[j] = fmaf(-mean[g] * rs, ga, beta[c]);
                dg[j] = 0.f; db[j] = 0.f;
            }
            auto accum = [&](const float* fx, const float* fd) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    float du = fd[j];
                    if (silu) {
                        const float u = fmaf(fx[j], P[j], Q[j]);
                        const float sg = sigmoid_t<T>(u);
                        du *= sg * fmaf(u, 1.f - sg, 1.f);
                    }
                    dg[j] = fmaf(du, fx[j], dg[j]); db[j] += du;
                }
            };
            int p = p0 + m.r;
            for (; p + m.rpi < p1; p += 2 * m.rpi) {
                typename Ld<T, VEC>::Raw rx[2], rd[2];   // raw (packed) loads in flight, unpacked one row at a time
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    Ld<T, VEC>::load_raw(xn + (long long)(p + u * m.rpi) * C + v * VEC, rx[u]);
                    Ld<T, VEC>::load_raw(dn + (long long)(p + u * m.rpi) * C + v * VEC, rd[u]);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    float fx[VEC], fd[VEC];
                    Ld<T, VEC>::unpack(rx[u], fx);
                    Ld<T, VEC>::unpack(rd[u], fd);
                    accum(fx, fd);
                }
            }
            for (; p < p1; p += m.rpi) {
                float fx[VEC], fd[VEC];
                Ld<T, VEC>::load(xn + (long long)p * C + v * VEC, fx);
                Ld<T, VEC>::load(dn + (long long)p * C + v * VEC, fd);
                accum(fx, fd);
            }
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const int g = (v * VEC + j) / cpg;
                const float rs = rstd[g];
                float* o = part + (size_t)(m.r * C + v * VEC + j) * 2;
                o[0] = fmaf(rs, dg[j], -mean[g] * rs * db[j]); o[1] = db[j];
            }
        }
    }
    __syncthreads();
    const long long cta = (long long)n * S + sp;
    float* chan = chan_scratch + cta * 2 * C;
    for (int c = threadIdx.x; c < C; c += GN_THREADS) {
        float dg = 0.f, db = 0.f;
        for (int r = 0; r < m.rpi; ++r) { dg += part[(size_t)(r * C + c) * 2]; db += part[(size_t)(r * C + c) * 2 + 1]; }
        chan[c] = dg; chan[C + c] = db;
    }
    __syncthreads();   // chan[] written by this CTA is read back below (same CTA: visible after the barrier)
    if (threadIdx.x < G) {
        float s1 = 0.f, s2 = 0.f;
        for (int c = threadIdx.x * cpg; c < (threadIdx.x + 1) * cpg; ++c) {
            const float ga = gamma[c];
            s1 = fmaf(ga, chan[C + c], s1); s2 = fmaf(ga, chan[c], s2);
        }
        float* o = scratch + (cta * G + threadIdx.x) * 2;
        o[0] = s1; o[1] = s2;
    }
}

// dgamma[c] = sum over the B*S CTA rows of chan[.][0][c], dbeta likewise; 32 columns x 8 row lanes
// per CTA, fixed summation order.
__global__ void __launch_bounds__(256) gn_param_reduce_kernel(const float* __restrict__ chan, float* __restrict__ dgamma,
                                                              float* __restrict__ dbeta, int rows, int C) {
    __shared__ float sh[8][33];
    const int col = blockIdx.x * 32 + (threadIdx.x & 31);   // column of the [rows][2C] matrix
    const int lane = threadIdx.x >> 5;
    float acc = 0.f;
    if (col < 2 * C)
        for (int r = lane; r < rows; r += 8) acc += chan[(long long)r * 2 * C + col];
    sh[lane][threadIdx.x & 31] = acc;
    __syncthreads();
    if (lane == 0 && col < 2 * C) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x];
        if (col < C) dgamma[col] = t; else dbeta[col - C] = t;
    }
}

// ------------------------------------------------------------------ backward: dx
// dx = rstd * (du*gamma - m1 - xhat*m2) + dx_add  =  du*P - K1 - xhat*K2 + dx_add
template <typename T, int VEC>
__global__ void __launch_bounds__(GN_THREADS, 3) gn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                                     const float* __restrict__ gamma,
                                                                     const float* __restrict__ beta,
                                                                     const float* __restrict__ stats,
                                                                     const float* __restrict__ scratch,
                                                                     const T* __restrict__ dx_add, T* __restrict__ dx,
                                                                     int HW, int C, int G, int S, int silu) {
    __shared__ float mean[MAX_G], rstd[MAX_G], k1[MAX_G], k2[MAX_G];
    const int n = blockIdx.y, sp = blockIdx.x;
    const int cpg = C / G;
    if (threadIdx.x < G) {
        float s1 = 0.f, s2 = 0.f;
        for (int k = 0; k < S; ++k) {
            const float* o = scratch + (((long long)n * S + k) * G + threadIdx.x) * 2;
            s1 += o[0]; s2 += o[1];
        }
        const float cnt = (float)HW * (float)cpg;
        const float rs = stats[((long long)n * G + threadIdx.x) * 2 + 1];
        mean[threadIdx.x] = stats[((long long)n * G + threadIdx.x) * 2];
        rstd[threadIdx.x] = rs;
        k1[threadIdx.x] = rs * (s1 / cnt); k2[threadIdx.x] = rs * (s2 / cnt);
    }
    __syncthreads();
    const long long base = (long long)n * HW * C;
    const int p0 = (int)(((long long)HW * sp) / S), p1 = (int)(((long long)HW * (sp + 1)) / S);
    const RowMap m = row_map<VEC>(C);
    for (int vb = 0; vb < C / VEC; vb += GN_THREADS) {
        const int v = vb + m.v;
        if (m.active && v < m.nvec) {
            // dx = du*P - K1 - xhat*K2 = du*P - A - x*Bc  with  A = K1 + M*K2, Bc = R*K2
            float P[VEC], Q[VEC], A[VEC], Bc[VEC];
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const int c = v * VEC + j, g = c / cpg;
                const float rs = rstd[g], ga = gamma[c], mm = -mean[g] * rs;
                P[j] = rs * ga; Q[j] = fmaf(mm, ga, beta[c]);
                A[j] = fmaf(mm, k2[g], k1[g]); Bc[j] = rs * k2[g];
            }
            auto one = [&](float* fx, const float* fd, const float* fa) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    float du = fd[j];
                    if (silu) {
                        const float u = fmaf(fx[j], P[j], Q[j]);
                        const float sg = sigmoid_t<T>(u);
                        du *= sg * fmaf(u, 1.f - sg, 1.f);
                    }
                    const float r = fmaf(du, P[j], -fmaf(fx[j], Bc[j], A[j]));
                    fx[j] = dx_add ? r + fa[j] : r;
                }
            };
            int p = p0 + m.r;
            for (; p + m.rpi < p1; p += 2 * m.rpi) {
                typename Ld<T, VEC>::Raw rx[2], rd[2], ra[2];
                long long off[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    off[u] = base + (long long)(p + u * m.rpi) * C + v * VEC;
                    Ld<T, VEC>::load_raw(x + off[u], rx[u]);
                    Ld<T, VEC>::load_raw(dy + off[u], rd[u]);
                    if (dx_add) Ld<T, VEC>::load_raw(dx_add + off[u], ra[u]);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    float fx[VEC], fd[VEC], fa[VEC];
                    Ld<T, VEC>::unpack(rx[u], fx);
                    Ld<T, VEC>::unpack(rd[u], fd);
                    if (dx_add) Ld<T, VEC>::unpack(ra[u], fa);
                    one(fx, fd, fa);
                    Ld<T, VEC>::store(dx + off[u], fx);
                }
            }
            for (; p < p1; p += m.rpi) {
                const long long off = base + (long long)p * C + v * VEC;
                float fx[VEC], fd[VEC], fa[VEC];
                Ld<T, VEC>::load(x + off, fx);
                Ld<T, VEC>::load(dy + off, fd);
                if (dx_add) Ld<T, VEC>::load(dx_add + off, fa);
                one(fx, fd, fa);
                Ld<T, VEC>::store(dx + off, fx);
            }
        }
    }
}

// ================================================================== one-pass variants
// When the conv kernel that PRODUCED x has already accumulated the per-(sample, group) sums in its
// epilogue (GnHook, conv_lower.h), GroupNorm is a single streaming pass in each direction:
//   forward :  y = [silu](x*a + b)                      reads x, writes y            (4 B/elem in bf16)
//   backward:  dx = du*P - K1 - xhat*K2 (+ dx_add)      reads du, x (, dx_add), writes dx
//              and, from the same registers, dgamma = sum du*xhat, dbeta = sum du and the column
//              sums of dx (= the bias gradient of the conv that produced x).
// Work items are (sample, row chunk) pairs dealt round-robin to a persistent grid sized to the
// machine (no tail wave); a thread owns 8 fixed channels, so per-channel coefficients live in
// registers and are rebuilt only when the sample changes.  Needs C % 8 == 0, C / 8 <= 256.
constexpr int GN1_THREADS = 256;

struct Gn1Map {
    int tpr, rpi, v, r;
    bool active;
};
__device__ __forceinline__ Gn1Map gn1_map(int C) {
    Gn1Map m;
    m.tpr = C / 8;
    m.rpi = GN1_THREADS / m.tpr;
    m.v = threadIdx.x % m.tpr;
    m.r = threadIdx.x / m.tpr;
    m.active = m.r < m.rpi;
    return m;
}

template <typename T>
__global__ void __launch_bounds__(GN1_THREADS, 3)
gn_apply_pre_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                    T* __restrict__ y, const float* __restrict__ sums, float* __restrict__ stats, int B, int HW, int C,
                    int G, int chunks, int rows_per, float eps, int silu) {
    const Gn1Map m = gn1_map(C);
    if (!m.active) return;
    const int cpg = C / G;
    const float inv_cnt = 1.f / ((float)HW * (float)cpg);
    const int items = B * chunks;
    const int c0 = m.v * 8;
    float a[8], b[8];
    int last_n = -1;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int n = item / chunks, ch = item - n * chunks;
        if (n != last_n) {
            last_n = n;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = c0 + j, g = c / cpg;
                const float2 sq = __ldg(reinterpret_cast<const float2*>(sums) + (long long)n * G + g);
                const float mu = sq.x * inv_cnt;
                float var = sq.y * inv_cnt - mu * mu;
                if (var < 0.f) var = 0.f;
                const float rs = rsqrtf(var + eps);
                a[j] = rs * __ldg(gamma + c);
                b[j] = __ldg(beta + c) - mu * a[j];
                if (ch == 0 && m.r == 0 && c == g * cpg)
                    *reinterpret_cast<float2*>(stats + ((long long)n * G + g) * 2) = make_float2(mu, rs);
            }
        }
        const int p0 = ch * rows_per;
        int p1 = p0 + rows_per;
        if (p1 > HW) p1 = HW;
        const T* xn = x + ((long long)n * HW) * C + c0;
        T* yn = y + ((long long)n * HW) * C + c0;
        int p = p0 + m.r;
        for (; p + 3 * m.rpi < p1; p += 4 * m.rpi) {
            Vec8<T> raw[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) raw[u].load(xn + (long long)(p + u * m.rpi) * C);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                float f[8];
                raw[u].unpack(f);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float t = fmaf(f[j], a[j], b[j]);
                    f[j] = silu ? t * sigmoid_t<T>(t) : t;
                }
                raw[u].pack(f);
                raw[u].store(yn + (long long)(p + u * m.rpi) * C);
            }
        }
        for (; p < p1; p += m.rpi) {
            Vec8<T> raw;
            raw.load(xn + (long long)p * C);
            float f[8];
            raw.unpack(f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float t = fmaf(f[j], a[j], b[j]);
                f[j] = silu ? t * sigmoid_t<T>(t) : t;
            }
            raw.pack(f);
            raw.store(yn + (long long)p * C);
        }
    }
}

// GPV = groups per 8-channel vector (1: C/G % 8 == 0;  2: C/G == 4;  4: C/G == 2)
template <typename T, int GPV>
__global__ void __launch_bounds__(GN1_THREADS, 2)
gn_bwd_pre_kernel(const T* __restrict__ du, const T* __restrict__ x, const float* __restrict__ gamma,
                  const float* __restrict__ stats, const float* __restrict__ s12, const T* __restrict__ dx_add,
                  T* __restrict__ dx, float* __restrict__ chan_scratch, int B, int HW, int C, int G, int chunks,
                  int rows_per) {
    extern __shared__ float part[];   // [rpi][3][C]
    const Gn1Map m = gn1_map(C);
    const int cpg = C / G;
    constexpr int CPV = 8 / GPV;      // channels of one group inside the vector
    const float inv_cnt = 1.f / ((float)HW * (float)cpg);
    const int items = B * chunks;
    const int c0 = m.v * 8;
    float dg[8], db[8], cs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[j] = 0.f; db[j] = 0.f; cs[j] = 0.f; }
    if (m.active) {
        float ga[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) ga[j] = __ldg(gamma + c0 + j);
        float P[8], R[GPV], M[GPV], K1[GPV], K2[GPV];
        int last_n = -1;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int n = item / chunks, ch = item - n * chunks;
            if (n != last_n) {
                last_n = n;
#pragma unroll
                for (int q = 0; q < GPV; ++q) {
                    const int g = (c0 + q * CPV) / cpg;
                    const float2 mr = __ldg(reinterpret_cast<const float2*>(stats) + (long long)n * G + g);
                    const float2 ss = __ldg(reinterpret_cast<const float2*>(s12) + (long long)n * G + g);
                    R[q] = mr.y; M[q] = -mr.x * mr.y;
                    K1[q] = mr.y * ss.x * inv_cnt; K2[q] = mr.y * ss.y * inv_cnt;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) P[j] = R[j / CPV] * ga[j];
            }
            const int p0 = ch * rows_per;
            int p1 = p0 + rows_per;
            if (p1 > HW) p1 = HW;
            const long long base = ((long long)n * HW) * C + c0;
            auto one = [&](const Vec8<T>& rx, const Vec8<T>& rd, const Vec8<T>& ra, long long off) {
                float fx[8], fd[8], fa[8];
                rx.unpack(fx); rd.unpack(fd);
                if (dx_add) ra.unpack(fa);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int q = j / CPV;
                    const float xh = fmaf(fx[j], R[q], M[q]);
                    dg[j] = fmaf(fd[j], xh, dg[j]);
                    db[j] += fd[j];
                    float r = fmaf(fd[j], P[j], -fmaf(xh, K2[q], K1[q]));
                    if (dx_add) r += fa[j];
                    cs[j] += r;
                    fx[j] = r;
                }
                Vec8<T> o; o.pack(fx); o.store(dx + off);
            };
            int p = p0 + m.r;
            for (; p + m.rpi < p1; p += 2 * m.rpi) {
                Vec8<T> rx[2], rd[2], ra[2];
                long long off[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    off[u] = base + (long long)(p + u * m.rpi) * C;
                    rx[u].load(x + off[u]); rd[u].load(du + off[u]);
                    if (dx_add) ra[u].load(dx_add + off[u]);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) one(rx[u], rd[u], ra[u], off[u]);
            }
            for (; p < p1; p += m.rpi) {
                Vec8<T> rx, rd, ra;
                const long long off = base + (long long)p * C;
                rx.load(x + off); rd.load(du + off);
                if (dx_add) ra.load(dx_add + off);
                one(rx, rd, ra, off);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            part[(m.r * 3 + 0) * C + c0 + j] = dg[j];
            part[(m.r * 3 + 1) * C + c0 + j] = db[j];
            part[(m.r * 3 + 2) * C + c0 + j] = cs[j];
        }
    }
    __syncthreads();
    float* dst = chan_scratch + (long long)blockIdx.x * 3 * C;
    for (int i = threadIdx.x; i < 3 * C; i += GN1_THREADS) {
        float t = 0.f;
        for (int r = 0; r < m.rpi; ++r) t += part[r * 3 * C + i];
        dst[i] = t;
    }
}

// out[col] = sum over rows of chan[row][col]; 32 columns x 8 row lanes per CTA, fixed order
__global__ void __launch_bounds__(256) gn_cols_reduce_kernel(const float* __restrict__ chan, float* __restrict__ out,
                                                             int rows, int ncols) {
    __shared__ float sh[8][33];
    const int col = blockIdx.x * 32 + (threadIdx.x & 31);
    const int lane = threadIdx.x >> 5;
    float acc = 0.f;
    if (col < ncols)
        for (int r = lane; r < rows; r += 8) acc += chan[(long long)r * ncols + col];
    sh[lane][threadIdx.x & 31] = acc;
    __syncthreads();
    if (lane == 0 && col < ncols) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x];
        out[col] = t;
    }
}

static int gn1_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}
// (sample, row chunk) items of ~elems elements; rows per chunk a multiple of rows-per-iteration
static void gn1_items(int HW, int C, long long elems, int& chunks, int& rows_per) {
    const int rpi = GN1_THREADS / (C / 8);
    long long rows = elems / C;
    if (rows < 4LL * rpi) rows = 4LL * rpi;
    if (rows > HW) rows = HW;
    chunks = (int)((HW + rows - 1) / rows);
    rows_per = (HW + chunks - 1) / chunks;
    rows_per = (rows_per + rpi - 1) / rpi * rpi;
    chunks = (HW + rows_per - 1) / rows_per;
}
bool gn_pre_ok(int C, int G) {
    if (G < 1 || G > MAX_G || C % G != 0 || C % 8 != 0 || C / 8 > GN1_THREADS || GN1_THREADS % (C / 8) != 0) return false;
    const int cpg = C / G;
    return cpg % 8 == 0 || cpg == 4 || cpg == 2;
}
int gn_bwd_pre_grid(int B, int HW, int C) {
    int chunks, rows_per;
    gn1_items(HW, C, 32768, chunks, rows_per);
    long long items = (long long)B * chunks;
    long long grid = 2LL * gn1_sms();
    return (int)(items < grid ? items : grid);
}

int gn_apply_pre(const void* x, const float* gamma, const float* beta, void* y, const float* sums, float* stats, int B,
                 int HW, int C, int G, float eps, int silu, int dtype, cudaStream_t st) {
    MVAE_CHECK(gn_pre_ok(C, G), "groupnorm_apply_fwd: unsupported C=%d G=%d", C, G);
    int chunks, rows_per;
    gn1_items(HW, C, 32768, chunks, rows_per);
    long long items = (long long)B * chunks;
    long long grid = 3LL * gn1_sms();
    if (items < grid) grid = items;
    if (dtype == MVAE_F32)
        gn_apply_pre_kernel<float><<<(unsigned)grid, GN1_THREADS, 0, st>>>((const float*)x, gamma, beta, (float*)y, sums,
                                                                            stats, B, HW, C, G, chunks, rows_per, eps, silu);
    else
        gn_apply_pre_kernel<bf16><<<(unsigned)grid, GN1_THREADS, 0, st>>>((const bf16*)x, gamma, beta, (bf16*)y, sums,
                                                                           stats, B, HW, C, G, chunks, rows_per, eps, silu);
    MVAE_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int gn_bwd_pre_t(const T* du, const T* x, const float* gamma, const float* stats, const float* s12,
                        const T* dx_add, T* dx, float* out3, float* scratch, int B, int HW, int C, int G,
                        cudaStream_t st) {
    int chunks, rows_per;
    gn1_items(HW, C, 32768, chunks, rows_per);
    const int grid = gn_bwd_pre_grid(B, HW, C);
    const int rpi = GN1_THREADS / (C / 8);
    const size_t sh = sizeof(float) * (size_t)rpi * 3 * C;
    const int cpg = C / G;
    if (cpg % 8 == 0)
        gn_bwd_pre_kernel<T, 1><<<grid, GN1_THREADS, sh, st>>>(du, x, gamma, stats, s12, dx_add, dx, scratch, B, HW, C, G,
                                                               chunks, rows_per);
    else if (cpg == 4)
        gn_bwd_pre_kernel<T, 2><<<grid, GN1_THREADS, sh, st>>>(du, x, gamma, stats, s12, dx_add, dx, scratch, B, HW, C, G,
                                                               chunks, rows_per);
    else
        gn_bwd_pre_kernel<T, 4><<<grid, GN1_THREADS, sh, st>>>(du, x, gamma, stats, s12, dx_add, dx, scratch, B, HW, C, G,
                                                               chunks, rows_per);
    MVAE_LAUNCH_CHECK();
    gn_cols_reduce_kernel<<<(3 * C + 31) / 32, 256, 0, st>>>(scratch, out3, grid, 3 * C);
    MVAE_LAUNCH_CHECK();
    return 0;
}
// out3: [3][C] fp32 = dgamma, dbeta, column sums of dx.  scratch: gn_bwd_pre_grid() * 3 * C floats.
int gn_bwd_pre(const void* du, const void* x, const float* gamma, const float* stats, const float* s12,
               const void* dx_add, void* dx, float* out3, float* scratch, int B, int HW, int C, int G, int dtype,
               cudaStream_t st) {
    MVAE_CHECK(gn_pre_ok(C, G), "groupnorm_apply_bwd: unsupported C=%d G=%d", C, G);
    if (dtype == MVAE_F32)
        return gn_bwd_pre_t<float>((const float*)du, (const float*)x, gamma, stats, s12, (const float*)dx_add, (float*)dx,
                                   out3, scratch, B, HW, C, G, st);
    return gn_bwd_pre_t<bf16>((const bf16*)du, (const bf16*)x, gamma, stats, s12, (const bf16*)dx_add, (bf16*)dx, out3,
                              scratch, B, HW, C, G, st);
}

// dynamic shared memory of the stats kernels: [rows per iteration][C][vals] floats
template <int VEC>
static size_t gn_part_bytes(int C, int vals) {
    int nvec = C / VEC;
    int tpr = nvec < GN_THREADS ? nvec : GN_THREADS;
    int rpi = GN_THREADS / tpr;
    return sizeof(float) * (size_t)rpi * C * vals;
}

template <typename T, int VEC>
static int gn_fwd_t(const T* x, const float* gamma, const float* beta, T* y, float* stats, float* scratch, int B,
                    int HW, int C, int G, float eps, int silu, cudaStream_t st) {
    const int nb = gn_chunk(B, HW, C, 1);
    const int S = gn_splits_for(nb, HW, C);
    for (int n0 = 0; n0 < B; n0 += nb) {
        const int n = (B - n0 < nb) ? (B - n0) : nb;
        const long long off = (long long)n0 * HW * C;
        dim3 grid(S, n);
        gn_stats_kernel<T, VEC><<<grid, GN_THREADS, gn_part_bytes<VEC>(C, 2), st>>>(x + off, scratch, HW, C, G, S);
        MVAE_LAUNCH_CHECK();
        gn_apply_kernel<T, VEC><<<grid, GN_THREADS, 0, st>>>(x + off, gamma, beta, y + off, stats + (long long)n0 * G * 2,
                                                              scratch, HW, C, G, S, eps, silu);
        MVAE_LAUNCH_CHECK();
    }
    return 0;
}
template <typename T, int VEC>
static int gn_bwd_t(const T* dy, const T* x, const float* gamma, const float* beta, const float* stats,
                    const T* dx_add, T* dx, float* dgamma, float* dbeta, float* scratch, int B, int HW, int C, int G,
                    int silu, cudaStream_t st) {
    const int nb = gn_chunk(B, HW, C, 2);
    const int S = gn_splits_for(nb, HW, C);
    float* chan = scratch + (long long)nb * S * G * 2;
    long long rows = 0;
    for (int n0 = 0; n0 < B; n0 += nb) {
        const int n = (B - n0 < nb) ? (B - n0) : nb;
        const long long off = (long long)n0 * HW * C;
        const float* st_n = stats + (long long)n0 * G * 2;
        dim3 grid(S, n);
        gn_bwd_stats_kernel<T, VEC><<<grid, GN_THREADS, gn_part_bytes<VEC>(C, 2), st>>>(
            dy + off, x + off, gamma, beta, st_n, scratch, chan + rows * 2 * C, HW, C, G, S, silu);
        MVAE_LAUNCH_CHECK();
        gn_bwd_apply_kernel<T, VEC><<<grid, GN_THREADS, 0, st>>>(dy + off, x + off, gamma, beta, st_n, scratch,
                                                                  dx_add ? dx_add + off : nullptr, dx + off, HW, C, G,
                                                                  S, silu);
        MVAE_LAUNCH_CHECK();
        rows += (long long)n * S;
    }
    gn_param_reduce_kernel<<<(2 * C + 31) / 32, 256, 0, st>>>(chan, dgamma, dbeta, (int)rows, C);
    MVAE_LAUNCH_CHECK();
    return 0;
}

int gn_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, float* scratch, int B, int HW,
           int C, int G, float eps, int silu, int dtype, cudaStream_t st) {
    MVAE_CHECK(G >= 1 && G <= MAX_G && C % G == 0, "groupnorm: bad groups %d for C=%d", G, C);
    const bool vec8 = (C % 8 == 0) && (C / 8 <= GN_THREADS || (C / 8) % GN_THREADS == 0);
    MVAE_CHECK(vec8 || C <= GN_THREADS, "groupnorm: unsupported channel count %d", C);
    if (dtype == MVAE_F32)
        return vec8 ? gn_fwd_t<float, 8>((const float*)x, gamma, beta, (float*)y, stats, scratch, B, HW, C, G, eps, silu, st)
                    : gn_fwd_t<float, 1>((const float*)x, gamma, beta, (float*)y, stats, scratch, B, HW, C, G, eps, silu, st);
    return vec8 ? gn_fwd_t<bf16, 8>((const bf16*)x, gamma, beta, (bf16*)y, stats, scratch, B, HW, C, G, eps, silu, st)
                : gn_fwd_t<bf16, 1>((const bf16*)x, gamma, beta, (bf16*)y, stats, scratch, B, HW, C, G, eps, silu, st);
}
int gn_bwd(const void* dy, const void* x, const float* gamma, const float* beta, const float* stats,
           const void* dx_add, void* dx, float* dgamma, float* dbeta, float* scratch, int B, int HW, int C, int G,
           int silu, int dtype, cudaStream_t st) {
    MVAE_CHECK(G >= 1 && G <= MAX_G && C % G == 0, "groupnorm: bad groups %d for C=%d", G, C);
    const bool vec8 = (C % 8 == 0) && (C / 8 <= GN_THREADS || (C / 8) % GN_THREADS == 0);
    MVAE_CHECK(vec8 || C <= GN_THREADS, "groupnorm: unsupported channel count %d", C);
    if (dtype == MVAE_F32)
        return vec8 ? gn_bwd_t<float, 8>((const float*)dy, (const float*)x, gamma, beta, stats, (const float*)dx_add,
                                         (float*)dx, dgamma, dbeta, scratch, B, HW, C, G, silu, st)
                    : gn_bwd_t<float, 1>((const float*)dy, (const float*)x, gamma, beta, stats, (const float*)dx_add,
                                         (float*)dx, dgamma, dbeta, scratch, B, HW, C, G, silu, st);
    return vec8 ? gn_bwd_t<bf16, 8>((const bf16*)dy, (const bf16*)x, gamma, beta, stats, (const bf16*)dx_add, (bf16*)dx,
                                    dgamma, dbeta, scratch, B, HW, C, G, silu, st)
                : gn_bwd_t<bf16, 1>((const bf16*)dy, (const bf16*)x, gamma, beta, stats, (const bf16*)dx_add, (bf16*)dx,
                                    dgamma, dbeta, scratch, B, HW, C, G, silu, st);
}

}  // namespace mvae
