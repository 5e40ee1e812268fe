// AttnBlock core (encoder_decoder.py:91-103): per image, single head of width C,
//   S = q k^T * C^-0.5 ; P = softmax_j(S) ; o_i = sum_j P_ij v_j
// and its backward.  Generic path: strided batched SIMT GEMMs + row softmax kernels with the
// probabilities kept in fp32 (exact enough for the fp32 verification mode).
#include "../../include/medvae_b200.h"
#include "common.cuh"

namespace mvae {

// C[b](m,n) = alpha * sum_k A[b](m,k) * B[b](k,n); fully strided operands (element strides)
struct BgemmArgs {
    int M, N, K, batch;
    long long a_b, a_m, a_k;
    long long b_b, b_k, b_n;
    long long c_b, c_m;     // c_n == 1
    float alpha;
};

template <typename TA, typename TB, typename TC>
__global__ void __launch_bounds__(256) bgemm_simt_kernel(const TA* __restrict__ A, const TB* __restrict__ Bm,
                                                         TC* __restrict__ Cm, const BgemmArgs g) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int b = blockIdx.z;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const TA* Ab = A + (long long)b * g.a_b;
    const TB* Bb = Bm + (long long)b * g.b_b;
    const int tid = threadIdx.x, ty = tid / 16, tx = tid % 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    // loader mapping: pick the fast-varying smem index along the operand's contiguous dimension
    const bool a_kfast = g.a_k == 1, b_kfast = g.b_k == 1;
    for (int k0 = 0; k0 < g.K; k0 += BK) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int e = tid + j * 256;   // 0..1023
            int am, ak;
            if (a_kfast) { ak = e % BK; am = e / BK; } else { am = e % BM; ak = e / BM; }
            float av = 0.f;
            if (m0 + am < g.M && k0 + ak < g.K) av = to_f<TA>(Ab[(long long)(m0 + am) * g.a_m + (long long)(k0 + ak) * g.a_k]);
            As[ak][am] = av;
            int bn, bk;
            if (b_kfast) { bk = e % BK; bn = e / BK; } else { bn = e % BN; bk = e / BN; }
            float bv = 0.f;
            if (n0 + bn < g.N && k0 + bk < g.K) bv = to_f<TB>(Bb[(long long)(k0 + bk) * g.b_k + (long long)(n0 + bn) * g.b_n]);
            Bs[bk][bn] = bv;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    TC* Cb = Cm + (long long)b * g.c_b;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int m = m0 + ty * 4 + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int n = n0 + tx * 4 + j;
            if (n < g.N) Cb[(long long)m * g.c_m + n] = from_f<TC>(acc[i][j] * g.alpha);
        }
    }
}

template <typename TA, typename TB, typename TC>
static int bgemm(const TA* A, const TB* B, TC* C, const BgemmArgs& g, cudaStream_t st) {
    dim3 grid((g.N + 63) / 64, (g.M + 63) / 64, g.batch);
    bgemm_simt_kernel<TA, TB, TC><<<grid, 256, 0, st>>>(A, B, C, g);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// in-place row softmax over [rows][T] fp32; one warp per row
__global__ void softmax_rows_kernel(float* __restrict__ s, long long rows, int T) {
    long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float* r = s + row * T;
    float mx = -INFINITY;
    for (int j = lane; j < T; j += 32) mx = fmaxf(mx, r[j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < T; j += 32) { float e = __expf(r[j] - mx); r[j] = e; sum += e; }
    sum = warp_sum(sum);
    float inv = 1.f / sum;
    for (int j = lane; j < T; j += 32) r[j] *= inv;
}
// dS = P * (dP - sum_j dP*P) * scale, in place on dP
__global__ void softmax_bwd_rows_kernel(float* __restrict__ dp, const float* __restrict__ p, long long rows, int T,
                                        float scale) {
    long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float* d = dp + row * T;
    const float* pr = p + row * T;
    float dot = 0.f;
    for (int j = lane; j < T; j += 32) dot = fmaf(d[j], pr[j], dot);
    dot = warp_sum(dot);
    for (int j = lane; j < T; j += 32) d[j] = pr[j] * (d[j] - dot) * scale;
}

template <typename T>
static int attn_fwd_t(const T* qkv, T* out, float* probs, int B, int Tn, int C, cudaStream_t st) {
    const float scale = 1.0f / sqrtf((float)C);   // int(c) ** (-0.5)
    BgemmArgs g;
    // S = scale * Q K^T
    g.M = Tn; g.N = Tn; g.K = C; g.batch = B;
    g.a_b = (long long)Tn * 3 * C; g.a_m = 3 * C; g.a_k = 1;
    g.b_b = (long long)Tn * 3 * C; g.b_k = 1; g.b_n = 3 * C;
    g.c_b = (long long)Tn * Tn; g.c_m = Tn; g.alpha = scale;
    if (bgemm<T, T, float>(qkv, qkv + C, probs, g, st)) return 1;
    long long rows = (long long)B * Tn;
    softmax_rows_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(probs, rows, Tn);
    MVAE_LAUNCH_CHECK();
    // O = P V
    g.M = Tn; g.N = C; g.K = Tn;
    g.a_b = (long long)Tn * Tn; g.a_m = Tn; g.a_k = 1;
    g.b_b = (long long)Tn * 3 * C; g.b_k = 3 * C; g.b_n = 1;
    g.c_b = (long long)Tn * C; g.c_m = C; g.alpha = 1.f;
    return bgemm<float, T, T>(probs, qkv + 2 * C, out, g, st);
}

template <typename T>
static int attn_bwd_t(const T* qkv, const float* probs, const T* dout, T* dqkv, float* ds, int B, int Tn, int C,
                      cudaStream_t st) {
    const float scale = 1.0f / sqrtf((float)C);
    const long long qs = (long long)Tn * 3 * C, ps = (long long)Tn * Tn, os = (long long)Tn * C;
    BgemmArgs g;
    g.batch = B;
    // dV = P^T dO          (m=j, k=i, n=c)
    g.M = Tn; g.N = C; g.K = Tn;
    g.a_b = ps; g.a_m = 1; g.a_k = Tn;
    g.b_b = os; g.b_k = C; g.b_n = 1;
    g.c_b = qs; g.c_m = 3 * C; g.alpha = 1.f;
    if (bgemm<float, T, T>(probs, dout, dqkv + 2 * C, g, st)) return 1;
    // dP = dO V^T          (m=i, k=c, n=j)
    g.M = Tn; g.N = Tn; g.K = C;
    g.a_b = os; g.a_m = C; g.a_k = 1;
    g.b_b = qs; g.b_k = 1; g.b_n = 3 * C;
    g.c_b = ps; g.c_m = Tn; g.alpha = 1.f;
    if (bgemm<T, T, float>(dout, qkv + 2 * C, ds, g, st)) return 1;
    long long rows = (long long)B * Tn;
    softmax_bwd_rows_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(ds, probs, rows, Tn, scale);
    MVAE_LAUNCH_CHECK();
    // dQ = dS K            (m=i, k=j, n=c)
    g.M = Tn; g.N = C; g.K = Tn;
    g.a_b = ps; g.a_m = Tn; g.a_k = 1;
    g.b_b = qs; g.b_k = 3 * C; g.b_n = 1;
    g.c_b = qs; g.c_m = 3 * C; g.alpha = 1.f;
    if (bgemm<float, T, T>(ds, qkv + C, dqkv, g, st)) return 1;
    // dK = dS^T Q          (m=j, k=i, n=c)
    g.a_b = ps; g.a_m = 1; g.a_k = Tn;
    return bgemm<float, T, T>(ds, qkv, dqkv + C, g, st);
}

// ------------------------------------------------------------------ LinearAttention (encoder_decoder.py:36-65)
// heads = 1, dim_head = C (LinAttnBlock): per image, on the fused NHWC qkv tensor [T][3C]
//   ks = softmax over the TOKENS of k (per channel) ; ctx[d][e] = sum_t ks[t][d] v[t][e] ; out[t][e] = sum_d q[t][d] ctx[d][e]
// No scale factor, no GroupNorm, no residual (the reference's block has none).  Generic SIMT path only: no
// reference config enables attn_type="linear".
template <typename T>
__global__ void __launch_bounds__(256) colsoftmax_kernel(const T* __restrict__ qkv, float* __restrict__ ks, int B, int Tn, int C) {
    const long long n = (long long)B * C;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / C), d = (int)(i % C);
        const T* k = qkv + (long long)b * Tn * 3 * C + C + d;
        float* o = ks + (long long)b * Tn * C + d;
        float mx = -INFINITY;
        for (int t = 0; t < Tn; ++t) mx = fmaxf(mx, to_f<T>(k[(long long)t * 3 * C]));
        float sum = 0.f;
        for (int t = 0; t < Tn; ++t) {
            float e = expf(to_f<T>(k[(long long)t * 3 * C]) - mx);
            o[(long long)t * C] = e;
            sum += e;
        }
        const float inv = 1.f / sum;
        for (int t = 0; t < Tn; ++t) o[(long long)t * C] *= inv;
    }
}
// dk[t][d] = ks[t][d] * (dks[t][d] - sum_t' dks[t'][d] ks[t'][d]) -> dqkv[:, C + d]
template <typename T>
__global__ void __launch_bounds__(256) colsoftmax_bwd_kernel(const float* __restrict__ dks, const float* __restrict__ ks,
                                                             T* __restrict__ dqkv, int B, int Tn, int C) {
    const long long n = (long long)B * C;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / C), d = (int)(i % C);
        const float* g = dks + (long long)b * Tn * C + d;
        const float* p = ks + (long long)b * Tn * C + d;
        T* o = dqkv + (long long)b * Tn * 3 * C + C + d;
        float dot = 0.f;
        for (int t = 0; t < Tn; ++t) dot = fmaf(g[(long long)t * C], p[(long long)t * C], dot);
        for (int t = 0; t < Tn; ++t) o[(long long)t * 3 * C] = from_f<T>(p[(long long)t * C] * (g[(long long)t * C] - dot));
    }
}
static int col_grid(long long n) {
    long long g = (n + 255) / 256;
    return (int)(g < 1 ? 1 : (g > 148 * 8 ? 148 * 8 : g));
}

// ks: fp32 [B][T][C], ctx: fp32 [B][C][C] (both kept for the backward)
template <typename T>
static int linattn_fwd_t(const T* qkv, T* out, float* ks, float* ctx, int B, int Tn, int C, cudaStream_t st) {
    const long long qs = (long long)Tn * 3 * C, ts = (long long)Tn * C, cs = (long long)C * C;
    colsoftmax_kernel<T><<<col_grid((long long)B * C), 256, 0, st>>>(qkv, ks, B, Tn, C);
    MVAE_LAUNCH_CHECK();
    BgemmArgs g;
    g.batch = B; g.alpha = 1.f;
    // ctx[d][e] = sum_t ks[t][d] v[t][e]
    g.M = C; g.N = C; g.K = Tn;
    g.a_b = ts; g.a_m = 1; g.a_k = C;
    g.b_b = qs; g.b_k = 3 * C; g.b_n = 1;
    g.c_b = cs; g.c_m = C;
    if (bgemm<float, T, float>(ks, qkv + 2 * C, ctx, g, st)) return 1;
    // out[t][e] = sum_d q[t][d] ctx[d][e]
    g.M = Tn; g.N = C; g.K = C;
    g.a_b = qs; g.a_m = 3 * C; g.a_k = 1;
    g.b_b = cs; g.b_k = C; g.b_n = 1;
    g.c_b = ts; g.c_m = C;
    return bgemm<T, float, T>(qkv, ctx, out, g, st);
}

// dctx: fp32 [B][C][C] scratch, dks: fp32 [B][T][C] scratch
template <typename T>
static int linattn_bwd_t(const T* qkv, const float* ks, const float* ctx, const T* dout, T* dqkv, float* dctx, float* dks,
                         int B, int Tn, int C, cudaStream_t st) {
    const long long qs = (long long)Tn * 3 * C, ts = (long long)Tn * C, cs = (long long)C * C;
    BgemmArgs g;
    g.batch = B; g.alpha = 1.f;
    // dq[t][d] = sum_e dout[t][e] ctx[d][e]
    g.M = Tn; g.N = C; g.K = C;
    g.a_b = ts; g.a_m = C; g.a_k = 1;
    g.b_b = cs; g.b_k = 1; g.b_n = C;
    g.c_b = qs; g.c_m = 3 * C;
    if (bgemm<T, float, T>(dout, ctx, dqkv, g, st)) return 1;
    // dctx[d][e] = sum_t q[t][d] dout[t][e]
    g.M = C; g.N = C; g.K = Tn;
    g.a_b = qs; g.a_m = 1; g.a_k = 3 * C;
    g.b_b = ts; g.b_k = C; g.b_n = 1;
    g.c_b = cs; g.c_m = C;
    if (bgemm<T, T, float>(qkv, dout, dctx, g, st)) return 1;
    // dv[t][e] = sum_d ks[t][d] dctx[d][e]
    g.M = Tn; g.N = C; g.K = C;
    g.a_b = ts; g.a_m = C; g.a_k = 1;
    g.b_b = cs; g.b_k = C; g.b_n = 1;
    g.c_b = qs; g.c_m = 3 * C;
    if (bgemm<float, float, T>(ks, dctx, dqkv + 2 * C, g, st)) return 1;
    // dks[t][d] = sum_e v[t][e] dctx[d][e]
    g.a_b = qs; g.a_m = 3 * C; g.a_k = 1;
    g.b_b = cs; g.b_k = 1; g.b_n = C;
    g.c_b = ts; g.c_m = C;
    if (bgemm<T, float, float>(qkv + 2 * C, dctx, dks, g, st)) return 1;
    colsoftmax_bwd_kernel<T><<<col_grid((long long)B * C), 256, 0, st>>>(dks, ks, dqkv, B, Tn, C);
    MVAE_LAUNCH_CHECK();
    return 0;
}

int linattn_fwd(const void* qkv, void* out, float* ks, float* ctx, int B, int Tn, int C, int dtype, cudaStream_t st) {
    if (dtype == MVAE_F32) return linattn_fwd_t<float>((const float*)qkv, (float*)out, ks, ctx, B, Tn, C, st);
    return linattn_fwd_t<bf16>((const bf16*)qkv, (bf16*)out, ks, ctx, B, Tn, C, st);
}
int linattn_bwd(const void* qkv, const float* ks, const float* ctx, const void* dout, void* dqkv, float* dctx, float* dks,
                int B, int Tn, int C, int dtype, cudaStream_t st) {
    if (dtype == MVAE_F32)
        return linattn_bwd_t<float>((const float*)qkv, ks, ctx, (const float*)dout, (float*)dqkv, dctx, dks, B, Tn, C, st);
    return linattn_bwd_t<bf16>((const bf16*)qkv, ks, ctx, (const bf16*)dout, (bf16*)dqkv, dctx, dks, B, Tn, C, st);
}

// attn_umma.cu: tensor-core path (bf16, C % 64 == 0, T <= 256)
bool attn_umma_ok(int Tn, int C);
int attn_fwd_umma(const bf16* qkv, bf16* out, void* probs, int B, int Tn, int C, cudaStream_t st);
int attn_bwd_umma(const bf16* qkv, const void* probs, const bf16* dout, bf16* dqkv, void* ds, int B, int Tn, int C,
                  cudaStream_t st);

int attn_fwd(const void* qkv, void* out, float* probs, int B, int Tn, int C, int dtype, cudaStream_t st) {
    if (dtype == MVAE_BF16 && attn_umma_ok(Tn, C) && debug_knob(2) == 0)
        return attn_fwd_umma((const bf16*)qkv, (bf16*)out, probs, B, Tn, C, st);
    if (dtype == MVAE_F32) return attn_fwd_t<float>((const float*)qkv, (float*)out, probs, B, Tn, C, st);
    return attn_fwd_t<bf16>((const bf16*)qkv, (bf16*)out, probs, B, Tn, C, st);
}
int attn_bwd(const void* qkv, const float* probs, const void* dout, void* dqkv, float* ds, int B, int Tn, int C,
             int dtype, cudaStream_t st) {
    if (dtype == MVAE_BF16 && attn_umma_ok(Tn, C) && debug_knob(2) == 0)
        return attn_bwd_umma((const bf16*)qkv, probs, (const bf16*)dout, (bf16*)dqkv, ds, B, Tn, C, st);
    if (dtype == MVAE_F32)
        return attn_bwd_t<float>((const float*)qkv, probs, (const float*)dout, (float*)dqkv, ds, B, Tn, C, st);
    return attn_bwd_t<bf16>((const bf16*)qkv, probs, (const bf16*)dout, (bf16*)dqkv, ds, B, Tn, C, st);
}

}  // namespace mvae
