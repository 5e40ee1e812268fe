// Callers either side of the hot path:
//  * batch feed: what MedMNISTDataset.__getitem__ + Normalize + mixed_modality_collate_fn do per sample
//    on the host (src/data/medmnist_data.py:186-251, 372-380, 16-72), as one HBM-bound kernel over a
//    staged uint8 batch: ToTensor (u8/255), RGB->gray (0.299, 0.587, 0.114) or gray->RGB (repeat) to the
//    modality's channel count, (v-0.5)/0.5, zero-pad to the batch's channel count, one-hot modality.
//    The arithmetic uses the same fp32 operations in the same order, so the result is bit-identical.
//  * dropout of ResnetBlock (encoder_decoder.py:164, nn.Dropout between swish(norm2(h)) and conv2):
//    counter-based Philox4x32-10 keep mask, regenerated in the backward instead of stored.
//  * FiLMLayer (conditional_vae.py:191-203): features * (W_s c + b_s) + (W_t c + b_t), fwd + bwd.
#include "../../include/medvae_b200.h"
#include "common.cuh"
#include "philox.cuh"

namespace mvae {

// ------------------------------------------------------------------ batch feed
__global__ void __launch_bounds__(256) batch_feed_u8_kernel(const uint8_t* __restrict__ src, const int* __restrict__ src_c,
                                                            const int* __restrict__ tgt_c, const long long* __restrict__ idx,
                                                            float* __restrict__ x, float* __restrict__ onehot, int B, int HW,
                                                            int SC, int OC, int n_mod) {
    const long long total = (long long)B * HW;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / HW), p = (int)(i % HW);
        const uint8_t* s = src + i * SC;
        const int cur = src_c[b] < SC ? src_c[b] : SC, tgt = tgt_c[b];
        float v[3];
        // ToTensor: uint8 -> float32 / 255 (IEEE division, as torch's div)
        const float c0 = __fdiv_rn((float)s[0], 255.f);
        if (cur >= 3) {
            const float c1 = __fdiv_rn((float)s[1], 255.f), c2 = __fdiv_rn((float)s[2], 255.f);
            if (tgt == 1) {   // 0.299*R + 0.587*G + 0.114*B, each product and sum rounded (no FMA contraction)
                v[0] = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, c0), __fmul_rn(0.587f, c1)), __fmul_rn(0.114f, c2));
            } else {
                v[0] = c0; v[1] = c1; v[2] = c2;
            }
        } else {
            v[0] = c0; v[1] = c0; v[2] = c0;   // image.repeat(3, 1, 1) when the modality wants RGB
        }
        float* o = x + (long long)b * OC * HW + p;
        for (int c = 0; c < OC; ++c) {
            // Normalize(mean 0.5, std 0.5): sub then div; channels beyond the modality's own are the collate's zero padding
            o[(long long)c * HW] = c < tgt ? __fdiv_rn(__fsub_rn(v[c], 0.5f), 0.5f) : 0.f;
        }
    }
    if (onehot) {   // modality one-hot (medmnist_data.py:243-245)
        const long long n1 = (long long)B * n_mod;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n1; i += (long long)gridDim.x * blockDim.x)
            onehot[i] = (idx[i / n_mod] == i % n_mod) ? 1.f : 0.f;
    }
}

int batch_feed_u8(const uint8_t* src, const int* src_c, const int* tgt_c, const long long* idx, float* x, float* onehot, int B,
                  int H, int W, int SC, int OC, int n_mod, cudaStream_t st) {
    const long long total = (long long)B * H * W;
    long long g = (total + 255) / 256;
    const int grid = (int)(g > 148 * 16 ? 148 * 16 : g);
    batch_feed_u8_kernel<<<grid, 256, 0, st>>>(src, src_c, tgt_c, idx, x, onehot, B, H * W, SC, OC, n_mod);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ dropout
// y[e] = x[e] * keep(e) / (1 - p); keep(e) = u(e) >= p with u from word (g & 3) of Philox(seed; counter g / 4, offset),
// g = elem_base + e the element's index in the GLOBAL batch (elem_base = first sample x elements per sample: a sample's
// mask does not depend on the rank count or the micro-batch split), or mask_in[e].  The backward is the same call on dy.
// 4 elements per thread (one Philox block, two when elem_base is not a multiple of 4).
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, float p,
                                                      uint64_t seed, uint64_t offset, uint64_t elem_base,
                                                      const double* __restrict__ dyn,
                                                      const uint8_t* __restrict__ mask_in, uint8_t* __restrict__ mask_out) {
    if (dyn) offset += ((uint64_t)dyn[2]) << 20;   // device-side stream position (CUDA-graph replays advance it)
    const float scale = 1.f / (1.f - p);
    const Philox ph(seed);
    const long long quads = (n + 3) / 4;
    const int shift = (int)(elem_base & 3);
    for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < quads; q += (long long)gridDim.x * blockDim.x) {
        uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        if (!mask_in) {
            const uint64_t blk = (elem_base + (uint64_t)q * 4) >> 2;
            const PhiloxWords r = ph(blk, offset);
            w[0] = r.x; w[1] = r.y; w[2] = r.z; w[3] = r.w;
            if (shift) {
                const PhiloxWords r2 = ph(blk + 1, offset);
                w[4] = r2.x; w[5] = r2.y; w[6] = r2.z; w[7] = r2.w;
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long e = q * 4 + j;
            if (e >= n) break;
            bool keep;
            if (mask_in) keep = mask_in[e] != 0;
            else keep = philox_u01_grid(w[shift + j]) >= p;
            if (mask_out) mask_out[e] = keep ? 1 : 0;
            y[e] = keep ? from_f<T>(to_f<T>(x[e]) * scale) : from_f<T>(0.f);
        }
    }
}

int dropout_apply(const void* x, void* y, long long n, float p, uint64_t seed, uint64_t offset, uint64_t elem_base,
                  const double* dyn, const uint8_t* mask_in, uint8_t* mask_out, int dtype, cudaStream_t st) {
    long long g = ((n + 3) / 4 + 255) / 256;
    const int grid = (int)(g > 148 * 16 ? 148 * 16 : (g < 1 ? 1 : g));
    if (dtype == MVAE_BF16)
        dropout_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x, (bf16*)y, n, p, seed, offset, elem_base, dyn, mask_in,
                                                   mask_out);
    else
        dropout_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, n, p, seed, offset, elem_base, dyn, mask_in,
                                                    mask_out);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ FiLM
// one CTA per (b, c) plane: scale/shift from the K-wide condition, then the plane.  st[0][b][c] = scale, st[1][b][c] = shift
__global__ void __launch_bounds__(256) film_fwd_kernel(const float* __restrict__ f, const float* __restrict__ cond,
                                                       const float* __restrict__ ws, const float* __restrict__ bs,
                                                       const float* __restrict__ wt, const float* __restrict__ bt,
                                                       float* __restrict__ out, float* __restrict__ st, int B, int C, int HW, int K) {
    const int bc = blockIdx.x, b = bc / C, c = bc % C;
    float s = bs[c], t = bt[c];
    for (int k = 0; k < K; ++k) {
        const float cv = cond[(long long)b * K + k];
        s = fmaf(ws[(long long)c * K + k], cv, s);
        t = fmaf(wt[(long long)c * K + k], cv, t);
    }
    if (threadIdx.x == 0) {
        st[bc] = s;
        st[(long long)B * C + bc] = t;
    }
    const float* fp = f + (long long)bc * HW;
    float* op = out + (long long)bc * HW;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) op[i] = __fadd_rn(__fmul_rn(fp[i], s), t);
}

// d_f = dy * scale; dst[0][b][c] = sum dy * f, dst[1][b][c] = sum dy
__global__ void __launch_bounds__(256) film_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ f,
                                                       const float* __restrict__ st, float* __restrict__ df,
                                                       float* __restrict__ dst, int B, int C, int HW) {
    __shared__ float sh[32];
    const int bc = blockIdx.x;
    const float s = st[bc];
    const float* dp = dy + (long long)bc * HW;
    const float* fp = f + (long long)bc * HW;
    float* op = df + (long long)bc * HW;
    float a = 0.f, b = 0.f;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) {
        const float d = dp[i];
        op[i] = d * s;
        a = fmaf(d, fp[i], a);
        b += d;
    }
    a = block_sum(a, sh);
    b = block_sum(b, sh);
    if (threadIdx.x == 0) {
        dst[bc] = a;
        dst[(long long)B * C + bc] = b;
    }
}

// parameter / condition gradients from dst [2][B][C]: dWs[c][k] = sum_b ds[b][c] cond[b][k] (same for Wt with dt),
// dbs[c] = sum_b ds[b][c], dbt likewise, d_cond[b][k] = sum_c ds[b][c] Ws[c][k] + dt[b][c] Wt[c][k]
__global__ void __launch_bounds__(256) film_param_grads_kernel(const float* __restrict__ dst, const float* __restrict__ cond,
                                                               const float* __restrict__ ws, const float* __restrict__ wt,
                                                               float* __restrict__ dws, float* __restrict__ dbs,
                                                               float* __restrict__ dwt, float* __restrict__ dbt,
                                                               float* __restrict__ dcond, int B, int C, int K) {
    const float* ds = dst;
    const float* dt = dst + (long long)B * C;
    const long long n_w = (long long)C * K, n_c = (long long)B * K, total = n_w + C + n_c;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        if (i < n_w) {
            const int c = (int)(i / K), k = (int)(i % K);
            float a = 0.f, b = 0.f;
            for (int bb = 0; bb < B; ++bb) {
                const float cv = cond[(long long)bb * K + k];
                a = fmaf(ds[(long long)bb * C + c], cv, a);
                b = fmaf(dt[(long long)bb * C + c], cv, b);
            }
            dws[i] = a; dwt[i] = b;
        } else if (i < n_w + C) {
            const int c = (int)(i - n_w);
            float a = 0.f, b = 0.f;
            for (int bb = 0; bb < B; ++bb) { a += ds[(long long)bb * C + c]; b += dt[(long long)bb * C + c]; }
            dbs[c] = a; dbt[c] = b;
        } else if (dcond) {
            const long long j = i - n_w - C;
            const int bb = (int)(j / K), k = (int)(j % K);
            float a = 0.f;
            for (int c = 0; c < C; ++c)
                a = fmaf(ds[(long long)bb * C + c], ws[(long long)c * K + k], fmaf(dt[(long long)bb * C + c], wt[(long long)c * K + k], a));
            dcond[j] = a;
        }
    }
}

int film_fwd(const float* f, const float* cond, const float* ws, const float* bs, const float* wt, const float* bt, float* out,
             float* st, int B, int C, int HW, int K, cudaStream_t s) {
    film_fwd_kernel<<<B * C, 256, 0, s>>>(f, cond, ws, bs, wt, bt, out, st, B, C, HW, K);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int film_bwd(const float* dy, const float* f, const float* cond, const float* ws, const float* wt, const float* st, float* df,
             float* dst, float* dws, float* dbs, float* dwt, float* dbt, float* dcond, int B, int C, int HW, int K,
             cudaStream_t s) {
    film_bwd_kernel<<<B * C, 256, 0, s>>>(dy, f, st, df, dst, B, C, HW);
    MVAE_LAUNCH_CHECK();
    const long long total = (long long)C * K + C + (long long)B * K;
    film_param_grads_kernel<<<(int)((total + 255) / 256), 256, 0, s>>>(dst, cond, ws, wt, dws, dbs, dwt, dbt, dcond, B, C, K);
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
