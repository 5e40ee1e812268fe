// Validation / metric step: MSE, MAE, PSNR and SSIM of (original, reconstruction)
// in ONE pass over the two images, and the KL / latent statistics of src/utils/metrics.py, each with
// a fixed-order two-stage reduction (bitwise reproducible) and device-resident results: the caller
// reads one small float vector instead of the reference's one `.item()` per metric
// (src/utils/metrics.py:14-102, src/lightning_module.py:248-257).
//
// SSIM follows torchmetrics' structural_similarity_index_measure with the arguments the reference
// passes (gaussian window 11x11, sigma 1.5, k1 0.01, k2 0.03, data_range 1.0): the SSIM map is
// cropped by 5 on every side before it is averaged, so only windows that lie fully inside the image
// count and the reflect padding never reaches the result.  The window is separable: a horizontal
// and a vertical 11-tap pass over x, r, x^2, r^2, x*r held in shared memory.
#include "../../include/medvae_b200.h"
#include "common.cuh"

namespace mvae {

constexpr int SSIM_K = 11, SSIM_PAD = 5, MT = 32, MTH = MT + 2 * SSIM_PAD;   // 32x32 pixel tile, 42x42 with halo

struct GaussTaps {
    float w[SSIM_K];
};

// grid: (tiles_x * tiles_y, planes); partial[(plane * tiles + tile) * 3 + {0,1,2}] = sum sq err, sum abs err, sum ssim
__global__ void __launch_bounds__(256) recon_metrics_kernel(const float* __restrict__ x, const float* __restrict__ r,
                                                            double* __restrict__ partial, int H, int W, int tiles_x,
                                                            GaussTaps g, float c1, float c2) {
    __shared__ float sx[MTH][MTH + 1], sr[MTH][MTH + 1];
    __shared__ float hx[5][MTH][MT + 1];
    __shared__ double shd[32];
    const int tile = blockIdx.x, plane = blockIdx.y;
    const int ty0 = (tile / tiles_x) * MT, tx0 = (tile % tiles_x) * MT;
    const float* px = x + (long long)plane * H * W;
    const float* pr = r + (long long)plane * H * W;
    // tile + halo, zero outside the image (those entries only feed SSIM positions that are cropped away)
    for (int i = threadIdx.x; i < MTH * MTH; i += blockDim.x) {
        int ly = i / MTH, lx = i % MTH, gy = ty0 + ly - SSIM_PAD, gx = tx0 + lx - SSIM_PAD;
        bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        sx[ly][lx] = in ? px[(long long)gy * W + gx] : 0.f;
        sr[ly][lx] = in ? pr[(long long)gy * W + gx] : 0.f;
    }
    __syncthreads();
    double sse = 0.0, sae = 0.0, ssim = 0.0;
    for (int i = threadIdx.x; i < MT * MT; i += blockDim.x) {
        int ly = i / MT, lx = i % MT;
        if (ty0 + ly < H && tx0 + lx < W) {
            float d = sr[ly + SSIM_PAD][lx + SSIM_PAD] - sx[ly + SSIM_PAD][lx + SSIM_PAD];
            sse += (double)(d * d);
            sae += (double)fabsf(d);
        }
    }
    // horizontal pass: rows of the halo tile, the MT centre columns
    for (int i = threadIdx.x; i < MTH * MT; i += blockDim.x) {
        int ly = i / MT, lx = i % MT;
        float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
        for (int k = 0; k < SSIM_K; ++k) {
            float xv = sx[ly][lx + k], rv = sr[ly][lx + k], w = g.w[k];
            a = fmaf(w, xv, a);
            b = fmaf(w, rv, b);
            aa = fmaf(w, xv * xv, aa);
            bb = fmaf(w, rv * rv, bb);
            ab = fmaf(w, xv * rv, ab);
        }
        hx[0][ly][lx] = a; hx[1][ly][lx] = b; hx[2][ly][lx] = aa; hx[3][ly][lx] = bb; hx[4][ly][lx] = ab;
    }
    __syncthreads();
    // vertical pass + SSIM map on positions whose window lies inside the image
    for (int i = threadIdx.x; i < MT * MT; i += blockDim.x) {
        int ly = i / MT, lx = i % MT, gy = ty0 + ly, gx = tx0 + lx;
        if (gy < SSIM_PAD || gy >= H - SSIM_PAD || gx < SSIM_PAD || gx >= W - SSIM_PAD) continue;
        float m[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < SSIM_K; ++k) {
            float w = g.w[k];
#pragma unroll
            for (int q = 0; q < 5; ++q) m[q] = fmaf(w, hx[q][ly + k][lx], m[q]);
        }
        float mu_x = m[0], mu_r = m[1];
        float var_x = fmaxf(m[2] - mu_x * mu_x, 0.f), var_r = fmaxf(m[3] - mu_r * mu_r, 0.f);
        float cov = m[4] - mu_x * mu_r;
        float s = ((2.f * mu_x * mu_r + c1) * (2.f * cov + c2)) / ((mu_x * mu_x + mu_r * mu_r + c1) * (var_x + var_r + c2));
        ssim += (double)s;
    }
    sse = block_sum_d(sse, shd);
    sae = block_sum_d(sae, shd);
    ssim = block_sum_d(ssim, shd);
    if (threadIdx.x == 0) {
        double* p = partial + ((long long)plane * gridDim.x + tile) * 3;
        p[0] = sse; p[1] = sae; p[2] = ssim;
    }
}

// out[0..3] = mse, mae, psnr, ssim
__global__ void __launch_bounds__(256) recon_metrics_final_kernel(const double* __restrict__ partial, long long n_partial,
                                                                  double n_elem, double n_ssim, float data_range,
                                                                  float* __restrict__ out) {
    __shared__ double shd[32];
    double a = 0.0, b = 0.0, c = 0.0;
    for (long long i = threadIdx.x; i < n_partial; i += blockDim.x) {
        a += partial[i * 3]; b += partial[i * 3 + 1]; c += partial[i * 3 + 2];
    }
    a = block_sum_d(a, shd);
    b = block_sum_d(b, shd);
    c = block_sum_d(c, shd);
    if (threadIdx.x == 0) {
        double mse = a / n_elem;
        out[0] = (float)mse;
        out[1] = (float)(b / n_elem);
        out[2] = (float)(10.0 * log10((double)data_range * (double)data_range / mse));
        out[3] = n_ssim > 0.0 ? (float)(c / n_ssim) : __int_as_float(0x7fc00000);
    }
}

// mean/logvar (B, C, S) fp32 (S = product of the trailing dims): one thread per (b, s) sums over C, which is
// what `kl_per_dim.sum(dim=1)` does.  partial[blk*2 + {0,1}] = sum of per-sample KL, sum of its square.
__global__ void __launch_bounds__(256) kl_metrics_kernel(const float* __restrict__ mean, const float* __restrict__ logvar,
                                                         double* __restrict__ partial, int B, int C, int S) {
    __shared__ double shd[32];
    const long long n = (long long)B * S;
    double s1 = 0.0, s2 = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long b = i / S, s = i % S;
        const float* m = mean + b * C * S + s;
        const float* lv = logvar + b * C * S + s;
        double acc = 0.0;
        for (int c = 0; c < C; ++c) {
            float mu = m[(long long)c * S], l = lv[(long long)c * S];
            acc += (double)(0.5f * (mu * mu + expf(l) - l - 1.f));
        }
        s1 += acc;
        s2 += acc * acc;
    }
    s1 = block_sum_d(s1, shd);
    s2 = block_sum_d(s2, shd);
    if (threadIdx.x == 0) {
        partial[blockIdx.x * 2] = s1;
        partial[blockIdx.x * 2 + 1] = s2;
    }
}

// out[0..3] = kl_total, kl_mean, kl_std (unbiased), kl_per_dim_mean
__global__ void __launch_bounds__(256) kl_metrics_final_kernel(const double* __restrict__ partial, int n_partial, double n_samples,
                                                               double n_elem, float* __restrict__ out) {
    __shared__ double shd[32];
    double s1 = 0.0, s2 = 0.0;
    for (int i = threadIdx.x; i < n_partial; i += blockDim.x) {
        s1 += partial[i * 2]; s2 += partial[i * 2 + 1];
    }
    s1 = block_sum_d(s1, shd);
    s2 = block_sum_d(s2, shd);
    if (threadIdx.x == 0) {
        double mean = s1 / n_samples;
        double var = n_samples > 1.0 ? fmax(s2 - n_samples * mean * mean, 0.0) / (n_samples - 1.0) : nan("");
        out[0] = (float)s1;
        out[1] = (float)mean;
        out[2] = (float)sqrt(var);
        out[3] = (float)(s1 / n_elem);
    }
}

// latents (B, D) fp32: one thread per feature d (coalesced over d), loop over the batch.
// partial[blk*3 + {0,1,2}] = sum_d mean_d, sum_d std_d (unbiased over the batch), count(|z| < thr)
__global__ void __launch_bounds__(256) latent_metrics_kernel(const float* __restrict__ z, double* __restrict__ partial, int B,
                                                             long long D, float thr) {
    __shared__ double shd[32];
    double sm = 0.0, ss = 0.0, cnt = 0.0;
    for (long long d = blockIdx.x * (long long)blockDim.x + threadIdx.x; d < D; d += (long long)gridDim.x * blockDim.x) {
        double s1 = 0.0, s2 = 0.0;
        for (int b = 0; b < B; ++b) {
            float v = z[(long long)b * D + d];
            s1 += (double)v;
            s2 += (double)v * (double)v;
            cnt += fabsf(v) < thr ? 1.0 : 0.0;
        }
        double mean = s1 / B;
        sm += mean;
        ss += B > 1 ? sqrt(fmax(s2 - B * mean * mean, 0.0) / (B - 1)) : nan("");
    }
    sm = block_sum_d(sm, shd);
    ss = block_sum_d(ss, shd);
    cnt = block_sum_d(cnt, shd);
    if (threadIdx.x == 0) {
        partial[blockIdx.x * 3] = sm; partial[blockIdx.x * 3 + 1] = ss; partial[blockIdx.x * 3 + 2] = cnt;
    }
}

// out[0..2] = latent_mean_avg, latent_std_avg, latent_sparsity
__global__ void __launch_bounds__(256) latent_metrics_final_kernel(const double* __restrict__ partial, int n_partial, double B,
                                                                   double D, float* __restrict__ out) {
    __shared__ double shd[32];
    double a = 0.0, b = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < n_partial; i += blockDim.x) {
        a += partial[i * 3]; b += partial[i * 3 + 1]; c += partial[i * 3 + 2];
    }
    a = block_sum_d(a, shd);
    b = block_sum_d(b, shd);
    c = block_sum_d(c, shd);
    if (threadIdx.x == 0) {
        out[0] = (float)(a / D);
        out[1] = (float)(b / D);
        out[2] = (float)(c / (B * D));
    }
}

// ------------------------------------------------------------------ host side
static int small_grid(long long items) {
    long long g = (items + 255) / 256;
    return (int)(g < 1 ? 1 : (g > 592 ? 592 : g));   // <= 4 CTAs per SM on 148 SMs
}

long long recon_metrics_ws_doubles(int planes, int H, int W) {
    long long tiles = (long long)((H + MT - 1) / MT) * ((W + MT - 1) / MT);
    return tiles * planes * 3;
}
long long stat_metrics_ws_doubles() { return 592 * 3; }

int recon_metrics(const float* x, const float* r, int planes, int H, int W, float data_range, double* ws, float* out,
                  cudaStream_t st) {
    const int tx = (W + MT - 1) / MT, ty = (H + MT - 1) / MT;
    GaussTaps g;
    {   // torchmetrics' _gaussian(): fp32 exp of -(d/sigma)^2/2, normalised
        float raw[SSIM_K], sum = 0.f;
        for (int k = 0; k < SSIM_K; ++k) {
            float d = (float)(k - SSIM_PAD) / 1.5f;
            raw[k] = expf(-d * d / 2.f);
            sum += raw[k];
        }
        for (int k = 0; k < SSIM_K; ++k) g.w[k] = raw[k] / sum;
    }
    const float c1 = (0.01f * data_range) * (0.01f * data_range), c2 = (0.03f * data_range) * (0.03f * data_range);
    // planes go on grid.y in slabs of 65535
    for (int p0 = 0; p0 < planes; p0 += 65535) {
        int np = planes - p0 < 65535 ? planes - p0 : 65535;
        dim3 grid(tx * ty, np);
        recon_metrics_kernel<<<grid, 256, 0, st>>>(x + (long long)p0 * H * W, r + (long long)p0 * H * W,
                                                   ws + (long long)p0 * tx * ty * 3, H, W, tx, g, c1, c2);
        MVAE_LAUNCH_CHECK();
    }
    const double n_ssim = (H > 2 * SSIM_PAD && W > 2 * SSIM_PAD) ? (double)planes * (H - 2 * SSIM_PAD) * (W - 2 * SSIM_PAD) : 0.0;
    recon_metrics_final_kernel<<<1, 256, 0, st>>>(ws, (long long)planes * tx * ty, (double)planes * H * W, n_ssim, data_range, out);
    MVAE_LAUNCH_CHECK();
    return 0;
}

int kl_metrics(const float* mean, const float* logvar, int B, int C, int S, double* ws, float* out, cudaStream_t st) {
    const int grid = small_grid((long long)B * S);
    kl_metrics_kernel<<<grid, 256, 0, st>>>(mean, logvar, ws, B, C, S);
    MVAE_LAUNCH_CHECK();
    kl_metrics_final_kernel<<<1, 256, 0, st>>>(ws, grid, (double)B * S, (double)B * C * S, out);
    MVAE_LAUNCH_CHECK();
    return 0;
}

int latent_metrics(const float* z, int B, long long D, float thr, double* ws, float* out, cudaStream_t st) {
    const int grid = small_grid(D);
    latent_metrics_kernel<<<grid, 256, 0, st>>>(z, ws, B, D, thr);
    MVAE_LAUNCH_CHECK();
    latent_metrics_final_kernel<<<1, 256, 0, st>>>(ws, grid, (double)B, (double)D, out);
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
