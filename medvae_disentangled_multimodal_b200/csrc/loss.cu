// Latent split + reparameterisation (Philox4x32-10) + KL and reconstruction losses, with their
// gradients.  Reference: base_vae.py:76,83-87; src/losses/vae_losses.py:41-58;
// disentangled_conditional_vae.py:409-412 (clamps), :524-525 (KL sum).
#include "../../include/medvae_b200.h"
#include "common.cuh"
#include "philox.cuh"

namespace mvae {

// ------------------------------------------------------------------ Philox4x32-10 noise (philox.cuh)
// one standard normal for element `e` of stream (seed, offset): Box-Muller on words (0,1) or (2,3)
__device__ __forceinline__ float philox_normal(uint64_t seed, uint64_t offset, uint64_t e) {
    uint32_t a, b;
    philox_normal_words(seed, offset, e, a, b);
    const float u1 = philox_u01(a), u2 = philox_u01(b);
    // u1 can round to 1 - 2^-24: the clamp keeps an approximate logarithm that lands on the wrong side of 0 from
    // turning into a NaN under the root
    return sqrtf(fmaxf(-2.0f * __logf(u1), 0.f)) * __cosf(6.283185307179586f * u2);
}

// ------------------------------------------------------------------ reparam + KL forward
// The encoder hands over NHWC moments [n][p][2zc] (channels innermost), the model's contract wants NCHW fp32 mean /
// logvar / z / std / eps [n][c][p] (positions innermost): a CTA moves one 32-position x 32-channel tile through shared
// memory, so both sides see full 64 / 128-byte segments (a thread per NCHW element with strided NHWC accesses, the
// first version, ran at 19-25 % of HBM bandwidth: profiles/r2a_ncu_full.csv).
constexpr int RP_T = 32;
template <typename T>
__global__ void __launch_bounds__(256) reparam_kl_fwd_kernel(const T* __restrict__ mom, const float* __restrict__ eps_in,
                                                             float* __restrict__ eps_out, float* __restrict__ mean,
                                                             float* __restrict__ logvar, float* __restrict__ stdv,
                                                             float* __restrict__ z_nchw,
                                                             T* __restrict__ z_nhwc, double* __restrict__ kl_sum,
                                                             int B, int zc, int RR, float clampv, uint64_t seed,
                                                             uint64_t offset, uint64_t elem_base,
                                                             const double* __restrict__ dyn) {
    __shared__ double shd[32];
    __shared__ float s_mu[RP_T][RP_T + 1], s_lv[RP_T][RP_T + 1], s_z[RP_T][RP_T + 1];
    if (dyn) offset += (uint64_t)dyn[2];   // device-side Philox stream position (CUDA-graph replays advance it)
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
    const int tiles_p = (RR + RP_T - 1) / RP_T, tiles_c = (zc + RP_T - 1) / RP_T;
    const long long tiles = (long long)B * tiles_p * tiles_c;
    float acc = 0.f;
    for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int tc = (int)(t % tiles_c);
        const int tp = (int)((t / tiles_c) % tiles_p);
        const int n = (int)(t / ((long long)tiles_c * tiles_p));
        const int p0 = tp * RP_T, c0 = tc * RP_T;
        // NHWC side: tx = channel, rows = positions
        for (int i = ty; i < RP_T; i += 8) {
            const int p = p0 + i, c = c0 + tx;
            if (p < RR && c < zc) {
                const T* m = mom + ((long long)n * RR + p) * (2 * zc);
                s_mu[i][tx] = to_f<T>(m[c]);
                s_lv[i][tx] = to_f<T>(m[zc + c]);
            }
        }
        __syncthreads();
        // NCHW side: tx = position, rows = channels
        for (int i = ty; i < RP_T; i += 8) {
            const int p = p0 + tx, c = c0 + i;
            if (p < RR && c < zc) {
                float mu = s_mu[tx][i], lv = s_lv[tx][i];
                if (clampv > 0.f) {
                    mu = fminf(fmaxf(mu, -clampv), clampv);
                    lv = fminf(fmaxf(lv, -clampv), clampv);
                }
                const long long e = ((long long)n * zc + c) * RR + p;
                const float ep = eps_in ? eps_in[e] : philox_normal(seed, offset, elem_base + (uint64_t)e);
                if (eps_out) eps_out[e] = ep;
                const float sd = expf(0.5f * lv);
                const float zz = fmaf(ep, sd, mu);
                mean[e] = mu; logvar[e] = lv; z_nchw[e] = zz;
                // the disentangled model (the only clamping caller) also clamps std to [1e-6, 10] for its Normal
                // objects (disentangled_conditional_vae.py:426-427); z above uses the unclamped std as the reference does
                if (stdv) stdv[e] = clampv > 0.f ? fminf(fmaxf(sd, 1e-6f), 10.f) : sd;
                s_z[tx][i] = zz;
                acc += 0.5f * (sd * sd + mu * mu - 1.0f - lv);      // 4 terms per thread per tile in fp32, fp64 across
            }
        }
        if (z_nhwc) {
            __syncthreads();
            for (int i = ty; i < RP_T; i += 8) {
                const int p = p0 + i, c = c0 + tx;
                if (p < RR && c < zc) z_nhwc[((long long)n * RR + p) * zc + c] = from_f<T>(s_z[i][tx]);
            }
        }
        __syncthreads();
    }
    const double tot = block_sum_d((double)acc, shd);
    if (threadIdx.x == 0 && kl_sum) atomicAdd(kl_sum, tot);
}

// d_moments[n][p][c]      = d_mean + d_z                       (masked by the clamp)
// d_moments[n][p][zc + c] = d_logvar + d_z * eps * 0.5 * std   (masked by the clamp)
template <typename T>
__global__ void __launch_bounds__(256) reparam_kl_bwd_kernel(const float* __restrict__ d_mean,
                                                             const float* __restrict__ d_logvar,
                                                             const float* __restrict__ d_z, const float* __restrict__ logvar,
                                                             const float* __restrict__ eps, const T* __restrict__ mom,
                                                             T* __restrict__ d_mom, int B, int zc, int RR, float clampv) {
    __shared__ float s_gm[RP_T][RP_T + 1], s_gl[RP_T][RP_T + 1];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int tiles_p = (RR + RP_T - 1) / RP_T, tiles_c = (zc + RP_T - 1) / RP_T;
    const long long tiles = (long long)B * tiles_p * tiles_c;
    for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int tc = (int)(t % tiles_c);
        const int tp = (int)((t / tiles_c) % tiles_p);
        const int n = (int)(t / ((long long)tiles_c * tiles_p));
        const int p0 = tp * RP_T, c0 = tc * RP_T;
        for (int i = ty; i < RP_T; i += 8) {          // NCHW side: tx = position
            const int p = p0 + tx, c = c0 + i;
            if (p < RR && c < zc) {
                const long long e = ((long long)n * zc + c) * RR + p;
                const float dz = d_z ? d_z[e] : 0.f;
                s_gm[tx][i] = (d_mean ? d_mean[e] : 0.f) + dz;
                s_gl[tx][i] = (d_logvar ? d_logvar[e] : 0.f) + dz * eps[e] * 0.5f * expf(0.5f * logvar[e]);
            }
        }
        __syncthreads();
        for (int i = ty; i < RP_T; i += 8) {          // NHWC side: tx = channel
            const int p = p0 + i, c = c0 + tx;
            if (p < RR && c < zc) {
                const long long o = ((long long)n * RR + p) * (2 * zc);
                float gm = s_gm[i][tx], gl = s_gl[i][tx];
                if (clampv > 0.f) {
                    const float mu_raw = to_f<T>(mom[o + c]), lv_raw = to_f<T>(mom[o + zc + c]);
                    if (mu_raw < -clampv || mu_raw > clampv) gm = 0.f;
                    if (lv_raw < -clampv || lv_raw > clampv) gl = 0.f;
                }
                d_mom[o + c] = from_f<T>(gm);
                d_mom[o + zc + c] = from_f<T>(gl);
            }
        }
        __syncthreads();
    }
}

static int rp_blocks(int B, int zc, int RR) {
    long long t = (long long)B * ((RR + RP_T - 1) / RP_T) * ((zc + RP_T - 1) / RP_T);
    if (t > 148LL * 8) t = 148LL * 8;
    return (int)(t < 1 ? 1 : t);
}

static int ew_blocks(long long total) {
    long long b = (total + 255) / 256;
    if (b > 148LL * 8) b = 148LL * 8;
    if (b < 1) b = 1;
    return (int)b;
}

int reparam_fwd(const void* mom, const float* eps, float* eps_out, float* mean, float* logvar, float* stdv,
                float* z_nchw, void* z_nhwc, double* kl_sum, int B, int zc, int R, float clampv, uint64_t seed, uint64_t offset,
                uint64_t elem_base, const double* dyn, int dtype, cudaStream_t st) {
    long long total = (long long)B * zc * R * R;
    if (dtype == MVAE_F32)
        reparam_kl_fwd_kernel<float><<<rp_blocks(B, zc, R * R), 256, 0, st>>>((const float*)mom, eps, eps_out, mean, logvar, stdv, z_nchw,
                                                                        (float*)z_nhwc, kl_sum, B, zc, R * R, clampv, seed, offset, elem_base, dyn);
    else
        reparam_kl_fwd_kernel<bf16><<<rp_blocks(B, zc, R * R), 256, 0, st>>>((const bf16*)mom, eps, eps_out, mean, logvar, stdv, z_nchw,
                                                                       (bf16*)z_nhwc, kl_sum, B, zc, R * R, clampv, seed, offset, elem_base, dyn);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int reparam_bwd(const float* d_mean, const float* d_logvar, const float* d_z, const float* logvar, const float* eps,
                const void* mom, void* d_mom, int B, int zc, int R, float clampv, int dtype, cudaStream_t st) {
    long long total = (long long)B * zc * R * R;
    if (dtype == MVAE_F32)
        reparam_kl_bwd_kernel<float><<<rp_blocks(B, zc, R * R), 256, 0, st>>>(d_mean, d_logvar, d_z, logvar, eps, (const float*)mom,
                                                                        (float*)d_mom, B, zc, R * R, clampv);
    else
        reparam_kl_bwd_kernel<bf16><<<rp_blocks(B, zc, R * R), 256, 0, st>>>(d_mean, d_logvar, d_z, logvar, eps, (const bf16*)mom,
                                                                       (bf16*)d_mom, B, zc, R * R, clampv);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ reconstruction loss (+ unit grad)
__global__ void __launch_bounds__(256) recon_loss_kernel(const float* __restrict__ r, const float* __restrict__ x,
                                                         double* __restrict__ loss_sum, float* __restrict__ d_unit,
                                                         long long n, int kind) {
    __shared__ double shd[32];
    const float inv_n = 1.0f / (float)n;
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float rv = r[i], xv = x[i], l, g;
        if (kind == 0) { float d = rv - xv; l = d * d; g = 2.f * d; }
        else if (kind == 1) { float d = rv - xv; l = fabsf(d); g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); }
        else {
            // binary_cross_entropy_with_logits: max(r,0) - r*x + log1p(exp(-|r|)); d/dr = sigmoid(r) - x
            l = fmaxf(rv, 0.f) - rv * xv + log1pf(__expf(-fabsf(rv)));
            g = 1.f / (1.f + __expf(-rv)) - xv;
        }
        acc += (double)l;
        if (d_unit) d_unit[i] = g * inv_n;
    }
    acc = block_sum_d(acc, shd);
    if (threadIdx.x == 0) atomicAdd(loss_sum, acc);
}
int recon_loss(const float* r, const float* x, double* loss_sum, float* d_unit, long long n, int kind, cudaStream_t st) {
    recon_loss_kernel<<<ew_blocks(n), 256, 0, st>>>(r, x, loss_sum, d_unit, n, kind);
    MVAE_LAUNCH_CHECK();
    return 0;
}

__global__ void __launch_bounds__(256) kl_fwd_kernel(const float* __restrict__ mean, const float* __restrict__ logvar,
                                                     double* __restrict__ kl_sum, long long n) {
    __shared__ double shd[32];
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float mu = mean[i], lv = logvar[i];
        acc += 0.5 * ((double)__expf(lv) + (double)mu * mu - 1.0 - (double)lv);
    }
    acc = block_sum_d(acc, shd);
    if (threadIdx.x == 0) atomicAdd(kl_sum, acc);
}
__global__ void kl_bwd_kernel(const float* __restrict__ mean, const float* __restrict__ logvar,
                              const float* __restrict__ coef, float* __restrict__ d_mean, float* __restrict__ d_logvar,
                              long long n) {
    const float c = *coef;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        d_mean[i] = c * mean[i];
        d_logvar[i] = c * 0.5f * (__expf(logvar[i]) - 1.f);
    }
}
__global__ void scale_kernel(const float* __restrict__ in, const float* __restrict__ s, float* __restrict__ out, long long n) {
    const float c = *s;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = in[i] * c;
}
int kl_fwd(const float* mean, const float* logvar, double* kl_sum, long long n, cudaStream_t st) {
    kl_fwd_kernel<<<ew_blocks(n), 256, 0, st>>>(mean, logvar, kl_sum, n);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int kl_bwd(const float* mean, const float* logvar, const float* coef, float* d_mean, float* d_logvar, long long n,
           cudaStream_t st) {
    kl_bwd_kernel<<<ew_blocks(n), 256, 0, st>>>(mean, logvar, coef, d_mean, d_logvar, n);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int scale_dev(const float* in, const float* s, float* out, long long n, cudaStream_t st) {
    scale_kernel<<<ew_blocks(n), 256, 0, st>>>(in, s, out, n);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ optimizer
// pass 1: sum of squares of finite gradient entries (non-finite entries are treated as zero,
// as lightning_module.py:468-477 zeroes them before clipping)
__global__ void __launch_bounds__(256) grad_sqnorm_kernel(const float* const* __restrict__ grads,
                                                          const long long* __restrict__ sizes, int nt,
                                                          double* __restrict__ out) {
    __shared__ double shd[32];
    double acc = 0.0;
    for (int t = blockIdx.y; t < nt; t += gridDim.y) {
        const float* g = grads[t];
        long long n = sizes[t];
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            float v = g[i];
            if (isfinite(v)) acc += (double)v * v;
        }
    }
    acc = block_sum_d(acc, shd);
    if (threadIdx.x == 0 && acc != 0.0) atomicAdd(out, acc);
}
__global__ void __launch_bounds__(256) adamw_kernel(float* const* __restrict__ params, const float* const* __restrict__ grads,
                                                    float* const* __restrict__ m1, float* const* __restrict__ m2,
                                                    const long long* __restrict__ sizes, int nt,
                                                    const double* __restrict__ sqnorm, float max_norm, float lr, float b1,
                                                    float b2, float eps, float wd, int decoupled, float bc1, float bc2) {
    float clip = 1.f;
    if (max_norm > 0.f && sqnorm) {
        float norm = (float)sqrt(*sqnorm);
        float c = max_norm / (norm + 1e-6f);
        clip = c < 1.f ? c : 1.f;
    }
    for (int t = blockIdx.y; t < nt; t += gridDim.y) {
        float* p = params[t];
        const float* g = grads[t];
        float* a = m1[t];
        float* b = m2[t];
        long long n = sizes[t];
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            float gv = g[i];
            gv = isfinite(gv) ? gv * clip : 0.f;
            float pv = p[i];
            if (decoupled) pv *= (1.f - lr * wd); else gv = fmaf(wd, pv, gv);
            float m = fmaf(b1, a[i], (1.f - b1) * gv);
            float v = fmaf(b2, b[i], (1.f - b2) * gv * gv);
            a[i] = m; b[i] = v;
            float denom = sqrtf(v) / sqrtf(bc2) + eps;
            p[i] = pv - (lr / bc1) * (m / denom);
        }
    }
}
// ---- chunked variants: a table of (tensor, offset, length) chunks of <= 64 Ki elements gives every CTA
// the same amount of work whatever the tensor sizes (the 17 9.4 M-element conv weights of C4 next to
// 128-element biases), and 16-byte accesses whenever the four pointers allow it.
// Pass 1 keeps one sum of squares and one "non-finite entry seen" flag PER TENSOR: the reference zeroes the WHOLE gradient
// of a parameter that holds any NaN / Inf before it clips (src/lightning_module.py:468-477), so a poisoned tensor must
// drop out of the global norm entirely — its finite entries included.
__global__ void __launch_bounds__(256) grad_sqnorm_chunked_kernel(const float* const* __restrict__ grads,
                                                                  const int* __restrict__ ch_tensor,
                                                                  const long long* __restrict__ ch_off,
                                                                  const int* __restrict__ ch_len, int n_chunks,
                                                                  double* __restrict__ tensor_sq,
                                                                  int* __restrict__ tensor_bad) {
    __shared__ double shd[32];
    __shared__ int sh_bad;
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const int t = ch_tensor[c];
        const float* g = grads[t] + ch_off[c];
        const int n = ch_len[c];
        if (threadIdx.x == 0) sh_bad = 0;
        __syncthreads();
        float part = 0.f;
        bool bad = false;
        if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
            const float4* g4 = reinterpret_cast<const float4*>(g);
            for (int i = threadIdx.x; i < n / 4; i += 256) {
                const float4 v = g4[i];
                bad |= !(isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w));
                part = fmaf(v.x, v.x, part); part = fmaf(v.y, v.y, part);
                part = fmaf(v.z, v.z, part); part = fmaf(v.w, v.w, part);
            }
            for (int i = (n / 4) * 4 + threadIdx.x; i < n; i += 256) {
                const float v = g[i];
                bad |= !isfinite(v);
                part = fmaf(v, v, part);
            }
        } else {
            for (int i = threadIdx.x; i < n; i += 256) {
                const float v = g[i];
                bad |= !isfinite(v);
                part = fmaf(v, v, part);
            }
        }
        if (bad) sh_bad = 1;                                  // (benign race: every writer stores 1)
        const double acc = block_sum_d(bad ? 0.0 : (double)part, shd);   // <= 256 fp32 terms per thread, then fp64
        __syncthreads();
        if (threadIdx.x == 0) {
            if (sh_bad) atomicOr(&tensor_bad[t], 1);
            else if (acc != 0.0) atomicAdd(&tensor_sq[t], acc);
        }
        __syncthreads();
    }
}

struct AdamArgs {
    float max_norm, lr, b1, b2, eps, wd, bc1, bc2;
    int decoupled;
};
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float clip, const AdamArgs& a) {
    g = clip == 0.f ? 0.f : g * clip;      // clip == 0: a tensor whose gradient held a NaN / Inf counts as zero as a whole
    if (a.decoupled) p *= (1.f - a.lr * a.wd); else g = fmaf(a.wd, p, g);
    m = fmaf(a.b1, m, (1.f - a.b1) * g);
    v = fmaf(a.b2, v, (1.f - a.b2) * g * g);
    const float denom = sqrtf(v) / sqrtf(a.bc2) + a.eps;
    p -= (a.lr / a.bc1) * (m / denom);
}
__global__ void __launch_bounds__(256) adamw_chunked_kernel(float* const* __restrict__ params,
                                                            const float* const* __restrict__ grads,
                                                            float* const* __restrict__ m1, float* const* __restrict__ m2,
                                                            const int* __restrict__ ch_tensor,
                                                            const long long* __restrict__ ch_off,
                                                            const int* __restrict__ ch_len, int n_chunks,
                                                            const double* __restrict__ tensor_sq,
                                                            const int* __restrict__ tensor_bad,
                                                            const int* __restrict__ tensor_skip, int n_tensors,
                                                            double* __restrict__ sqnorm_out, AdamArgs a,
                                                            const double* __restrict__ dyn) {
    if (dyn) {   // step count and learning rate live on the device (CUDA-graph replays): dyn = {step, lr, rng}
        const float t = (float)dyn[0];
        a.lr = (float)dyn[1];
        a.bc1 = 1.f - powf(a.b1, t);
        a.bc2 = 1.f - powf(a.b2, t);
    }
    // global norm over the tensors that take part: not skipped (never produced a gradient this step) and not poisoned
    __shared__ double shd[32];
    __shared__ double sh_total;
    double part = 0.0;
    for (int t = threadIdx.x; t < n_tensors; t += 256)
        if (!tensor_bad[t] && !(tensor_skip && tensor_skip[t])) part += tensor_sq[t];
    part = block_sum_d(part, shd);
    if (threadIdx.x == 0) {
        sh_total = part;
        if (blockIdx.x == 0 && sqnorm_out) *sqnorm_out = part;
    }
    __syncthreads();
    float clip_all = 1.f;
    if (a.max_norm > 0.f) {
        const float norm = (float)sqrt(sh_total);
        const float c = a.max_norm / (norm + 1e-6f);
        clip_all = c < 1.f ? c : 1.f;
    }
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const int t = ch_tensor[c];
        if (tensor_skip && tensor_skip[t]) continue;       // torch.optim skips a parameter whose .grad is None
        const float clip = tensor_bad[t] ? 0.f : clip_all;
        const long long off = ch_off[c];
        const int n = ch_len[c];
        float* p = params[t] + off;
        const float* g = grads[t] + off;
        float* m = m1[t] + off;
        float* v = m2[t] + off;
        const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                           reinterpret_cast<uintptr_t>(v)) & 15) == 0;
        int done = 0;
        if (vec) {
            for (int i = threadIdx.x; i < n / 4; i += 256) {
                float4 pv = reinterpret_cast<float4*>(p)[i];
                const float4 gv = reinterpret_cast<const float4*>(g)[i];
                float4 mv = reinterpret_cast<float4*>(m)[i];
                float4 vv = reinterpret_cast<float4*>(v)[i];
                adam_one(pv.x, gv.x, mv.x, vv.x, clip, a);
                adam_one(pv.y, gv.y, mv.y, vv.y, clip, a);
                adam_one(pv.z, gv.z, mv.z, vv.z, clip, a);
                adam_one(pv.w, gv.w, mv.w, vv.w, clip, a);
                reinterpret_cast<float4*>(p)[i] = pv;
                reinterpret_cast<float4*>(m)[i] = mv;
                reinterpret_cast<float4*>(v)[i] = vv;
            }
            done = (n / 4) * 4;
        }
        for (int i = done + threadIdx.x; i < n; i += 256) {
            float pv = p[i], mv = m[i], vv = v[i];
            adam_one(pv, g[i], mv, vv, clip, a);
            p[i] = pv; m[i] = mv; v[i] = vv;
        }
    }
}
static int chunk_grid(int n_chunks) {
    int g = 148 * 8;
    return n_chunks < g ? n_chunks : g;
}
int grad_sqnorm_chunked(const float* const* grads, const int* ch_tensor, const long long* ch_off, const int* ch_len,
                        int n_chunks, double* tensor_sq, int* tensor_bad, cudaStream_t st) {
    grad_sqnorm_chunked_kernel<<<chunk_grid(n_chunks), 256, 0, st>>>(grads, ch_tensor, ch_off, ch_len, n_chunks, tensor_sq,
                                                                      tensor_bad);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int adamw_step_chunked(float* const* params, const float* const* grads, float* const* m1, float* const* m2,
                       const int* ch_tensor, const long long* ch_off, const int* ch_len, int n_chunks,
                       const double* tensor_sq, const int* tensor_bad, const int* tensor_skip, int n_tensors,
                       double* sqnorm_out, float max_norm, float lr, float b1, float b2, float eps, float wd,
                       int decoupled, int step, const double* dyn, cudaStream_t st) {
    AdamArgs a;
    a.max_norm = max_norm; a.lr = lr; a.b1 = b1; a.b2 = b2; a.eps = eps; a.wd = wd; a.decoupled = decoupled;
    a.bc1 = 1.f - powf(b1, (float)step); a.bc2 = 1.f - powf(b2, (float)step);
    adamw_chunked_kernel<<<chunk_grid(n_chunks), 256, 0, st>>>(params, grads, m1, m2, ch_tensor, ch_off, ch_len, n_chunks,
                                                                tensor_sq, tensor_bad, tensor_skip, n_tensors, sqnorm_out,
                                                                a, dyn);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int grad_sqnorm(const float* const* grads, const long long* sizes, int nt, double* out, cudaStream_t st) {
    dim3 grid(32, nt < 64 ? nt : 64);
    grad_sqnorm_kernel<<<grid, 256, 0, st>>>(grads, sizes, nt, out);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int adamw_step(float* const* params, const float* const* grads, float* const* m1, float* const* m2,
               const long long* sizes, int nt, const double* sqnorm, float max_norm, float lr, float b1, float b2,
               float eps, float wd, int decoupled, int step, cudaStream_t st) {
    float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
    dim3 grid(32, nt < 64 ? nt : 64);
    adamw_kernel<<<grid, 256, 0, st>>>(params, grads, m1, m2, sizes, nt, sqnorm, max_norm, lr, b1, b2, eps, wd, decoupled, bc1, bc2);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// dyn = {optimizer step, learning rate, Philox stream position}: one training step forward
__global__ void step_state_advance_kernel(double* dyn, double rng_inc) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        dyn[0] += 1.0;
        dyn[2] += rng_inc;
    }
}
int step_state_advance(double* dyn, double rng_inc, cudaStream_t st) {
    step_state_advance_kernel<<<1, 32, 0, st>>>(dyn, rng_inc);
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
