// DisentangledConditionalVAE extras (src/models/disentangled_conditional_vae.py):
//   * per-modality decoder heads  Conv2d(3,3,3,1,1) -> ReLU -> Conv2d(3,3,3,1,1) [-> Conv2d(3,1,1)]
//     with the zero padding back to the batch's widest sample (:241-303), forward and backward,
//     batched over the samples with the weights indexed by modality (no per-sample Python loop,
//     no .item() syncs);
//   * modality_separation_loss (:305-349) and contrastive_loss (:351-386) on the partitioned latent
//     (partition_latent :195-206 slices the FLATTENED (C,R,R) latent), forward and backward;
//   * the guarded total of DisentangledVAELoss (:527-565).
// All 3-channel stencil / B x B work: tiny next to the encoder/decoder, written for clarity and
// fixed-order (deterministic) reductions except the per-modality weight-gradient atomics.
#include "../../include/medvae_b200.h"
#include "common.cuh"

namespace mvae {

constexpr int HEAD_W = 81, HEAD_PER = 2 * (HEAD_W + 3) + 4;   // w1 b1 w2 b2 pw(3) pb(1) per modality

__device__ __forceinline__ int clamp_mod(long long m, int n_mod) {
    if (m >= n_mod) m = n_mod - 1;   // :142-146, :260-265: out-of-range indices use the last entry
    if (m < 0) m = 0;
    return (int)m;
}

// ------------------------------------------------------------------ head: 3 -> 3 conv3x3 (+ReLU)
// in: [B][H][W][in_stride] (first 3 channels used), out: fp32 [B][H][W][3]
template <typename T>
__global__ void __launch_bounds__(256) head_conv3_kernel(const T* __restrict__ in, int in_stride,
                                                         const long long* __restrict__ idx, const float* __restrict__ w,
                                                         const float* __restrict__ b, float* __restrict__ out, int H, int W,
                                                         int n_mod, int relu) {
    __shared__ float sw[HEAD_W + 3];
    const int n = blockIdx.y;
    const int m = clamp_mod(idx[n], n_mod);
    if (threadIdx.x < HEAD_W) sw[threadIdx.x] = w[m * HEAD_W + threadIdx.x];
    if (threadIdx.x < 3) sw[HEAD_W + threadIdx.x] = b[m * 3 + threadIdx.x];
    __syncthreads();
    const int HW = H * W;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const int y = p / W, x = p % W;
        float acc[3] = {sw[HEAD_W], sw[HEAD_W + 1], sw[HEAD_W + 2]};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int yy = y + r - 1;
            if (yy < 0 || yy >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int xx = x + s - 1;
                if (xx < 0 || xx >= W) continue;
                const T* ip = in + ((long long)n * HW + yy * W + xx) * in_stride;
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const float v = to_f<T>(ip[i]);
#pragma unroll
                    for (int o = 0; o < 3; ++o) acc[o] = fmaf(sw[((o * 3 + i) * 3 + r) * 3 + s], v, acc[o]);
                }
            }
        }
        float* op = out + ((long long)n * HW + p) * 3;
#pragma unroll
        for (int o = 0; o < 3; ++o) op[o] = relu ? fmaxf(acc[o], 0.f) : acc[o];
    }
}

// y: fp32 [B][H][W][3] -> out NCHW fp32 [B][out_c][H][W]: gray modalities: out[0] = pw . y + pb, rest 0
__global__ void __launch_bounds__(256) head_out_kernel(const float* __restrict__ y, const long long* __restrict__ idx,
                                                       const float* __restrict__ pw, const float* __restrict__ pb,
                                                       const int* __restrict__ is_gray, float* __restrict__ out, int HW,
                                                       int n_mod, int out_c) {
    const int n = blockIdx.y;
    const int m = clamp_mod(idx[n], n_mod);
    const bool gray = is_gray[m] != 0;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const float* yp = y + ((long long)n * HW + p) * 3;
        float* op = out + (long long)n * out_c * HW + p;
        if (gray) {
            op[0] = fmaf(pw[m * 3], yp[0], fmaf(pw[m * 3 + 1], yp[1], fmaf(pw[m * 3 + 2], yp[2], pb[m])));
            for (int c = 1; c < out_c; ++c) op[(long long)c * HW] = 0.f;
        } else {
            for (int c = 0; c < out_c; ++c) op[(long long)c * HW] = c < 3 ? yp[c] : 0.f;
        }
    }
}

// per-block sum of `v` added to *dst (one atomic per block); sh: 8 floats
__device__ __forceinline__ void block_atomic_add(float v, float* dst, float* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += sh[i];
        if (t != 0.f) atomicAdd(dst, t);
    }
}

// d_out NCHW fp32 -> dy [B][HW][3]; gray: dy[c] = pw[c]*d_out[0], dpw[c] += d_out[0]*y[c], dpb += d_out[0]
__global__ void __launch_bounds__(256) head_out_bwd_kernel(const float* __restrict__ d_out, const float* __restrict__ y,
                                                           const long long* __restrict__ idx, const float* __restrict__ pw,
                                                           const int* __restrict__ is_gray, float* __restrict__ dy,
                                                           float* __restrict__ gflat, int HW, int n_mod, int out_c) {
    __shared__ float sh[8];
    const int n = blockIdx.y;
    const int m = clamp_mod(idx[n], n_mod);
    const bool gray = is_gray[m] != 0;
    float aw[3] = {0.f, 0.f, 0.f}, ab = 0.f;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const float* gp = d_out + (long long)n * out_c * HW + p;
        float* dp = dy + ((long long)n * HW + p) * 3;
        if (gray) {
            const float g0 = gp[0];
            const float* yp = y + ((long long)n * HW + p) * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c) { dp[c] = pw[m * 3 + c] * g0; aw[c] = fmaf(g0, yp[c], aw[c]); }
            ab += g0;
        } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) dp[c] = c < out_c ? gp[(long long)c * HW] : 0.f;
        }
    }
    if (gray) {   // block-uniform branch
        float* g = gflat + (long long)m * HEAD_PER;
        for (int c = 0; c < 3; ++c) block_atomic_add(aw[c], g + 2 * (HEAD_W + 3) + c, sh);
        block_atomic_add(ab, g + 2 * (HEAD_W + 3) + 3, sh);
    }
}

// backward of out = conv3x3(in) (3 -> 3, pad 1): dy fp32 [B][HW][3] ->
//   d_in[i](q) = sum dy[o](q - off) w[o][i][r][s]   (x 1[in > 0] when relu_mask: `in` is a ReLU output)
//   dW[o][i][r][s] += sum_p dy[o](p) in[i](p + off),  db[o] += sum_p dy[o](p)   -> gflat[m][goff ...]
template <typename T>
__global__ void __launch_bounds__(256) head_conv3_bwd_kernel(const float* __restrict__ dy, const T* __restrict__ in,
                                                             int in_stride, const long long* __restrict__ idx,
                                                             const float* __restrict__ w, T* __restrict__ d_in,
                                                             int din_stride, float* __restrict__ gflat, int goff, int H,
                                                             int W, int n_mod, int relu_mask) {
    __shared__ float sw[HEAD_W];
    __shared__ float sg[HEAD_W + 3];
    const int n = blockIdx.y;
    const int m = clamp_mod(idx[n], n_mod);
    if (threadIdx.x < HEAD_W) sw[threadIdx.x] = w[m * HEAD_W + threadIdx.x];
    if (threadIdx.x < HEAD_W + 3) sg[threadIdx.x] = 0.f;
    __syncthreads();
    const int HW = H * W;
    float gw[HEAD_W], gb[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < HEAD_W; ++i) gw[i] = 0.f;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const int y = p / W, x = p % W;
        const float* dp = dy + ((long long)n * HW + p) * 3;
        const float d0 = dp[0], d1 = dp[1], d2 = dp[2];
        gb[0] += d0; gb[1] += d1; gb[2] += d2;
        float di[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                // weight gradient: this output pixel p with input pixel p + off
                const int yy = y + r - 1, xx = x + s - 1;
                if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
                    const T* ip = in + ((long long)n * HW + yy * W + xx) * in_stride;
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        const float v = to_f<T>(ip[i]);
                        gw[((0 * 3 + i) * 3 + r) * 3 + s] = fmaf(d0, v, gw[((0 * 3 + i) * 3 + r) * 3 + s]);
                        gw[((1 * 3 + i) * 3 + r) * 3 + s] = fmaf(d1, v, gw[((1 * 3 + i) * 3 + r) * 3 + s]);
                        gw[((2 * 3 + i) * 3 + r) * 3 + s] = fmaf(d2, v, gw[((2 * 3 + i) * 3 + r) * 3 + s]);
                    }
                }
                // input gradient at q = p: contributions of output pixels q - off
                const int py = y - (r - 1), px = x - (s - 1);
                if (py >= 0 && py < H && px >= 0 && px < W) {
                    const float* ep = dy + ((long long)n * HW + py * W + px) * 3;
#pragma unroll
                    for (int i = 0; i < 3; ++i)
#pragma unroll
                        for (int o = 0; o < 3; ++o) di[i] = fmaf(ep[o], sw[((o * 3 + i) * 3 + r) * 3 + s], di[i]);
                }
            }
        }
        if (d_in) {
            const T* ip = in + ((long long)n * HW + p) * in_stride;
            T* op = d_in + ((long long)n * HW + p) * din_stride;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                float v = di[i];
                if (relu_mask && !(to_f<T>(ip[i]) > 0.f)) v = 0.f;
                op[i] = from_f<T>(v);
            }
        }
    }
    // block reduction: warp shuffles, then shared-memory atomics, then one global atomic per entry
#pragma unroll
    for (int i = 0; i < HEAD_W; ++i) {
        const float t = warp_sum(gw[i]);
        if ((threadIdx.x & 31) == 0 && t != 0.f) atomicAdd(&sg[i], t);
    }
#pragma unroll
    for (int o = 0; o < 3; ++o) {
        const float t = warp_sum(gb[o]);
        if ((threadIdx.x & 31) == 0 && t != 0.f) atomicAdd(&sg[HEAD_W + o], t);
    }
    __syncthreads();
    if (threadIdx.x < HEAD_W + 3 && sg[threadIdx.x] != 0.f)
        atomicAdd(gflat + (long long)m * HEAD_PER + goff + threadIdx.x, sg[threadIdx.x]);
}

static dim3 head_grid(int B, int HW) {
    int bx = (HW + 255) / 256;
    if (bx > 64) bx = 64;
    return dim3(bx, B);
}

// r: decoder output NHWC (dtype, r_stride channels, first 3 used) -> hid, y (fp32 [B][HW][3], kept for
// the backward) and out NCHW fp32 [B][out_c][H][W]
int dis_head_fwd(const void* r, int r_stride, const long long* idx, const float* w1, const float* b1, const float* w2,
                 const float* b2, const float* pw, const float* pb, const int* is_gray, float* hid, float* y, float* out,
                 int B, int H, int W, int n_mod, int out_c, int dtype, cudaStream_t st) {
    dim3 grid = head_grid(B, H * W);
    if (dtype == MVAE_F32)
        head_conv3_kernel<float><<<grid, 256, 0, st>>>((const float*)r, r_stride, idx, w1, b1, hid, H, W, n_mod, 1);
    else
        head_conv3_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)r, r_stride, idx, w1, b1, hid, H, W, n_mod, 1);
    MVAE_LAUNCH_CHECK();
    head_conv3_kernel<float><<<grid, 256, 0, st>>>(hid, 3, idx, w2, b2, y, H, W, n_mod, 0);
    MVAE_LAUNCH_CHECK();
    head_out_kernel<<<grid, 256, 0, st>>>(y, idx, pw, pb, is_gray, out, H * W, n_mod, out_c);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// d_out NCHW fp32 -> d_r (dtype, r_stride channels; only channels 0..2 are written: caller zero-fills) and
// gflat fp32 [n_mod][HEAD_PER] = (dw1 81, db1 3, dw2 81, db2 3, dpw 3, dpb 1), zeroed here.
// scratch: 2 * B*HW*3 floats.
int dis_head_bwd(const float* d_out, const void* r, int r_stride, const float* hid, const float* y, const long long* idx,
                 const float* w1, const float* w2, const float* pw, const int* is_gray, void* d_r, float* gflat,
                 float* scratch, int B, int H, int W, int n_mod, int out_c, int dtype, cudaStream_t st) {
    const int HW = H * W;
    dim3 grid = head_grid(B, HW);
    float* dy = scratch;
    float* dhid = scratch + (size_t)B * HW * 3;
    MVAE_CUDA(cudaMemsetAsync(gflat, 0, sizeof(float) * (size_t)n_mod * HEAD_PER, st));
    head_out_bwd_kernel<<<grid, 256, 0, st>>>(d_out, y, idx, pw, is_gray, dy, gflat, HW, n_mod, out_c);
    MVAE_LAUNCH_CHECK();
    head_conv3_bwd_kernel<float><<<grid, 256, 0, st>>>(dy, hid, 3, idx, w2, dhid, 3, gflat, HEAD_W + 3, H, W, n_mod, 1);
    MVAE_LAUNCH_CHECK();
    if (dtype == MVAE_F32)
        head_conv3_bwd_kernel<float><<<grid, 256, 0, st>>>(dhid, (const float*)r, r_stride, idx, w1, (float*)d_r, r_stride,
                                                            gflat, 0, H, W, n_mod, 0);
    else
        head_conv3_bwd_kernel<bf16><<<grid, 256, 0, st>>>(dhid, (const bf16*)r, r_stride, idx, w1, (bf16*)d_r, r_stride,
                                                           gflat, 0, H, W, n_mod, 0);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ================================================================== auxiliary latent losses
// zm[i][d] = z_flat[i][s0 + d], d < m  (partition_latent :195-206).  tau = contrastive temperature.
// aux layout (fp32): [0] separation loss, [1] contrastive loss, [2] rows with a positive, [3] centroid
// pairs; then inv_norm[B], pos[B], all[B], cents[n_mod][m], dcent[n_mod][m] (unit gradient of the
// separation loss w.r.t. each centroid), count[n_mod].
struct AuxLayout {
    int B, m, n_mod;
    __host__ __device__ long long inv() const { return 4; }
    __host__ __device__ long long pos() const { return 4 + (long long)B; }
    __host__ __device__ long long all() const { return 4 + 2LL * B; }
    __host__ __device__ long long cents() const { return 4 + 3LL * B; }
    __host__ __device__ long long dcent() const { return cents() + (long long)n_mod * m; }
    __host__ __device__ long long count() const { return dcent() + (long long)n_mod * m; }
    __host__ __device__ long long total() const { return count() + n_mod; }
};

__global__ void __launch_bounds__(128) aux_norm_kernel(const float* __restrict__ z, long long D, int s0, AuxLayout L,
                                                       float* __restrict__ aux) {
    __shared__ float sh[32];
    const int i = blockIdx.x;
    float a = 0.f;
    for (int d = threadIdx.x; d < L.m; d += blockDim.x) {
        const float v = z[(long long)i * D + s0 + d];
        a = fmaf(v, v, a);
    }
    a = block_sum(a, sh);
    if (threadIdx.x == 0) aux[L.inv() + i] = 1.f / fmaxf(sqrtf(a), 1e-12f);   // F.normalize eps
}

// row i: pos_i = sum_{j != i, same modality} exp(S_ij), all_i = sum_{j != i} exp(S_ij)
__global__ void __launch_bounds__(128) aux_rows_kernel(const float* __restrict__ z, const long long* __restrict__ idx,
                                                       long long D, int s0, float inv_tau, AuxLayout L,
                                                       float* __restrict__ aux) {
    extern __shared__ float zi[];   // [m] normalised row i
    __shared__ float sh[32];
    const int i = blockIdx.x;
    const float inv_i = aux[L.inv() + i];
    for (int d = threadIdx.x; d < L.m; d += blockDim.x) zi[d] = z[(long long)i * D + s0 + d] * inv_i;
    __syncthreads();
    const long long mi = idx[i];
    float pos = 0.f, all = 0.f;
    for (int j = threadIdx.x; j < L.B; j += blockDim.x) {
        if (j == i) continue;
        const float* zj = z + (long long)j * D + s0;
        float dot = 0.f;
        for (int d = 0; d < L.m; ++d) dot = fmaf(zi[d], zj[d], dot);
        const float e = expf(dot * aux[L.inv() + j] * inv_tau);
        all += e;
        if (idx[j] == mi) pos += e;
    }
    pos = block_sum(pos, sh);
    all = block_sum(all, sh);
    if (threadIdx.x == 0) { aux[L.pos() + i] = pos; aux[L.all() + i] = all; }
}

// one CTA: centroids, pdist, both loss values, unit centroid gradients
__global__ void __launch_bounds__(256) aux_final_kernel(const float* __restrict__ z, const long long* __restrict__ idx,
                                                        long long D, int s0, AuxLayout L, float* __restrict__ aux) {
    __shared__ float sh[32];
    __shared__ float dist[128];   // pairwise distances (n_mod <= 16 -> 120 pairs)
    float* cents = aux + L.cents();
    float* dcent = aux + L.dcent();
    float* count = aux + L.count();
    for (int k = threadIdx.x; k < L.n_mod; k += blockDim.x) {
        int c = 0;
        for (int i = 0; i < L.B; ++i) c += (idx[i] == k);
        count[k] = (float)c;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < L.n_mod * L.m; t += blockDim.x) {
        const int k = t / L.m, d = t % L.m;
        float a = 0.f;
        for (int i = 0; i < L.B; ++i)
            if (idx[i] == k) a += z[(long long)i * D + s0 + d];
        cents[t] = count[k] > 0.f ? a / count[k] : 0.f;
        dcent[t] = 0.f;
    }
    __syncthreads();
    // pairwise distances between PRESENT modalities (torch.unique order = ascending index)
    int npairs = 0;
    for (int a = 0; a < L.n_mod; ++a)
        for (int b = a + 1; b < L.n_mod; ++b) {
            if (count[a] > 0.f && count[b] > 0.f) {
                float s = 0.f;
                for (int d = threadIdx.x; d < L.m; d += blockDim.x) {
                    const float df = cents[a * L.m + d] - cents[b * L.m + d];
                    s = fmaf(df, df, s);
                }
                s = block_sum(s, sh);
                if (threadIdx.x == 0) dist[npairs] = sqrtf(s);
                ++npairs;
            }
        }
    __syncthreads();
    float sep = 0.f;
    if (npairs > 0) {
        for (int q = 0; q < npairs; ++q) sep += dist[q];
        sep = -sep / (float)npairs;
        int q = 0;
        for (int a = 0; a < L.n_mod; ++a)
            for (int b = a + 1; b < L.n_mod; ++b) {
                if (count[a] > 0.f && count[b] > 0.f) {
                    const float dd = dist[q++];
                    if (dd > 0.f) {   // torch.pdist backward yields 0 at zero distance
                        const float c = -1.f / ((float)npairs * dd);
                        for (int d = threadIdx.x; d < L.m; d += blockDim.x) {
                            const float df = cents[a * L.m + d] - cents[b * L.m + d];
                            dcent[a * L.m + d] += c * df;
                            dcent[b * L.m + d] -= c * df;
                        }
                    }
                }
            }
    }
    // contrastive: mean over rows with a positive of -log(pos/all + 1e-8)
    float ls = 0.f, nv = 0.f;
    for (int i = threadIdx.x; i < L.B; i += blockDim.x) {
        const float p = aux[L.pos() + i];
        if (p > 0.f) { ls += -logf(p / aux[L.all() + i] + 1e-8f); nv += 1.f; }
    }
    ls = block_sum(ls, sh);
    nv = block_sum(nv, sh);
    if (threadIdx.x == 0) {
        aux[0] = sep;
        aux[1] = nv > 0.f ? ls / nv : 0.f;
        aux[2] = nv;
        aux[3] = (float)npairs;
    }
}

// dz[k][s0 + d] = g_sep * dcent[mod(k)][d] / count + (normalize backward of) sum_j w_kj zn_j[d]
__global__ void __launch_bounds__(128) aux_bwd_kernel(const float* __restrict__ z, const long long* __restrict__ idx,
                                                      long long D, int s0, float inv_tau, AuxLayout L,
                                                      const float* __restrict__ aux, const float* __restrict__ g_sep,
                                                      const float* __restrict__ g_con, float* __restrict__ dz) {
    extern __shared__ float smem[];   // zk[m], w[B]
    __shared__ float sh[32];
    float* zk = smem;
    float* w = smem + L.m;
    const int k = blockIdx.x;
    const float inv_k = aux[L.inv() + k];
    for (int d = threadIdx.x; d < L.m; d += blockDim.x) zk[d] = z[(long long)k * D + s0 + d] * inv_k;
    __syncthreads();
    const long long mk = idx[k];
    const float nv = aux[2];
    const float gc = (g_con && nv > 0.f) ? (*g_con) / nv : 0.f;
    const float pos_k = aux[L.pos() + k], all_k = aux[L.all() + k];
    const float coef_k = pos_k > 0.f ? -gc / (pos_k / all_k + 1e-8f) : 0.f;
    for (int j = threadIdx.x; j < L.B; j += blockDim.x) {
        float wj = 0.f;
        if (j != k && gc != 0.f) {
            const float* zj = z + (long long)j * D + s0;
            float dot = 0.f;
            for (int d = 0; d < L.m; ++d) dot = fmaf(zk[d], zj[d], dot);
            const float e = expf(dot * aux[L.inv() + j] * inv_tau);
            const float same = (idx[j] == mk) ? 1.f : 0.f;
            const float pos_j = aux[L.pos() + j], all_j = aux[L.all() + j];
            const float coef_j = pos_j > 0.f ? -gc / (pos_j / all_j + 1e-8f) : 0.f;
            const float a_kj = coef_k * (same / all_k - pos_k / (all_k * all_k));
            const float a_jk = coef_j * (same / all_j - pos_j / (all_j * all_j));
            wj = e * inv_tau * (a_kj + a_jk) * aux[L.inv() + j];   // inv_j folded in: zn_j = inv_j * z_j
        }
        w[j] = wj;
    }
    __syncthreads();
    // dzn[d] for the dims this thread owns (m <= 4 * 128)
    float dzn[4] = {0.f, 0.f, 0.f, 0.f};
    float dotp = 0.f;
    for (int q = 0; q < 4; ++q) {
        const int d = threadIdx.x + q * blockDim.x;
        if (d < L.m) {
            float a = 0.f;
            for (int j = 0; j < L.B; ++j) a = fmaf(w[j], z[(long long)j * D + s0 + d], a);
            dzn[q] = a;
            dotp = fmaf(a, zk[d], dotp);
        }
    }
    dotp = block_sum(dotp, sh);
    const int mc = clamp_mod(mk, L.n_mod);
    const bool in_range = mk >= 0 && mk < L.n_mod;
    const float gs = g_sep ? *g_sep : 0.f;
    for (int q = 0; q < 4; ++q) {
        const int d = threadIdx.x + q * blockDim.x;
        if (d < L.m) {
            float g = inv_k * (dzn[q] - zk[d] * dotp);
            if (in_range && aux[L.count() + mc] > 0.f) g = fmaf(gs, aux[L.dcent() + mc * L.m + d] / aux[L.count() + mc], g);
            dz[(long long)k * D + s0 + d] = g;
        }
    }
}

long long dis_aux_floats(int B, int m, int n_mod) {
    AuxLayout L{B, m, n_mod};
    return L.total();
}

// z: fp32 [B][D] (flattened NCHW latent); aux: dis_aux_floats() floats; aux[0] = separation loss,
// aux[1] = contrastive loss afterwards
int dis_aux_fwd(const float* z, const long long* idx, float* aux, int B, long long D, int s0, int m, int n_mod, float tau,
                cudaStream_t st) {
    MVAE_CHECK(m >= 1 && m <= 512 && s0 >= 0 && s0 + m <= D && n_mod >= 1 && n_mod <= 16 && B >= 1,
               "dis_aux: unsupported sizes (m=%d, n_mod=%d)", m, n_mod);
    AuxLayout L{B, m, n_mod};
    aux_norm_kernel<<<B, 128, 0, st>>>(z, D, s0, L, aux);
    MVAE_LAUNCH_CHECK();
    aux_rows_kernel<<<B, 128, sizeof(float) * m, st>>>(z, idx, D, s0, 1.f / tau, L, aux);
    MVAE_LAUNCH_CHECK();
    aux_final_kernel<<<1, 256, 0, st>>>(z, idx, D, s0, L, aux);
    MVAE_LAUNCH_CHECK();
    return 0;
}
// dz: fp32 [B][D], columns [s0, s0+m) are written (caller zero-fills the rest)
int dis_aux_bwd(const float* z, const long long* idx, const float* aux, const float* g_sep, const float* g_con, float* dz,
                int B, long long D, int s0, int m, int n_mod, float tau, cudaStream_t st) {
    MVAE_CHECK(m >= 1 && m <= 512 && n_mod <= 16, "dis_aux: unsupported sizes");
    AuxLayout L{B, m, n_mod};
    const size_t sh = sizeof(float) * ((size_t)m + B);
    MVAE_CHECK(sh <= 48 * 1024, "dis_aux_bwd: batch %d too large for one CTA's shared memory", B);
    aux_bwd_kernel<<<B, 128, sh, st>>>(z, idx, D, s0, 1.f / tau, L, aux, g_sep, g_con, dz);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ================================================================== guarded total (:527-565)
// terms[4] = recon, kl, separation, contrastive (device scalars, fp32); out[5] = total + the four terms
// with NaN/Inf terms replaced by 0 and a NaN/Inf total by 1e6; mask[4] = 1 where the term (and the
// total) was finite, i.e. where the gradient flows.
__global__ void dis_total_kernel(const float* t0, const float* t1, const float* t2, const float* t3, float w0, float w1,
                                 float w2, float w3, float* __restrict__ out, float* __restrict__ mask) {
    const float w[4] = {w0, w1, w2, w3};
    const float* terms[4] = {t0, t1, t2, t3};
    float t[4], ok[4];
    float total = 0.f;
    for (int i = 0; i < 4; ++i) {
        const float v = *terms[i];
        ok[i] = isfinite(v) ? 1.f : 0.f;
        t[i] = isfinite(v) ? v : 0.f;
        total = fmaf(w[i], t[i], total);
    }
    const bool tok = isfinite(total);
    out[0] = tok ? total : 1e6f;
    for (int i = 0; i < 4; ++i) { out[1 + i] = t[i]; mask[i] = tok ? ok[i] : 0.f; }
}
int dis_total(const float* t0, const float* t1, const float* t2, const float* t3, float w0, float w1, float w2, float w3,
              float* out, float* mask, cudaStream_t st) {
    dis_total_kernel<<<1, 1, 0, st>>>(t0, t1, t2, t3, w0, w1, w2, w3, out, mask);
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
