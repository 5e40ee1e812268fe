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
