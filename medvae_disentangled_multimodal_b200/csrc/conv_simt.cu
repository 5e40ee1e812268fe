// SIMT (CUDA-core, fp32 accumulate) execution of conv tap programs.
// Used for (1) the fp32 verification mode, (2) the skinny first/last layers whose K or N is
// too small for a tensor-core tile (encoder conv_in: K = 9*Cin <= 54; decoder conv_out:
// N <= 3 — SURVEY.md 7.2), and (3) as an on-device cross-check of the tcgen05 kernels.
#include "conv_lower.h"

namespace mvae {

constexpr int TM = 64, TN = 64, TK = 16;

template <typename T>
struct GatherArgs {
    GatherProblem p;
    const T* in;
    const T* w;          // [slab][N][K]
    const float* bias;   // [N] or null
    const T* res;        // addressed like out, or null
    T* out;
};

// 64 pixels x 64 output channels per CTA, 256 threads, 4x4 micro-tile.
template <typename T>
__global__ void __launch_bounds__(256) conv_gather_simt_kernel(const GatherArgs<T> a) {
    const GatherProblem& p = a.p;
    __shared__ float As[TK][TM + 4];
    __shared__ float Bs[TK][TN + 4];
    __shared__ int rn[TM], ry[TM], rx[TM];

    const int tid = threadIdx.x;
    const long long M = (long long)p.B * p.OH * p.OW;
    const long long m0 = (long long)blockIdx.x * TM;
    const int n0 = blockIdx.y * TN;
    if (tid < TM) {
        long long m = m0 + tid;
        if (m < M) {
            int x = (int)(m % p.OW);
            long long r = m / p.OW;
            ry[tid] = (int)(r % p.OH);
            rn[tid] = (int)(r / p.OH);
            rx[tid] = x;
        } else {
            rn[tid] = -1; ry[tid] = 0; rx[tid] = 0;
        }
    }
    __syncthreads();

    const int ty = tid / 16, tx = tid % 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int lrow = tid / 4;         // 0..63
    const int lk = (tid % 4) * 4;     // 0,4,8,12

    for (int t = 0; t < p.taps.ntaps; ++t) {
        const Tap tp = p.taps.t[t];
        const View v = p.in_views[tp.view];
        // per-thread source pixel for the A rows it loads
        const int n = rn[lrow];
        const int yy = ry[lrow] + tp.dy, xx = rx[lrow] + tp.dx;
        const bool pix_ok = (n >= 0) && yy >= 0 && yy < v.Hs && xx >= 0 && xx < v.Ws;
        const T* src = a.in + v.off + (long long)(n < 0 ? 0 : n) * v.sn + (long long)yy * v.sy + (long long)xx * v.sx;
        const T* wsl = a.w + (long long)tp.slab * p.N * p.K;
        for (int k0 = 0; k0 < p.K; k0 += TK) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int k = k0 + lk + j;
                float av = 0.f;
                if (pix_ok && k < p.K) av = to_f<T>(src[k]);
                As[lk + j][lrow] = av;
                float bv = 0.f;
                int nn = n0 + lrow;
                if (nn < p.N && k < p.K) bv = to_f<T>(wsl[(long long)nn * p.K + k]);
                Bs[lk + j][lrow] = bv;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
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
    }

    const View& ov = p.out_view;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int r = ty * 4 + i;
        int n = rn[r];
        if (n < 0) continue;
        long long base = ov.off + (long long)n * ov.sn + (long long)ry[r] * ov.sy + (long long)rx[r] * ov.sx;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int c = n0 + tx * 4 + j;
            if (c >= p.N) continue;
            float v = acc[i][j];
            if (a.bias) v += a.bias[c];
            if (a.res) v += to_f<T>(a.res[base + c]);
            a.out[base + c] = from_f<T>(v);
        }
    }
}

int simt_gather(const GatherProblem& p, const void* in, const void* w, const float* bias,
                const void* residual, void* out, int dtype, cudaStream_t st) {
    long long M = (long long)p.B * p.OH * p.OW;
    if (M == 0) return 0;
    dim3 grid((unsigned)((M + TM - 1) / TM), (unsigned)((p.N + TN - 1) / TN));
    if (dtype == MVAE_F32) {
        GatherArgs<float> a{p, (const float*)in, (const float*)w, bias, (const float*)residual, (float*)out};
        conv_gather_simt_kernel<float><<<grid, 256, 0, st>>>(a);
    } else {
        GatherArgs<bf16> a{p, (const bf16*)in, (const bf16*)w, bias, (const bf16*)residual, (bf16*)out};
        conv_gather_simt_kernel<bf16><<<grid, 256, 0, st>>>(a);
    }
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------
// wgrad: partial[slab][co][ci] += sum over a pixel range of dy(pix, co) * in(pix + tap, ci)
// grid: (co_tiles*ci_tiles, nslabs, ksplits); fp32 atomics into `partial` (caller zeroes).
template <typename T>
struct ReduceArgs {
    ReduceProblem p;
    const T* in;
    const T* dy;
    float* partial;   // [nslabs][Cout][Cin]
    int ci_tiles;
    long long pix_per_split;
};

template <typename T>
__global__ void __launch_bounds__(256) conv_reduce_simt_kernel(const ReduceArgs<T> a) {
    const ReduceProblem& p = a.p;
    __shared__ float Ds[TK][TM + 4];   // dy  [pix][co]
    __shared__ float Xs[TK][TN + 4];   // in  [pix][ci]
    const int tid = threadIdx.x;
    const int co0 = (blockIdx.x / a.ci_tiles) * TM;
    const int ci0 = (blockIdx.x % a.ci_tiles) * TN;
    const Tap tp = p.taps.t[blockIdx.y];
    const View vin = p.in_views[tp.view];
    const View vdy = p.dy_views[tp.oview];
    const long long M = (long long)p.B * p.GH * p.GW;
    const long long mbeg = (long long)blockIdx.z * a.pix_per_split;
    long long mend = mbeg + a.pix_per_split;
    if (mend > M) mend = M;

    const int ty = tid / 16, tx = tid % 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int lp = tid / 16;          // pixel within chunk 0..15
    const int lc = (tid % 16) * 4;    // channel offset 0..60

    for (long long mb = mbeg; mb < mend; mb += TK) {
        long long m = mb + lp;
        bool ok = m < mend;
        int x = 0, y = 0, n = 0;
        if (ok) {
            x = (int)(m % p.GW);
            long long r = m / p.GW;
            y = (int)(r % p.GH);
            n = (int)(r / p.GH);
        }
        const T* dsrc = a.dy + vdy.off + (long long)n * vdy.sn + (long long)y * vdy.sy + (long long)x * vdy.sx;
        int yy = y + tp.dy, xx = x + tp.dx;
        bool in_ok = ok && yy >= 0 && yy < vin.Hs && xx >= 0 && xx < vin.Ws;
        const T* xsrc = a.in + vin.off + (long long)n * vin.sn + (long long)yy * vin.sy + (long long)xx * vin.sx;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int co = co0 + lc + j, ci = ci0 + lc + j;
            Ds[lp][lc + j] = (ok && co < p.Cout) ? to_f<T>(dsrc[co]) : 0.f;
            Xs[lp][lc + j] = (in_ok && ci < p.Cin) ? to_f<T>(xsrc[ci]) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
            float dv[4], xv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) dv[i] = Ds[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) xv[j] = Xs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(dv[i], xv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* dst = a.partial + (long long)tp.slab * p.Cout * p.Cin;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int co = co0 + ty * 4 + i;
        if (co >= p.Cout) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int ci = ci0 + tx * 4 + j;
            if (ci >= p.Cin) continue;
            atomicAdd(dst + (long long)co * p.Cin + ci, acc[i][j]);
        }
    }
}

int simt_reduce(const ReduceProblem& p, const void* in, const void* dy, float* partial, int dtype,
                cudaStream_t st) {
    const int nslabs = p.taps.ntaps;
    MVAE_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * (size_t)nslabs * p.Cout * p.Cin, st));
    long long M = (long long)p.B * p.GH * p.GW;
    if (M == 0) return 0;
    int co_tiles = (p.Cout + TM - 1) / TM, ci_tiles = (p.Cin + TN - 1) / TN;
    long long base_ctas = (long long)co_tiles * ci_tiles * nslabs;
    long long want = (148LL * 8 + base_ctas - 1) / base_ctas;   // ~8 CTAs per SM
    long long max_splits = (M + 255) / 256;
    long long splits = want < 1 ? 1 : (want > max_splits ? max_splits : want);
    if (splits > 65535) splits = 65535;
    long long pps = (M + splits - 1) / splits;
    pps = (pps + TK - 1) / TK * TK;
    splits = (M + pps - 1) / pps;
    dim3 grid((unsigned)(co_tiles * ci_tiles), (unsigned)nslabs, (unsigned)splits);
    if (dtype == MVAE_F32) {
        ReduceArgs<float> a{p, (const float*)in, (const float*)dy, partial, ci_tiles, pps};
        conv_reduce_simt_kernel<float><<<grid, 256, 0, st>>>(a);
    } else {
        ReduceArgs<bf16> a{p, (const bf16*)in, (const bf16*)dy, partial, ci_tiles, pps};
        conv_reduce_simt_kernel<bf16><<<grid, 256, 0, st>>>(a);
    }
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
