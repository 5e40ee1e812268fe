// Operand packing, wgrad finalisation, bias-gradient column sums, layout conversion and the
// input stage (NCHW -> NHWC + modality-conditioning map).
#include "conv_lower.h"

namespace mvae {

// ------------------------------------------------------------------ weight packing
// nn.Conv2d weight [Cout][Cin][R][R] fp32 -> fprop operand [slab][Ctot][Cin] (rows co_off..)
// and dgrad operand [slab][Cin][Ctot] (cols co_off..).  UP mode folds the nearest-2x upsample
// into 16 slabs of 2x2 taps: Weff[a,b,ui,vi] = sum_{r in R(a,ui)} sum_{s in R(b,vi)} W[r,s].
__device__ __forceinline__ void up_rset(int a, int ui, int& lo, int& hi) {
    // R(0,0)={0}  R(0,1)={1,2}  R(1,0)={0,1}  R(1,1)={2}
    if (a == 0) { lo = ui == 0 ? 0 : 1; hi = ui == 0 ? 0 : 2; }
    else        { lo = ui == 0 ? 0 : 2; hi = ui == 0 ? 1 : 2; }
}

// A CTA takes a 32 (co) x 32 (ci) tile with all R*R taps: the tile's rows are contiguous spans of the OIHW tensor
// (coalesced reads into shared memory), and each slab of both operands is written from there in 64-byte row
// pieces — ci-major for the fprop operand, co-major for the dgrad operand (row pitch 289 and tap stride 9 are odd:
// both transposed reads are bank-conflict free).  8 bytes of HBM traffic per weight, once per optimizer step.
constexpr int PK_T = 32, PK_PITCH = PK_T * 9 + 1;

template <typename T>
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ w, T* __restrict__ wf,
                                                           T* __restrict__ wd, int Cout, int Cin, int R, int mode,
                                                           int nslabs, int Ctot, int co_off) {
    __shared__ float s[PK_T][PK_PITCH];
    const int RR = R * R;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_ci = (Cin + PK_T - 1) / PK_T, tiles_co = (Cout + PK_T - 1) / PK_T;
    for (int tile = blockIdx.x; tile < tiles_ci * tiles_co; tile += gridDim.x) {
        const int co0 = (tile / tiles_ci) * PK_T, ci0 = (tile % tiles_ci) * PK_T;
        const int nco = Cout - co0 < PK_T ? Cout - co0 : PK_T, nci = Cin - ci0 < PK_T ? Cin - ci0 : PK_T;
        for (int r = warp; r < nco; r += 8) {
            const float* src = w + ((long long)(co0 + r) * Cin + ci0) * RR;
            for (int k = lane; k < nci * RR; k += 32) s[r][k] = src[k];
        }
        __syncthreads();
        for (int slab = 0; slab < nslabs; ++slab) {
            int rlo = 0, rhi = 0, slo = 0, shi = 0;
            if (mode == MVAE_CONV_UP) {
                const int vi = slab & 1, ui = (slab >> 1) & 1, b = (slab >> 2) & 1, a = (slab >> 3) & 1;
                up_rset(a, ui, rlo, rhi);
                up_rset(b, vi, slo, shi);
            }
            auto value = [&](int co, int ci) -> float {
                const float* e = &s[co][ci * RR];
                if (mode != MVAE_CONV_UP) return e[slab];
                float v = 0.f;
                for (int r = rlo; r <= rhi; ++r)
                    for (int q = slo; q <= shi; ++q) v += e[r * 3 + q];
                return v;
            };
            if (wf && lane < nci)
                for (int co = warp; co < nco; co += 8)
                    wf[((long long)slab * Ctot + co_off + co0 + co) * Cin + ci0 + lane] = from_f<T>(value(co, lane));
            if (wd && lane < nco)
                for (int ci = warp; ci < nci; ci += 8)
                    wd[((long long)slab * Cin + ci0 + ci) * Ctot + co_off + co0 + lane] = from_f<T>(value(lane, ci));
        }
        __syncthreads();
    }
}

int pack_weights(const ConvGeom& g, const float* w, void* wf, void* wd, int Ctot, int co_off, int dtype,
                 cudaStream_t st) {
    int nslabs = conv_num_slabs(g);
    long long tiles = (long long)((g.Cout + PK_T - 1) / PK_T) * ((g.Cin + PK_T - 1) / PK_T);
    int blocks = (int)(tiles < 148 * 8 ? tiles : 148 * 8);
    if (dtype == MVAE_F32)
        pack_weights_kernel<float><<<blocks, 256, 0, st>>>(w, (float*)wf, (float*)wd, g.Cout, g.Cin, g.R, g.mode,
                                                            nslabs, Ctot, co_off);
    else
        pack_weights_kernel<bf16><<<blocks, 256, 0, st>>>(w, (bf16*)wf, (bf16*)wd, g.Cout, g.Cin, g.R, g.mode,
                                                           nslabs, Ctot, co_off);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ---- all convs of a model in ONE launch (once per optimizer step): a device table of jobs, each job owning a contiguous
// range of the launch's 32 x 32 tiles; a job may write into a window of a wider / zero-padded packed operand (q,k,v fused
// along Cout; conv_in / conv_out padded to one tensor-core tile: cin_pitch / ctot are the operand's dimensions) and also
// refreshes the fp32 bias vector the conv kernels read.
// two consecutive elements in one store (the destination is 4-byte aligned: even column, even row pitch), else two
template <typename T> __device__ __forceinline__ void store2(T* dst, float a, float b);
template <> __device__ __forceinline__ void store2<bf16>(bf16* dst, float a, float b) {
    if ((reinterpret_cast<uintptr_t>(dst) & 3) == 0) *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(a, b);
    else { dst[0] = __float2bfloat16_rn(a); dst[1] = __float2bfloat16_rn(b); }
}
template <> __device__ __forceinline__ void store2<float>(float* dst, float a, float b) {
    if ((reinterpret_cast<uintptr_t>(dst) & 7) == 0) *reinterpret_cast<float2*>(dst) = make_float2(a, b);
    else { dst[0] = a; dst[1] = b; }
}

template <typename T>
__global__ void __launch_bounds__(256) pack_weights_batched_kernel(const mvae_pack_job* __restrict__ jobs, int n_jobs,
                                                                   int total_tiles) {
    __shared__ float s[PK_T][PK_PITCH];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int gt = blockIdx.x; gt < total_tiles; gt += gridDim.x) {
        int lo = 0, hi = n_jobs - 1;
        while (lo < hi) {                       // last job whose first tile is <= gt
            const int mid = (lo + hi + 1) >> 1;
            if (jobs[mid].tile_begin <= gt) lo = mid; else hi = mid - 1;
        }
        const mvae_pack_job j = jobs[lo];
        const int tile = gt - j.tile_begin;
        const int Cout = j.out_c, Cin = j.in_c, RR = j.ksize * j.ksize, Ctot = j.out_c_total, co_off = j.out_c_offset;
        const int CinP = j.in_c_pitch;
        const float* w = j.w_oihw;
        T* wf = reinterpret_cast<T*>(j.w_fprop);
        T* wd = reinterpret_cast<T*>(j.w_dgrad);
        if (tile == 0 && j.bias_dst)
            for (int c = threadIdx.x; c < Cout; c += 256) j.bias_dst[co_off + c] = j.bias_src ? j.bias_src[c] : 0.f;
        const int tiles_ci = (Cin + PK_T - 1) / PK_T;
        const int co0 = (tile / tiles_ci) * PK_T, ci0 = (tile % tiles_ci) * PK_T;
        const int nco = Cout - co0 < PK_T ? Cout - co0 : PK_T, nci = Cin - ci0 < PK_T ? Cin - ci0 : PK_T;
        {
            // all (up to 4 rows x 9) loads of a thread are issued before the first shared-memory store: a loop with a
            // runtime trip count exposed one DRAM round trip per iteration (29 us per tile, profiles/r2k_ncu_full.csv)
            float v[4][9];
            const int n = nci * RR;
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const int r = warp + 8 * rr;
                const float* src = w + ((long long)(co0 + r) * Cin + ci0) * RR;
#pragma unroll
                for (int i = 0; i < 9; ++i) {
                    const int k = lane + 32 * i;
                    v[rr][i] = (r < nco && k < n) ? __ldg(src + k) : 0.f;
                }
            }
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const int r = warp + 8 * rr;
#pragma unroll
                for (int i = 0; i < 9; ++i) {
                    const int k = lane + 32 * i;
                    if (r < nco && k < n) s[r][k] = v[rr][i];
                }
            }
        }
        __syncthreads();
        for (int slab = 0; slab < j.n_slabs; ++slab) {
            int rlo = 0, rhi = 0, slo = 0, shi = 0;
            if (j.mode == MVAE_CONV_UP) {
                const int vi = slab & 1, ui = (slab >> 1) & 1, b = (slab >> 2) & 1, a = (slab >> 3) & 1;
                up_rset(a, ui, rlo, rhi);
                up_rset(b, vi, slo, shi);
            }
            auto value = [&](int co, int ci) -> float {
                const float* e = &s[co][ci * RR];
                if (j.mode != MVAE_CONV_UP) return e[slab];
                float v = 0.f;
                for (int r = rlo; r <= rhi; ++r)
                    for (int q = slo; q <= shi; ++q) v += e[r * 3 + q];
                return v;
            };
            // two adjacent elements per lane, two rows per warp instruction: 4-byte stores (the LSU handles 2-byte lanes
            // at a fraction of its rate: the one-element-per-lane version ran at 1.2 TB/s, profiles/r2k_ncu_full.csv)
            const int half = lane >> 4, e2 = (lane & 15) * 2;
            if (wf)
                for (int co = warp * 2 + half; co < nco; co += 16) {
                    T* dst = wf + ((long long)slab * Ctot + co_off + co0 + co) * CinP + ci0 + e2;
                    if (e2 + 1 < nci) store2<T>(dst, value(co, e2), value(co, e2 + 1));
                    else if (e2 < nci) dst[0] = from_f<T>(value(co, e2));
                }
            if (wd)
                for (int ci = warp * 2 + half; ci < nci; ci += 16) {
                    T* dst = wd + ((long long)slab * CinP + ci0 + ci) * Ctot + co_off + co0 + e2;
                    if (e2 + 1 < nco) store2<T>(dst, value(e2, ci), value(e2 + 1, ci));
                    else if (e2 < nco) dst[0] = from_f<T>(value(e2, ci));
                }
        }
        __syncthreads();
    }
}

int pack_weights_batched(const mvae_pack_job* jobs_dev, int n_jobs, int total_tiles, int dtype, cudaStream_t st) {
    const int blocks = total_tiles < 148 * 8 ? total_tiles : 148 * 8;
    if (dtype == MVAE_F32)
        pack_weights_batched_kernel<float><<<blocks, 256, 0, st>>>(jobs_dev, n_jobs, total_tiles);
    else
        pack_weights_batched_kernel<bf16><<<blocks, 256, 0, st>>>(jobs_dev, n_jobs, total_tiles);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ wgrad finalisation
// partial [splits][nslabs][Cout][Cin] fp32 -> dW [Cout][Cin][R][R] fp32 (sum over splits; UP
// mode: each (r,s) receives the 4 slabs whose R-sets contain it).  dW = beta * dW + sum: beta = 1
// accumulates straight into the training step's flat gradient buffer (no temporary, no add pass).
// A CTA owns 256 consecutive (co, ci) pairs: slab reads are coalesced along ci, the R*R values of a
// pair are transposed through shared memory (stride 9: conflict-free) and the CTA's contiguous
// 256*R*R-float span of OIHW is written with fully coalesced stores.
// RR = taps per (co, ci) pair known at compile time: the RR x splits partial reads of a thread are independent loads
// issued back to back (the first version walked them in a runtime-bounded nested loop, one L2 round trip after the
// other: 22 us per launch for the 14x14 / 7x7 layers of the 28x28 workloads, profiles/r2d_launches_c3.csv).
template <int RR, bool UP>
__global__ void __launch_bounds__(256) wgrad_finalize_kernel(const float* __restrict__ partial, float* __restrict__ dw,
                                                             int splits, int nslabs, int Cout, int Cin, float beta) {
    __shared__ float tile[256 * RR];
    const long long pairs = (long long)Cout * Cin;
    const long long slab_stride = pairs;
    const long long split_stride = slab_stride * nslabs;
    for (long long i0 = blockIdx.x * 256LL; i0 < pairs; i0 += (long long)gridDim.x * 256) {
        const long long i = i0 + threadIdx.x;
        if (i < pairs) {
            float acc[RR];
#pragma unroll
            for (int rs = 0; rs < RR; ++rs) acc[rs] = 0.f;
            if (UP) {
                // tap (r, s) of the 3x3 kernel receives the four parity slabs whose 2x2 footprint contains it
                for (int k = 0; k < splits; ++k) {
                    const float* pk = partial + k * split_stride + i;
                    float v[16];
#pragma unroll
                    for (int sl = 0; sl < 16; ++sl) v[sl] = pk[sl * slab_stride];
#pragma unroll
                    for (int rs = 0; rs < 9; ++rs) {
                        const int r = rs / 3, s_ = rs % 3;
#pragma unroll
                        for (int a = 0; a < 2; ++a) {
                            const int ui = (a == 0) ? (r == 0 ? 0 : 1) : (r == 2 ? 1 : 0);
#pragma unroll
                            for (int b = 0; b < 2; ++b) {
                                const int vi = (b == 0) ? (s_ == 0 ? 0 : 1) : (s_ == 2 ? 1 : 0);
                                acc[rs] += v[((a * 2 + b) * 2 + ui) * 2 + vi];
                            }
                        }
                    }
                }
            } else {
                int k = 0;
                for (; k + 2 <= splits; k += 2) {
                    const float* p0 = partial + k * split_stride + i;
                    const float* p1 = p0 + split_stride;
                    float v0[RR], v1[RR];
#pragma unroll
                    for (int rs = 0; rs < RR; ++rs) { v0[rs] = p0[rs * slab_stride]; v1[rs] = p1[rs * slab_stride]; }
#pragma unroll
                    for (int rs = 0; rs < RR; ++rs) acc[rs] += v0[rs] + v1[rs];
                }
                if (k < splits) {
                    const float* p0 = partial + k * split_stride + i;
#pragma unroll
                    for (int rs = 0; rs < RR; ++rs) acc[rs] += p0[rs * slab_stride];
                }
            }
#pragma unroll
            for (int rs = 0; rs < RR; ++rs) tile[threadIdx.x * RR + rs] = acc[rs];
        }
        __syncthreads();
        const long long span = (pairs - i0 < 256 ? pairs - i0 : 256) * RR;
        float* o = dw + i0 * RR;
        for (int k = threadIdx.x; k < span; k += 256) o[k] = (beta != 0.f) ? fmaf(beta, o[k], tile[k]) : tile[k];
        __syncthreads();
    }
}

int wgrad_finalize(const ConvGeom& g, const float* partial, float* dw, int splits, float beta, cudaStream_t st) {
    long long total = (long long)g.Cout * g.Cin;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    const int ns = conv_num_slabs(g);
    if (g.mode == MVAE_CONV_UP)
        wgrad_finalize_kernel<9, true><<<blocks, 256, 0, st>>>(partial, dw, splits, ns, g.Cout, g.Cin, beta);
    else if (g.R == 3)
        wgrad_finalize_kernel<9, false><<<blocks, 256, 0, st>>>(partial, dw, splits, ns, g.Cout, g.Cin, beta);
    else
        wgrad_finalize_kernel<1, false><<<blocks, 256, 0, st>>>(partial, dw, splits, ns, g.Cout, g.Cin, beta);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ bias gradient: column sums
// db[c] = sum_rows dy[row][c];  dy dense [M][C].  Two stages (deterministic): per-CTA partial
// rows in `scratch` ([gridDim.x][C]), then a fold.
template <typename T>
__global__ void __launch_bounds__(256) colsum_partial_kernel(const T* __restrict__ dy, float* __restrict__ scratch,
                                                             long long M, int C, long long rows_per_cta) {
    __shared__ float sh[256];
    const long long r0 = blockIdx.x * rows_per_cta;
    long long r1 = r0 + rows_per_cta;
    if (r1 > M) r1 = M;
    // threads: tc = channel lane, tr = row lane
    const int tpr = C < 256 ? C : 256;       // threads per row
    const int rpi = 256 / tpr;               // rows per iteration
    const int tc = threadIdx.x % tpr, tr = threadIdx.x / tpr;
    for (int cb = 0; cb < C; cb += tpr) {
        float acc = 0.f;
        int c = cb + tc;
        if (tr < rpi && c < C)
            for (long long r = r0 + tr; r < r1; r += rpi) acc += to_f<T>(dy[r * C + c]);
        sh[threadIdx.x] = (tr < rpi) ? acc : 0.f;
        __syncthreads();
        if (tr == 0 && c < C) {
            float s = 0.f;
            for (int j = 0; j < rpi; ++j) s += sh[j * tpr + tc];
            scratch[(long long)blockIdx.x * C + c] = s;
        }
        __syncthreads();
    }
}
// vectorised variant (C % 8 == 0): 16-byte (bf16) / 32-byte (fp32) loads, 4 independent rows in
// flight per thread; HBM-bound (one read of dy).
template <typename T>
__global__ void __launch_bounds__(256) colsum_partial_vec_kernel(const T* __restrict__ dy, float* __restrict__ scratch,
                                                                 long long M, int C, long long rows_per_cta) {
    __shared__ float sh[256 * 8];
    const long long r0 = blockIdx.x * rows_per_cta;
    long long r1 = r0 + rows_per_cta;
    if (r1 > M) r1 = M;
    const int nvec = C / 8;
    const int tpr = nvec < 256 ? nvec : 256;
    const int rpi = 256 / tpr;
    const int tv = threadIdx.x % tpr, tr = threadIdx.x / tpr;
    for (int vb = 0; vb < nvec; vb += tpr) {
        const int v = vb + tv;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        if (tr < rpi && v < nvec) {
            const T* base = dy + (long long)v * 8;
            long long r = r0 + tr;
            for (; r + 3LL * rpi < r1; r += 4LL * rpi) {
                Vec8<T> a, b, c, d;
                a.load(base + r * C); b.load(base + (r + rpi) * C);
                c.load(base + (r + 2LL * rpi) * C); d.load(base + (r + 3LL * rpi) * C);
                float fa[8], fb[8], fc[8], fd[8];
                a.unpack(fa); b.unpack(fb); c.unpack(fc); d.unpack(fd);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += (fa[j] + fb[j]) + (fc[j] + fd[j]);
            }
            for (; r < r1; r += rpi) {
                Vec8<T> a; a.load(base + r * C);
                float fa[8]; a.unpack(fa);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += fa[j];
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) sh[threadIdx.x * 8 + j] = acc[j];
        __syncthreads();
        if (tr == 0 && v < nvec) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float t = 0.f;
                for (int q = 0; q < rpi; ++q) t += sh[(q * tpr + tv) * 8 + j];
                scratch[(long long)blockIdx.x * C + v * 8 + j] = t;
            }
        }
        __syncthreads();
    }
}
__global__ void colsum_fold_kernel(const float* __restrict__ scratch, float* __restrict__ db, int parts, int C, float beta) {
    __shared__ float sh[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    float s = 0.f;
    if (c < C)
        for (int p = threadIdx.y; p < parts; p += 8) s += scratch[(long long)p * C + c];
    sh[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x];
        db[c] = (beta != 0.f) ? fmaf(beta, db[c], t) : t;
    }
}

int colsum_parts(long long M) {
    long long parts = (M + 127) / 128;
    if (parts > 1184) parts = 1184;   // 8 CTAs per SM
    if (parts < 1) parts = 1;
    return (int)parts;
}

int colsum(const void* dy, float* db, float* scratch, long long M, int C, int dtype, float beta, cudaStream_t st) {
    int parts = colsum_parts(M);
    long long rpc = (M + parts - 1) / parts;
    parts = (int)((M + rpc - 1) / rpc);
    if (C % 8 == 0) {
        if (dtype == MVAE_F32)
            colsum_partial_vec_kernel<float><<<parts, 256, 0, st>>>((const float*)dy, scratch, M, C, rpc);
        else
            colsum_partial_vec_kernel<bf16><<<parts, 256, 0, st>>>((const bf16*)dy, scratch, M, C, rpc);
    } else if (dtype == MVAE_F32)
        colsum_partial_kernel<float><<<parts, 256, 0, st>>>((const float*)dy, scratch, M, C, rpc);
    else
        colsum_partial_kernel<bf16><<<parts, 256, 0, st>>>((const bf16*)dy, scratch, M, C, rpc);
    MVAE_LAUNCH_CHECK();
    colsum_fold_kernel<<<(C + 31) / 32, dim3(32, 8), 0, st>>>(scratch, db, parts, C, beta);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ layout conversion
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, float* __restrict__ dst, int B, int HW, int Cs,
                                    int c_off, int Cd) {
    long long total = (long long)B * Cd * HW;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        int p = (int)(i % HW);
        long long r = i / HW;
        int c = (int)(r % Cd);
        int n = (int)(r / Cd);
        dst[i] = to_f<T>(src[((long long)n * HW + p) * Cs + c_off + c]);
    }
}
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, T* __restrict__ dst, int B, int HW, int C) {
    long long total = (long long)B * C * HW;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % C);
        long long r = i / C;
        int p = (int)(r % HW);
        int n = (int)(r / HW);
        dst[i] = from_f<T>(src[((long long)n * C + c) * HW + p]);
    }
}

static int ew_blocks(long long total) {
    long long b = (total + 255) / 256;
    if (b > 148LL * 32) b = 148LL * 32;
    if (b < 1) b = 1;
    return (int)b;
}

int nhwc_to_nchw(const void* src, float* dst, int B, int HW, int Cs, int c_off, int Cd, int dtype, cudaStream_t st) {
    long long total = (long long)B * Cd * HW;
    if (dtype == MVAE_F32)
        nhwc_to_nchw_kernel<float><<<ew_blocks(total), 256, 0, st>>>((const float*)src, dst, B, HW, Cs, c_off, Cd);
    else
        nhwc_to_nchw_kernel<bf16><<<ew_blocks(total), 256, 0, st>>>((const bf16*)src, dst, B, HW, Cs, c_off, Cd);
    MVAE_LAUNCH_CHECK();
    return 0;
}
int nchw_to_nhwc(const float* src, void* dst, int B, int HW, int C, int dtype, cudaStream_t st) {
    long long total = (long long)B * C * HW;
    if (dtype == MVAE_F32)
        nchw_to_nhwc_kernel<float><<<ew_blocks(total), 256, 0, st>>>(src, (float*)dst, B, HW, C);
    else
        nchw_to_nhwc_kernel<bf16><<<ew_blocks(total), 256, 0, st>>>(src, (bf16*)dst, B, HW, C);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ input stage
// F.interpolate(mode="bilinear", align_corners=False) source coordinates for an 8 -> size resize
__device__ __forceinline__ void bilinear_src(int o, int out_size, int& i0, int& i1, float& l1) {
    float scale = 8.0f / (float)out_size;
    float s = ((float)o + 0.5f) * scale - 0.5f;
    if (s < 0.f) s = 0.f;
    i0 = (int)s;
    if (i0 > 7) i0 = 7;
    i1 = i0 + (i0 < 7 ? 1 : 0);
    l1 = s - (float)i0;
}

// out[n,y,x, 0:C) = x ; out[n,y,x, C:2C) = bilinear(ReLU(W cond_n + b)) ; rest zero
template <typename T>
__global__ void input_pack_kernel(const float* __restrict__ x, const float* __restrict__ cond,
                                  const float* __restrict__ w, const float* __restrict__ b, T* __restrict__ out,
                                  int B, int C, int H, int W, int cond_dim, int out_c) {
    extern __shared__ float cmap[];   // [C][64] low-res condition map of this image
    const int n = blockIdx.y;
    if (w) {
        for (int i = threadIdx.x; i < C * 64; i += blockDim.x) {
            float acc = b[i];
            for (int k = 0; k < cond_dim; ++k) acc = fmaf(w[(long long)i * cond_dim + k], cond[n * cond_dim + k], acc);
            cmap[i] = acc > 0.f ? acc : 0.f;
        }
    }
    __syncthreads();
    const int HW = H * W;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        int yy = p / W, xx = p % W;
        T* o = out + ((long long)n * HW + p) * out_c;
        int y0 = 0, y1 = 0, x0 = 0, x1 = 0; float ly = 0.f, lx = 0.f;
        if (w) {
            bilinear_src(yy, H, y0, y1, ly);
            bilinear_src(xx, W, x0, x1, lx);
        }
        auto value = [&](int c) -> float {
            if (c < C) return x[((long long)n * C + c) * HW + p];
            if (w && c < 2 * C) {
                const float* m = cmap + (c - C) * 64;
                return (1.f - ly) * ((1.f - lx) * m[y0 * 8 + x0] + lx * m[y0 * 8 + x1]) +
                       ly * ((1.f - lx) * m[y1 * 8 + x0] + lx * m[y1 * 8 + x1]);
            }
            return 0.f;
        };
        if (out_c % 8 == 0) {
            for (int c0 = 0; c0 < out_c; c0 += 8) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = (c0 < 2 * C) ? value(c0 + j) : 0.f;
                Vec8<T> v8; v8.pack(f); v8.store(o + c0);
            }
        } else {
            for (int c = 0; c < out_c; ++c) o[c] = from_f<T>(value(c));
        }
    }
}

int input_pack_fwd(const float* x, const float* cond, const float* w, const float* b, void* out, int B, int C,
                   int H, int W, int cond_dim, int out_c, int dtype, cudaStream_t st) {
    int HW = H * W;
    dim3 grid((HW + 255) / 256 < 64 ? (HW + 255) / 256 : 64, B);
    size_t sh = w ? sizeof(float) * C * 64 : 0;
    if (dtype == MVAE_F32)
        input_pack_kernel<float><<<grid, 256, sh, st>>>(x, cond, w, b, (float*)out, B, C, H, W, cond_dim, out_c);
    else
        input_pack_kernel<bf16><<<grid, 256, sh, st>>>(x, cond, w, b, (bf16*)out, B, C, H, W, cond_dim, out_c);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// backward of the condition branch.  One CTA per (channel c, image-chunk); it first reduces
// d_map (H x W) of each image onto the 8x8 grid (transpose of the bilinear resize), masks by
// ReLU, and accumulates dW[c*64+q][k] += g * cond[n][k], db[c*64+q] += g with fp32 atomics.
template <typename T>
__global__ void __launch_bounds__(256) input_pack_bwd_kernel(const T* __restrict__ d_in, const float* __restrict__ cond,
                                                             const float* __restrict__ w, const float* __restrict__ b,
                                                             float* __restrict__ dw, float* __restrict__ db, int B,
                                                             int C, int H, int W, int cond_dim, int out_c) {
    __shared__ float g8[64];
    const int c = blockIdx.x;
    const int HW = H * W;
    for (int n = blockIdx.y; n < B; n += gridDim.y) {
        if (threadIdx.x < 64) g8[threadIdx.x] = 0.f;
        __syncthreads();
        for (int p = threadIdx.x; p < HW; p += blockDim.x) {
            int yy = p / W, xx = p % W;
            float g = to_f<T>(d_in[((long long)n * HW + p) * out_c + C + c]);
            int y0, y1, x0, x1; float ly, lx;
            bilinear_src(yy, H, y0, y1, ly);
            bilinear_src(xx, W, x0, x1, lx);
            atomicAdd(&g8[y0 * 8 + x0], g * (1.f - ly) * (1.f - lx));
            atomicAdd(&g8[y0 * 8 + x1], g * (1.f - ly) * lx);
            atomicAdd(&g8[y1 * 8 + x0], g * ly * (1.f - lx));
            atomicAdd(&g8[y1 * 8 + x1], g * ly * lx);
        }
        __syncthreads();
        if (threadIdx.x < 64) {
            int row = c * 64 + threadIdx.x;
            float pre = b[row];
            for (int k = 0; k < cond_dim; ++k) pre = fmaf(w[(long long)row * cond_dim + k], cond[n * cond_dim + k], pre);
            float g = pre > 0.f ? g8[threadIdx.x] : 0.f;
            if (g != 0.f) {
                atomicAdd(&db[row], g);
                for (int k = 0; k < cond_dim; ++k) {
                    float cv = cond[n * cond_dim + k];
                    if (cv != 0.f) atomicAdd(&dw[(long long)row * cond_dim + k], g * cv);
                }
            }
        }
        __syncthreads();
    }
}

int input_pack_bwd(const void* d_in, const float* cond, const float* w, const float* b, float* dw, float* db,
                   int B, int C, int H, int W, int cond_dim, int out_c, int dtype, cudaStream_t st) {
    MVAE_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)C * 64 * cond_dim, st));
    MVAE_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * (size_t)C * 64, st));
    dim3 grid(C, B < 128 ? B : 128);
    if (dtype == MVAE_F32)
        input_pack_bwd_kernel<float><<<grid, 256, 0, st>>>((const float*)d_in, cond, w, b, dw, db, B, C, H, W,
                                                            cond_dim, out_c);
    else
        input_pack_bwd_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)d_in, cond, w, b, dw, db, B, C, H, W,
                                                           cond_dim, out_c);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ disentangled input projector
template <typename T>
__global__ void dis_input_kernel(const float* __restrict__ x, const long long* __restrict__ idx,
                                 const float* __restrict__ pw, const float* __restrict__ pb,
                                 const int* __restrict__ is_gray, T* __restrict__ out, int B, int xc, int HW,
                                 int n_mod, int out_c) {
    long long total = (long long)B * HW;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        int p = (int)(i % HW);
        int n = (int)(i / HW);
        long long m = idx[n];
        if (m >= n_mod) m = n_mod - 1;
        if (m < 0) m = 0;
        T* o = out + i * out_c;
        if (is_gray[m]) {
            float v = x[((long long)n * xc) * HW + p];
            for (int c = 0; c < 3; ++c) o[c] = from_f<T>(fmaf(pw[m * 3 + c], v, pb[m * 3 + c]));
        } else {
            for (int c = 0; c < 3; ++c) o[c] = from_f<T>(c < xc ? x[((long long)n * xc + c) * HW + p] : 0.f);
        }
        for (int c = 3; c < out_c; ++c) o[c] = from_f<T>(0.f);
    }
}
int dis_input_fwd(const float* x, const long long* idx, const float* pw, const float* pb, const int* is_gray,
                  void* out, int B, int xc, int HW, int n_mod, int out_c, int dtype, cudaStream_t st) {
    long long total = (long long)B * HW;
    if (dtype == MVAE_F32)
        dis_input_kernel<float><<<ew_blocks(total), 256, 0, st>>>(x, idx, pw, pb, is_gray, (float*)out, B, xc, HW, n_mod, out_c);
    else
        dis_input_kernel<bf16><<<ew_blocks(total), 256, 0, st>>>(x, idx, pw, pb, is_gray, (bf16*)out, B, xc, HW, n_mod, out_c);
    MVAE_LAUNCH_CHECK();
    return 0;
}
// d_pw[m][c] += sum_p d_in[n,p,c] * x[n,0,p];  d_pb[m][c] += sum_p d_in[n,p,c]  for gray samples
template <typename T>
__global__ void __launch_bounds__(256) dis_input_bwd_kernel(const T* __restrict__ d_in, const float* __restrict__ x,
                                                            const long long* __restrict__ idx,
                                                            const int* __restrict__ is_gray, float* __restrict__ dpw,
                                                            float* __restrict__ dpb, int B, int xc, int HW, int n_mod,
                                                            int out_c) {
    __shared__ float sh[32];
    const int n = blockIdx.x;
    long long m = idx[n];
    if (m >= n_mod) m = n_mod - 1;
    if (m < 0) m = 0;
    if (!is_gray[m]) return;
    float aw[3] = {0.f, 0.f, 0.f}, ab[3] = {0.f, 0.f, 0.f};
    for (int p = threadIdx.x; p < HW; p += blockDim.x) {
        float v = x[((long long)n * xc) * HW + p];
        const T* g = d_in + ((long long)n * HW + p) * out_c;
        for (int c = 0; c < 3; ++c) {
            float gg = to_f<T>(g[c]);
            aw[c] = fmaf(gg, v, aw[c]);
            ab[c] += gg;
        }
    }
    for (int c = 0; c < 3; ++c) {
        float s = block_sum(aw[c], sh);
        float t = block_sum(ab[c], sh);
        if (threadIdx.x == 0) {
            atomicAdd(&dpw[m * 3 + c], s);
            atomicAdd(&dpb[m * 3 + c], t);
        }
    }
}
int dis_input_bwd(const void* d_in, const float* x, const long long* idx, const int* is_gray, float* dpw,
                  float* dpb, int B, int xc, int HW, int n_mod, int out_c, int dtype, cudaStream_t st) {
    MVAE_CUDA(cudaMemsetAsync(dpw, 0, sizeof(float) * n_mod * 3, st));
    MVAE_CUDA(cudaMemsetAsync(dpb, 0, sizeof(float) * n_mod * 3, st));
    if (dtype == MVAE_F32)
        dis_input_bwd_kernel<float><<<B, 256, 0, st>>>((const float*)d_in, x, idx, is_gray, dpw, dpb, B, xc, HW, n_mod, out_c);
    else
        dis_input_bwd_kernel<bf16><<<B, 256, 0, st>>>((const bf16*)d_in, x, idx, is_gray, dpw, dpb, B, xc, HW, n_mod, out_c);
    MVAE_LAUNCH_CHECK();
    return 0;
}

}  // namespace mvae
