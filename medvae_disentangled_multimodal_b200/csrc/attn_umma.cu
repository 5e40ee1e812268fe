// tcgen05 batched GEMM with fused softmax epilogues: the tensor-core path of the AttnBlock core
// (encoder_decoder.py:91-103) and its backward.  bf16 operands, fp32 accumulation in TMEM.
//
//   D[b](m, n) = sum_k A[b](m, k) * B[b](n, k)            one CTA per (image, 128-row M tile, N tile)
//
// Operands are read straight out of the NHWC qkv / dout / P tensors by 3-D TMA maps
// {inner, rows, image}; an operand whose reduction index is NOT its contiguous index is fed to the
// MMA as an "MN-major" tile (the same trick the wgrad kernel uses), so no transposes are ever
// materialised:
//     S  = Q K^T        A = Q  (K-major)   B = K  (K-major)   epilogue: row softmax  -> P  (bf16)
//     O  = P V          A = P  (K-major)   B = V  (N-major)
//     dV = P^T dO       A = P  (M-major)   B = dO (N-major)
//     dP = dO V^T       A = dO (K-major)   B = V  (K-major)   epilogue: softmax bwd  -> dS (bf16)
//     dQ = dS K         A = dS (K-major)   B = K  (N-major)
//     dK = dS^T Q       A = dS (M-major)   B = Q  (N-major)
// In the two softmax epilogues the N tile spans the whole row (T <= 256 keys), and because TMEM lane
// == accumulator row, one thread owns one row of scores: max / sum / normalise are plain loops
// over tcgen05.ld chunks, no cross-thread reduction and no T x T fp32 round trip through HBM.
// Out-of-range rows / reduction indices (T = 49, 196 are not tile multiples) are TMA zero fill.
#include "conv_lower.h"
#include "umma.cuh"

namespace mvae {
using namespace sm100;

int encode_tmap_3d_bf16(CUtensorMap* m, const void* base, unsigned long long inner, unsigned long long rows,
                        unsigned long long batch, unsigned long long pitch_elems, unsigned long long batch_elems,
                        unsigned box_inner, unsigned box_rows);   // conv_umma.cu

enum { EPI_STORE = 0, EPI_SOFTMAX = 1, EPI_SOFTMAX_BWD = 2 };

struct alignas(64) BgemmTmaps {
    CUtensorMap a, b;
};
struct BgemmDev {
    int M, N, K;            // per image
    int BN;                 // N tile (multiple of 16, <= 256)
    int m_tiles, n_tiles;
    int a_mmajor, b_nmajor;
    int mode;
    float alpha;
    bf16* out; long long o_pitch, o_batch; int out_cols;   // columns [N, out_cols) are written as zero
    const bf16* P; long long p_pitch, p_batch;             // EPI_SOFTMAX_BWD
    uint32_t idesc, tmem_cols;
};

constexpr int BG_THREADS = 192;
constexpr int BG_STAGES = 3;
constexpr int BG_A_BYTES = 128 * 128;   // 128 m x 64 k bf16 (either major-ness)

__global__ void __launch_bounds__(BG_THREADS)
bgemm_umma_kernel(const __grid_constant__ BgemmTmaps tm, const __grid_constant__ BgemmDev p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int b_bytes = p.BN * 128;
    uint8_t* sA = smem;
    uint8_t* sB = sA + BG_STAGES * BG_A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(sB + BG_STAGES * b_bytes);
    uint64_t* empty = full + BG_STAGES;
    uint64_t* tfull = empty + BG_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    int w = blockIdx.x;
    const int nt = w % p.n_tiles; w /= p.n_tiles;
    const int mt = w % p.m_tiles;
    const int img = w / p.m_tiles;
    const int m0 = mt * 128, n0 = nt * p.BN;
    const int kiters = (p.K + 63) / 64;

    if (warp == 0 && lane == 0) { prefetch_tmap(&tm.a); prefetch_tmap(&tm.b); }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < BG_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
            mbar_init(&tfull[0], 1);
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, p.tmem_cols);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t stage_bytes = (uint32_t)(BG_A_BYTES + b_bytes);

    if (warp == 0) {
        if (elect_one_sync()) {
            int stage = 0; uint32_t phase = 0;
            for (int kb = 0; kb < kiters; ++kb) {
                mbar_wait(&empty[stage], phase ^ 1u);
                mbar_expect_tx(&full[stage], stage_bytes);
                uint8_t* a = sA + stage * BG_A_BYTES;
                uint8_t* b = sB + stage * b_bytes;
                if (!p.a_mmajor) {
                    tma_load_3d(a, &tm.a, &full[stage], kb * 64, m0, img);
                } else {
                    tma_load_3d(a, &tm.a, &full[stage], m0, kb * 64, img);
                    tma_load_3d(a + 8192, &tm.a, &full[stage], m0 + 64, kb * 64, img);
                }
                if (!p.b_nmajor) {
                    tma_load_3d(b, &tm.b, &full[stage], kb * 64, n0, img);
                } else {
                    for (int c = 0; c < p.BN / 64; ++c)
                        tma_load_3d(b + c * 8192, &tm.b, &full[stage], n0 + c * 64, kb * 64, img);
                }
                if (++stage == BG_STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (elect_one_sync()) {
            int stage = 0; uint32_t phase = 0;
            for (int kb = 0; kb < kiters; ++kb) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t a0 = smem_u32(sA + stage * BG_A_BYTES);
                const uint32_t b0 = smem_u32(sB + stage * b_bytes);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint64_t ad = p.a_mmajor ? smem_desc_sw128(a0 + k * 2048, 8192, 1024)
                                                   : smem_desc_sw128(a0 + k * 32, 16, 1024);
                    const uint64_t bd = p.b_nmajor ? smem_desc_sw128(b0 + k * 2048, 8192, 1024)
                                                   : smem_desc_sw128(b0 + k * 32, 16, 1024);
                    umma_bf16(tmem_base, ad, bd, p.idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit(&empty[stage]);
                if (++stage == BG_STAGES) { stage = 0; phase ^= 1u; }
            }
            umma_commit(&tfull[0]);
        }
    } else {
        const int quarter = warp & 3;
        const int m = m0 + quarter * 32 + lane;
        const bool valid = m < p.M;
        mbar_wait(&tfull[0], 0);
        tc_fence_after();
        const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16);
        bf16* orow = p.out + (long long)img * p.o_batch + (long long)m * p.o_pitch;
        if (p.mode == EPI_STORE) {
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int n = n0 + c0 + g * 8;
                        if (n < p.N) {
                            float f[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g * 8 + j]) * p.alpha;
                            Vec8<bf16> o8; o8.pack(f);
                            o8.store(orow + n);
                        }
                    }
                }
            }
        } else if (p.mode == EPI_SOFTMAX) {
            float mx = -INFINITY;
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (c0 + i < p.N) mx = fmaxf(mx, __uint_as_float(v[i]) * p.alpha);
            }
            float sum = 0.f;
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (c0 + i < p.N) sum += __expf(__uint_as_float(v[i]) * p.alpha - mx);
            }
            const float inv = 1.f / sum;
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int n = c0 + g * 8;
                        if (n < p.out_cols) {
                            float f[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                f[j] = (n + j < p.N) ? __expf(__uint_as_float(v[g * 8 + j]) * p.alpha - mx) * inv : 0.f;
                            Vec8<bf16> o8; o8.pack(f);
                            o8.store(orow + n);
                        }
                    }
                }
            }
        } else {
            const bf16* prow = p.P + (long long)img * p.p_batch + (long long)m * p.p_pitch;
            float dot = 0.f;
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int n = c0 + g * 8;
                        if (n < p.out_cols) {   // P's pad columns hold zeros
                            Vec8<bf16> p8; p8.load(prow + n);
                            float pf[8]; p8.unpack(pf);
#pragma unroll
                            for (int j = 0; j < 8; ++j) dot = fmaf(__uint_as_float(v[g * 8 + j]), pf[j], dot);
                        }
                    }
                }
            }
#pragma unroll 1
            for (int c0 = 0; c0 < p.BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t0 + (uint32_t)c0, v);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int n = c0 + g * 8;
                        if (n < p.out_cols) {
                            Vec8<bf16> p8; p8.load(prow + n);
                            float pf[8]; p8.unpack(pf);
                            float f[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                f[j] = (n + j < p.N) ? pf[j] * (__uint_as_float(v[g * 8 + j]) - dot) * p.alpha : 0.f;
                            Vec8<bf16> o8; o8.pack(f);
                            o8.store(orow + n);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

struct BgemmArgsU {
    int batch, M, N, K;
    const bf16* A; long long a_pitch, a_batch; int a_mmajor;   // K-major: [M][K]; M-major: [K][M]
    const bf16* B; long long b_pitch, b_batch; int b_nmajor;   // K-major: [N][K]; N-major: [K][N]
    int mode; float alpha;
    bf16* out; long long o_pitch, o_batch; int out_cols;
    const bf16* P; long long p_pitch, p_batch;
};

static int bgemm_umma(const BgemmArgsU& g, cudaStream_t st) {
    BgemmDev d;
    d.M = g.M; d.N = g.N; d.K = g.K;
    if (g.mode == EPI_STORE) {
        d.BN = (g.N % 256 == 0) ? 256 : (g.N % 128 == 0 ? 128 : 64);
        MVAE_CHECK(g.N % 64 == 0, "bgemm_umma: N=%d must be a multiple of 64", g.N);
        d.n_tiles = g.N / d.BN;
    } else {
        d.BN = (g.N + 15) / 16 * 16;
        MVAE_CHECK(d.BN <= 256, "bgemm_umma: softmax row of %d keys does not fit one N tile", g.N);
        d.n_tiles = 1;
    }
    MVAE_CHECK(!g.b_nmajor || d.BN % 64 == 0, "bgemm_umma: N-major B needs a 64-multiple N tile");
    d.m_tiles = (g.M + 127) / 128;
    d.a_mmajor = g.a_mmajor; d.b_nmajor = g.b_nmajor; d.mode = g.mode; d.alpha = g.alpha;
    d.out = g.out; d.o_pitch = g.o_pitch; d.o_batch = g.o_batch; d.out_cols = g.out_cols;
    d.P = g.P; d.p_pitch = g.p_pitch; d.p_batch = g.p_batch;
    d.idesc = make_idesc_bf16(128, d.BN, g.a_mmajor, g.b_nmajor);
    uint32_t cols = 32;
    while ((int)cols < d.BN) cols <<= 1;
    d.tmem_cols = cols;
    BgemmTmaps tm;
    memset(&tm, 0, sizeof(tm));
    if (!g.a_mmajor) {
        if (encode_tmap_3d_bf16(&tm.a, g.A, g.K, g.M, g.batch, g.a_pitch, g.a_batch, 64, 128)) return 1;
    } else {
        if (encode_tmap_3d_bf16(&tm.a, g.A, g.M, g.K, g.batch, g.a_pitch, g.a_batch, 64, 64)) return 1;
    }
    if (!g.b_nmajor) {
        if (encode_tmap_3d_bf16(&tm.b, g.B, g.K, g.N, g.batch, g.b_pitch, g.b_batch, 64, d.BN)) return 1;
    } else {
        if (encode_tmap_3d_bf16(&tm.b, g.B, g.N, g.K, g.batch, g.b_pitch, g.b_batch, 64, 64)) return 1;
    }
    const int smem = 1024 + BG_STAGES * (BG_A_BYTES + d.BN * 128) + 256;
    static bool attr_set = false;
    if (!attr_set) {
        MVAE_CUDA(cudaFuncSetAttribute(bgemm_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       1024 + BG_STAGES * (BG_A_BYTES + 256 * 128) + 256));
        attr_set = true;
    }
    long long grid = (long long)g.batch * d.m_tiles * d.n_tiles;
    bgemm_umma_kernel<<<(unsigned)grid, BG_THREADS, smem, st>>>(tm, d);
    MVAE_LAUNCH_CHECK();
    return 0;
}

bool attn_umma_ok(int Tn, int C) { return C % 64 == 0 && Tn <= 256 && Tn >= 8; }

static inline int pad8(int t) { return (t + 7) / 8 * 8; }

// qkv [B][T][3C] bf16 -> out [B][T][C]; probs (>= B*T*pad8(T) bf16) keeps P for the backward
int attn_fwd_umma(const bf16* qkv, bf16* out, void* probs, int B, int Tn, int C, cudaStream_t st) {
    const int Tp = pad8(Tn);
    const float scale = 1.0f / sqrtf((float)C);
    bf16* P = (bf16*)probs;
    BgemmArgsU g;
    memset(&g, 0, sizeof(g));
    g.batch = B;
    // P = softmax(scale * Q K^T)
    g.M = Tn; g.N = Tn; g.K = C;
    g.A = qkv; g.a_pitch = 3LL * C; g.a_batch = (long long)Tn * 3 * C; g.a_mmajor = 0;
    g.B = qkv + C; g.b_pitch = 3LL * C; g.b_batch = (long long)Tn * 3 * C; g.b_nmajor = 0;
    g.mode = EPI_SOFTMAX; g.alpha = scale;
    g.out = P; g.o_pitch = Tp; g.o_batch = (long long)Tn * Tp; g.out_cols = Tp;
    if (bgemm_umma(g, st)) return 1;
    // O = P V
    g.M = Tn; g.N = C; g.K = Tn;
    g.A = P; g.a_pitch = Tp; g.a_batch = (long long)Tn * Tp; g.a_mmajor = 0;
    g.B = qkv + 2 * C; g.b_pitch = 3LL * C; g.b_batch = (long long)Tn * 3 * C; g.b_nmajor = 1;
    g.mode = EPI_STORE; g.alpha = 1.f;
    g.out = out; g.o_pitch = C; g.o_batch = (long long)Tn * C; g.out_cols = C;
    return bgemm_umma(g, st);
}

// dqkv [B][T][3C] from dout [B][T][C]; ds: scratch >= B*T*pad8(T) bf16
int attn_bwd_umma(const bf16* qkv, const void* probs, const bf16* dout, bf16* dqkv, void* ds, int B, int Tn, int C,
                  cudaStream_t st) {
    const int Tp = pad8(Tn);
    const float scale = 1.0f / sqrtf((float)C);
    const bf16* P = (const bf16*)probs;
    bf16* dS = (bf16*)ds;
    const long long qs = (long long)Tn * 3 * C, ps = (long long)Tn * Tp, os = (long long)Tn * C;
    BgemmArgsU g;
    memset(&g, 0, sizeof(g));
    g.batch = B;
    // dV = P^T dO
    g.M = Tn; g.N = C; g.K = Tn;
    g.A = P; g.a_pitch = Tp; g.a_batch = ps; g.a_mmajor = 1;
    g.B = dout; g.b_pitch = C; g.b_batch = os; g.b_nmajor = 1;
    g.mode = EPI_STORE; g.alpha = 1.f;
    g.out = dqkv + 2 * C; g.o_pitch = 3LL * C; g.o_batch = qs; g.out_cols = C;
    if (bgemm_umma(g, st)) return 1;
    // dS = P * (dP - rowsum(dP * P)) * scale,  dP = dO V^T
    g.M = Tn; g.N = Tn; g.K = C;
    g.A = dout; g.a_pitch = C; g.a_batch = os; g.a_mmajor = 0;
    g.B = qkv + 2 * C; g.b_pitch = 3LL * C; g.b_batch = qs; g.b_nmajor = 0;
    g.mode = EPI_SOFTMAX_BWD; g.alpha = scale;
    g.out = dS; g.o_pitch = Tp; g.o_batch = ps; g.out_cols = Tp;
    g.P = P; g.p_pitch = Tp; g.p_batch = ps;
    if (bgemm_umma(g, st)) return 1;
    // dQ = dS K
    g.M = Tn; g.N = C; g.K = Tn;
    g.A = dS; g.a_pitch = Tp; g.a_batch = ps; g.a_mmajor = 0;
    g.B = qkv + C; g.b_pitch = 3LL * C; g.b_batch = qs; g.b_nmajor = 1;
    g.mode = EPI_STORE; g.alpha = 1.f;
    g.out = dqkv; g.o_pitch = 3LL * C; g.o_batch = qs; g.out_cols = C;
    g.P = nullptr;
    if (bgemm_umma(g, st)) return 1;
    // dK = dS^T Q
    g.A = dS; g.a_mmajor = 1;
    g.B = qkv; g.b_nmajor = 1;
    g.out = dqkv + C;
    return bgemm_umma(g, st);
}

}  // namespace mvae
