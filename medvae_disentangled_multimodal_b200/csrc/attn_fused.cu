// Fused AttnBlock core (encoder_decoder.py:91-103) on tcgen05 / TMEM: ONE persistent kernel per direction, the scores of
// an image never leave the SM.
//
//   forward   item = (image, 128-query tile):   S = Q K^T  (TMEM)  ->  softmax on TMEM lanes (one thread = one query row)
//                                               ->  P~ = exp2(S' - max) as bf16 pairs written back IN PLACE into the score
//                                                   columns of tensor memory: the A operand of the second GEMM comes
//                                                   from TMEM (tcgen05.mma [d], [a_tmem], b_desc); knob 15 keeps the
//                                                   first version, a swizzled shared-memory tile
//                                               ->  O = (P~ V) / rowsum in 128-channel chunks (TMEM, double buffered)
//             saved for the backward: ONE float per query row, lse2 = max + log2(rowsum)  (no [B][T][T] tensor)
//   backward  recomputes P from lse2.  Two kinds of work item, both (image, 128-row tile):
//     Q-side  rows = queries:  S = Q K^T, dP = dO V^T (TMEM)  ->  dS = P (dP - D) alpha  (TMEM, in place)  ->  dQ = dS K
//     K-side  rows = keys:     S^T = K Q^T, dP^T = V dO^T     ->  P^T, dS^T (TMEM, in place) ->  dV = P^T dO, dK = dS^T Q
//     D[q] = sum_c dO[q][c] O[q][c] comes from a small row-dot kernel launched in front.
//   A key tile needs the contributions of ALL queries (and vice versa), which is why the backward has the two transposed
//   kinds instead of atomics: every output element is written exactly once, deterministically.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM allocation), warps 2-5 = epilogue (TMEM
// lane quarter = warp & 3).  Operands are read in place from the fused NHWC qkv tensor [B][T][3C] and from dO / written to
// dqkv through ONE 3-D tensor map per tensor with a {64 channels, 64 tokens} box: a K-major operand (reduction over
// channels) and an N-major operand (reduction over tokens) are the same box seen through different UMMA descriptors.
// Rows / tokens beyond T are TMA zero fill; the score columns beyond T are masked in the epilogue.
#include <mutex>
#include "../../include/medvae_b200.h"
#include "umma.cuh"

namespace mvae {
using namespace sm100;

int encode_tmap_3d_bf16(CUtensorMap* m, const void* base, unsigned long long inner, unsigned long long rows,
                        unsigned long long batch, unsigned long long pitch_elems, unsigned long long batch_elems,
                        unsigned box_inner, unsigned box_rows);   // conv_umma.cu

enum { AF_FWD = 0, AF_BWD_Q = 1, AF_BWD_K = 2 };

// Three boxes per tensor ([0] = qkv, [1] = dO): the kernel is bound by the rate at which ONE thread can issue TMA
// instructions (~0.28 us each, measured: every shape cost 0.28-0.29 us per {64 x 64} box), so every operand moves in the
// largest box its layout allows and the lanes of the producer warp issue the boxes of a stage in parallel.
struct alignas(64) AfTmaps {
    CUtensorMap a[2];   // {64 channels, 128 tokens}: the A rows of a score GEMM
    CUtensorMap b[2];   // {64 channels, kboxes * 64 tokens}: ALL tokens, the B operand of a score GEMM
    CUtensorMap n[2];   // {64 channels, kg * 64 tokens}: a token group, the N-major B operand of an output GEMM
    CUtensorMap o;      // {64 channels, 32 tokens} of the OUTPUT (out, or dqkv): one epilogue warp's rows, TMA-stored
};
struct AfDev {
    int B, T, C;
    int Tp;            // N of the score GEMMs: T rounded up to 16
    int kboxes;        // 64-token boxes covering T
    int mtiles;        // 128-row tiles per image
    int bwd;
    int stages;
    int bno;           // channel chunk of the output GEMMs (128, or 64 when C is not a multiple of 128)
    int kg;            // 64-token boxes per output-GEMM stage (a divisor of kboxes)
    int n_items;
    int o_alias;       // output accumulators alias the score columns (backward, T > 128)
    uint32_t stage_bytes, tile_bytes;
    uint32_t col_s, col_dp, col_o0, col_o1;
    int ts;            // P / dS go back to TENSOR memory as the A operand of the output GEMMs (no shared-memory tiles)
    uint32_t col_p, col_ds;   // their first (packed bf16 pair) columns
    uint32_t idesc_s, idesc_o;
    float alpha, alpha_log2e;
    bf16* out;         // forward [B][T][C]
    float* lse2;       // [B][T]
    const float* D;    // [B][T]
    bf16* dqkv;        // backward [B][T][3C]
    int dbg;           // knob 19 = 1: block 0's MMA thread prints where its cycles went
    int so;            // 1: output rows leave as TMA stores of swizzled shared-memory boxes (0: no room for them, one row per thread)
};

constexpr int AF_THREADS = 192;
constexpr int AF_MAX_STAGES = 8;

// one GEMM whose reduction runs over channels (scores): A rows r0.. of (map_a, channel offset ca), B all tokens of (map_b, cb)
struct AfScoreJob { int a_dy, ca, b_dy, cb; uint32_t col; };
// one GEMM whose reduction runs over tokens (outputs): A = shared tile `tile`, B = (map, channel offset cb) N-major
struct AfOutJob { int tile, b_dy, cb; int out_off; };

__device__ __forceinline__ void af_jobs(const AfDev& p, int kind, AfScoreJob* sj, int& ns, AfOutJob* oj, int& no) {
    const int C = p.C;
    if (kind == AF_FWD) {
        ns = 1; sj[0] = {0, 0, 0, C, p.col_s};
        no = 1; oj[0] = {0, 0, 2 * C, 0};
    } else if (kind == AF_BWD_Q) {
        ns = 2; sj[0] = {0, 0, 0, C, p.col_s}; sj[1] = {1, 0, 0, 2 * C, p.col_dp};
        no = 1; oj[0] = {1, 0, C, 0};
    } else {
        ns = 2; sj[0] = {0, C, 0, 0, p.col_s}; sj[1] = {0, 2 * C, 1, 0, p.col_dp};
        no = 2; oj[0] = {0, 1, 0, 2 * C}; oj[1] = {1, 0, 0, C};
    }
}
__device__ __forceinline__ void af_item(const AfDev& p, int it, int& kind, int& img, int& r0) {
    if (!p.bwd) {
        kind = AF_FWD;
    } else {
        // the two kinds of one (image, row tile) are items 2k and 2k+1; a K-side item costs ~1.5x a Q-side one, and with
        // an even grid `it & 1 == blockIdx & 1` would give every odd CTA nothing but K-side items.  Odd rounds swap
        // the pair (grid and n_items are even: a permutation within each round).
        it ^= (it / (int)gridDim.x) & 1;
        kind = (it & 1) ? AF_BWD_K : AF_BWD_Q;
        it >>= 1;
    }
    img = it / p.mtiles;
    r0 = (it % p.mtiles) * 128;
}
// 16-byte store of 8 bf16 into a K-major SWIZZLE_128B tile: rows of 128 B (64 elements), 16 KB per 64-column chunk
__device__ __forceinline__ void af_tile_store8(uint8_t* tile, int row, int col, const float* f) {
    Vec8<bf16> v;
    v.pack(f);
    const int chunk = col >> 6, unit = (col & 63) >> 3;
    uint8_t* dst = tile + chunk * 16384 + row * 128 + ((unit ^ (row & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = v.raw;
}

// 32 packed bf16 (16 words: columns col .. col+31 of one row) into the same tile: four 16-byte stores
__device__ __forceinline__ void af_tile_store32(uint8_t* tile, int row, int col, const uint32_t* pk) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int c = col + g * 8;
        const int chunk = c >> 6, unit = (c & 63) >> 3;
        uint8_t* dst = tile + chunk * 16384 + row * 128 + ((unit ^ (row & 7)) << 4);
        *reinterpret_cast<uint4*>(dst) = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
    }
}
// 2^x in one MUFU (flushes results below 2^-126 to zero; the arguments here are <= 0 up to rounding)
__device__ __forceinline__ float af_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// 8 floats -> 4 packed bf16 pairs
__device__ __forceinline__ void af_pack8(uint32_t* dst, const float* f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
        dst[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
}

template <bool TS>
__global__ void __launch_bounds__(AF_THREADS, 1)
attn_fused_kernel(const __grid_constant__ AfTmaps tm, const __grid_constant__ AfDev p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sTile0 = smem;                                         // P~ (forward) / P^T (K-side)
    uint8_t* sTile1 = sTile0 + p.tile_bytes;                        // dS / dS^T (backward only)
    uint8_t* sStage = sTile1 + (p.bwd ? p.tile_bytes : 0u);
    // [4 epilogue warps][2 buffers][32 rows x 128 B]: output boxes on their way to global (1024-byte aligned: swizzled)
    uint8_t* sOut = sStage + (uint32_t)p.stages * p.stage_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(sOut + (p.so ? 32768u : 0u));
    uint64_t* empty = full + AF_MAX_STAGES;
    uint64_t* tfull_s = empty + AF_MAX_STAGES;    // scores (S and dP) complete in TMEM
    uint64_t* pready = tfull_s + 1;               // shared-memory tiles written by the epilogue warps
    uint64_t* tfull_o = pready + 1;               // [2] output accumulator complete
    uint64_t* tempty_o = tfull_o + 2;             // [2] output accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_o + 2);
    float* sStat = reinterpret_cast<float*>(tmem_slot + 4);   // [2][256]: lse2, D per query (K-side items)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if ((smem_u32(smem) & 1023u) != 0u) {   // SWIZZLE_128B tiles need 1024-byte alignment
        if (threadIdx.x == 0) printf("mvae: attn_fused dynamic shared memory is not 1024-byte aligned\n");
        __trap();
    }
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm.a[0]); prefetch_tmap(&tm.b[0]); prefetch_tmap(&tm.n[0]); prefetch_tmap(&tm.o);
        if (p.bwd) { prefetch_tmap(&tm.a[1]); prefetch_tmap(&tm.b[1]); prefetch_tmap(&tm.n[1]); }
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < AF_MAX_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
            mbar_init(tfull_s, 1);
            mbar_init(pready, 4);
            for (int i = 0; i < 2; ++i) { mbar_init(&tfull_o[i], 1); mbar_init(&tempty_o[i], 4); }
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tmem_base != 0u) __trap();          // all 512 columns: the allocation starts at column 0 (the MMA issuer relies on it)
    const int kiters = p.C / 64;            // channel chunks of a score GEMM
    const int nchunks = p.C / p.bno;        // channel chunks of an output GEMM
    const uint32_t score_stage_bytes = 16384u + (uint32_t)p.kboxes * 8192u;
    const uint32_t out_stage_bytes = (uint32_t)p.bno * 128u * (uint32_t)p.kg;
    const int kgroups = p.kboxes / p.kg;

    if (warp == 0) {
        // ===================================================== TMA producer: one elected thread (elect.sync: the assembler
        // then emits each UTMALDG once, without the any-number-of-lanes loop a `lane == 0` region gets)
        if (elect_one_sync()) {
            int stage = 0; uint32_t phase = 0;
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
                int kind, img, r0, ns, no;
                AfScoreJob sj[2]; AfOutJob oj[2];
                af_item(p, it, kind, img, r0);
                af_jobs(p, kind, sj, ns, oj, no);
                for (int j = 0; j < ns; ++j) {
                    for (int kb = 0; kb < kiters; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1u);
                        mbar_expect_tx(&full[stage], score_stage_bytes);
                        uint8_t* a = sStage + (uint32_t)stage * p.stage_bytes;
                        tma_load_3d(a, &tm.a[sj[j].a_dy], &full[stage], sj[j].ca + kb * 64, r0, img);
                        tma_load_3d(a + 16384, &tm.b[sj[j].b_dy], &full[stage], sj[j].cb + kb * 64, 0, img);
                        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
                    }
                }
                for (int j = 0; j < no; ++j) {
                    for (int nc = 0; nc < nchunks; ++nc) {
                        for (int g = 0; g < kgroups; ++g) {
                            mbar_wait(&empty[stage], phase ^ 1u);
                            mbar_expect_tx(&full[stage], out_stage_bytes);
                            uint8_t* b = sStage + (uint32_t)stage * p.stage_bytes;
                            for (int c = 0; c < p.bno / 64; ++c)
                                tma_load_3d(b + c * (p.kg * 8192), &tm.n[oj[j].b_dy], &full[stage],
                                            oj[j].cb + nc * p.bno + c * 64, g * p.kg * 64, img);
                            if (++stage == p.stages) { stage = 0; phase ^= 1u; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one elected thread; accumulator and operand
        // addresses in tensor memory are plain column numbers: the allocation starts at 0)
        if (elect_one_sync()) {
            constexpr uint32_t dhi = smem_desc_hi_sw128(1024);
            int stage = 0; uint32_t phase = 0;
            uint32_t n_local = 0;            // items done by this CTA (parity of tfull_s / pready)
            uint32_t uses[2] = {0, 0};       // uses of each output accumulator so far
            long long w_ls = 0, w_pr = 0, w_te = 0, w_lo = 0, c0 = 0;
            const long long t_begin = clock64();
#define AF_TIMED(acc, stmt) do { if (p.dbg) c0 = clock64(); stmt; if (p.dbg) acc += clock64() - c0; } while (0)
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, ++n_local) {
                int kind, img, r0, ns, no;
                AfScoreJob sj[2]; AfOutJob oj[2];
                af_item(p, it, kind, img, r0);
                af_jobs(p, kind, sj, ns, oj, no);
                if (p.o_alias) {   // the score columns double as output accumulators: the last item's must be drained
                    AF_TIMED(w_te, mbar_wait(&tempty_o[0], (uses[0] & 1u) ^ 1u); mbar_wait(&tempty_o[1], (uses[1] & 1u) ^ 1u));
                    tc_fence_after();
                }
                for (int j = 0; j < ns; ++j) {
                    for (int kb = 0; kb < kiters; ++kb) {
                        AF_TIMED(w_ls, mbar_wait(&full[stage], phase));
                        tc_fence_after();
                        const uint32_t a_lo = smem_desc_lo(smem_u32(sStage + (uint32_t)stage * p.stage_bytes));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_lohi(sj[j].col, a_lo + (uint32_t)(k * 2), dhi, a_lo + (uint32_t)(1024 + k * 2), dhi, p.idesc_s,
                                           (kb | k) != 0 ? 1u : 0u);
                        umma_commit(&empty[stage]);
                        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
                    }
                }
                umma_commit(tfull_s);
                AF_TIMED(w_pr, mbar_wait(pready, n_local & 1u));
                tc_fence_after();
                for (int j = 0; j < no; ++j) {
                    const uint32_t t0 = smem_u32(oj[j].tile ? sTile1 : sTile0);
                    for (int nc = 0; nc < nchunks; ++nc) {
                        const int buf = (int)((uses[0] + uses[1]) & 1u);
                        AF_TIMED(w_te, mbar_wait(&tempty_o[buf], (uses[buf] & 1u) ^ 1u));
                        tc_fence_after();
                        const uint32_t dcol = buf ? p.col_o1 : p.col_o0;
                        for (int g = 0; g < kgroups; ++g) {
                            AF_TIMED(w_lo, mbar_wait(&full[stage], phase));
                            tc_fence_after();
                            const uint32_t b_lo = smem_desc_lo(smem_u32(sStage + (uint32_t)stage * p.stage_bytes), (uint32_t)p.kg * 8192u);
                            const uint32_t pcol = oj[j].tile ? p.col_ds : p.col_p;
                            for (int kc = 0; kc < p.kg; ++kc) {
                                const int kt = g * p.kg + kc;          // 64-token box of the reduction
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const uint32_t bl = b_lo + (uint32_t)(kc * 512 + k * 128);
                                    if constexpr (TS)
                                        umma_bf16_ts(dcol, pcol + (uint32_t)(kt * 32 + k * 8),
                                                     ((uint64_t)dhi << 32) | bl, p.idesc_o, (kt | k) != 0 ? 1u : 0u);
                                    else
                                        umma_bf16_lohi(dcol, smem_desc_lo(t0 + kt * 16384) + (uint32_t)(k * 2), dhi, bl, dhi, p.idesc_o,
                                                       (kt | k) != 0 ? 1u : 0u);
                                }
                            }
                            umma_commit(&empty[stage]);
                            if (++stage == p.stages) { stage = 0; phase ^= 1u; }
                        }
                        umma_commit(&tfull_o[buf]);
                        ++uses[buf];
                    }
                }
            }
#undef AF_TIMED
            if (p.dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x / 2 || blockIdx.x == gridDim.x - 1))
                printf("mvae attn dbg (bwd %d T %d C %d): %u items, %lld clk issuing; waited %lld on score operand boxes, %lld on the "
                       "softmax / dS warps, %lld on drained output accumulators, %lld on output operand boxes\n", p.bwd, p.T, p.C,
                       n_local, clock64() - t_begin, w_ls, w_pr, w_te, w_lo);
        }
    } else {
        // ===================================================== epilogue warps: one thread = one tile row (TMEM lane)
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const int et = threadIdx.x - 64;     // 0..127
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const int cols_tile = p.kboxes * 64;
        uint32_t n_local = 0;
        uint32_t uses[2] = {0, 0};
        uint32_t obuf = 0;               // which of the warp's two output boxes the next TMA store takes
        long long e_ws = 0, e_cs = 0, e_wo = 0, e_co = 0, e0 = 0;   // knob 19: waits / work of the score and the output phases
        const long long e_begin = clock64();
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, ++n_local) {
            int kind, img, r0, ns, no;
            AfScoreJob sj[2]; AfOutJob oj[2];
            af_item(p, it, kind, img, r0);
            af_jobs(p, kind, sj, ns, oj, no);
            const int grow = r0 + row;
            const bool valid = grow < p.T;
            const long long tok0 = (long long)img * p.T;
            float row_lse = 0.f, row_d = 0.f, inv_l = 1.f;
            if (kind == AF_BWD_K) {
                // per-QUERY statistics: the score columns of this kind are queries
                asm volatile("bar.sync 1, 128;" ::: "memory");     // the previous item's readers are done
                for (int q = et; q < cols_tile; q += 128) {
                    sStat[q] = q < p.T ? p.lse2[tok0 + q] : 0.f;
                    sStat[256 + q] = q < p.T ? p.D[tok0 + q] * p.alpha : 0.f;    // D * alpha: the dS loop folds the scale
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            } else if (kind == AF_BWD_Q && valid) {
                row_lse = p.lse2[tok0 + grow];
                row_d = p.D[tok0 + grow];
            }
            if (p.dbg) e0 = clock64();
            mbar_wait(tfull_s, n_local & 1u);
            if (p.dbg) { const long long t = clock64(); e_ws += t - e0; e0 = t; }
            tc_fence_after();
            // P / dS of 32 score columns at a time.  One warp per scheduler does this, so the instruction count per element
            // is what matters (scripts/attn_stall_probe.py): full chunks run without per-element predicates (only the
            // chunk that holds column T selects), exp2 is one MUFU (ex2.approx.ftz), constants are folded (D * alpha).
            const float a2 = p.alpha_log2e, alpha = p.alpha;
            const int T = p.T;
            // a warp whose 32 rows all lie beyond T (the upper warps of a T <= 64 or T = 196 tile) has nothing to compute
            // or store: its rows of P / dS stay whatever tensor memory holds — operand rows only reach their own,
            // never-stored, output rows — and it only keeps the barrier protocol going
            const bool wlive = r0 + quarter * 32 < T;
            if (!wlive) {
            } else if (kind == AF_FWD) {
                float mx = -INFINITY;
#pragma unroll 1
                for (int c0 = 0; c0 < p.Tp; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + p.col_s + (uint32_t)c0, v);
                    tmem_ld_wait();
                    if (c0 + 32 <= T) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c0 + i < T) mx = fmaxf(mx, __uint_as_float(v[i]));
                    }
                }
                mx *= a2;                       // alpha > 0: the maximum commutes with the scale
                float sum = 0.f;
#pragma unroll 1
                for (int c0 = 0; c0 < cols_tile; c0 += 32) {
                    uint32_t v[32], pk[16];
                    const bool live = c0 < p.Tp;
                    if (live) { tmem_ld_32x32(lane_addr + p.col_s + (uint32_t)c0, v); tmem_ld_wait(); }
                    if (live && c0 + 32 <= T) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const __nv_bfloat162 h = __floats2bfloat162_rn(af_ex2(fmaf(__uint_as_float(v[2 * i]), a2, -mx)),
                                                                           af_ex2(fmaf(__uint_as_float(v[2 * i + 1]), a2, -mx)));
                            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                            sum += __low2float(h) + __high2float(h);      // the values the PV GEMM sees
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float e0 = (live && c0 + 2 * i < T) ? af_ex2(fmaf(__uint_as_float(v[2 * i]), a2, -mx)) : 0.f;
                            const float e1 = (live && c0 + 2 * i + 1 < T) ? af_ex2(fmaf(__uint_as_float(v[2 * i + 1]), a2, -mx)) : 0.f;
                            const __nv_bfloat162 h = __floats2bfloat162_rn(e0, e1);
                            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                            sum += __low2float(h) + __high2float(h);
                        }
                    }
                    if constexpr (TS) tmem_st_32x16(lane_addr + p.col_p + (uint32_t)(c0 >> 1), pk);
                    else af_tile_store32(sTile0, row, c0, pk);
                }
                inv_l = 1.f / sum;
                if (valid) p.lse2[tok0 + grow] = mx + log2f(sum);
            } else if (kind == AF_BWD_Q) {
                const float row_da = row_d * alpha;
#pragma unroll 1
                for (int c0 = 0; c0 < cols_tile; c0 += 32) {
                    uint32_t sv[32], dv[32], pk[16];
                    const bool live = c0 < p.Tp;
                    if (live) {
                        tmem_ld_32x32(lane_addr + p.col_s + (uint32_t)c0, sv);
                        tmem_ld_32x32(lane_addr + p.col_dp + (uint32_t)c0, dv);
                        tmem_ld_wait();
                    }
                    if (live && c0 + 32 <= T) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float f0 = af_ex2(fmaf(__uint_as_float(sv[2 * i]), a2, -row_lse)) * fmaf(__uint_as_float(dv[2 * i]), alpha, -row_da);
                            const float f1 = af_ex2(fmaf(__uint_as_float(sv[2 * i + 1]), a2, -row_lse)) * fmaf(__uint_as_float(dv[2 * i + 1]), alpha, -row_da);
                            const __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
                            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            float f[2];
#pragma unroll
                            for (int u = 0; u < 2; ++u)
                                f[u] = (live && c0 + 2 * i + u < T)
                                           ? af_ex2(fmaf(__uint_as_float(sv[2 * i + u]), a2, -row_lse)) * fmaf(__uint_as_float(dv[2 * i + u]), alpha, -row_da)
                                           : 0.f;
                            const __nv_bfloat162 h = __floats2bfloat162_rn(f[0], f[1]);
                            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                        }
                    }
                    if constexpr (TS) tmem_st_32x16(lane_addr + p.col_ds + (uint32_t)(c0 >> 1), pk);
                    else af_tile_store32(sTile1, row, c0, pk);
                }
            } else {
#pragma unroll 1
                for (int c0 = 0; c0 < cols_tile; c0 += 32) {
                    uint32_t sv[32], dv[32], pk[16], pk2[16];
                    const bool live = c0 < p.Tp;
                    if (live) {
                        tmem_ld_32x32(lane_addr + p.col_s + (uint32_t)c0, sv);
                        tmem_ld_32x32(lane_addr + p.col_dp + (uint32_t)c0, dv);
                        tmem_ld_wait();
                    }
                    if (live && c0 + 32 <= T) {
                        const float4* ls4 = reinterpret_cast<const float4*>(sStat + c0);          // lse2 of the queries
                        const float4* da4 = reinterpret_cast<const float4*>(sStat + 256 + c0);    // D * alpha of the queries
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const float4 l4 = ls4[g], d4 = da4[g];
                            const float ls[4] = {l4.x, l4.y, l4.z, l4.w}, da[4] = {d4.x, d4.y, d4.z, d4.w};
                            float pr[4], fd[4];
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                pr[u] = af_ex2(fmaf(__uint_as_float(sv[4 * g + u]), a2, -ls[u]));
                                fd[u] = pr[u] * fmaf(__uint_as_float(dv[4 * g + u]), alpha, -da[u]);
                            }
#pragma unroll
                            for (int u = 0; u < 2; ++u) {
                                const __nv_bfloat162 h = __floats2bfloat162_rn(pr[2 * u], pr[2 * u + 1]);
                                const __nv_bfloat162 h2 = __floats2bfloat162_rn(fd[2 * u], fd[2 * u + 1]);
                                pk[2 * g + u] = *reinterpret_cast<const uint32_t*>(&h);
                                pk2[2 * g + u] = *reinterpret_cast<const uint32_t*>(&h2);
                            }
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            float pr[2], fd[2];
#pragma unroll
                            for (int u = 0; u < 2; ++u) {
                                const int c = c0 + 2 * i + u;      // query index
                                const bool on = live && c < T;
                                pr[u] = on ? af_ex2(fmaf(__uint_as_float(sv[2 * i + u]), a2, -sStat[c])) : 0.f;
                                fd[u] = on ? pr[u] * fmaf(__uint_as_float(dv[2 * i + u]), alpha, -sStat[256 + c]) : 0.f;
                            }
                            const __nv_bfloat162 h = __floats2bfloat162_rn(pr[0], pr[1]);
                            const __nv_bfloat162 h2 = __floats2bfloat162_rn(fd[0], fd[1]);
                            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                            pk2[i] = *reinterpret_cast<const uint32_t*>(&h2);
                        }
                    }
                    if constexpr (TS) {
                        tmem_st_32x16(lane_addr + p.col_p + (uint32_t)(c0 >> 1), pk);
                        tmem_st_32x16(lane_addr + p.col_ds + (uint32_t)(c0 >> 1), pk2);
                    } else {
                        af_tile_store32(sTile0, row, c0, pk);
                        af_tile_store32(sTile1, row, c0, pk2);
                    }
                }
            }
            // the tiles go to the tensor core through the async proxy; the score columns may be overwritten next
            if constexpr (TS) tmem_st_wait();
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(pready);
            if (p.dbg) e_cs += clock64() - e0;

            for (int j = 0; j < no; ++j) {
                bf16* obase;
                long long pitch;
                if (kind == AF_FWD) { obase = p.out; pitch = p.C; }
                else { obase = p.dqkv + oj[j].out_off; pitch = 3LL * p.C; }
                bf16* orow = obase + (tok0 + grow) * pitch;
                for (int nc = 0; nc < nchunks; ++nc) {
                    const int buf = (int)((uses[0] + uses[1]) & 1u);
                    if (p.dbg) e0 = clock64();
                    mbar_wait(&tfull_o[buf], uses[buf] & 1u);
                    if (p.dbg) { const long long t = clock64(); e_wo += t - e0; e0 = t; }
                    tc_fence_after();
                    const uint32_t ocol = lane_addr + (buf ? p.col_o1 : p.col_o0);
                    // A thread owns one accumulator row; stored as it is, every 16-byte store instruction of a warp touches 32
                    // different 128-byte lines (32 LSU wavefronts: the drain was wavefront-bound, ~1 clk per 16 B per SM).  The
                    // warp's 32 rows x 64 columns are written as a SWIZZLE_128B box into shared memory instead (conflict-free
                    // 16-byte stores) and one elected lane hands the box to TMA; two boxes per warp alternate, a box is reused
                    // once the store before last has finished reading (rows beyond T are clipped by the tensor map).
                    const uint32_t so_w = smem_u32(sOut) + (uint32_t)((warp - 2) * 8192 + lane * 128);
                    const int col0 = (kind == AF_FWD ? 0 : (int)oj[j].out_off) + nc * p.bno;
#pragma unroll 1
                    for (int c0 = 0; wlive && c0 < p.bno; c0 += 64) {      // bno is 64 or 128: two loads in flight per wait
                        uint32_t v[64];
                        tmem_ld_32x32(ocol + (uint32_t)c0, v);
                        tmem_ld_32x32(ocol + (uint32_t)(c0 + 32), v + 32);
                        if (p.so) {
                            if (elect_one_sync()) tma_store_wait_read<1>();
                            __syncwarp();
                        }
                        tmem_ld_wait();
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            uint32_t w[4];
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[g * 8 + 2 * u]) * inv_l,
                                                                               __uint_as_float(v[g * 8 + 2 * u + 1]) * inv_l);
                                w[u] = *reinterpret_cast<const uint32_t*>(&h);
                            }
                            if (p.so)
                                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(so_w + obuf * 4096u + (uint32_t)((g ^ (lane & 7)) << 4)),
                                             "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
                            else if (valid)
                                *reinterpret_cast<uint4*>(orow + nc * p.bno + c0 + g * 8) = make_uint4(w[0], w[1], w[2], w[3]);
                        }
                        if (!p.so) continue;
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (elect_one_sync())
                            tma_store_3d(&tm.o, sOut + (warp - 2) * 8192 + obuf * 4096u, col0 + c0, r0 + quarter * 32, img);
                        obuf ^= 1u;
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_o[buf]);
                    ++uses[buf];
                    if (p.dbg) e_co += clock64() - e0;
                }
            }
        }
        if (p.so) {      // the shared-memory boxes must outlive their stores
            if (elect_one_sync()) tma_store_wait_all();
            __syncwarp();
        }
        if (p.dbg && blockIdx.x == 0 && threadIdx.x == 64)
            printf("mvae attn dbg, epilogue warp: %lld clk; waited %lld for scores, %lld computing P / dS, waited %lld for output "
                   "accumulators, %lld draining them\n", clock64() - e_begin, e_ws, e_cs, e_wo, e_co);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// D[row] = sum_c dO[row][c] * O[row][c]  (rows = B*T tokens, one warp per row, 16-byte loads)
__global__ void __launch_bounds__(256) attn_rowdot_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ out,
                                                          float* __restrict__ D, long long rows, int C) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long r = warp0; r < rows; r += nwarps) {
        const bf16* a = dout + r * C;
        const bf16* b = out + r * C;
        float acc = 0.f;
        for (int c = lane * 8; c < C; c += 256) {
            Vec8<bf16> va, vb;
            va.load(a + c); vb.load(b + c);
            float fa[8], fb[8];
            va.unpack(fa); vb.unpack(fb);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc = fmaf(fa[j], fb[j], acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) D[r] = acc;
    }
}

bool attn_fused_ok(int Tn, int C) { return C % 64 == 0 && C >= 64 && Tn >= 1 && Tn <= 256; }

static int af_num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    int r = debug_knob(12);       // SMs left free for a concurrent collective (see conv_umma.cu:num_sms)
    int m = n - (r > 0 ? r : 0);
    return m < 8 ? 8 : m;
}

static int af_launch(const bf16* qkv, const bf16* dout, AfDev& d, cudaStream_t st) {
    const int Tn = d.T, C = d.C;
    d.Tp = (Tn + 15) / 16 * 16;
    d.kboxes = (Tn + 63) / 64;
    d.mtiles = (Tn + 127) / 128;
    d.bno = (C % 128 == 0) ? 128 : 64;
    d.tile_bytes = (uint32_t)d.kboxes * 16384u;
    d.stage_bytes = 16384u + (uint32_t)d.kboxes * 8192u;
    // token boxes per output-GEMM stage: the largest divisor of kboxes whose (bno channels x kg * 64 tokens) fits a stage
    d.kg = 1;
    for (int k = d.kboxes; k >= 1; --k)
        if (d.kboxes % k == 0 && (uint32_t)(d.bno * 128 * k) <= d.stage_bytes) { d.kg = k; break; }
    AfTmaps tm;
    memset(&tm, 0, sizeof(tm));
    for (int t = 0; t < (d.bwd ? 2 : 1); ++t) {
        const bf16* base = t ? dout : qkv;
        const unsigned long long inner = t ? (unsigned long long)C : 3ULL * C;
        const unsigned long long img = (unsigned long long)Tn * inner;
        if (encode_tmap_3d_bf16(&tm.a[t], base, inner, Tn, d.B, inner, img, 64, 128)) return 1;
        if (encode_tmap_3d_bf16(&tm.b[t], base, inner, Tn, d.B, inner, img, 64, (unsigned)(d.kboxes * 64))) return 1;
        if (encode_tmap_3d_bf16(&tm.n[t], base, inner, Tn, d.B, inner, img, 64, (unsigned)(d.kg * 64))) return 1;
    }
    {   // the output: out [B][T][C] (forward) or dqkv [B][T][3C] (backward), boxes of 64 channels x 32 tokens
        const void* obase = d.bwd ? (const void*)d.dqkv : (const void*)d.out;
        const unsigned long long inner = d.bwd ? 3ULL * C : (unsigned long long)C;
        if (encode_tmap_3d_bf16(&tm.o, obase, inner, Tn, d.B, inner, (unsigned long long)Tn * inner, 64, 32)) return 1;
    }
    d.alpha = 1.0f / sqrtf((float)C);                         // int(c) ** (-0.5), encoder_decoder.py:94
    d.alpha_log2e = d.alpha * 1.4426950408889634f;
    d.idesc_s = make_idesc_bf16(128, d.Tp, 0, 0);
    d.idesc_o = make_idesc_bf16(128, d.bno, 0, 1);
    d.col_s = 0;
    d.dbg = debug_knob(19) == 1 ? 1 : 0;
    d.ts = debug_knob(15) == 0 ? 1 : 0;       // knob 15 = 1: the shared-memory-tile variant (A/B: 12.3 vs 11.6 ms backward at C4)
    d.col_p = 0; d.col_ds = 256;
    if (d.ts) {
        // scores at [0, Tp) and [256, 256 + Tp); P / dS are packed IN PLACE into the lower halves of those ranges, the
        // output accumulators take the upper halves (still holding score columns when Tp > 128: o_alias)
        d.tile_bytes = 0;
        d.col_dp = 256;
        if (!d.bwd) { d.col_o0 = 256; d.col_o1 = 384; d.o_alias = 0; }
        else { d.col_o0 = 128; d.col_o1 = 384; d.o_alias = d.Tp > 128 ? 1 : 0; }
    } else if (!d.bwd) {
        d.col_dp = 0; d.col_o0 = 256; d.col_o1 = 384; d.o_alias = 0;
    } else if (d.Tp <= 128) {
        d.col_dp = 128; d.col_o0 = 256; d.col_o1 = 384; d.o_alias = 0;
    } else {
        d.col_dp = 256; d.col_o0 = 0; d.col_o1 = 128; d.o_alias = 1;
    }
    d.n_items = d.B * d.mtiles * (d.bwd ? 2 : 1);
    const long long budget = 232448;    // 227 KB of dynamic shared memory per CTA on sm_100
    long long fixed = (long long)d.tile_bytes * (d.bwd ? 2 : 1) + 2048 /* sStat */ + 256 /* barriers */;
    // sOut (4 warps x 2 boxes x 32 rows x 128 B) unless it would cost the pipeline its second stage (the shared-memory-tile
    // variant's backward at T > 192)
    d.so = (budget - fixed - 32768) / d.stage_bytes >= 2 ? 1 : 0;
    if (d.so) fixed += 32768;
    int stages = (int)((budget - fixed) / d.stage_bytes);
    if (stages > AF_MAX_STAGES) stages = AF_MAX_STAGES;
    MVAE_CHECK(stages >= 2, "attn_fused: T=%d does not leave room for a 2-stage pipeline", Tn);
    d.stages = stages;
    long long smem = fixed + (long long)stages * d.stage_bytes;
    if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM: every CTA allocates all 512 TMEM columns
    static bool attr_set = false;
    if (!attr_set) {
        MVAE_CUDA(cudaFuncSetAttribute(attn_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        MVAE_CUDA(cudaFuncSetAttribute(attn_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        attr_set = true;
    }
    int grid = d.n_items < af_num_sms() ? d.n_items : af_num_sms();
    if (d.bwd && grid > 1) grid &= ~1;      // af_item's pair swap
    if (d.ts) attn_fused_kernel<true><<<grid, AF_THREADS, (size_t)smem, st>>>(tm, d);
    else attn_fused_kernel<false><<<grid, AF_THREADS, (size_t)smem, st>>>(tm, d);
    MVAE_LAUNCH_CHECK();
    return 0;
}

// qkv [B][T][3C] bf16 -> out [B][T][C] bf16, lse2 [B][T] fp32 (log2-domain log-sum-exp of the scaled scores, kept for
// the backward)
int attn_fused_fwd(const bf16* qkv, bf16* out, float* lse2, int B, int Tn, int C, cudaStream_t st) {
    MVAE_CHECK(attn_fused_ok(Tn, C), "attn_fused_fwd: unsupported shape T=%d C=%d (needs C %% 64 == 0, T <= 256)", Tn, C);
    AfDev d;
    memset(&d, 0, sizeof(d));
    d.B = B; d.T = Tn; d.C = C; d.bwd = 0;
    d.out = out; d.lse2 = lse2;
    return af_launch(qkv, nullptr, d, st);
}

// dqkv [B][T][3C] from dout [B][T][C], the forward's out and lse2; dscratch: B*T floats
int attn_fused_bwd(const bf16* qkv, const bf16* out, const float* lse2, const bf16* dout, bf16* dqkv, float* dscratch,
                   int B, int Tn, int C, cudaStream_t st) {
    MVAE_CHECK(attn_fused_ok(Tn, C), "attn_fused_bwd: unsupported shape T=%d C=%d (needs C %% 64 == 0, T <= 256)", Tn, C);
    const long long rows = (long long)B * Tn;
    long long blocks = (rows + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    attn_rowdot_kernel<<<(unsigned)blocks, 256, 0, st>>>(dout, out, dscratch, rows, C);
    MVAE_LAUNCH_CHECK();
    AfDev d;
    memset(&d, 0, sizeof(d));
    d.B = B; d.T = Tn; d.C = C; d.bwd = 1;
    d.lse2 = const_cast<float*>(lse2); d.D = dscratch; d.dqkv = dqkv;
    return af_launch(qkv, dout, d, st);
}

// ------------------------------------------------------------------ no-allocation variant of the backward
// For callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
// it): the row-dot scratch D[B*T] of the backward lives in a file-scope device array of a fixed, modest size instead of a
// caller-provided buffer.  Every GPU that loads the library holds these 4 MiB until exit.
//   * `g_af_static_rowdot` in HOST code is a host-side stand-in for the symbol, not a device address: the address the
//     kernels get comes from cudaGetSymbolAddress, resolved once per device with that device current.
//   * Every call shares the array and it keeps the last call's contents.  Nothing has to start at zero: the row-dot
//     kernel writes D[0, B*T) completely before the fused kernel reads exactly that range on the same stream.
//   * Two calls that overlap on the GPU (different streams, or graph branches without an edge between them) RACE on it.
static __device__ __align__(256) float g_af_static_rowdot[MVAE_ATTN_STATIC_MAX_ROWS];

static int af_static_rowdot(float** out) {
    constexpr int MAX_DEV = 64;
    static float* ptrs[MAX_DEV];
    static std::mutex mu;
    int dev = -1;
    MVAE_CUDA(cudaGetDevice(&dev));
    MVAE_CHECK(dev >= 0 && dev < MAX_DEV, "attn_fused_bwd_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        MVAE_CUDA(cudaGetSymbolAddress(&p, g_af_static_rowdot));
        MVAE_CHECK(p != nullptr, "attn_fused_bwd_static: cudaGetSymbolAddress returned a null address");
        ptrs[dev] = static_cast<float*>(p);
    }
    *out = ptrs[dev];
    return 0;
}

}  // namespace mvae

// include/medvae_b200.h: no counterpart in the reference
extern "C" int mvae_attn_fused_bwd_static(const void* qkv, const void* out, const float* lse2, const void* dout, void* dqkv,
                                          int32_t batch, int32_t tokens, int32_t c, int32_t dtype, void* stream) {
    using namespace mvae;
    MVAE_CHECK(qkv && out && lse2 && dout && dqkv && batch > 0 && tokens > 0 && c > 0, "attn_fused_bwd_static: bad argument");
    MVAE_CHECK(dtype == MVAE_BF16, "attn_fused_bwd_static: bf16 only");
    const long long rows = (long long)batch * tokens;
    MVAE_CHECK(rows <= (long long)MVAE_ATTN_STATIC_MAX_ROWS,
               "attn_fused_bwd_static: batch * tokens = %lld exceeds the static scratch of %d rows "
               "(MVAE_ATTN_STATIC_MAX_ROWS); use mvae_attn_fused_bwd with a caller-provided buffer",
               rows, (int)MVAE_ATTN_STATIC_MAX_ROWS);
    float* scratch = nullptr;
    if (af_static_rowdot(&scratch)) return 1;
    return attn_fused_bwd((const bf16*)qkv, (const bf16*)out, lse2, (const bf16*)dout, (bf16*)dqkv, scratch, batch, tokens, c,
                          (cudaStream_t)stream);
}
