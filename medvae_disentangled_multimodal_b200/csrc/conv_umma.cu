// tcgen05 / TMEM / TMA implicit-GEMM execution of conv tap programs (bf16 in, fp32 accumulate).
//
//  gather kernel (fprop and dgrad):  D[128 pixels, BN channels] = sum_taps A_tap[128, K] * W_tap[BN, K]^T
//      A tiles: TMA 4-D boxes {64 ch, bw, bh, bn} over an NHWC view, shifted per tap; the conv's
//      zero padding (and the Downsample pad) is TMA out-of-bounds zero fill.  K-major, 128B swizzle.
//      B tiles: TMA 3-D boxes {64 ch, BN, 1 slab} of the packed weights.  K-major, 128B swizzle.
//      warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM alloc) — one thread each, chosen with elect.sync so that
//      the assembler emits every UTMALDG / UTCHMMA once (umma.cuh) — warps 2-5 = epilogue
//      (tcgen05.ld -> +bias (+residual) -> bf16 -> global, 32 bytes per access).  Persistent CTAs, 2 TMEM accumulator
//      stages so the epilogue of tile i overlaps the MMAs of tile i+1.
//      Halo variant (HL = 1, 128-channel stride-1 3x3 layers): 8 x 16 pixel tiles, ONE box {64 ch, 16, 18} per tile and
//      64-channel chunk, the nine taps are nine descriptor views of it; weights through their own ring (see HaloCfg).
//      Epilogue variants (template EPI): the GroupNorm either side of the conv rides here —
//      EPI_GN_STATS accumulates the consumer GroupNorm's per-(sample, group) sums of the output,
//      EPI_GN_BWD turns dL/d[silu(GN(x))] into dL/du and accumulates the GroupNorm-backward group
//      sums (warps 2..: CS = 2 or 4 warps per TMEM lane quarter, one column strip each; x / residual
//      strips and the per-channel constants fetched before the accumulator is ready).
//
//  reduce kernel (wgrad):  D[128 co, BN ci] (per slab, per K-split) = sum_pixels dy[pix, co] * x[pix+tap, ci]
//      Both operands are "MN-major" (the channel index is contiguous in NHWC), loaded by the same
//      4-D boxes; 128 pixels per pipeline stage, 8 MMAs (K=16 pixels) per stage.
#include <type_traits>
#include "conv_lower.h"
#include "umma.cuh"

namespace mvae {
using namespace sm100;

// ------------------------------------------------------------------ tensor-map creation (host)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = (EncodeTiledFn)p;
    }
    return fn;
}

static int encode_tmap(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims,
                       const cuuint64_t* strides_bytes, const cuuint32_t* box) {
    EncodeTiledFn fn = get_encode_fn();
    MVAE_CHECK(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims,
                    strides_bytes, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MVAE_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu %llu %llu)",
               (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
               (unsigned long long)dims[2]);
    return 0;
}

// 4-D map {C, Ws, Hs, B} over a strided NHWC view; box {64, bw, bh, bn}
static int view_tmap(CUtensorMap* m, const bf16* base, const View& v, int C, int B, int bw, int bh, int bn) {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)v.Ws, (cuuint64_t)v.Hs, (cuuint64_t)B};
    cuuint64_t st[3] = {(cuuint64_t)v.sx * 2, (cuuint64_t)v.sy * 2, (cuuint64_t)v.sn * 2};
    cuuint32_t box[4] = {64u, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
    return encode_tmap(m, base + v.off, 4, dims, st, box);
}

// 3-D map {inner, rows, batch} over a pitched bf16 matrix stack (attn_umma.cu)
int encode_tmap_3d_bf16(CUtensorMap* m, const void* base, unsigned long long inner, unsigned long long rows,
                        unsigned long long batch, unsigned long long pitch_elems, unsigned long long batch_elems,
                        unsigned box_inner, unsigned box_rows) {
    cuuint64_t dims[3] = {inner, rows, batch};
    cuuint64_t st[2] = {pitch_elems * 2, batch_elems * 2};
    cuuint32_t box[3] = {box_inner, box_rows, 1u};
    return encode_tmap(m, base, 3, dims, st, box);
}

// pixel-box geometry: 128 GEMM rows = bw x bh x bn pixels of the iteration grid
struct BoxGeom {
    int bw, bh, bn, tiles_x, tiles_y, tiles_n, rows;
};
static BoxGeom box_geom(int B, int GH, int GW, int budget = 128) {
    BoxGeom g;
    g.bw = GW < budget ? GW : budget;
    int rem = budget / g.bw;
    g.bh = GH < rem ? GH : rem;
    if (g.bh < 1) g.bh = 1;
    rem = budget / (g.bw * g.bh);
    g.bn = B < rem ? B : rem;
    if (g.bn < 1) g.bn = 1;
    g.tiles_x = (GW + g.bw - 1) / g.bw;
    g.tiles_y = (GH + g.bh - 1) / g.bh;
    g.tiles_n = (B + g.bn - 1) / g.bn;
    g.rows = g.bw * g.bh * g.bn;
    return g;
}

static int num_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    // knob 12: SMs left free for a concurrent collective (NCCL's CTAs cannot start while a persistent one-CTA-per-SM
    // kernel holds every SM, and vice versa: data-parallel runs reserve a few, train.TrainStep sets the knob)
    int r = debug_knob(12);
    int m = n - (r > 0 ? r : 0);
    if (m < 8) m = 8;
    return m & ~1;
}

// ================================================================== gather (fprop / dgrad)
struct alignas(64) GatherTmaps {
    CUtensorMap a[MAX_VIEWS];
    CUtensorMap b;
    CUtensorMap h;   // halo variant: box {64 ch, 16, 18, 1} over view 0
};
struct GatherDev {
    int bw, bh, bn, rows;
    int OW, OH, B;
    int tiles_x, tiles_y, tiles_n, n_tiles, kchunks, total_tiles;
    int N;
    long long o_off, o_sn, o_sy, o_sx;
    const float* bias;
    const bf16* res;
    bf16* out;
    TapProg taps;
    // GroupNorm fusion (GnHook, conv_lower.h); which one is compiled in is the EPI template argument
    int gn_G, gn_cpg_log2, gn_silu;
    float* gn_sums;
    const bf16* gn_x;
    const float* gn_stat;
    const float* gn_gamma;
    const float* gn_beta;
    int halo_bo;   // halo variant: 1 = put (start >> 7) & 7 into the descriptor's base-offset field (knob 18)
    int dbg;       // knob 19 = 1: block 0's MMA thread prints where its cycles went (waits on loads / on the epilogue)
};

constexpr int G_THREADS = 192;
enum { EPI_PLAIN = 0, EPI_GN_STATS = 1, EPI_GN_BWD = 2 };
// The GroupNorm epilogues do 8-18 instructions and a 2-byte global read per element: eight epilogue
// warps (two per TMEM lane quarter, each taking half of the tile's columns) keep them hidden under
// the MMAs of the next tile; the plain epilogue keeps four.
// CS = column strips of the tile = epilogue warps per TMEM lane quarter.  The GroupNorm-backward epilogue (a tanh
// and ~18 dependent instructions per element) is latency-bound with two warps per scheduler; where the MMAs of a
// tile are too short to cover it (K = 9 x 128) CS = 4, i.e. 16 epilogue warps, hides it.
__host__ __device__ constexpr int gather_csplit(int epi) { return epi == EPI_PLAIN ? 1 : 2; }
__host__ __device__ constexpr int gather_threads_cs(int cs) { return 64 + 128 * cs; }

// Sum N per-lane values over the 32 lanes of a warp with a transposing butterfly (N-1 + log2(32/N)
// shuffles instead of 5N): afterwards lane l holds in v[0] the total of value index l / (32/N).
template <int N>
struct WarpReduceScatter {
    static __device__ __forceinline__ void run(float* v, int lane, int off) {
        constexpr int H = N / 2;
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < H; ++i) {
            const float send = up ? v[i] : v[i + H];
            const float keep = up ? v[i + H] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
        WarpReduceScatter<H>::run(v, lane, off >> 1);
    }
};
template <>
struct WarpReduceScatter<1> {
    static __device__ __forceinline__ void run(float* v, int lane, int off) {
        for (; off > 0; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
};

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}

// sa / sb: the two per-row partial sums of the eight 4-channel sub-blocks of one 32-channel chunk.
// NG = GroupNorm groups inside the chunk.  Rows of one warp that all belong to image n_ref are
// summed in registers first (one atomic per value per warp); otherwise every valid row adds its own.
// dst -> element (n, first group of the chunk, 0) of the [B][G][2] accumulator, for n = n_ref
// (uniform) or the thread's own image.
template <int NG>
__device__ __forceinline__ void gn_flush(const float* sa, const float* sb, float* dst, int lane, bool uniform,
                                         bool valid, float scale) {
    constexpr int PER = 8 / NG;
    float v[2 * NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int i = 0; i < PER; ++i) { a += sa[g * PER + i]; b += sb[g * PER + i]; }
        v[2 * g] = a; v[2 * g + 1] = b;
    }
    if (uniform) {
        WarpReduceScatter<2 * NG>::run(v, lane, 16);
        constexpr int LOW = 32 / (2 * NG);
        if ((lane & (LOW - 1)) == 0) atomicAdd(dst + lane / LOW, v[0] * scale);
    } else if (valid) {
#pragma unroll
        for (int i = 0; i < 2 * NG; ++i) atomicAdd(dst + i, v[i] * scale);
    }
}
__device__ __forceinline__ void gn_flush_any(int cpg_log2, const float* sa, const float* sb, float* dst, int lane,
                                             bool uniform, bool valid, float scale = 1.f) {
    if (cpg_log2 == 2) gn_flush<8>(sa, sb, dst, lane, uniform, valid, scale);
    else if (cpg_log2 == 3) gn_flush<4>(sa, sb, dst, lane, uniform, valid, scale);
    else if (cpg_log2 == 4) gn_flush<2>(sa, sb, dst, lane, uniform, valid, scale);
    else gn_flush<1>(sa, sb, dst, lane, uniform, valid, scale);
}

// MT pixel tiles (128 rows each) share every weight stage: a CTA tile is (MT x 128) pixels x BN channels.
// What bounds this kernel is the bytes an SM pulls from L2 per MMA cycle: 128 x 128 tiles need
// (16 + 16) KB per 256 cycles = 128 B/clk, twice what 128 x 256 tiles need per FLOP of B traffic; pairing two
// pixel tiles on one 128-channel weight stage brings N = 128 layers to (32 + 16) KB per 512 cycles = 96 B/clk,
// the same as the 128 x 256 configuration.
constexpr int A_TILE_BYTES = 128 * 128;                 // 128 rows x 64 bf16
// CG = 2: a CTA PAIR (cluster of two, cta_group::2) runs one 256-pixel x BN MMA; each CTA holds its own 128 pixel
// rows and HALF of the weight tile, so a 128 x 256 accumulator per SM costs (16 + 16) KB per 512 MMA cycles =
// 64 B/clk of ingest instead of 96, with the accumulators still double-buffered.
template <int BN, int MT, int CG = 1>
struct GatherCfg {
    static constexpr int A_BYTES = MT * A_TILE_BYTES;
    static constexpr int B_BYTES = BN * 128 / CG;
    // as many stages as fit 200 KB (the GroupNorm tables, barriers and alignment slack take the rest of the 227 KB), <= 8:
    // 64 KB stages -> 3, 48 KB -> 4, 40 KB -> 5, 32 KB -> 6, 24 KB -> 8
    static constexpr int STAGES = (204800 / (A_BYTES + B_BYTES)) > 8 ? 8 : (204800 / (A_BYTES + B_BYTES));
    // accumulators are double-buffered (the epilogue of one CTA tile runs under the MMAs of the next) while they fit
    // the 512 TMEM columns; the 256 x 256 tile (MT = 2, BN = 256) uses all 512 for one buffer and pays an exposed
    // epilogue for halving the bytes pulled per FLOP (64 KB per 1024 MMA cycles = 64 B/clk, the SM's ingest rate)
    static constexpr int ACC_BUFS = (2 * MT * BN <= 512) ? 2 : 1;
    static constexpr int TMEM_COLS = (ACC_BUFS * MT * BN < 32) ? 32 : ACC_BUFS * MT * BN;   // powers of two
    static constexpr int SMEM = 1024 + STAGES * (A_BYTES + B_BYTES) + 256;
};
// EPI_GN_BWD keeps two small tables behind the pipeline stages: {gamma/2, beta/2} of every output channel (filled
// once per CTA) and, per epilogue warp and pixel tile, {rstd, -mean*rstd} of each 4-channel sub-block of the warp's
// column strip (refilled per tile, before the accumulator wait) — the epilogue then reads its per-channel constants
// with shared-memory broadcasts at compile-time offsets instead of dependent global loads in the inner loop.
constexpr int GN_COEF_MAX = 1024;                                   // channels
constexpr int GN_BWD_SMEM = GN_COEF_MAX * 8 + 8 * 2 * 32 * 8;       // coefficient table + per-warp statistics slots
__host__ __device__ constexpr int gather_smem(int base, int epi) { return base + (epi == EPI_GN_BWD ? GN_BWD_SMEM : 0); }

// HALO variant (HL = 1; stride-1 3x3 convs, output grid a multiple of 8 x 16).  Everything above pulls each input
// pixel from L2 once per TAP, and what bounds these kernels is L2 -> SM throughput (~6300 B/clk chip-wide = 42.6
// B/clk/SM: every gather launch in profiles/r2k_ncu_full.csv sits at 11-13 TB/s of l1tex__m_xbar2l1tex_read_bytes).
// Here a pixel tile is 8 wide x 16 tall and ONE TMA box {64 ch, W, 18} brings its (1-pixel) halo per 64-channel
// chunk; the nine taps are nine views of that box: descriptor start = ((dy+1) * W + (dx+1)) * 128 B into it, stride
// between 8-row groups (= 8 pixels of one image row) = the box pitch W * 128 B.  The 128B swizzle is a function
// of the absolute shared-memory address bits, so a start that is not 1024-aligned reads what TMA wrote.  Activation
// bytes per pixel tile and chunk: 22.5 KB instead of 9 x 16 KB; the weights still stream per (chunk, tap), through their
// own ring.  The k loop runs chunk-major (chunk outer, tap inner).
// Box = the 8 x 16 pixel tile plus one pixel all round: 10 x 18 pixels, dense in shared memory (pitch 10 pixels = 1280 B:
// the stride between 8-row groups need not be a multiple of the 1024-byte swizzle atom, for the same reason the start
// need not be).  The first version padded the pitch to 16 pixels (36 KB per box); 22.5 KB leave room for a weight ring of
// seven stages instead of four, and the issuing thread waited on weight boxes for 15 % of its cycles.
constexpr int HALO_TW = 8, HALO_TH = 16;                 // pixel tile
constexpr int HALO_W = HALO_TW + 2, HALO_H = HALO_TH + 2;
constexpr int HALO_BOX_BYTES = HALO_W * HALO_H * 128;    // 23040: what one TMA box delivers
constexpr int HALO_TILE_BYTES = (HALO_BOX_BYTES + 1023) / 1024 * 1024;   // its 1024-byte aligned slot
template <int BN, int MT>
struct HaloCfg {
    static constexpr int A_STAGE = MT * HALO_TILE_BYTES;
    static constexpr int B_BYTES = BN * 128;
    static constexpr int AS = 2;
    static constexpr int BS_RAW = (212992 - AS * A_STAGE) / B_BYTES;
    static constexpr int BS = BS_RAW > 8 ? 8 : BS_RAW;
    static constexpr int SMEM = 1024 + AS * A_STAGE + BS * B_BYTES + 256;
};

template <int BN, int EPI, int MT, int CG, int CS, int HL = 0>
__global__ void __launch_bounds__(gather_threads_cs(CS), 1)
conv_gather_umma_kernel(const __grid_constant__ GatherTmaps tm, const __grid_constant__ GatherDev p) {
    using Cfg = GatherCfg<BN, MT, CG>;
    using HCfg = HaloCfg<BN, MT>;
    static_assert(HL == 0 || CG == 1, "the halo variant is single-CTA");
    // a scheduled tile is CG * MT pixel tiles: MT per CTA, the CTAs of a pair side by side (rank * MT + j)
    // CTA pair: rank within the pair, pair index and number of pairs replace blockIdx / gridDim in the tile loop
    const uint32_t rank = (CG == 2) ? cluster_ctarank() : 0u;
    const int sched_id = (CG == 2) ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int sched_n = (CG == 2) ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    // halo variant: ring of AS halo stages (MT boxes each), ring of BS weight stages; `full` / `empty` = fullA, emptyA
    // (AS each) followed by fullB, emptyB (BS each)
    constexpr int PIPE_BYTES = HL ? (HCfg::AS * HCfg::A_STAGE + HCfg::BS * HCfg::B_BYTES) : STAGES * (Cfg::A_BYTES + Cfg::B_BYTES);
    constexpr int NBAR = HL ? (HCfg::AS + HCfg::BS) : STAGES;
    uint8_t* sA = smem;
    uint8_t* sB = sA + (HL ? HCfg::AS * HCfg::A_STAGE : STAGES * Cfg::A_BYTES);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + PIPE_BYTES);
    uint64_t* empty = full + NBAR;
    uint64_t* tfull = empty + NBAR;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    static_assert((2 * NBAR + 4) * 8 + 4 <= 256, "barrier block");
    float2* sCoef = reinterpret_cast<float2*>(smem + PIPE_BYTES + 256);   // EPI_GN_BWD only
    float2* sStat = sCoef + GN_COEF_MAX;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (EPI == EPI_GN_BWD) {
        for (int c = threadIdx.x; c < p.N; c += gather_threads_cs(CS))
            sCoef[c] = make_float2(0.5f * __ldg(p.gn_gamma + c), 0.5f * __ldg(p.gn_beta + c));
    }
    if (!HL && p.rows < 128) {
        // rows the TMA box never writes feed accumulator rows nobody stores; keep them finite
        for (int i = threadIdx.x; i < STAGES * Cfg::A_BYTES / 16; i += gather_threads_cs(CS))
            reinterpret_cast<uint4*>(sA)[i] = make_uint4(0, 0, 0, 0);
        fence_proxy_async_smem();
    }
    if (warp == 0 && lane == 0) {
        if (HL) prefetch_tmap(&tm.h);
        else for (int v = 0; v < MAX_VIEWS; ++v) prefetch_tmap(&tm.a[v]);
        prefetch_tmap(&tm.b);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < NBAR; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
            // pair: the epilogue warps of BOTH CTAs release the leader's accumulator barrier
            for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], CG * 32 * (4 * CS)); }
            fence_barrier_init();
        }
        __syncwarp();
        if (CG == 2) tmem_alloc2(tmem_slot, Cfg::TMEM_COLS);
        else tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();   // the peer's barriers exist before anything arrives on them remotely
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // An allocation of all 512 columns can only start at column 0: the MMA issuer then forms its accumulator addresses
    // from loop counters alone (uniform registers; a value read from shared memory costs an ELECT / R2UR.BROADCAST
    // loop around every tcgen05.mma)
    if (Cfg::TMEM_COLS == 512 && tmem_base != 0u) __trap();
    const uint32_t tmem_mma = (Cfg::TMEM_COLS == 512) ? 0u : tmem_base;

    const int kiters = p.taps.ntaps * p.kchunks;
    // bytes that complete on one "full" barrier: both CTAs' loads land on the leader's barrier in a pair
    const uint32_t stage_bytes = (uint32_t)CG * ((uint32_t)(MT * p.rows) * 128u + (uint32_t)Cfg::B_BYTES);

    if (HL && warp == 0) {
        if (elect_one_sync()) {
            // one issuing thread, two rings: whichever has a free slot goes next (neither cursor ever blocks the other)
            uint64_t* fullA = full; uint64_t* emptyA = empty;
            uint64_t* fullB = full + HCfg::AS; uint64_t* emptyB = empty + HCfg::AS;
            int ta = sched_id, kca = 0, sa = 0; uint32_t pa = 0;              // halo cursor: (tile, chunk)
            int tb = sched_id, kcb = 0, tpb = 0, sb = 0; uint32_t pb = 0;     // weight cursor: (tile, chunk, tap)
            bool a_live = ta < p.total_tiles, b_live = a_live;
            long long t_idle = clock64();   // bounded like mbar_wait: a protocol bug traps instead of hanging the GPU
            while (a_live || b_live) {
                if (clock64() - t_idle > 4000000000LL) {
                    printf("mvae: halo producer stalled (block %d)\n", (int)blockIdx.x);
                    __trap();
                }
                if (a_live && mbar_try_wait(&emptyA[sa], pa ^ 1u)) {
                    t_idle = clock64();
                    mbar_expect_tx(&fullA[sa], (uint32_t)(MT * HALO_BOX_BYTES));
#pragma unroll
                    for (int j = 0; j < MT; ++j) {
                        int mt = (ta / p.n_tiles) * MT + j;   // one past the last: n0 >= B, the box is zero-filled
                        const int tx = mt % p.tiles_x; mt /= p.tiles_x;
                        tma_load_4d(sA + sa * HCfg::A_STAGE + j * HALO_TILE_BYTES, &tm.h, &fullA[sa], kca * 64,
                                    tx * p.bw - 1, (mt % p.tiles_y) * p.bh - 1, mt / p.tiles_y);
                    }
                    if (++sa == HCfg::AS) { sa = 0; pa ^= 1u; }
                    if (++kca == p.kchunks) { kca = 0; ta += sched_n; a_live = ta < p.total_tiles; }
                }
                if (b_live && mbar_try_wait(&emptyB[sb], pb ^ 1u)) {
                    t_idle = clock64();
                    mbar_expect_tx(&fullB[sb], (uint32_t)HCfg::B_BYTES);
                    tma_load_3d(sB + sb * HCfg::B_BYTES, &tm.b, &fullB[sb], kcb * 64, (tb % p.n_tiles) * BN,
                                p.taps.t[tpb].slab);
                    if (++sb == HCfg::BS) { sb = 0; pb ^= 1u; }
                    if (++tpb == p.taps.ntaps) {
                        tpb = 0;
                        if (++kcb == p.kchunks) { kcb = 0; tb += sched_n; b_live = tb < p.total_tiles; }
                    }
                }
            }
        }
    } else if (HL && warp == 1) {
        if (elect_one_sync()) {
            uint64_t* fullA = full; uint64_t* emptyA = empty;
            uint64_t* fullB = full + HCfg::AS; uint64_t* emptyB = empty + HCfg::AS;
            constexpr uint32_t idesc = make_idesc_bf16(128, BN, 0, 0);
            int sa = 0; uint32_t pa = 0; int sb = 0; uint32_t pb = 0;
            int acc = 0; uint32_t acc_phase = 0;
            long long w_acc = 0, w_a = 0, w_b = 0, c0 = 0; int ntl = 0;
            const long long t_begin = clock64();
            for (int tile = sched_id; tile < p.total_tiles; tile += sched_n) {
                if (p.dbg) c0 = clock64();
                mbar_wait(&tempty[acc], acc_phase ^ 1u);
                if (p.dbg) { w_acc += clock64() - c0; ++ntl; }
                tc_fence_after();
                const uint32_t d_tmem = tmem_mma + (uint32_t)(acc * MT * BN);
                for (int kc = 0; kc < p.kchunks; ++kc) {
                    if (p.dbg) c0 = clock64();
                    mbar_wait(&fullA[sa], pa);
                    if (p.dbg) w_a += clock64() - c0;
                    const uint32_t h0 = smem_u32(sA + sa * HCfg::A_STAGE);
                    for (int t = 0; t < p.taps.ntaps; ++t) {
                        if (p.dbg) c0 = clock64();
                        mbar_wait(&fullB[sb], pb);
                        if (p.dbg) w_b += clock64() - c0;
                        tc_fence_after();
                        const Tap tp = p.taps.t[t];
                        const uint32_t a0 = h0 + (uint32_t)(((tp.dy + 1) * HALO_W + (tp.dx + 1)) * 128);
                        const uint32_t a_lo = smem_desc_lo(a0), b_lo = smem_desc_lo(smem_u32(sB + sb * HCfg::B_BYTES));
                        // knob 18 (probe only): the descriptor's base-offset field set from the start address — wrong on
                        // this hardware (the swizzle follows the absolute address bits), kept to show it
                        const uint32_t a_hi = smem_desc_hi_sw128(HALO_W * 128) | (p.halo_bo ? (((a0 >> 7) & 7u) << 17) : 0u);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
#pragma unroll
                            for (int j = 0; j < MT; ++j)
                                umma_bf16_lohi(d_tmem + (uint32_t)(j * BN), a_lo + (uint32_t)((j * HALO_TILE_BYTES + k * 32) >> 4), a_hi,
                                               b_lo + (uint32_t)(k * 2), smem_desc_hi_sw128(1024), idesc, (kc | t | k) != 0 ? 1u : 0u);
                        }
                        umma_commit(&emptyB[sb]);
                        if (++sb == HCfg::BS) { sb = 0; pb ^= 1u; }
                    }
                    umma_commit(&emptyA[sa]);
                    if (++sa == HCfg::AS) { sa = 0; pa ^= 1u; }
                }
                umma_commit(&tfull[acc]);
                if (++acc == Cfg::ACC_BUFS) { acc = 0; acc_phase ^= 1u; }
            }
            if (p.dbg && blockIdx.x == 0)
                printf("mvae gather dbg (halo BN %d MT %d EPI %d): %d tiles, %lld clk issuing; waited %lld on the epilogue, "
                       "%lld on halo boxes, %lld on weight boxes\n", BN, MT, EPI, ntl, clock64() - t_begin, w_acc, w_a, w_b);
        }
    } else if (warp == 0) {
        if (elect_one_sync()) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = sched_id; tile < p.total_tiles; tile += sched_n) {
                const int nt = tile % p.n_tiles;
                int x0[MT], y0[MT], n0[MT];
#pragma unroll
                for (int j = 0; j < MT; ++j) {
                    // pixel tile (tile / n_tiles) * MT + j; one past the last decodes to n0 >= B: the box is
                    // entirely out of bounds, TMA zero-fills it and the epilogue stores nothing
                    int mt = (tile / p.n_tiles) * (CG * MT) + (int)rank * MT + j;
                    const int tx = mt % p.tiles_x; mt /= p.tiles_x;
                    x0[j] = tx * p.bw; y0[j] = (mt % p.tiles_y) * p.bh; n0[j] = (mt / p.tiles_y) * p.bn;
                }
                for (int t = 0; t < p.taps.ntaps; ++t) {
                    const Tap tp = p.taps.t[t];
                    for (int kc = 0; kc < p.kchunks; ++kc) {
                        mbar_wait(&empty[stage], phase ^ 1u);
                        if (CG == 2) {
                            if (rank == 0) mbar_expect_tx(&full[stage], stage_bytes);
                            const uint32_t lead_full = mapa_u32(smem_u32(&full[stage]), 0);
#pragma unroll
                            for (int j = 0; j < MT; ++j)
                                tma2_load_4d(sA + stage * Cfg::A_BYTES + j * A_TILE_BYTES, &tm.a[tp.view], lead_full, kc * 64,
                                             x0[j] + tp.dx, y0[j] + tp.dy, n0[j]);
                            tma2_load_3d(sB + stage * Cfg::B_BYTES, &tm.b, lead_full, kc * 64,
                                         nt * BN + (int)rank * (BN / 2), tp.slab);
                            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                            continue;
                        }
                        mbar_expect_tx(&full[stage], stage_bytes);
#pragma unroll
                        for (int j = 0; j < MT; ++j)
                            tma_load_4d(sA + stage * Cfg::A_BYTES + j * A_TILE_BYTES, &tm.a[tp.view], &full[stage], kc * 64,
                                        x0[j] + tp.dx, y0[j] + tp.dy, n0[j]);
                        tma_load_3d(sB + stage * Cfg::B_BYTES, &tm.b, &full[stage], kc * 64, nt * BN, tp.slab);
                        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0 && elect_one_sync()) {
            constexpr uint32_t idesc = make_idesc_bf16(128 * CG, BN, 0, 0);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            long long w_acc = 0, w_f = 0, c0 = 0; int ntl = 0;
            const long long t_begin = clock64();
            for (int tile = sched_id; tile < p.total_tiles; tile += sched_n) {
                if (p.dbg) c0 = clock64();
                mbar_wait(&tempty[acc], acc_phase ^ 1u);
                if (p.dbg) { w_acc += clock64() - c0; ++ntl; }
                tc_fence_after();
                const uint32_t d_tmem = tmem_mma + (uint32_t)(acc * MT * BN);
                for (int it = 0; it < kiters; ++it) {
                    if (p.dbg) c0 = clock64();
                    mbar_wait(&full[stage], phase);
                    if (p.dbg) w_f += clock64() - c0;
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(sA + stage * Cfg::A_BYTES);
                    const uint32_t b0 = smem_u32(sB + stage * Cfg::B_BYTES);
                    const uint32_t a_lo = smem_desc_lo(a0), b_lo = smem_desc_lo(b0);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int j = 0; j < MT; ++j) {
                            if (CG == 2) {
                                const uint64_t bd = smem_desc_sw128(b0 + k * 32, 16, 1024);
                                const uint64_t ad = smem_desc_sw128(a0 + j * A_TILE_BYTES + k * 32, 16, 1024);
                                umma2_bf16(d_tmem + (uint32_t)(j * BN), ad, bd, idesc, (it | k) != 0 ? 1u : 0u);
                            } else {
                                umma_bf16_lohi(d_tmem + (uint32_t)(j * BN), a_lo + (uint32_t)((j * A_TILE_BYTES + k * 32) >> 4),
                                               smem_desc_hi_sw128(1024), b_lo + (uint32_t)(k * 2), smem_desc_hi_sw128(1024), idesc,
                                               (it | k) != 0 ? 1u : 0u);
                            }
                        }
                    }
                    if (CG == 2) umma2_commit(&empty[stage]);
                    else umma_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                if (CG == 2) umma2_commit(&tfull[acc]);
                else umma_commit(&tfull[acc]);
                if (++acc == Cfg::ACC_BUFS) { acc = 0; acc_phase ^= 1u; }
            }
            if (p.dbg && blockIdx.x == 0)
                printf("mvae gather dbg (BN %d MT %d CG %d EPI %d): %d tiles, %lld clk issuing; waited %lld on the epilogue, "
                       "%lld on operand boxes\n", BN, MT, CG, EPI, ntl, clock64() - t_begin, w_acc, w_f);
        }
    } else {
        const int quarter = warp & 3;            // TMEM lane quarter this warp may read
        const int row = quarter * 32 + lane;     // GEMM row == pixel within the box
        const int ix = row % p.bw;
        const int iy = (row / p.bw) % p.bh;
        const int in = row / (p.bw * p.bh);
        int acc = 0; uint32_t acc_phase = 0;
        for (int tile = sched_id; tile < p.total_tiles; tile += sched_n) {
            const int nt = tile % p.n_tiles;
          if constexpr (EPI == EPI_GN_BWD) {
            // dgrad + GroupNorm backward.  Everything that does not depend on the accumulator — the GN input x of
            // this thread's pixel row for all MT pixel tiles, the statistics of the row's image — is fetched before
            // the accumulator wait; the inner loop then runs on registers and shared-memory broadcasts only.
            constexpr int CW = BN / CS, XCH = CW / 32, NSUB = CW / 4;
            static_assert(4 * CS * MT * NSUB * 8 <= 8 * 2 * 32 * 8, "statistics slots");
            const int chalf = (warp - 2) >> 2;                // which column strip of the tile
            const int cw0 = nt * BN + chalf * CW;             // first channel of this warp's column strip
            float2* wstat = sStat + (warp - 2) * (MT * NSUB);         // [MT][NSUB] {rstd, -mean*rstd} per 4-channel sub-block
            const float2* stat2 = reinterpret_cast<const float2*>(p.gn_stat);
            uint4 aux[4];                                     // GN input x of the chunk being worked on (32 channels)
            long long pixj[MT];
            int nj[MT], slotn[MT];
            unsigned vmj[MT];
            bool validj[MT];
            __syncwarp();                                     // every lane is done with the previous tile's slots
#pragma unroll
            for (int j = 0; j < MT; ++j) {
                int mt = (tile / p.n_tiles) * (CG * MT) + (int)rank * MT + j;
                const int tx = mt % p.tiles_x; mt /= p.tiles_x;
                const int ty = mt % p.tiles_y;
                const int tn = mt / p.tiles_y;
                const int x = tx * p.bw + ix, y = ty * p.bh + iy, n = tn * p.bn + in;
                const bool valid = (row < p.rows) && x < p.OW && y < p.OH && n < p.B;
                const long long pix = p.o_off + (long long)n * p.o_sn + (long long)y * p.o_sy + (long long)x * p.o_sx;
                const unsigned vm = __ballot_sync(0xffffffffu, valid);
                // image of the warp's first valid row: the slots start out holding its statistics
                const int n_ref = vm ? __shfl_sync(0xffffffffu, n, __ffs(vm) - 1) : -1;
                pixj[j] = pix; nj[j] = n; validj[j] = valid; vmj[j] = vm; slotn[j] = n_ref;
                if (vm && lane < NSUB) {
                    const float2 mr = __ldg(stat2 + (long long)n_ref * p.gn_G + ((cw0 + 4 * lane) >> p.gn_cpg_log2));
                    wstat[j * NSUB + lane] = make_float2(mr.y, -mr.x * mr.y);
                }
            }
            // chunk (j, ci)'s x is fetched while the chunk before it is reduced (the first one before the wait)
            auto load_aux = [&](int j, int ci) {
                if (validj[j]) {
                    const uint4* ap = reinterpret_cast<const uint4*>(p.gn_x + pixj[j] + cw0 + ci * 32);
#pragma unroll
                    for (int i = 0; i < 4; i += 2) ldg256_nc(ap + i, aux[i], aux[i + 1]);
                }
            };
            load_aux(0, 0);
            __syncwarp();
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < MT; ++j) {
                const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * MT * BN + j * BN + chalf * CW);
#pragma unroll
                for (int ci = 0; ci < XCH; ++ci) {
                    const int cbase = cw0 + ci * 32;
                    // one pass per image among the warp's valid rows: exactly one unless the images are smaller than
                    // a warp's 32 pixel rows; rows of other images sit a pass out (their sums stay zero), so every
                    // pass ends in one warp-reduced atomic per value
                    unsigned todo = vmj[j];
                    int ncur = slotn[j];
                    while (todo) {
                        const bool active = validj[j] && nj[j] == ncur;
                        if (slotn[j] != ncur) {
                            __syncwarp();
                            if (lane < NSUB) {
                                const float2 mr = __ldg(stat2 + (long long)ncur * p.gn_G + ((cw0 + 4 * lane) >> p.gn_cpg_log2));
                                wstat[j * NSUB + lane] = make_float2(mr.y, -mr.x * mr.y);
                            }
                            slotn[j] = ncur;
                            __syncwarp();
                        }
                        float sa[8], sb[8];   // GN partial sums of the eight 4-channel sub-blocks (of gamma/2 * du)
#pragma unroll
                        for (int k = 0; k < 8; ++k) { sa[k] = 0.f; sb[k] = 0.f; }
                        const uint32_t cf = smem_u32(sCoef + cbase), sf = smem_u32(wstat + j * NSUB + ci * 8);
#pragma unroll
                        for (int hf = 0; hf < 2; ++hf) {     // 16 accumulator columns at a time
                            uint32_t v[16];
                            tmem_ld_32x16(t0 + (uint32_t)(ci * 32 + hf * 16), v);
                            tmem_ld_wait();
                            if (active) {
                                float f[16];
#pragma unroll
                                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
                                if (p.bias) {
                                    const float4* bp = reinterpret_cast<const float4*>(p.bias + cbase + hf * 16);
#pragma unroll
                                    for (int i = 0; i < 4; ++i) {
                                        float4 b4 = __ldg(bp + i);
                                        f[4 * i] += b4.x; f[4 * i + 1] += b4.y; f[4 * i + 2] += b4.z; f[4 * i + 3] += b4.w;
                                    }
                                }
                                // f = dL/da  ->  du = f * silu'(u), u = xhat*gamma + beta, xhat = (x - mean)*rstd.  With
                                // h = u/2 and t = tanh(h):  silu'(u) = (1 + t + h*(1 - t*t)) / 2.
#pragma unroll
                                for (int i = 0; i < 2; ++i) {
                                    Vec8<bf16> x8; x8.raw = aux[2 * hf + i];
                                    float xf[8]; x8.unpack(xf);
#pragma unroll
                                    for (int hh = 0; hh < 2; ++hh) {
                                        const int kl = 2 * i + hh, k = 4 * hf + kl;   // sub-block within the half / the chunk
                                        const float2 s2 = lds_f2(sf + k * 8);
                                        const float rs = s2.x, nm = s2.y;
                                        const float4 c01 = lds_f4(cf + k * 32), c23 = lds_f4(cf + k * 32 + 16);   // {gamma/2, beta/2} x 4
                                        const float gh[4] = {c01.x, c01.z, c23.x, c23.z}, bh[4] = {c01.y, c01.w, c23.y, c23.w};
#pragma unroll
                                        for (int q = 0; q < 4; ++q) {
                                            const float xh = fmaf(xf[4 * hh + q], rs, nm);
                                            float d = f[4 * kl + q];
                                            if (p.gn_silu) {
                                                const float h = fmaf(xh, gh[q], bh[q]);
                                                const float t = tanh_approx(h);
                                                const float r = fmaf(h, fmaf(-t, t, 1.f), t);
                                                d *= fmaf(r, 0.5f, 0.5f);
                                                f[4 * kl + q] = d;
                                            }
                                            const float gd = d * gh[q];
                                            sa[k] += gd; sb[k] = fmaf(gd, xh, sb[k]);
                                        }
                                    }
                                }
                                Vec8<bf16> o8a, o8b;
                                o8a.pack(f); o8b.pack(f + 8);
                                stg256(p.out + pixj[j] + cbase + hf * 16, o8a.raw, o8b.raw);
                            }
                        }
                        todo &= ~__ballot_sync(0xffffffffu, active);
                        const int ndone = ncur;
                        if (todo) ncur = __shfl_sync(0xffffffffu, nj[j], __ffs(todo) - 1);
                        else if (ci + 1 < XCH) load_aux(j, ci + 1);      // x of the next chunk, under this chunk's reduction
                        else if (j + 1 < MT) load_aux(j + 1, 0);
                        float* dst = p.gn_sums + ((long long)ndone * p.gn_G + (cbase >> p.gn_cpg_log2)) * 2;
                        gn_flush_any(p.gn_cpg_log2, sa, sb, dst, lane, true, active, 2.f);   // the sums carried gamma/2
                    }
                    if (vmj[j] == 0u) {                                  // no valid row in this pixel tile: nothing ran
                        if (ci + 1 < XCH) load_aux(j, ci + 1);
                        else if (j + 1 < MT) load_aux(j + 1, 0);
                    }
                }
            }
          } else {
#pragma unroll 1
          for (int j = 0; j < MT; ++j) {
            const uint32_t acc_col = (uint32_t)(acc * MT * BN + j * BN);
            int mt = (tile / p.n_tiles) * (CG * MT) + (int)rank * MT + j;
            const int tx = mt % p.tiles_x; mt /= p.tiles_x;
            const int ty = mt % p.tiles_y;
            const int tn = mt / p.tiles_y;
            const int x = tx * p.bw + ix, y = ty * p.bh + iy, n = tn * p.bn + in;
            const bool valid = (row < p.rows) && x < p.OW && y < p.OH && n < p.B;
            const long long pix = p.o_off + (long long)n * p.o_sn + (long long)y * p.o_sy + (long long)x * p.o_sx;
            if (EPI == EPI_PLAIN) {
                // the residual of a 32-channel chunk is fetched one chunk ahead (the first before the accumulator wait): a
                // dependent global load per chunk inside the loop left the MMA warp waiting on this epilogue for up to 40 %
                // of a 128-channel tile (scripts/gather_stall_probe.py)
                uint4 rcur[4], rnext[4];
                const bool have_res = valid && p.res != nullptr;
                const uint4* rp0 = reinterpret_cast<const uint4*>(p.res + pix + nt * BN);
                if (have_res) {
#pragma unroll
                    for (int i = 0; i < 4; i += 2) ldg256_nc(rp0 + i, rcur[i], rcur[i + 1]);
                }
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc_col;
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(t0 + (uint32_t)c0, v);
                    if (have_res && c0 + 32 < BN) {
#pragma unroll
                        for (int i = 0; i < 4; i += 2) ldg256_nc(rp0 + (c0 + 32) / 8 + i, rnext[i], rnext[i + 1]);
                    }
                    tmem_ld_wait();
                    if (valid) {
                        const int cbase = nt * BN + c0;
                        float f[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
                        if (p.bias) {
                            const float4* bp = reinterpret_cast<const float4*>(p.bias + cbase);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                float4 b4 = __ldg(bp + i);
                                f[4 * i] += b4.x; f[4 * i + 1] += b4.y; f[4 * i + 2] += b4.z; f[4 * i + 3] += b4.w;
                            }
                        }
                        if (p.res) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                Vec8<bf16> r8; r8.raw = rcur[i];
                                float rf[8]; r8.unpack(rf);
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[8 * i + j] += rf[j];
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 4; i += 2) {
                            Vec8<bf16> o8a, o8b;
                            o8a.pack(f + 8 * i); o8b.pack(f + 8 * i + 8);
                            stg256(p.out + pix + cbase + 8 * i, o8a.raw, o8b.raw);
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) rcur[i] = rnext[i];
                }
            } else {
                // this warp's half of the tile's columns: XCH chunks of 32
                constexpr int CW = BN / CS, XCH = CW / 32;
                const int chalf = (warp - 2) >> 2;
                // are all valid rows of this warp in one image?  (then one atomic per value per warp)
                const unsigned vm = __ballot_sync(0xffffffffu, valid);
                const bool any_valid = vm != 0u;
                const int n_ref = __shfl_sync(0xffffffffu, n, any_valid ? (__ffs(vm) - 1) : 0);
                const bool uniform = __all_sync(0xffffffffu, !valid || n == n_ref);
                // the per-element global operand (GN input x / residual) does not depend on the
                // accumulator: fetch the whole strip before waiting for the MMAs
                uint4 aux[XCH][4];
                const bf16* auxp = p.res;
                const bool have_aux = valid && auxp != nullptr;
                if (have_aux) {
#pragma unroll
                    for (int ci = 0; ci < XCH; ++ci) {
                        const uint4* ap = reinterpret_cast<const uint4*>(auxp + pix + nt * BN + chalf * CW + ci * 32);
#pragma unroll
                        for (int i = 0; i < 4; i += 2) ldg256_nc(ap + i, aux[ci][i], aux[ci][i + 1]);
                    }
                }
                mbar_wait(&tfull[acc], acc_phase);
                tc_fence_after();
                const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc_col + (uint32_t)(chalf * CW);
#pragma unroll
                for (int ci = 0; ci < XCH; ++ci) {
                    uint32_t v[32];
                    tmem_ld_32x32(t0 + (uint32_t)(ci * 32), v);
                    tmem_ld_wait();
                    const int cbase = nt * BN + chalf * CW + ci * 32;
                    float sa[8], sb[8];   // GN partial sums of the eight 4-channel sub-blocks of this chunk
#pragma unroll
                    for (int k = 0; k < 8; ++k) { sa[k] = 0.f; sb[k] = 0.f; }
                    if (valid) {
                        float f[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
                        if (p.bias) {
                            const float4* bp = reinterpret_cast<const float4*>(p.bias + cbase);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                float4 b4 = __ldg(bp + i);
                                f[4 * i] += b4.x; f[4 * i + 1] += b4.y; f[4 * i + 2] += b4.z; f[4 * i + 3] += b4.w;
                            }
                        }
                        if (EPI == EPI_GN_STATS && p.res) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                Vec8<bf16> r8; r8.raw = aux[ci][i];
                                float rf[8]; r8.unpack(rf);
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[8 * i + j] += rf[j];
                            }
                        }
                        uint4 ov[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            Vec8<bf16> o8; o8.pack(f + 8 * i);
                            ov[i] = o8.raw;
                            if (EPI == EPI_GN_STATS) {
                                float yf[8]; o8.unpack(yf);   // statistics of the values as stored
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const int k = 2 * i + (j >> 2);
                                    sa[k] += yf[j]; sb[k] = fmaf(yf[j], yf[j], sb[k]);
                                }
                            }
                        }
                        stg256(p.out + pix + cbase, ov[0], ov[1]);
                        stg256(p.out + pix + cbase + 16, ov[2], ov[3]);
                    }
                    if (any_valid) {
                        float* dst = p.gn_sums + ((long long)(uniform ? n_ref : n) * p.gn_G + (cbase >> p.gn_cpg_log2)) * 2;
                        gn_flush_any(p.gn_cpg_log2, sa, sb, dst, lane, uniform, valid);
                    }
                }
            }
          }   // pixel tiles of this CTA tile
          }
            tc_fence_before();
            if (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty[acc]), 0));   // the leader issues the next MMAs
            else mbar_arrive(&tempty[acc]);
            if (++acc == Cfg::ACC_BUFS) { acc = 0; acc_phase ^= 1u; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();   // neither CTA of a pair frees its TMEM / exits while the other still uses it
    if (warp == 1) {
        tc_fence_after();
        if (CG == 2) tmem_dealloc2(tmem_base, Cfg::TMEM_COLS);
        else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

bool umma_gather_ok(const GatherProblem& p) {
    if (p.K % 64 != 0 || p.N % 64 != 0) return false;
    if (p.OW > 128 && p.OW % 128 != 0) return false;
    BoxGeom g = box_geom(p.B, p.OH, p.OW);
    for (int v = 0; v < p.nviews; ++v)
        if (p.in_views[v].Ws < g.bw || p.in_views[v].Hs < g.bh) return false;
    return true;
}

template <int BN, int EPI, int MT, int CG, int CS = gather_csplit(EPI), int HL = 0>
static int launch_gather_epi(const GatherTmaps& tm, GatherDev& d, cudaStream_t st) {
    using Cfg = GatherCfg<BN, MT, CG>;
    constexpr int SMEM = gather_smem(HL ? HaloCfg<BN, MT>::SMEM : Cfg::SMEM, EPI);
    static_assert(SMEM <= 232448, "shared memory of one CTA");
    static bool attr_set = false;
    if (!attr_set) {
        MVAE_CUDA(cudaFuncSetAttribute(conv_gather_umma_kernel<BN, EPI, MT, CG, CS, HL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       SMEM));
        attr_set = true;
    }
    if (CG == 2) {
        // one CTA pair per TPC: a cluster of two, persistent over the tile pairs
        int pairs = num_sms() / 2;
        if (d.total_tiles < pairs) pairs = d.total_tiles;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(2 * pairs);
        cfg.blockDim = dim3(gather_threads_cs(CS));
        cfg.dynamicSmemBytes = SMEM;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        count_launch();
        MVAE_CUDA(cudaLaunchKernelEx(&cfg, conv_gather_umma_kernel<BN, EPI, MT, CG, CS, HL>, tm, d));
        MVAE_CUDA(cudaGetLastError());
        return 0;
    }
    int grid = d.total_tiles < num_sms() ? d.total_tiles : num_sms();
    conv_gather_umma_kernel<BN, EPI, MT, CG, CS, HL><<<grid, gather_threads_cs(CS), SMEM, st>>>(tm, d);
    MVAE_LAUNCH_CHECK();
    return 0;
}
template <int BN, int MT, int CG = 1, int HL = 0>
static int launch_gather(const GatherTmaps& tm, GatherDev& d, int epi, cudaStream_t st) {
    if (epi == EPI_GN_STATS) {
        if constexpr (BN == 128 && HL == 0) {
            // OPT-IN (knob 16 = 2): four column strips (16 epilogue warps) instead of two for the statistics epilogue of
            // the 64x64 layers.  Measured on one box (trip r2n): fprop 62.7 ms with, 62.6 ms without — unlike the
            // GroupNorm-backward epilogue this one is not what holds the 128-channel tiles at ~50 % tensor-pipe.
            if (debug_knob(16) == 2) return launch_gather_epi<BN, EPI_GN_STATS, MT, CG, 4>(tm, d, st);
        }
        return launch_gather_epi<BN, EPI_GN_STATS, MT, CG, gather_csplit(EPI_GN_STATS), HL>(tm, d, st);
    }
    if (epi == EPI_GN_BWD) {
        if constexpr (BN >= 128) {
            // MMA cycles of one CTA tile against the ~25k cycles the 8-warp epilogue takes for it (knob 9: 1 = never, 2 = always)
            const long long mma_cycles = (long long)d.taps.ntaps * d.kchunks * 4 * MT * (BN / 2);
            const int knob = debug_knob(9);
            if (knob == 2 || (knob == 0 && mma_cycles <= 12288)) return launch_gather_epi<BN, EPI_GN_BWD, MT, CG, 4, HL>(tm, d, st);
        }
        return launch_gather_epi<BN, EPI_GN_BWD, MT, CG, gather_csplit(EPI_GN_BWD), HL>(tm, d, st);
    }
    return launch_gather_epi<BN, EPI_PLAIN, MT, CG, gather_csplit(EPI_PLAIN), HL>(tm, d, st);
}

// halo variant: a stride-1 3 x 3 tap program over one view, output grid in whole 8 x 16 pixel tiles
static bool halo_ok(const GatherProblem& p) {
    if (p.nviews != 1 || p.taps.ntaps != 9 || p.OW % HALO_TW != 0 || p.OH % HALO_TH != 0) return false;
    if (p.in_views[0].Ws < HALO_W || p.in_views[0].Hs < HALO_H) return false;
    for (int t = 0; t < 9; ++t) {
        const Tap& tp = p.taps.t[t];
        if (tp.view != 0 || tp.dx < -1 || tp.dx > 1 || tp.dy < -1 || tp.dy > 1) return false;
    }
    return true;
}

int umma_gather(const GatherProblem& p, const void* in, const void* w, int nslabs, const float* bias,
                const void* residual, void* out, cudaStream_t st, const GnHook* hook) {
    MVAE_CHECK(umma_gather_ok(p), "umma_gather: unsupported shape K=%d N=%d OW=%d", p.K, p.N, p.OW);
    BoxGeom g = box_geom(p.B, p.OH, p.OW);
    const int BN = (p.N % 256 == 0) ? 256 : (p.N % 128 == 0 ? 128 : 64);
    // CTA pairs (cta_group::2) for 256-channel tiles when there are enough pixel-tile pairs for every TPC
    int CG = 1;
    {
        BoxGeom g0 = box_geom(p.B, p.OH, p.OW);
        const long long mt0 = (long long)g0.tiles_x * g0.tiles_y * g0.tiles_n;
        if (BN == 256 && debug_knob(8) == 0 && ((mt0 + 1) / 2) * (p.N / 256) >= num_sms() / 2) CG = 2;   // knob 8 = 1: off
        // 128-channel tiles: a pair with TWO pixel tiles per CTA (2 x 2 x 128 pixels x 128 channels per scheduled tile),
        // each CTA loading half of the weight tile: (32 + 8) KB per 512 MMA cycles = 80 B/clk of ingest instead of
        // the 96 B/clk of the single-CTA two-tile variant.  OPT-IN (knob 13 = 1): measured on C4 on one box (trip r2c)
        // it changes nothing — fprop 64.0 ms with, 63.1 ms without, step 4113 vs 4119 img/s — so the 128-wide tiles
        // are not bound by the weight half of the ingest; the single-CTA variant stays the default.
        if (BN == 128 && debug_knob(13) == 1 && debug_knob(5) == 0 && ((mt0 + 3) / 4) * (p.N / 128) >= num_sms() / 2) CG = 2;
    }
    GatherTmaps tm;
    memset(&tm, 0, sizeof(tm));
    for (int v = 0; v < MAX_VIEWS; ++v) {
        const View& vw = p.in_views[v < p.nviews ? v : 0];
        if (view_tmap(&tm.a[v], (const bf16*)in, vw, p.K, p.B, g.bw, g.bh, g.bn)) return 1;
    }
    {
        cuuint64_t dims[3] = {(cuuint64_t)p.K, (cuuint64_t)p.N, (cuuint64_t)nslabs};
        cuuint64_t stb[2] = {(cuuint64_t)p.K * 2, (cuuint64_t)p.K * p.N * 2};
        cuuint32_t box[3] = {64u, (cuuint32_t)(BN / CG), 1u};
        if (encode_tmap(&tm.b, w, 3, dims, stb, box)) return 1;
    }
    GatherDev d;
    d.bw = g.bw; d.bh = g.bh; d.bn = g.bn; d.rows = g.rows;
    d.OW = p.OW; d.OH = p.OH; d.B = p.B;
    d.tiles_x = g.tiles_x; d.tiles_y = g.tiles_y; d.tiles_n = g.tiles_n;
    d.n_tiles = p.N / BN; d.kchunks = p.K / 64;
    // 128-channel tiles take two pixel tiles per weight stage when that still leaves every SM a CTA tile
    long long m_tiles = (long long)g.tiles_x * g.tiles_y * g.tiles_n;
    int MT = (BN == 128 && debug_knob(5) == 0 && ((m_tiles + 1) / 2) * d.n_tiles >= num_sms()) ? 2 : 1;
    if (BN == 128 && CG == 2) MT = 2;
    if (CG == 1 && BN == 256 && debug_knob(7) == 1) {
        // OPT-IN experiment (knob 7): 256 x 256 CTA tiles, 64 instead of 96 B/clk of ingest but a single accumulator
        // buffer, i.e. an exposed epilogue.  Measured on C4 (r1e): fprop 16.7 -> 17.9 ms, dgrad 19.6 -> 21.6 ms per
        // 256-image micro-batch — slower, so the default stays 128 x 256 with double-buffered accumulators.
        const long long t1 = m_tiles * d.n_tiles, t2 = ((m_tiles + 1) / 2) * d.n_tiles, sms = num_sms();
        if (t2 >= sms && 3 * ((t2 + sms - 1) / sms) < 2 * ((t1 + sms - 1) / sms)) MT = 2;
    }
    // 128-channel tiles of stride-1 3x3 convs: 8 x 16 pixel tiles fed by one halo box per chunk (knob 17 = 1: off)
    const bool halo = BN == 128 && MT == 2 && CG == 1 && debug_knob(17) != 1 && halo_ok(p);
    d.halo_bo = debug_knob(18) == 1 ? 1 : 0;
    d.dbg = debug_knob(19) == 1 ? 1 : 0;
    if (halo) {
        g.bw = 8; g.bh = 16; g.bn = 1; g.rows = 128;
        g.tiles_x = p.OW / 8; g.tiles_y = p.OH / 16; g.tiles_n = p.B;
        d.bw = g.bw; d.bh = g.bh; d.bn = g.bn; d.rows = g.rows;
        d.tiles_x = g.tiles_x; d.tiles_y = g.tiles_y; d.tiles_n = g.tiles_n;
        m_tiles = (long long)g.tiles_x * g.tiles_y * g.tiles_n;
        if (view_tmap(&tm.h, (const bf16*)in, p.in_views[0], p.K, p.B, HALO_W, HALO_H, 1)) return 1;
    }
    const int m_per_tile = MT * CG;   // pixel tiles per scheduled tile: MT per CTA, one per CTA of a pair
    d.total_tiles = (int)((m_tiles + m_per_tile - 1) / m_per_tile) * d.n_tiles;
    d.N = p.N;
    d.o_off = p.out_view.off; d.o_sn = p.out_view.sn; d.o_sy = p.out_view.sy; d.o_sx = p.out_view.sx;
    d.bias = bias; d.res = (const bf16*)residual; d.out = (bf16*)out;
    d.taps = p.taps;
    int epi = EPI_PLAIN;
    d.gn_G = 0; d.gn_cpg_log2 = 0; d.gn_silu = 0;
    d.gn_sums = nullptr; d.gn_x = nullptr; d.gn_stat = nullptr; d.gn_gamma = nullptr; d.gn_beta = nullptr;
    if (hook && hook->mode != 0) {
        MVAE_CHECK(hook->cpg_log2 >= 2 && (p.N % (hook->G << hook->cpg_log2)) == 0 && hook->sums,
                   "umma_gather: GroupNorm hook needs >= 4 channels per group (power of two) over all N channels");
        MVAE_CHECK(hook->mode == EPI_GN_STATS || (hook->x && hook->stat && hook->gamma && hook->beta),
                   "umma_gather: GroupNorm backward hook needs x, stats, gamma, beta");
        MVAE_CHECK(hook->mode != EPI_GN_BWD || p.N <= GN_COEF_MAX, "umma_gather: GroupNorm backward hook takes <= %d channels",
                   GN_COEF_MAX);
        epi = hook->mode;
        d.gn_G = hook->G; d.gn_cpg_log2 = hook->cpg_log2; d.gn_silu = hook->silu;
        d.gn_sums = hook->sums; d.gn_x = (const bf16*)hook->x; d.gn_stat = hook->stat;
        d.gn_gamma = hook->gamma; d.gn_beta = hook->beta;
    }
    if (halo) return launch_gather<128, 2, 1, 1>(tm, d, epi, st);
    if (CG == 2 && BN == 128) return launch_gather<128, 2, 2>(tm, d, epi, st);
    if (CG == 2) return launch_gather<256, 1, 2>(tm, d, epi, st);
    if (BN == 256 && MT == 2) return launch_gather<256, 2>(tm, d, epi, st);
    if (BN == 256) return launch_gather<256, 1>(tm, d, epi, st);
    if (BN == 128 && MT == 2) return launch_gather<128, 2>(tm, d, epi, st);
    if (BN == 128) return launch_gather<128, 1>(tm, d, epi, st);
    return launch_gather<64, 1>(tm, d, epi, st);
}

// ================================================================== reduce (wgrad)
// D[co, ci] (per slab, per K-split) = sum_pixels dy[pix, co] * x[pix + tap, ci].  Both operands are
// MN-major (channels contiguous in NHWC); 64 pixels per pipeline stage = 4 MMAs of K = 16.
// CTA tile = (MT x 128 co) x (BN ci), MT in {1, 2}, BN in {128, 256}: MT * BN <= 512 TMEM columns.
// Bytes pulled from L2 per MMA cycle fall from 128 B/clk (128 x 128) to 64 B/clk (256 x 256), which
// is what bounds this kernel (SM ingest), see DESIGN.md.
struct alignas(64) ReduceTmaps {
    CUtensorMap dy[MAX_VIEWS];
    CUtensorMap in[MAX_VIEWS];
};
struct ReduceDev {
    int bw, bh, bn, rows;
    int tiles_x, tiles_y, tiles_n, ptiles;
    int co_tiles, ci_tiles, nslabs, splits;
    int Cout, Cin;
    float* partial;    // [splits][nslabs][Cout][Cin]
    TapProg taps;
};

constexpr int R_PIX = 64;            // pixels (K) per stage
constexpr int R_CHUNK = R_PIX * 128; // 64 pixels x 64 channels bf16 = 8 KB

// NT filter taps per CTA share every dy stage (NT = 3 for 128 x 128 channel tiles: 16 KB of dy + 3 x 16 KB of x
// per 768 MMA cycles = 85 B/clk instead of 128 B/clk with one tap per CTA).
template <int MT, int BN, int NT>
struct ReduceCfg {
    static constexpr int A_BYTES = MT * 2 * R_CHUNK;
    static constexpr int B_TAP = (BN / 64) * R_CHUNK;
    static constexpr int B_BYTES = NT * B_TAP;
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int STAGES = STAGE <= 32768 ? 6 : (STAGE <= 49152 ? 4 : 3);
    static constexpr int ACC_COLS = MT * BN * NT;
    static constexpr int TMEM_COLS = ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512);
    static constexpr int SMEM = 1024 + STAGES * STAGE + 256;
};

template <int MT, int BN, int NT>
__global__ void __launch_bounds__(G_THREADS, 1)
conv_reduce_umma_kernel(const __grid_constant__ ReduceTmaps tm, const __grid_constant__ ReduceDev p) {
    using Cfg = ReduceCfg<MT, BN, NT>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = sA + STAGES * Cfg::A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // work item -> (split, slab, co tile, ci tile)
    int w = blockIdx.x;
    const int cit = w % p.ci_tiles; w /= p.ci_tiles;
    const int cot = w % p.co_tiles; w /= p.co_tiles;
    const int n_groups = p.nslabs / NT;           // NT consecutive taps (same dy view) per CTA
    const int slab_i = (w % n_groups) * NT;
    const int split = w / n_groups;
    const Tap tp = p.taps.t[slab_i];
    const int pt_beg = (int)(((long long)p.ptiles * split) / p.splits);
    const int pt_end = (int)(((long long)p.ptiles * (split + 1)) / p.splits);
    const int kiters = pt_end - pt_beg;

    if (p.rows < R_PIX) {
        // pixel rows beyond the TMA box take part in the K reduction: they must be zero
        for (int i = threadIdx.x; i < STAGES * Cfg::STAGE / 16; i += G_THREADS)
            reinterpret_cast<uint4*>(sA)[i] = make_uint4(0, 0, 0, 0);
        fence_proxy_async_smem();
    }
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm.dy[tp.oview]);
        for (int j = 0; j < NT; ++j) prefetch_tmap(&tm.in[p.taps.t[slab_i + j].view]);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
            mbar_init(&tfull[0], 1);
            fence_barrier_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (Cfg::TMEM_COLS == 512 && tmem_base != 0u) __trap();   // all 512 columns start at column 0 (see the gather kernel)
    const uint32_t tmem_mma = (Cfg::TMEM_COLS == 512) ? 0u : tmem_base;
    const uint32_t stage_bytes = (uint32_t)p.rows * 128u * (MT * 2 + NT * (BN / 64));

    if (warp == 0) {
        if (elect_one_sync()) {
            int stage = 0; uint32_t phase = 0;
            for (int pt = pt_beg; pt < pt_end; ++pt) {
                int mt = pt;
                const int tx = mt % p.tiles_x; mt /= p.tiles_x;
                const int ty = mt % p.tiles_y;
                const int tn = mt / p.tiles_y;
                const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bn;
                mbar_wait(&empty[stage], phase ^ 1u);
                mbar_expect_tx(&full[stage], stage_bytes);
                uint8_t* a = sA + stage * Cfg::A_BYTES;
                uint8_t* b = sB + stage * Cfg::B_BYTES;
#pragma unroll
                for (int c = 0; c < MT * 2; ++c)
                    tma_load_4d(a + c * R_CHUNK, &tm.dy[tp.oview], &full[stage], cot * (MT * 128) + c * 64, x0, y0, n0);
#pragma unroll
                for (int j = 0; j < NT; ++j) {
                    const Tap tj = p.taps.t[slab_i + j];
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c)
                        tma_load_4d(b + j * Cfg::B_TAP + c * R_CHUNK, &tm.in[tj.view], &full[stage], cit * BN + c * 64,
                                    x0 + tj.dx, y0 + tj.dy, n0);
                }
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (elect_one_sync()) {
            constexpr uint32_t idesc = make_idesc_bf16(128, BN, 1, 1);
            constexpr uint32_t dhi = smem_desc_hi_sw128(1024);
            int stage = 0; uint32_t phase = 0;
            for (int it = 0; it < kiters; ++it) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t a_lo = smem_desc_lo(smem_u32(sA + stage * Cfg::A_BYTES), R_CHUNK);
                const uint32_t b_lo = smem_desc_lo(smem_u32(sB + stage * Cfg::B_BYTES), R_CHUNK);
#pragma unroll
                for (int m = 0; m < MT; ++m) {
#pragma unroll
                    for (int k = 0; k < R_PIX / 16; ++k) {   // 16 pixels (= 16 rows of 128 B) per MMA
#pragma unroll
                        for (int j = 0; j < NT; ++j)
                            umma_bf16_lohi(tmem_mma + (uint32_t)((m * NT + j) * BN), a_lo + (uint32_t)((m * 2 * R_CHUNK + k * 2048) >> 4), dhi,
                                           b_lo + (uint32_t)((j * Cfg::B_TAP + k * 2048) >> 4), dhi, idesc, (it | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&empty[stage]);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
            umma_commit(&tfull[0]);
        }
    } else {
        const int quarter = warp & 3;
        if (kiters > 0) {
            mbar_wait(&tfull[0], 0);
            tc_fence_after();
        }
#pragma unroll 1
        for (int mj = 0; mj < MT * NT; ++mj) {
            const int m = mj / NT, j = mj % NT;
            const int co = cot * (MT * 128) + m * 128 + quarter * 32 + lane;
            const bool co_ok = co < p.Cout;   // last co tile may be partly TMA zero fill
            float* dst = p.partial + (((long long)split * p.nslabs + p.taps.t[slab_i + j].slab) * p.Cout + co) * p.Cin + cit * BN;
            const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(mj * BN);
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                if (kiters > 0) {
                    tmem_ld_32x32(t0 + (uint32_t)c0, v);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = 0u;
                }
                if (co_ok && cit * BN + c0 < p.Cin) {
                    float4* o = reinterpret_cast<float4*>(dst + c0);
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        o[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                           __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

static void reduce_tile(const ReduceProblem& p, int& MT, int& BN, int& NT) {
    MT = p.Cout >= 256 ? 2 : 1;
    BN = p.Cin >= 256 ? 256 : 128;
    // 128 x 128 channel tiles: three taps per CTA on one dy stage, when all taps read the same dy view
    NT = 1;
    if (MT == 1 && BN == 128 && p.taps.ntaps % 3 == 0 && debug_knob(6) == 0) {
        NT = 3;
        for (int t = 1; t < p.taps.ntaps; ++t)
            if (p.taps.t[t].oview != p.taps.t[0].oview) NT = 1;
    }
}

bool umma_reduce_ok(const ReduceProblem& p) {
    if (p.Cout % 64 != 0 || p.Cin % 64 != 0) return false;
    if (p.GW > R_PIX && p.GW % R_PIX != 0) return false;
    BoxGeom g = box_geom(p.B, p.GH, p.GW, R_PIX);
    for (int v = 0; v < p.n_inviews; ++v)
        if (p.in_views[v].Ws < g.bw || p.in_views[v].Hs < g.bh) return false;
    for (int v = 0; v < p.n_dyviews; ++v)
        if (p.dy_views[v].Ws < g.bw || p.dy_views[v].Hs < g.bh) return false;
    return true;
}

int umma_reduce_splits(const ReduceProblem& p) {
    int MT, BN, NT;
    reduce_tile(p, MT, BN, NT);
    BoxGeom g = box_geom(p.B, p.GH, p.GW, R_PIX);
    long long ptiles = (long long)g.tiles_x * g.tiles_y * g.tiles_n;
    long long items = (long long)(p.taps.ntaps / NT) * ((p.Cout + MT * 128 - 1) / (MT * 128)) * ((p.Cin + BN - 1) / BN);
    // one wave of CTAs (one per SM): as many K-splits as fit, each with at least ~8 pixel tiles
    long long want = num_sms() / items;
    if (want < 1) want = 1;
    long long cap = (ptiles + 7) / 8;
    if (cap < 1) cap = 1;
    if (want > cap) want = cap;
    if (want > 64) want = 64;
    return (int)want;
}

template <int MT, int BN, int NT>
static int launch_reduce(const ReduceTmaps& tm, const ReduceDev& d, long long grid, cudaStream_t st) {
    using Cfg = ReduceCfg<MT, BN, NT>;
    static bool attr_set = false;
    if (!attr_set) {
        MVAE_CUDA(cudaFuncSetAttribute(conv_reduce_umma_kernel<MT, BN, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM));
        attr_set = true;
    }
    conv_reduce_umma_kernel<MT, BN, NT><<<(unsigned)grid, G_THREADS, Cfg::SMEM, st>>>(tm, d);
    MVAE_LAUNCH_CHECK();
    return 0;
}

int umma_reduce(const ReduceProblem& p, const void* in, const void* dy, float* partial, int splits,
                cudaStream_t st) {
    MVAE_CHECK(umma_reduce_ok(p), "umma_reduce: unsupported shape Cout=%d Cin=%d", p.Cout, p.Cin);
    int MT, BN, NT;
    reduce_tile(p, MT, BN, NT);
    BoxGeom g = box_geom(p.B, p.GH, p.GW, R_PIX);
    ReduceTmaps tm;
    memset(&tm, 0, sizeof(tm));
    for (int v = 0; v < MAX_VIEWS; ++v) {
        const View& vd = p.dy_views[v < p.n_dyviews ? v : 0];
        const View& vi = p.in_views[v < p.n_inviews ? v : 0];
        MVAE_CHECK(vd.Ws >= g.bw && vd.Hs >= g.bh && vi.Ws >= g.bw && vi.Hs >= g.bh,
                   "umma_reduce: view %d smaller than the pixel box", v);
        if (view_tmap(&tm.dy[v], (const bf16*)dy, vd, p.Cout, p.B, g.bw, g.bh, g.bn)) return 1;
        if (view_tmap(&tm.in[v], (const bf16*)in, vi, p.Cin, p.B, g.bw, g.bh, g.bn)) return 1;
    }
    ReduceDev d;
    d.bw = g.bw; d.bh = g.bh; d.bn = g.bn; d.rows = g.rows;
    d.tiles_x = g.tiles_x; d.tiles_y = g.tiles_y; d.tiles_n = g.tiles_n;
    d.ptiles = g.tiles_x * g.tiles_y * g.tiles_n;
    d.co_tiles = (p.Cout + MT * 128 - 1) / (MT * 128); d.ci_tiles = (p.Cin + BN - 1) / BN;
    d.nslabs = p.taps.ntaps; d.splits = splits;
    d.Cout = p.Cout; d.Cin = p.Cin; d.partial = partial;
    d.taps = p.taps;
    long long grid = (long long)d.co_tiles * d.ci_tiles * (d.nslabs / NT) * splits;
    if (MT == 2 && BN == 256) return launch_reduce<2, 256, 1>(tm, d, grid, st);
    if (MT == 2) return launch_reduce<2, 128, 1>(tm, d, grid, st);
    if (BN == 256) return launch_reduce<1, 256, 1>(tm, d, grid, st);
    if (NT == 3) return launch_reduce<1, 128, 3>(tm, d, grid, st);
    return launch_reduce<1, 128, 1>(tm, d, grid, st);
}

}  // namespace mvae
