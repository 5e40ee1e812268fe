// Cycles per tcgen05.mma instruction on one SM, operands resident in shared memory (no loads in flight):
//   SS  M=128 N in {64,128,256}, one or two alternating accumulators;  TS (A from tensor memory) N in {128,256}
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/umma_rate_probe scripts/umma_rate_probe.cu -I medvae_disentangled_multimodal_b200/csrc
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include "umma.cuh"
using namespace mvae::sm100;

struct Res { long long cyc; int n; };

template <int N, int TS, int NACC>
__global__ void __launch_bounds__(128, 1) probe(Res* out, int rounds) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
    for (int i = threadIdx.x; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        if (lane == 0) { mbar_init(&bar[0], 1); fence_barrier_init(); }
        __syncwarp();
        tmem_alloc(slot, 512);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
        // 4 stages of A (16 KB) and B (N x 128 B) like the conv kernel's ring
        const uint32_t a_base = smem_u32(smem), b_base = smem_u32(smem + 64 * 1024);
        const long long t0 = clock64();
        int n = 0;
        for (int r = 0; r < rounds; ++r) {
            const int st = r & 3;
            const uint32_t a0 = a_base + st * 16384, b0 = b_base + (st % (96 * 1024 / (N * 128))) * (N * 128);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint64_t bd = smem_desc_sw128(b0 + k * 32, 16, 1024);
#pragma unroll
                for (int j = 0; j < NACC; ++j) {
                    if (TS) umma_bf16_ts(tmem + (uint32_t)(j * N), tmem + 480u, bd, idesc, 1u);
                    else umma_bf16(tmem + (uint32_t)(j * N), smem_desc_sw128(a0 + k * 32, 16, 1024), bd, idesc, 1u);
                    ++n;
                }
            }
        }
        umma_commit(&bar[0]);
        mbar_wait(&bar[0], 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0) { out->cyc = t1 - t0; out->n = n; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

template <int N, int TS, int NACC>
static void run(const char* name, int grid) {
    Res* d; cudaMalloc(&d, sizeof(Res));
    auto k = probe<N, TS, NACC>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 170 * 1024);
    k<<<grid, 128, 170 * 1024>>>(d, 64);
    cudaDeviceSynchronize();
    k<<<grid, 128, 170 * 1024>>>(d, 2048);
    cudaError_t e = cudaDeviceSynchronize();
    Res h; cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const double cyc = (double)h.cyc / h.n;
    printf("%-44s grid %3d  %7.1f clk/MMA  -> %5.1f %% of 8192 flop/clk  (%s)\n", name, grid, cyc,
           100.0 * (2.0 * 128 * N * 16 / cyc) / 8192.0, cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    for (int grid : {1, 148}) {
        run<64, 0, 1>("SS M128 N64", grid);
        run<128, 0, 1>("SS M128 N128", grid);
        run<128, 0, 2>("SS M128 N128 two accumulators (MT=2)", grid);
        run<256, 0, 1>("SS M128 N256", grid);
        run<128, 1, 1>("TS M128 N128 (A in TMEM)", grid);
        run<128, 1, 2>("TS M128 N128 two accumulators", grid);
        run<256, 1, 1>("TS M128 N256 (A in TMEM)", grid);
    }
    return 0;
}
