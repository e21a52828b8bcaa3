// stream_bench.cu -- how fast can 148 CTAs pull data into shared memory with 1-D bulk async copies
// (cp.async.bulk + mbarrier), from HBM (buffer >> L2) and from L2 (buffer re-read while resident)?  Tuning aid.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mb_expect(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mb_wait(uint64_t* b, uint32_t par) {
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk(void* dst, const void* src, uint32_t n, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(src), "r"(n), "r"(s32(bar)) : "memory");
}

// every CTA streams [blockIdx.x * per_cta, +per_cta) of the buffer `reps` times through an ns-slot ring of `tile` bytes
__global__ void __launch_bounds__(64, 1) k(const uint8_t* buf, size_t per_cta, int tile, int ns, int reps, int pieces) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + 16;
    uint8_t* ring = smem + 256;
    if (threadIdx.x == 0) { for (int i = 0; i < ns; i++) { mb_init(&full[i], 1); mb_init(&empty[i], 1); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const int ntiles = (int)(per_cta / tile) * reps;
    const int per_rep = (int)(per_cta / tile);
    const uint8_t* base = buf + (size_t)blockIdx.x * per_cta;
    if (threadIdx.x == 0) {
        for (int it = 0; it < ntiles; it++) {
            const int s = it % ns;
            if (it >= ns) mb_wait(&empty[s], ((it / ns) - 1) & 1);
            mb_expect(&full[s], tile);
            const uint8_t* src = base + (size_t)(it % per_rep) * tile;
            for (int p = 0; p < pieces; p++) bulk(ring + (size_t)s * tile + p * (tile / pieces), src + p * (tile / pieces), tile / pieces, &full[s]);
        }
    } else if (threadIdx.x == 32) {
        for (int it = 0; it < ntiles; it++) {
            const int s = it % ns;
            mb_wait(&full[s], (it / ns) & 1);
            mb_arrive(&empty[s]);
        }
    }
}

int main() {
    const size_t cap = (size_t)4 << 30;
    uint8_t* buf; CK(cudaMalloc(&buf, cap)); CK(cudaMemset(buf, 1, cap));
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int grid : {132, 148}) for (int tile : {8 << 10, 16 << 10, 32 << 10, 64 << 10}) for (int pieces : {1, 4}) {
        const int ns = (200 * 1024) / tile > 16 ? 16 : (200 * 1024) / tile;
        for (size_t total : {(size_t)24 << 20, (size_t)48 << 20, (size_t)96 << 20, (size_t)3 << 30}) {
            size_t per_cta = total / grid / tile * tile;
            const int reps = (int)(((size_t)6 << 30) / (per_cta * grid)) + 1;
            k<<<grid, 64, 256 + ns * tile>>>(buf, per_cta, tile, ns, 2, pieces);      // warm (and load L2)
            CK(cudaEventRecord(e0));
            k<<<grid, 64, 256 + ns * tile>>>(buf, per_cta, tile, ns, reps, pieces);
            CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            printf("grid %3d tile %3d KB x%d slots, %d copies/tile, working set %5zu MB: %7.1f GB/s\n", grid, tile >> 10, ns, pieces, (per_cta * grid) >> 20,
                   (double)per_cta * grid * reps / (ms * 1e-3) / 1e9);
        }
    }
    return 0;
}
