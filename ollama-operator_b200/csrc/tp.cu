// tp.cu -- the exchange step of tensor parallelism: all-reduce of the K-split projections
// (attention output, FFN down) over NVLink peer memory, fused with the residual add.
//
// One process drives all shards (one per GPU, peer access enabled).  Each shard's GEMV leaves its
// UNROUNDED f64 partial sums in a local buffer (EPI_PARTIAL).  This kernel then
//   1. publishes "epoch e of shard r is complete" into every peer's flag array (st.release.sys),
//   2. waits until every peer has published epoch e into the local array (ld.acquire.sys),
//   3. reads all partial vectors (its own from HBM, the peers' over NVLink), adds them in device
//      order in f64, rounds once to f32 and adds the residual: x += (float)sum.
// Every shard computes the full vector redundantly, in the same order, so the residual stream stays
// bit-identical on all GPUs -- and bit-identical to a single-GPU run, because the canonical
// arithmetic adds block terms in f64 and rounds once either way.  No NCCL call,
// no host involvement: the kernel sits in each shard's CUDA graph like any other.
//
// Buffers alternate (attention output uses 0, FFN down uses 1): a shard can only overwrite buffer b
// again after it has passed the reduce of the other buffer, which in turn waited for every peer to
// have finished reading b (peers publish the next epoch only after their previous reduce kernel).
#include <algorithm>

#include "common.cuh"
#include "kernels.h"

namespace b2 {

namespace {

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double2 ld_relaxed_sys_f64x2(const double* p) {
    double2 v;
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(256) tp_allreduce_residual_kernel(const TpPeers p, const int buf, float* __restrict__ x, const int n) {
    pdl_trigger();       // the next GEMV may start streaming its weights; it waits for this kernel before touching x
    pdl_wait();          // the GEMV that wrote our partial sums has completed and flushed
    const unsigned e = *reinterpret_cast<volatile unsigned*>(p.epoch) + 1;     // the same value in every CTA: only the last CTA to leave bumps it
    if (blockIdx.x == 0 && threadIdx.x < p.n && (int)threadIdx.x != p.rank) {
        // the GEMV that wrote our partial sums finished before this kernel started (stream order)
        __threadfence_system();
        st_release_sys(p.flags[threadIdx.x] + p.rank, e);
    }
    if (threadIdx.x < p.n && (int)threadIdx.x != p.rank) {
        const unsigned* f = p.flags[p.rank] + threadIdx.x;
        SpinGuard guard;           // a peer that never publishes this epoch becomes a trap (B2_ERR_CUDA on the host), not a hung GPU
        while ((int)(ld_acquire_sys(f) - e) < 0) guard.tick();
    }
    __syncthreads();
    for (int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x); i < n; i += 2 * gridDim.x * blockDim.x) {
        // all peers' values in flight at once (one NVLink round trip, not one per peer), then the fixed-order sum
        double2 v[TP_MAX];
#pragma unroll
        for (int d = 0; d < TP_MAX; d++)
            if (d < p.n) v[d] = ld_relaxed_sys_f64x2(p.part[buf][d] + i);
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int d = 0; d < TP_MAX; d++)
            if (d < p.n) { s0 += v[d].x; s1 += v[d].y; }
        float2 r = *reinterpret_cast<float2*>(x + i);
        r.x = __fadd_rn(r.x, (float)s0);
        r.y = __fadd_rn(r.y, (float)s1);
        *reinterpret_cast<float2*>(x + i) = r;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(p.ticket, 1u);
        if (t == gridDim.x - 1) {
            *p.ticket = 0;
            __threadfence();
            *reinterpret_cast<volatile unsigned*>(p.epoch) = e;
        }
    }
}

}  // namespace

int tp_allreduce_residual_launch(const TpPeers& p, int buf, float* x, int n, cudaStream_t st, bool pdl) {
    if (p.n < 2 || p.n > TP_MAX || (n & 1) || buf < 0 || buf > 1) B2_FAIL(B2_ERR_ARG, "tp all-reduce: bad arguments");
    const int grid = std::max(1, std::min(32, n / 512));     // one pair of elements per thread
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, tp_allreduce_residual_kernel, p, buf, x, n));
    return B2_OK;
}

}  // namespace b2
