// rowloop_bench.cu -- micro-benchmark of the Q4_0 row-dot inner loop variants (tuning aid, not product code).
// Each consumer warp repeats the row loop over a shared-memory tile; cycles per iteration are reported
// for several numbers of active warps, so issue-bound and latency-bound regimes can be told apart.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t u4c(const uint4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) { int d; asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }

struct X { uint4 a, b; float d; int s8; };

__device__ __forceinline__ float unit_dot(const X& x, const uint8_t* q, const uint8_t* sc, int u) {
    const uint4 w = *reinterpret_cast<const uint4*>(q + u * 16);
    const float dw = __half2float(*reinterpret_cast<const __half*>(sc + u * 2));
    int lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t ww = u4c(w, i);
        lo = __dp4a((int)(ww & 0x0F0F0F0Fu), (int)u4c(x.a, i), lo);
        hi = dp4a_us(ww & 0xF0F0F0F0u, (int)u4c(x.b, i), hi);
    }
    return __fmul_rn(__fmul_rn((float)(lo + (hi >> 4) - x.s8), dw), x.d);
}

template <typename T> __device__ __forceinline__ T bfly2(T acc0, T acc1, int lane) {
    const bool upper = (lane & 16) != 0;
    T keep = upper ? acc1 : acc0;
    const T give = upper ? acc0 : acc1;
    keep += __shfl_xor_sync(0xffffffffu, give, 16);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
    return keep;
}
// four rows: lanes end with row ((lane>>4)&1)*2 + ((lane>>3)&1)
template <typename T> __device__ __forceinline__ T bfly4(const T (&a)[4], int lane) {
    const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0;
    T k[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const T keep = b4 ? a[2 + i] : a[i], send = b4 ? a[i] : a[2 + i];
        k[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    T r = (b3 ? k[1] : k[0]) + __shfl_xor_sync(0xffffffffu, b3 ? k[0] : k[1], 8);
    r += __shfl_xor_sync(0xffffffffu, r, 4);
    r += __shfl_xor_sync(0xffffffffu, r, 2);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

constexpr int NU = 128;                 // units per row (K = 4096)
constexpr int QS = NU * 16, SS = NU * 2;    // plane strides
constexpr int RT = 32;                  // rows in the tile

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int V>
__global__ void __launch_bounds__(512, 1) k(const uint8_t* gq, const uint8_t* gs, int iters, int nact, long long* out, double* sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t sbar[2];
    __shared__ volatile int sdone;
    uint8_t* q = smem; uint8_t* sc = smem + RT * QS;
    for (int i = threadIdx.x; i < RT * QS / 16; i += blockDim.x) reinterpret_cast<uint4*>(q)[i] = reinterpret_cast<const uint4*>(gq)[i];
    for (int i = threadIdx.x; i < RT * SS / 16; i += blockDim.x) reinterpret_cast<uint4*>(sc)[i] = reinterpret_cast<const uint4*>(gs)[i];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (V == 5) {
        if (threadIdx.x == 0) { sdone = 0; for (int i = 0; i < 2; i++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&sbar[i]))); asm volatile("fence.mbarrier_init.release.cluster;"); }
        __syncthreads();
        if (warp == 15) {      // streamer: 64 KB bulk copies from global into a scratch region of shared memory, back to back, while the others compute
            if (lane == 0) {
                uint8_t* dst = smem + RT * (QS + SS);
                for (int it = 0; !sdone; it++) {
                    const int b = it & 1;
                    if (it >= 2) { uint32_t ok = 0; while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(&sbar[b])), "r"(((it >> 1) - 1) & 1)); }
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&sbar[b])), "r"(32768));
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst + b * 32768)), "l"(gq + (size_t)((blockIdx.x * 7 + it) % 1024) * 32768), "r"(32768), "r"(s32(&sbar[b])));
                }
            }
            return;
        }
    }
    if (warp >= nact) { return; }
    asm volatile("bar.sync 1, %0;" :: "r"(nact * 32));
    X x[4]; int unit[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        unit[j] = j * 32 + lane;
        x[j].a = make_uint4(lane * 7 + j, lane + 3, j * 11, 5); x[j].b = make_uint4(lane, 9 * j, 77, lane * 3);
        x[j].d = 0.01f * (j + 1); x[j].s8 = 8 * (lane - 16);
    }
    double total = 0;
    const long long t0 = clock64();
    if (V == 0 || V == 1 || V == 5) {                     // current structure: two rows, butterfly each iteration (V1: f32 accumulation)
        for (int it = 0; it < iters; it++) {
            const int r = (2 * warp + 2 * it) % RT;
            const uint8_t *q0 = q + r * QS, *q1 = q0 + QS, *s0 = sc + r * SS, *s1 = s0 + SS;
            if (V == 0 || V == 5) {
                double a0 = 0, a1 = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) { a0 += (double)unit_dot(x[j], q0, s0, unit[j]); a1 += (double)unit_dot(x[j], q1, s1, unit[j]); }
                total += bfly2(a0, a1, lane);
            } else {
                float a0 = 0, a1 = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) { a0 += unit_dot(x[j], q0, s0, unit[j]); a1 += unit_dot(x[j], q1, s1, unit[j]); }
                total += (double)bfly2(a0, a1, lane);
            }
        }
    } else if (V == 2) {                        // butterfly of the previous iteration after this iteration's dots
        double p0 = 0, p1 = 0;
        for (int it = 0; it < iters; it++) {
            const int r = (2 * warp + 2 * it) % RT;
            const uint8_t *q0 = q + r * QS, *q1 = q0 + QS, *s0 = sc + r * SS, *s1 = s0 + SS;
            double a0 = 0, a1 = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) { a0 += (double)unit_dot(x[j], q0, s0, unit[j]); a1 += (double)unit_dot(x[j], q1, s1, unit[j]); }
            total += bfly2(p0, p1, lane);
            p0 = a0; p1 = a1;
        }
        total += bfly2(p0, p1, lane);
    } else if (V == 3) {                        // four rows per iteration, one transposed butterfly (iters counts row PAIRS)
        for (int it = 0; it < iters; it += 2) {
            const int r = (4 * warp + 2 * it) % RT;
            double a[4] = {0, 0, 0, 0};
#pragma unroll
            for (int j = 0; j < 4; j++)
#pragma unroll
                for (int rr = 0; rr < 4; rr++) a[rr] += (double)unit_dot(x[j], q + ((r + rr) % RT) * QS, sc + ((r + rr) % RT) * SS, unit[j]);
            total += bfly4(a, lane);
        }
    } else if (V == 4) {                        // four rows, f32 per-lane partial sums of the 4 units kept separate, f64 only in the butterfly
        for (int it = 0; it < iters; it += 2) {
            const int r = (4 * warp + 2 * it) % RT;
            double a[4];
#pragma unroll
            for (int rr = 0; rr < 4; rr++) {
                const uint8_t* qq = q + ((r + rr) % RT) * QS; const uint8_t* ss = sc + ((r + rr) % RT) * SS;
                const float t0 = unit_dot(x[0], qq, ss, unit[0]), t1 = unit_dot(x[1], qq, ss, unit[1]), t2 = unit_dot(x[2], qq, ss, unit[2]), t3 = unit_dot(x[3], qq, ss, unit[3]);
                a[rr] = ((double)t0 + (double)t1) + ((double)t2 + (double)t3);
            }
            total += bfly4(a, lane);
        }
    }
    const long long t1 = clock64();
    if (V == 5) { asm volatile("bar.sync 2, %0;" ::"r"(nact * 32)); if (threadIdx.x == 0) sdone = 1; }
    if (lane == 0) out[blockIdx.x * 16 + warp] = t1 - t0;
    if (total == 1.2345) sink[0] = total;
}

template <int V> void run(const char* name, const uint8_t* gq, const uint8_t* gs, long long* dout, double* sink, int iters) {
    const int smem = RT * (QS + SS) + (V == 5 ? 65536 : 0);
    CK(cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    static long long h[148 * 16];
    for (int nact : {8, 15}) {
        k<V><<<148, 512, smem>>>(gq, gs, iters, nact, dout, sink);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost));
        double mx = 0, mean = 0; int n = 0;
        for (int b = 0; b < 148; b++) for (int w = 0; w < nact; w++) { const double c = (double)h[b * 16 + w] / iters; mx = c > mx ? c : mx; mean += c; n++; }
        printf("%-28s warps %2d  cycles/row-pair: mean %7.1f max %7.1f   -> per-SM row pairs per us @1.9GHz: %6.2f\n", name, nact, mean / n, mx, nact * 1900.0 / (mean / n));
    }
}

int main() {
    const size_t nq = RT * QS, ns = RT * SS;
    uint8_t *hq = (uint8_t*)malloc(nq), *hs = (uint8_t*)malloc(ns);
    for (size_t i = 0; i < nq; i++) hq[i] = (uint8_t)(rand() & 0xFF);
    for (size_t i = 0; i < ns / 2; i++) reinterpret_cast<__half*>(hs)[i] = __float2half(0.01f + (rand() % 100) * 1e-4f);
    uint8_t *gq, *gs; long long* dout; double* sink;
    CK(cudaMalloc(&gq, (size_t)1024 * 32768 + nq)); CK(cudaMalloc(&gs, ns)); CK(cudaMalloc(&dout, 148 * 16 * 8)); CK(cudaMalloc(&sink, 8));
    CK(cudaMemcpy(gq, hq, nq, cudaMemcpyHostToDevice)); CK(cudaMemcpy(gs, hs, ns, cudaMemcpyHostToDevice));
    for (int it : {1, 2, 4, 6, 12}) { printf("-- short runs: %d iterations per warp (cold start after a block barrier)\n", it); run<0>("V0 short", gq, gs, dout, sink, it); run<2>("V2 short", gq, gs, dout, sink, it); }
    const int iters = 400;
    run<0>("V0 current (f64, 2 rows)", gq, gs, dout, sink, iters);
    run<1>("V1 f32 accumulate", gq, gs, dout, sink, iters);
    run<2>("V2 deferred butterfly", gq, gs, dout, sink, iters);
    run<3>("V3 four rows", gq, gs, dout, sink, iters);
    run<4>("V4 four rows, tree cvt", gq, gs, dout, sink, iters);
    run<5>("V5 = V0 + bulk copies into smem", gq, gs, dout, sink, iters);
    return 0;
}
