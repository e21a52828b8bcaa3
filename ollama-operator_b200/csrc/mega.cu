// mega.cu -- the persistent decode kernel: every layer of one decode step in ONE launch.
//
// Why.  At batch 1 a 7B layer is 113.8 MB of weights = 17 us of HBM time on B200, split over four
// matrices and an attention; run as five kernels, each boundary costs 4-5 us of drain, launch,
// prologue and pipeline refill (profiles/r01a_*), i.e. more than the streaming itself.  Here one
// CTA per SM stays resident for the whole step:
//   * a producer warp walks the step's weight matrices in order and streams this CTA's row tiles
//     into a shared-memory ring (1-D bulk async copies, mbarrier complete_tx, L2 evict-first); it
//     never waits for anything but a free ring slot, so HBM keeps streaming across matrix
//     boundaries, grid barriers and the attention stage;
//   * the consumer warps run the stages of a layer -- fused norm+QKV+RoPE+KV-append, attention,
//     wo+residual, norm+gate|up+SiLU, down+residual -- separated by grid-wide barriers (one
//     atomic counter in global memory), using the same prologue / integer-dot / epilogue code as
//     the per-matrix kernel (gemv_dev.cuh), hence bit-identical results;
//   * attention runs inside the same kernel: the grid is launched as thread-block clusters, each
//     cluster takes KV heads round-robin and exchanges max / sum / partial outputs through DSMEM
//     (attn_dev.cuh), signalling with remote mbarrier arrives so the producer warp is not involved.
// The kernel is launched cooperatively (all CTAs must be co-resident for the grid barrier).
#include "attn_dev.cuh"
#include "common.cuh"
#include "formats.cuh"
#include "gemv_dev.cuh"
#include "kernels.h"

namespace b2 {

namespace {

using namespace gemvdev;
using namespace attndev;

constexpr int MEGA_NS_MAX = 8;

struct MegaSmem {            // dynamic shared-memory carve-up (byte offsets), computed on the host
    uint32_t off_red, off_xq, off_xd, off_xs, off_att, off_ring;
    uint32_t stage_bytes;
    int ns;
};

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Grid-wide barrier among the consumer threads of all CTAs.  `count` is zeroed before the launch;
// the n-th barrier completes when it reaches n * gridDim.x.
__device__ __forceinline__ void grid_barrier(unsigned* count, unsigned& epoch, int tid, int ntc) {
    named_bar_sync(1, ntc);                    // every consumer of this CTA has issued its global writes
    if (tid == 0) {
        epoch++;
        __threadfence();
        atomicAdd(count, 1u);
        const unsigned target = epoch * gridDim.x;
        while (ld_acquire_gpu(count) < target) {
        }
        __threadfence();
    }
    named_bar_sync(1, ntc);
}

template <int TYPE, int NB, int NCW, int GRP, int HD>
__global__ void __launch_bounds__((NCW + 1) * 32, 1) mega_kernel(const MegaStage* __restrict__ stages, int s_begin, int s_end, unsigned* grid_count,
                                                                   const MegaSmem sm) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + MEGA_NS_MAX;
    uint64_t* xbar = empty + MEGA_NS_MAX;                                        // cluster exchange barrier (attention)
    double* dscratch = reinterpret_cast<double*>(smem + (2 * MEGA_NS_MAX + 2) * 8);   // 32 doubles
    double* red = reinterpret_cast<double*>(smem + sm.off_red);
    uint8_t* xq = smem + sm.off_xq;
    float* xd = reinterpret_cast<float*>(smem + sm.off_xd);
    int* xs = reinterpret_cast<int*>(smem + sm.off_xs);
    uint8_t* att_smem = smem + sm.off_att;
    uint8_t* ring = smem + sm.off_ring;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int ntc = NCW * 32;
    constexpr int NP = TYPE == T_Q4_0 || TYPE == T_Q8_0 ? 2 : TYPE == T_Q4_K ? 3 : 4;
    const int cluster_size = (int)cluster_nctarank();

    if (tid == 0) {
        for (int s = 0; s < sm.ns; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], NCW); }
        mbar_init(xbar, cluster_size);
        mbar_fence_init();
    }
    __syncthreads();
    cluster_sync_all();        // every CTA's exchange barrier exists before a peer can arrive on it

    if (warp == NCW) {
        // ===== producer: this CTA's tiles of every GEMV stage, in stage order =====
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            uint32_t it = 0;
            for (int st = s_begin; st < s_end; st++) {
                const MegaStage& S = stages[st];
                if (S.kind != 0) continue;
                const QMat& W = S.g.W;
                const int64_t pairs = W.rows >> 1;
                const int64_t r_begin = 2 * ((int64_t)blockIdx.x * pairs / gridDim.x), r_end = 2 * ((int64_t)(blockIdx.x + 1) * pairs / gridDim.x);
                for (int64_t row0 = r_begin; row0 < r_end; row0 += S.RT, it++) {
                    const uint32_t s = it % sm.ns;
                    if (it >= (uint32_t)sm.ns) mbar_wait(&empty[s], ((it / sm.ns) - 1) & 1);
                    const int rows_here = (int)min((int64_t)S.RT, r_end - row0);
                    uint32_t total = 0;
#pragma unroll
                    for (int p = 0; p < NP; p++) total += (uint32_t)rows_here * W.stride[p];
                    mbar_expect_tx(&full[s], total);
                    uint8_t* dst = ring + (size_t)s * sm.stage_bytes;
#pragma unroll
                    for (int p = 0; p < NP; p++) {
                        bulk_g2s(dst, W.plane[p] + (size_t)row0 * W.stride[p], (uint32_t)rows_here * W.stride[p], &full[s], pol);
                        dst += (size_t)S.RT * W.stride[p];
                    }
                }
            }
        }
        return;
    }

    // ===== consumers =====
    uint32_t it = 0, xphase = 0;
    unsigned epoch = 0;
    for (int st = s_begin; st < s_end; st++) {
        const MegaStage& S = stages[st];
        if (st > s_begin) grid_barrier(grid_count, epoch, tid, ntc);     // the previous stage's outputs are complete and visible

        if (S.kind != 0) {
            // ---- attention: clusters take KV heads round-robin ----
            const AttnArgs& a = S.at;
            AttnCfg ac;
            ac.cluster = cluster_size;
            ac.max_local = S.att_max_local;
            const int n_clusters = gridDim.x / cluster_size, my_cluster = blockIdx.x / cluster_size;
            for (int kvh = my_cluster; kvh < a.n_head_kv; kvh += n_clusters)
                attn_kv_head<GRP, HD, ntc>(a, ac, kvh, att_smem, tid, xbar, xphase, 1);
            continue;
        }

        // ---- GEMV stage ----
        const GemvArgs& a = S.g;
        const int Sw = S.S, P = S.P, G = S.G, RPG = S.RPG, RT = S.RT, nunits = S.nunits, nbeff = S.nbeff, WPG = Sw * P;
        {
            const ProStats ps = prologue_stats(a, dscratch, tid, ntc);
            if (TYPE == T_Q4_0 || TYPE == T_Q8_0) prologue_q8_0(a, ps, xq, xd, xs, tid, ntc);
            else prologue_q8_K(a, ps, xq, xd, xs, tid, ntc);
            named_bar_sync(1, ntc);
        }
        const bool working = warp < G * WPG;
        const int g = warp / WPG, ch = (warp % WPG) % Sw, rsel = (warp % WPG) / Sw;
        XRegs<TYPE> xr[NB];
        int unit[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) {
            const int u = (ch * nbeff + j) * 32 + lane;
            const bool valid = working && j < nbeff && u < nunits;
            unit[j] = u < nunits ? u : nunits - 1;
            xr[j].load(xq, xd, xs, unit[j], nunits, valid);
        }
        named_bar_sync(1, ntc);        // x is in registers: the next stage's prologue may overwrite xq/xd/xs

        const int64_t pairs = a.W.rows >> 1;
        const int64_t r_begin = 2 * ((int64_t)blockIdx.x * pairs / gridDim.x), r_end = 2 * ((int64_t)(blockIdx.x + 1) * pairs / gridDim.x);
        unsigned long long best = 0ull;
        for (int64_t row0 = r_begin; row0 < r_end; row0 += RT, it++) {
            const uint32_t s = it % sm.ns;
            mbar_wait(&full[s], (it / sm.ns) & 1);
            const int rows_here = (int)min((int64_t)RT, r_end - row0);
            const uint8_t* sp = ring + (size_t)s * sm.stage_bytes;
            double* redt = red + (size_t)(it & 1) * RT * Sw;
            if (working) {
                for (int rr = rsel; rr < RPG; rr += P) {
                    const int r = g * RPG + rr;
                    if (r >= rows_here) break;
                    RowPtrs rp;
                    const uint8_t* pb = sp;
#pragma unroll
                    for (int p = 0; p < NP; p++) { rp.p[p] = pb + (size_t)r * a.W.stride[p]; pb += (size_t)RT * a.W.stride[p]; }
                    double acc = 0.0;
#pragma unroll
                    for (int j = 0; j < NB; j++) acc += (double)unit_dot(xr[j], rp, unit[j], nunits);
                    acc = warp_sum_d(acc);
                    if (lane == 0) redt[r * Sw + ch] = acc;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
            if (working) {
                if (WPG > 1) named_bar_sync(2 + g, WPG * 32); else __syncwarp();
                if (ch == 0 && rsel == 0) {
                    const int r = g * RPG + 2 * lane;
                    if (2 * lane < RPG && r < rows_here) {
                        double d0 = 0.0, d1 = 0.0;
                        for (int k = 0; k < Sw; k++) { d0 += redt[r * Sw + k]; d1 += redt[(r + 1) * Sw + k]; }
                        epilogue_pair(a, a.epi_row0 + row0 + r, (float)d0, (float)d1, best);
                    }
                }
            }
        }
    }
}

int g_mega_sms = 0;

}  // namespace

// ---- host side ---------------------------------------------------------------------------------------
int mega_plan_stage(MegaStage& st, int ncw) {
    const QMat& W = st.g.W;
    const bool kq = W.type == T_Q4_K || W.type == T_Q6_K;
    const int nb_max = kq ? 1 : 4;
    st.nunits = (int)(kq ? W.k / 128 : W.k / 32);
    st.P = 2;
    // units per lane: the choice that leaves the fewest consumer warps idle, then the fewest idle lanes
    int best_idle = 1 << 30;
    for (int nbe = nb_max; nbe >= 1; nbe--) {
        const int S = (st.nunits + 32 * nbe - 1) / (32 * nbe);
        const int wpg = S * st.P;
        if (wpg > ncw || ncw / wpg > 13) continue;
        const int G = ncw / wpg;
        const int idle = (ncw - G * wpg) * 1024 + (S * 32 * nbe - st.nunits);
        if (idle < best_idle) { best_idle = idle; st.nbeff = nbe; st.S = S; st.G = G; }
    }
    if (best_idle == (1 << 30)) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: k=%lld does not map onto %d consumer warps", (long long)W.k, ncw);
    st.RPG = 2;
    st.RT = st.G * st.RPG;
    return B2_OK;
}

template <int TYPE, int NB, int NCW, int GRP, int HD>
static int mega_launch_t(const MegaLaunch& L, cudaStream_t stream) {
    auto kern = mega_kernel<TYPE, NB, NCW, GRP, HD>;
    static bool configured = false;
    if (!configured) {
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 0));
        configured = true;
    }
    MegaSmem sm;
    sm.off_red = L.off_red; sm.off_xq = L.off_xq; sm.off_xd = L.off_xd; sm.off_xs = L.off_xs; sm.off_att = L.off_att; sm.off_ring = L.off_ring;
    sm.stage_bytes = L.stage_bytes; sm.ns = L.ns;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(L.grid);
    cfg.blockDim = dim3((NCW + 1) * 32);
    cfg.dynamicSmemBytes = L.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = L.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative;
    at[1].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 2;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, L.stages, L.s_begin, L.s_end, L.grid_count, sm));
    return B2_OK;
}

template <int TYPE, int NB, int NCW, int GRP, int HD>
static int mega_occupancy_t(const MegaLaunch& L, int* max_clusters) {
    auto kern = mega_kernel<TYPE, NB, NCW, GRP, HD>;
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(L.cluster * 64);
    cfg.blockDim = dim3((NCW + 1) * 32);
    cfg.dynamicSmemBytes = L.smem_bytes;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = L.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B2_CUDA(cudaOccupancyMaxActiveClusters(max_clusters, kern, &cfg));
    return B2_OK;
}

#define B2_MEGA_DISPATCH(FN, ...)                                                                          \
    do {                                                                                                   \
        if (L.type == T_Q4_0 && L.grp == 1 && L.head_dim == 64) return FN<T_Q4_0, 4, MEGA_NCW, 1, 64>(__VA_ARGS__);   \
        if (L.type == T_Q4_0 && L.grp == 1 && L.head_dim == 128) return FN<T_Q4_0, 4, MEGA_NCW, 1, 128>(__VA_ARGS__); \
        if (L.type == T_Q4_0 && L.grp == 8 && L.head_dim == 128) return FN<T_Q4_0, 4, MEGA_NCW, 8, 128>(__VA_ARGS__); \
    } while (0)

bool mega_supported(int type, int grp, int head_dim) {
    return type == T_Q4_0 && ((grp == 1 && (head_dim == 64 || head_dim == 128)) || (grp == 8 && head_dim == 128));
}

// Fill in the shared-memory layout and the grid for a model: k_max = longest activation vector,
// row_bytes_max = bytes of the largest stage tile row set (RT rows).
int mega_configure(MegaLaunch& L, int64_t k_max, size_t tile_bytes_max, int rt_s_max, int max_ctx) {
    if (!mega_supported(L.type, L.grp, L.head_dim)) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: no kernel for type %d, %d query heads per KV head, head_dim %d", L.type, L.grp, L.head_dim);
    const bool kq = L.type == T_Q4_K || L.type == T_Q6_K;
    size_t off = (2 * MEGA_NS_MAX + 2) * 8 + 32 * 8;
    L.off_red = (uint32_t)off; off += (size_t)2 * rt_s_max * 8; off = (off + 15) / 16 * 16;
    L.off_xq = (uint32_t)off; off += (size_t)k_max; off = (off + 15) / 16 * 16;
    L.off_xd = (uint32_t)off; off += (size_t)(kq ? k_max / 256 : k_max / 32) * 4; off = (off + 15) / 16 * 16;
    L.off_xs = (uint32_t)off; off += (size_t)(kq ? k_max / 16 : k_max / 32) * 4; off = (off + 15) / 16 * 16;
    L.off_att = (uint32_t)off;
    L.att_max_local = ((max_ctx + L.cluster - 1) / L.cluster + 3) / 4 * 4;
    off += attndev::attn_smem_bytes(L.grp, L.head_dim, L.att_max_local, L.cluster, MEGA_NCW * 32);
    off = (off + 127) / 128 * 128;
    L.off_ring = (uint32_t)off;
    L.stage_bytes = (uint32_t)((tile_bytes_max + 127) / 128 * 128);
    const size_t cap = 227 * 1024;
    if (off + 2 * (size_t)L.stage_bytes > cap) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: shared memory does not fit (%zu B before the ring, %u B per stage)", off, L.stage_bytes);
    L.ns = (int)std::min<size_t>(MEGA_NS_MAX, (cap - off) / L.stage_bytes);
    L.smem_bytes = off + (size_t)L.ns * L.stage_bytes;
    int max_clusters = 0;
    {
        auto occ = [&](int* mc) -> int { B2_MEGA_DISPATCH(mega_occupancy_t, L, mc); B2_FAIL(B2_ERR_UNSUPPORTED, "mega: dispatch"); };
        B2_TRY(occ(&max_clusters));
    }
    if (max_clusters <= 0) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: the kernel does not fit on this device");
    L.grid = max_clusters * L.cluster;
    return B2_OK;
}

int mega_launch(const MegaLaunch& L, cudaStream_t stream) {
    B2_MEGA_DISPATCH(mega_launch_t, L, stream);
    B2_FAIL(B2_ERR_UNSUPPORTED, "mega: no kernel for type %d grp %d head_dim %d", L.type, L.grp, L.head_dim);
}

}  // namespace b2
