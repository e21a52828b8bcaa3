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
#include <type_traits>

#include "attn_dev.cuh"
#include "common.cuh"
#include "formats.cuh"
#include "gemv_dev.cuh"
#include "kernels.h"

namespace b2 {

namespace {

using namespace gemvdev;
using namespace attndev;

constexpr int MEGA_NS_MAX = 16;
// Arrival units of a ring slot's `empty` barrier: every WORKING warp of a stage (G * S of them, owner of a pair in the
// tile or not) arrives with MEGA_EMPTY_BASE / (G * S) units, so any stage shape completes the same barrier.  A slot must
// not be refilled before every worker has passed its wait on that tile's `full` barrier: a worker without a pair in the
// tile holds nothing, but if the producer laps it (refills the slot before it has looked), its parity wait matches the
// wrong phase and the kernel deadlocks.  (Seen at contexts beyond ~3800 with tiles of fewer pairs than workers: the KV
// prefetch issued before the row loop staggers the warps by more than a tile there.)
constexpr uint32_t MEGA_EMPTY_BASE = 360360;      // lcm(1..15)
constexpr int MEGA_MAX_PAGES = 256;
constexpr int MEGA_PF_LANES_MAX = 1536;           // contexts up to this many positions: KV prefetch from the producer warp's idle lanes

struct MegaSmem {            // dynamic shared-memory carve-up (byte offsets), computed on the host
    uint32_t off_red, off_xq, off_xd, off_xs, off_att, off_ring;
    uint32_t stage_bytes;
    int ns;
    uint32_t pf_bytes;       // how far (bytes per CTA) the L2 prefetch cursor runs ahead of the ring
    int debug;               // tuning experiments: 1 = consumers skip the arithmetic, 2 = consumers do not wait for data
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
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
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(count) : "memory");
        const unsigned target = epoch * gridDim.x;
        // (polling with relaxed loads plus one acquire fence was measured slower: 1.5-2.1 us per barrier against 1.2-1.5)
        SpinGuard guard;           // a CTA that never arrives (a fault elsewhere) becomes a trap, not a hung GPU
        while (ld_acquire_gpu(count) < target) guard.tick();
    }
    named_bar_sync(1, ntc);
}

// Tensor-parallel exchange, after the grid barrier that completed this GPU's partial sums: publish the
// epoch to the peers, wait for theirs, then every CTA reduces its slice of the vector (tp.cu has the protocol).
// Kept out of line: its registers (one value per peer in flight) must not weigh on the GEMV stages.
__device__ __noinline__ void tp_exchange(const TpPeers& tp, unsigned e, int buf, float* x, int n, int tid, int ntc) {
    if (tid < tp.n && tid != tp.rank) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(tp.flags[tid] + tp.rank), "r"(e) : "memory");
        }
        const unsigned* f = tp.flags[tp.rank] + tid;
        unsigned v;
        SpinGuard guard;           // a peer GPU that never publishes this epoch becomes a trap here
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            guard.tick();
        } while ((int)(v - e) < 0);
    }
    named_bar_sync(1, ntc);
    for (int i = 2 * (blockIdx.x * ntc + tid); i < n; i += 2 * gridDim.x * ntc) {
        double2 v[TP_MAX];
#pragma unroll
        for (int d = 0; d < TP_MAX; d++)
            if (d < tp.n) asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v[d].x), "=d"(v[d].y) : "l"(tp.part[buf][d] + i));
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int d = 0; d < TP_MAX; d++)
            if (d < tp.n) { s0 += v[d].x; s1 += v[d].y; }
        float2 r = *reinterpret_cast<float2*>(x + i);
        r.x = __fadd_rn(r.x, (float)s0);
        r.y = __fadd_rn(r.y, (float)s1);
        *reinterpret_cast<float2*>(x + i) = r;
    }
}

// TRACE: the per-stage timestamps of b2_mega_trace are compiled into a variant of their own, so that the kernel that serves
// requests carries neither their registers nor their branches (it sits at the 128-register limit of its 512 threads)
template <int TYPE, int NB, int NCW, int GRP, int HD, bool TP, bool TRACE = false>
__global__ void __launch_bounds__((NCW + 1) * 32, 1) mega_kernel(const MegaStage* __restrict__ stages, int s_begin, int s_end, unsigned* grid_count,
                                                                   const MegaSmem sm, unsigned long long* trace, const int* pos_ptr, const int* page_table, int n_pages,
                                                                   const __grid_constant__ TpPeers tp) {
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
    constexpr int NP = TYPE == T_Q4_0 || TYPE == T_Q8_0 ? 2 : TYPE == T_Q4_K ? 3 : 4;      // T_KQ: up to 4, the matrix says how many
    const int cluster_size = (int)cluster_nctarank();                            // 1, 2 or 4
    const int n_clusters = gridDim.x >> (__ffs(cluster_size) - 1), my_cluster = blockIdx.x >> (__ffs(cluster_size) - 1);
    // position and page table of this step: fetched once, used by every layer's KV append and attention
    __shared__ int s_pos;
    __shared__ int s_pt[MEGA_MAX_PAGES];
    // The stage table lives in global memory; a field read there costs an L2 round trip, and a stage reads a dozen
    // of them in dependent steps (barrier -> operands -> strides -> loop bounds).  Each stage's descriptor is
    // therefore copied into shared memory one stage ahead (bulk async copy, completion on dbar).
    __shared__ __align__(16) uint8_t s_desc[2][sizeof(MegaStage)];
    // per-thread state kept from layer to layer (gemvdev::EpiCache): thread p of the epilogue owns the same row pair in every layer
    constexpr int EPI_CACHE_PAIRS = 128;
    __shared__ float s_res[2 * EPI_CACHE_PAIRS];
    __shared__ float2 s_rope[EPI_CACHE_PAIRS];
    __shared__ uint64_t dbar[2];
    __shared__ int s_progress;            // stage the consumers have entered (read by the prefetch lanes of the producer warp)
    for (int i = tid; i < n_pages; i += blockDim.x) s_pt[i] = page_table[i];
    if (tid == 0) { s_pos = *pos_ptr; s_progress = s_begin - 1; }

    if (tid == 0) {
        for (int s = 0; s < sm.ns; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], MEGA_EMPTY_BASE); }
        mbar_init(xbar, cluster_size);
        mbar_init(&dbar[0], 1); mbar_init(&dbar[1], 1);
        mbar_fence_init();
        for (int i = 0; i < 2 && s_begin + i < s_end; i++) {
            mbar_expect_tx(&dbar[i], (uint32_t)sizeof(MegaStage));
            bulk_g2s_plain(s_desc[i], &stages[s_begin + i], (uint32_t)sizeof(MegaStage), &dbar[i]);
        }
    }
    __syncthreads();
    cluster_sync_all();        // every CTA's exchange barrier exists before a peer can arrive on it

    if (warp == NCW) {
        // ===== producer: this CTA's tiles of every GEMV stage, in stage order =====
        if (lane == 0 && !(sm.debug & 2)) {
            const uint64_t pol = (sm.debug & 4) ? l2_policy_evict_last() : (sm.debug & 8) ? l2_policy_evict_normal() : l2_policy_evict_first();
            // cursor over (stage, tile) pairs; two of them walk the same sequence: `f` feeds the
            // shared-memory ring, `p` runs up to sm.pf_bytes further ahead issuing L2 prefetches, so
            // that HBM keeps streaming while the consumers sit in a barrier, a prologue or the
            // attention (the ring alone covers ~3 us of HBM time, L2 holds tens of MB more)
            // Everything a tile needs sits in registers once a cursor has entered a stage: the stage table is in
            // global memory, and with tiles of a few KB a descriptor read per tile would bound the producer.
            struct Cursor {
                int st; int rows_left, RT;            // rows of this CTA still to issue in stage `st`
                const uint8_t* pl[NP]; uint32_t stride[NP];
            };
            auto seek = [&](Cursor& c) {         // position on the first tile of the next GEMV stage at or after c.st
                for (; c.st < s_end; c.st++) {
                    const MegaStage& S = stages[c.st];
                    if (S.kind != 0) continue;
                    // row pairs of this CTA: pairs_base each, the first pairs_rem CTAs one more (the consumers use the same split)
                    const uint32_t row0 = 2u * (blockIdx.x * (uint32_t)S.pairs_base + min(blockIdx.x, (uint32_t)S.pairs_rem));
                    c.rows_left = 2 * (S.pairs_base + (blockIdx.x < (uint32_t)S.pairs_rem ? 1 : 0));
                    if (c.rows_left <= 0) continue;
                    c.RT = S.RT;
#pragma unroll
                    for (int pl = 0; pl < NP; pl++) {
                        const bool on = TYPE != T_KQ || pl < S.g.W.n_planes;
                        c.stride[pl] = on ? S.g.W.stride[pl] : 0u;
                        c.pl[pl] = on ? S.g.W.plane[pl] + (size_t)row0 * S.g.W.stride[pl] : nullptr;
                    }
                    return;
                }
            };
            auto advance = [&](Cursor& c) {
                c.rows_left -= c.RT;
                if (c.rows_left <= 0) { c.st++; seek(c); return; }
#pragma unroll
                for (int pl = 0; pl < NP; pl++) c.pl[pl] += (size_t)c.RT * c.stride[pl];
            };
            Cursor f, p;
            f.st = p.st = s_begin; f.rows_left = p.rows_left = 0; f.RT = p.RT = 0;
            seek(f); seek(p);
            long long ahead = 0;                  // bytes prefetched to L2 but not yet requested for the ring
            uint32_t s = 0, lap = 0;              // ring slot of the next tile, times the ring has wrapped
            while (f.st < s_end) {
                while (p.st < s_end && ahead < (long long)sm.pf_bytes) {
                    const uint32_t rows_here = (uint32_t)min(p.RT, p.rows_left);
#pragma unroll
                    for (int pl = 0; pl < NP; pl++) {
                        if (p.stride[pl] == 0) continue;
                        prefetch_l2(p.pl[pl], rows_here * p.stride[pl]);
                        ahead += (long long)(rows_here * p.stride[pl]);
                    }
                    advance(p);
                }
                if (lap) mbar_wait(&empty[s], (lap - 1) & 1);
                const uint32_t rows_here = (uint32_t)min(f.RT, f.rows_left);
                uint32_t total = 0;
#pragma unroll
                for (int pl = 0; pl < NP; pl++) total += rows_here * f.stride[pl];
                mbar_expect_tx(&full[s], total);
                uint8_t* dst = ring + (size_t)s * sm.stage_bytes;
#pragma unroll
                for (int pl = 0; pl < NP; pl++) {
                    if (f.stride[pl] == 0) continue;
                    bulk_g2s(dst, f.pl[pl], rows_here * f.stride[pl], &full[s], pol);
                    dst += (size_t)f.RT * f.stride[pl];
                }
                ahead -= total;
                advance(f);
                if (++s == (uint32_t)sm.ns) { s = 0; lap++; }
            }
        }
        // Lanes 1-31: the L2 prefetch of the K/V rows an attention stage will read, issued when the consumers enter the GEMV
        // stage before it -- here, off the consumers' path (it used to sit between their prologue and their row loop, where
        // it cost 0.7 us per layer at 512 positions and staggered the warps at long contexts).
        // Up to MEGA_PF_LANES_MAX positions only: beyond, 31 lanes are too slow for the ~(pos + 1) * 1.5 lines of a CTA and the
        // consumers' 480 threads keep the job (measured: +1.3 % at 512 positions, equal at 2048, -7 % at 3800).
        if (lane >= 1 && s_pos < MEGA_PF_LANES_MAX && !(sm.debug & (16 | 64))) {
            const int pos_value = s_pos;
            for (int st = s_begin; st + 1 < s_end; st++) {
                if (stages[st].kind != 0 || stages[st + 1].kind != 1) continue;
                while (*reinterpret_cast<volatile int*>(&s_progress) < st - 1) __nanosleep(200);      // (one stage of lead: 31 lanes need a while at long contexts)
                const AttnArgs& na = stages[st + 1].at;
                for (int kvh = my_cluster; kvh < na.n_head_kv; kvh += n_clusters) attn_prefetch<GRP, HD, 31>(na, cluster_size, kvh, lane - 1, pos_value, s_pt);
            }
        }
        return;
    }

    // ===== consumers =====
    const int pos_now = s_pos;        // written before the __syncthreads above
    uint32_t ring_slot = 0, ring_phase = 0, xphase = 0;      // ring position of the next stage's first tile
    unsigned epoch = 0;
    unsigned tp_epoch = TP ? *reinterpret_cast<volatile unsigned*>(tp.epoch) : 0u;      // exchanges completed before this launch
    // Hand-over buffers (GemvArgs::x_ll): write number w of a vector goes to buffer w % NB of its rotation, and the thread
    // that stores row r of write w puts the poison back into row r of the buffer of write w - 1.
    //  * Nobody reads that buffer any more: its readers are the prologue of the stage after write w - 1 (for the q/k/v
    //    hand-over: the attention stage), and the writer of w is past its own prologue, which needed every CTA's output
    //    of a stage after that one.
    //  * The poison is visible before that buffer is polled again (for write w - 1 + NB): a grid barrier lies in between
    //    -- up -> down, the one barrier a layer keeps.  att, h and q/k/v are written once per layer (NB = 2 is enough); x
    //    is written twice (by wo and down), so its rotation has NB = 3 buffers to put that barrier between poison and poll.
    //  * The buffers of the first NB - 1 writes of a launch are poisoned by the kernel that precedes this one (step_begin).
    double* ysum = red;                                   // per-row partial sums of this CTA's rows: [row_local][S]
    bool res_valid = false, rope_valid = false;          // the caches above hold this launch's values
    for (int st = s_begin; st < s_end; st++) {
        const int db = (st - s_begin) & 1;
        mbar_wait(&dbar[db], (uint32_t)((st - s_begin) >> 1) & 1u);
        const MegaStage& S = *reinterpret_cast<const MegaStage*>(s_desc[db]);
        if (tid == 0) *reinterpret_cast<volatile int*>(&s_progress) = st;
        // after this stage's grid barrier nobody reads the previous stage's descriptor any more: fetch the next one over it
        auto fetch_next_desc = [&]() {
            if (tid == 0 && st > s_begin && st + 1 < s_end) {
                mbar_expect_tx(&dbar[db ^ 1], (uint32_t)sizeof(MegaStage));
                bulk_g2s_plain(s_desc[db ^ 1], &stages[st + 1], (uint32_t)sizeof(MegaStage), &dbar[db ^ 1]);
            }
        };
        // optional per-stage timestamps (ns): [0] stage entered, [1] barrier passed, [2] prologue done, [3] stage done
        unsigned long long* tr = TRACE && trace && tid == 0 ? trace + ((size_t)blockIdx.x * (s_end - s_begin) + (st - s_begin)) * 16 : nullptr;
        if (tr) tr[0] = global_ns();

        if (TP && S.kind == 2) {
            // ---- tensor-parallel exchange: x += (float)(sum over the GPUs of the K-split partial sums) ----
            // (tp.cu describes the protocol; here the publishing side is "all CTAs passed the grid barrier")
            grid_barrier(grid_count, epoch, tid, ntc);              // this GPU's partial sums are complete
            fetch_next_desc();
            if (tr) tr[1] = global_ns();
            tp_epoch++;
            tp_exchange(tp, tp_epoch, S.tp_buf, S.tp_x, S.tp_n, tid, ntc);
            if (tr) tr[2] = tr[3] = global_ns();
            continue;
        }
        if (S.kind != 0) {
            const bool att_ll = !TP && S.ll_in != 0;      // q and the new K/V rows come through a hand-over buffer (attn_dev.cuh)
            if (att_ll) named_bar_sync(1, ntc);           // (every thread of this CTA has left the previous stage: its descriptor may go)
            else if (st > s_begin) grid_barrier(grid_count, epoch, tid, ntc);
            fetch_next_desc();
            if (tr) tr[1] = global_ns();
            // ---- attention: clusters take KV heads round-robin ----
            const AttnArgs& a = S.at;
            AttnCfg ac;
            ac.cluster = cluster_size;
            ac.max_local = S.att_max_local;
            ac.ll_out = a.out_ll != nullptr;
            ac.ll_in = att_ll;
            if (GRP > 1 && 2 * a.n_head_kv <= n_clusters) {
                // few KV heads (GQA under tensor parallelism): one query head per cluster instead of one group
                for (int h = my_cluster; h < a.n_head; h += n_clusters)
                    attn_kv_head<1, HD, ntc>(a, ac, h, att_smem, tid, xbar, xphase, 1, pos_now, s_pt, tr ? tr + 4 : nullptr, h / GRP);
            } else {
                for (int kvh = my_cluster; kvh < a.n_head_kv; kvh += n_clusters)
                    attn_kv_head<GRP, HD, ntc>(a, ac, kvh, att_smem, tid, xbar, xphase, 1, pos_now, s_pt, tr ? tr + 4 : nullptr);
            }
            if (tr) tr[2] = tr[3] = global_ns();
            continue;
        }

        // ---- GEMV stage ----
        const GemvArgs& a = S.g;
        const int Sw = S.S, NW = S.G, RT = S.RT, nunits = S.nunits, nbeff = S.nbeff;
        const int K = (int)a.W.k, nchunks = K >> 2;
        constexpr int MAXC = 8;                               // float4 chunks of x per thread held in registers (K <= 4*ntc*MAXC)
        const bool fast = (TYPE == T_Q4_0 || TYPE == T_Q8_0) && a.pro != PRO_LAYERNORM && nchunks <= ntc * MAXC && (K & 255) == 0;
        // ll_in: the input vector carries the producing stage's tag next to every value; polling those replaces the barrier
        // (and the next descriptor is fetched further down, once every thread of this CTA is known to have left the previous stage)
        const bool ll_in = !TP && S.ll_in != 0;
        auto stage_barrier = [&]() {
            if (!ll_in) {
                if (st > s_begin) grid_barrier(grid_count, epoch, tid, ntc);     // the previous stage's outputs are complete and visible
                fetch_next_desc();
            }
            if (tr) { tr[1] = global_ns(); tr[10] = (unsigned long long)clock64(); }
        };
        // Fast prologue: x once into registers (one L2 round trip), RMS statistics, Q8_0 quantisation -- the arithmetic of
        // gemvdev::prologue_stats / prologue_q8_0.  A thread owns "octets" of 8 consecutive values (octet o = tid + k * ntc,
        // k < NO), four lanes share a 32-value block: per value that is half the shuffles, divisions and address arithmetic of
        // a 4-value mapping, and this phase is bound by the instructions the fifteen warps issue.
        auto fast_prologue = [&](auto no_tag, auto norm_tag) {
            constexpr int NO = decltype(no_tag)::value;
            constexpr bool NORM = decltype(norm_tag)::value;     // RMS norm (gains held in registers) or plain quantisation
            const int noct = K >> 3;
            // the norm gain does not depend on the previous stage: fetch it while waiting for the input
            float4 gv[NORM ? 2 * NO : 1];
            if constexpr (NORM) {
#pragma unroll
                for (int k = 0; k < NO; k++) {
                    const int o = tid + k * ntc;
                    gv[2 * k] = o < noct ? *reinterpret_cast<const float4*>(a.gain + 8 * o) : make_float4(0.f, 0.f, 0.f, 0.f);
                    gv[2 * k + 1] = o < noct ? *reinterpret_cast<const float4*>(a.gain + 8 * o + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            stage_barrier();
            float4 xv[2 * NO];
            if (ll_in) {
#pragma unroll
                for (int k = 0; k < NO; k++) {
                    const int o = tid + k * ntc;
                    xv[2 * k] = o < noct ? ll_load4(a.x_ll + 8 * o) : make_float4(0.f, 0.f, 0.f, 0.f);
                    xv[2 * k + 1] = o < noct ? ll_load4(a.x_ll + 8 * o + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int c = 0; c < 2 * NO; c++) {
                    const int o = tid + (c >> 1) * ntc;
                    SpinGuard guard;                        // a producer that never writes becomes a trap, not a hung GPU
                    while (!ll_ready(xv[c])) { xv[c] = ll_load4(a.x_ll + 8 * o + 4 * (c & 1)); guard.tick(); }      // (any __nanosleep here costs 10-20 %;
                }                                                                                                    //  polling all pending chunks at once: 2 % slower)
            } else {
#pragma unroll
                for (int k = 0; k < NO; k++) {
                    const int o = tid + k * ntc;
                    xv[2 * k] = o < noct ? __ldcg(reinterpret_cast<const float4*>(a.x + 8 * o)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    xv[2 * k + 1] = o < noct ? __ldcg(reinterpret_cast<const float4*>(a.x + 8 * o + 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            if (tr) tr[4] = (unsigned long long)clock64();          // (x loads issued)
            float scale = 1.f;
            if constexpr (NORM) {
                double ss = 0;
#pragma unroll
                for (int c = 0; c < 2 * NO; c++) {
                    if ((c >> 1) * ntc >= noct) break;          // (the octets past the vector are zeros: skipping them changes nothing)
                    ss += (double)__fmul_rn(xv[c].x, xv[c].x); ss += (double)__fmul_rn(xv[c].y, xv[c].y);
                    ss += (double)__fmul_rn(xv[c].z, xv[c].z); ss += (double)__fmul_rn(xv[c].w, xv[c].w);
                }
                if (tr) tr[5] = (unsigned long long)clock64();      // (x arrived, squares summed per thread)
                ss = consumers_sum(ss, dscratch, tid, ntc, false);      // (the barrier after the quantisation separates it from the next use)
                const float mean = (float)(ss / (double)K);
                scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, a.eps)));
            }
            if (tr) tr[6] = (unsigned long long)clock64();          // (statistics reduced)
            const int nbk = K >> 5;
            // K is a multiple of 256, so a warp's 32 octets are all inside or all outside the vector:
            // full-mask shuffles (xor 1, 2 stay inside the 4 lanes that own one 32-element block)
            const unsigned gmask = 0xffffffffu;
#pragma unroll
            for (int k = 0; k < NO; k++) {
                const int o = tid + k * ntc;
                if (o < noct) {
                    float4 v0 = xv[2 * k], v1 = xv[2 * k + 1];
                    if constexpr (NORM) {
                        v0.x = __fmul_rn(__fmul_rn(v0.x, scale), gv[2 * k].x); v0.y = __fmul_rn(__fmul_rn(v0.y, scale), gv[2 * k].y);
                        v0.z = __fmul_rn(__fmul_rn(v0.z, scale), gv[2 * k].z); v0.w = __fmul_rn(__fmul_rn(v0.w, scale), gv[2 * k].w);
                        v1.x = __fmul_rn(__fmul_rn(v1.x, scale), gv[2 * k + 1].x); v1.y = __fmul_rn(__fmul_rn(v1.y, scale), gv[2 * k + 1].y);
                        v1.z = __fmul_rn(__fmul_rn(v1.z, scale), gv[2 * k + 1].z); v1.w = __fmul_rn(__fmul_rn(v1.w, scale), gv[2 * k + 1].w);
                    }
                    float amax = fmaxf(fmaxf(fmaxf(fabsf(v0.x), fabsf(v0.y)), fmaxf(fabsf(v0.z), fabsf(v0.w))),
                                       fmaxf(fmaxf(fabsf(v1.x), fabsf(v1.y)), fmaxf(fabsf(v1.z), fabsf(v1.w))));
                    amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, 1));
                    amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, 2));
                    const float d = div127_rn(amax);
                    const float id = d != 0.0f ? __frcp_rn(d) : 0.0f;
                    const int q0 = round_away_i(__fmul_rn(v0.x, id)), q1 = round_away_i(__fmul_rn(v0.y, id));
                    const int q2 = round_away_i(__fmul_rn(v0.z, id)), q3 = round_away_i(__fmul_rn(v0.w, id));
                    const int q4 = round_away_i(__fmul_rn(v1.x, id)), q5 = round_away_i(__fmul_rn(v1.y, id));
                    const int q6 = round_away_i(__fmul_rn(v1.z, id)), q7 = round_away_i(__fmul_rn(v1.w, id));
                    int sum = ((q0 + q1) + (q2 + q3)) + ((q4 + q5) + (q6 + q7));
                    sum += __shfl_xor_sync(gmask, sum, 1);
                    sum += __shfl_xor_sync(gmask, sum, 2);
                    const int bq = o >> 2, e0 = (o & 3) * 8;         // block, first element of the octet inside it
                    uint8_t* dst = xq + (e0 < 16 ? 0 : nbk * 16) + bq * 16 + (e0 & 15);
                    *reinterpret_cast<uint2*>(dst) = make_uint2(pack4(q0, q1, q2, q3), pack4(q4, q5, q6, q7));
                    if (e0 == 0) {
                        xd[bq] = __half2float(__float2half_rn(d));
                        xs[bq] = 8 * sum;
                    }
                }
            }
        };
        if (fast) {
            const bool norm = a.pro == PRO_RMSNORM;
            const int noct = K >> 3;
            if (noct <= 2 * ntc) {
                if (norm) fast_prologue(std::integral_constant<int, 2>{}, std::true_type{});
                else fast_prologue(std::integral_constant<int, 2>{}, std::false_type{});
            } else if (noct <= 3 * ntc) {
                if (norm) fast_prologue(std::integral_constant<int, 3>{}, std::true_type{});
                else fast_prologue(std::integral_constant<int, 3>{}, std::false_type{});
            } else {
                if (norm) fast_prologue(std::integral_constant<int, MAXC / 2>{}, std::true_type{});
                else fast_prologue(std::integral_constant<int, MAXC / 2>{}, std::false_type{});
            }
        } else {
            stage_barrier();
            if (ll_in) {
                // generic prologue over a hand-over buffer: first see every chunk final (each thread polls a share), then
                // run the usual passes, whose loads of x go to L2 (gemvdev::load_x4)
                for (int i = tid * 4; i < K; i += ntc * 4) {
                    float4 v = ll_load4(a.x_ll + i);
                    SpinGuard guard;
                    while (!ll_ready(v)) { v = ll_load4(a.x_ll + i); guard.tick(); }
                }
                named_bar_sync(1, ntc);
            }
            const ProStats ps = prologue_stats(a, dscratch, tid, ntc);
            if (TYPE == T_Q4_0 || TYPE == T_Q8_0) prologue_q8_0(a, ps, xq, xd, xs, tid, ntc);
            else prologue_q8_K(a, ps, xq, xd, xs, tid, ntc);
        }
        if (tr) tr[7] = (unsigned long long)clock64();              // (quantised)
        named_bar_sync(1, ntc);
        if (ll_in) fetch_next_desc();                               // every thread of this CTA is past the previous stage now
        if (tr) tr[8] = (unsigned long long)clock64();              // (block barrier)

        const int worker = (warp * S.inv_S) >> 16, ch = warp - worker * Sw;      // warp / Sw, warp % Sw (reciprocal from the host)
        const bool working = worker < NW;
        XRegs<TYPE> xr[NB];
        int unit[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) {
            const int u = (ch * nbeff + j) * 32 + lane;
            const bool valid = working && j < nbeff && u < nunits;
            unit[j] = u < nunits ? u : nunits - 1;
            xr[j].load(xq, xd, xs, unit[j], nunits, valid, a.W.type);
        }
        if (tr) { tr[2] = global_ns(); tr[9] = (unsigned long long)clock64(); }
        // the attention stage that follows will read K/V rows that already exist: pull them into L2
        // now, while this stage streams its weights
        if (S.next_attn && !(sm.debug & 16) && (pos_now >= MEGA_PF_LANES_MAX || (sm.debug & 64))) {      // (else: the producer warp's lanes 1-31 do it)
            const AttnArgs& na = S.at;                              // the host copied the next stage's arguments here
            for (int kvh = my_cluster; kvh < na.n_head_kv; kvh += n_clusters) attn_prefetch<GRP, HD, ntc>(na, cluster_size, kvh, tid, pos_now, s_pt);
        }

        // rows [r_begin, r_end) of this CTA: whole pairs, split evenly (32-bit: pairs * grid < 2^31 for every supported matrix)
        const int r_begin = (int)(2u * (blockIdx.x * (uint32_t)S.pairs_base + min(blockIdx.x, (uint32_t)S.pairs_rem)));
        const int r_end = r_begin + 2 * (S.pairs_base + (blockIdx.x < (uint32_t)S.pairs_rem ? 1 : 0));
        // the stage descriptor lives in global memory and the waits below are compiler barriers: keep what
        // the row loop needs in registers
        uint32_t wstride[NP];
#pragma unroll
        for (int p = 0; p < NP; p++) wstride[p] = TYPE != T_KQ || p < a.W.n_planes ? (uint32_t)a.W.stride[p] : 0u;
        // Row pairs of this CTA are handed out as "pair slots" j = 0, 1, ...: slot j belongs to tile j / PT (one ring
        // slot, PT pairs, a few KB) and to row worker j % NW.  Small tiles keep the ring densely packed and let the
        // producer refill as soon as PT owners are done instead of waiting for all fifteen warps.
        const int PT = S.PT;
        const int npairs_cta = (r_end - r_begin) >> 1;
        const int ntiles = ((npairs_cta + PT - 1) * S.inv_PT) >> 16;             // ceil(npairs_cta / PT)
        const uint32_t arrive_units = S.arrive_units;                            // MEGA_EMPTY_BASE / (NW * Sw)
        long long c_wait = 0, c_work = 0;      // (trace only) cycles warp 0 spends waiting for tiles / computing
        if (tr) tr[14] = (unsigned long long)clock64();           // (stage set-up done, row loop starts)
        if (working) {
            // Every worker walks every tile of the stage (a parity wait must not skip a lap of its barrier), and computes in
            // the tiles where it owns a pair slot: slot j = t * PT + i belongs to worker j % NW, and PT <= NW, so at most one.
            int jn = worker;                                   // this worker's next pair slot
            uint32_t slot = ring_slot, phase = ring_phase;
            for (int t = 0; t < ntiles; t++) {
                const int i = jn - t * PT;                     // pair inside tile t, if 0 <= i < PT
                const bool own = i < PT;
                const int j = jn;
                const long long c0 = tr ? clock64() : 0;
                if (!(sm.debug & 2)) mbar_wait(&full[slot], phase);
                const long long c1 = tr ? clock64() : 0;
                if (own && j < npairs_cta && !(sm.debug & 1)) {
                    // two rows at a time: twice the independent dp4a chains, and one transposed f64 butterfly
                    // (lanes 0-15 end up with row r, lanes 16-31 with row r+1) instead of two
                    const uint8_t* pb = ring + slot * sm.stage_bytes;
                    const uint32_t r = 2u * (uint32_t)i;
                    RowPtrs rp0, rp1;
#pragma unroll
                    for (int p = 0; p < NP; p++) {
                        rp0.p[p] = pb + r * wstride[p];
                        rp1.p[p] = rp0.p[p] + wstride[p];
                        pb += (uint32_t)RT * wstride[p];
                    }
                    double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
                    for (int jj = 0; jj < NB; jj++) {
                        acc0 += (double)unit_dot(xr[jj], rp0, unit[jj], nunits);
                        acc1 += (double)unit_dot(xr[jj], rp1, unit[jj], nunits);
                    }
                    const bool upper = (lane & 16) != 0;
                    double keep = upper ? acc1 : acc0;
                    const double give = upper ? acc0 : acc1;
                    keep += __shfl_xor_sync(0xffffffffu, give, 16);
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
                    if ((lane & 15) == 0) ysum[(2 * j + (upper ? 1 : 0)) * Sw + ch] = keep;
                }
                __syncwarp();
                if (lane == 0 && !(sm.debug & 2)) mbar_arrive_units(&empty[slot], arrive_units);   // this warp is past the slot (owner or not)
                if (own) jn += NW;
                if (tr) { const long long c2 = clock64(); c_wait += c1 - c0; c_work += c2 - c1; }
                if (++slot == (uint32_t)sm.ns) { slot = 0; phase ^= 1u; }
            }
        }
        if (tr) { tr[12] = (unsigned long long)c_wait; tr[13] = (unsigned long long)c_work; }
        ring_slot += (uint32_t)ntiles;
        while (ring_slot >= (uint32_t)sm.ns) { ring_slot -= sm.ns; ring_phase ^= 1u; }
        if (tr) tr[11] = (unsigned long long)clock64();           // (this warp's row loop done)
        // ---- deferred epilogue: all row sums of this CTA are in shared memory; one thread per row pair,
        //      so the long-latency parts (f64 exp of SiLU, RoPE table, KV page lookup) are paid once ----
        named_bar_sync(1, ntc);
        if (tr) tr[15] = (unsigned long long)clock64();           // (every warp's row loop done)
        {
            unsigned long long best = 0ull;
            const int npairs = (int)((r_end - r_begin) >> 1);
            const bool ll_out = !TP && a.y_ll != nullptr;
            for (int p = tid; p < npairs; p += ntc) {
                double d0 = 0.0, d1 = 0.0;
                for (int k = 0; k < Sw; k++) { d0 += ysum[(size_t)(2 * p) * Sw + k]; d1 += ysum[(size_t)(2 * p + 1) * Sw + k]; }
                if (TP && a.epi == EPI_PARTIAL) {          // tensor-parallel K slice: unrounded, the exchange stage adds the shards
                    a.y64[a.epi_row0 + r_begin + 2 * p] = d0;
                    a.y64[a.epi_row0 + r_begin + 2 * p + 1] = d1;
                } else {
                    EpiCache ec;
                    if (!TP && p < EPI_CACHE_PAIRS) {
                        if (S.epi_cache == 1) { ec.res = s_res + 2 * p; ec.res_valid = res_valid; }
                        else if (S.epi_cache == 2) { ec.rope = s_rope + p; ec.rope_valid = rope_valid; }
                    }
                    epilogue_pair(a, a.epi_row0 + r_begin + 2 * p, (float)d0, (float)d1, best, pos_now, s_pt, ll_out, ec);
                }
            }
            if (S.epi_cache == 1) res_valid = true;
            else if (S.epi_cache == 2) rope_valid = true;
        }
        if (tr) tr[3] = global_ns();
    }
    // every CTA read the exchange counter before its first grid barrier, CTA 0 is past the last one
    if (TP && blockIdx.x == 0 && tid == 0) *reinterpret_cast<volatile unsigned*>(tp.epoch) = tp_epoch;
}

int g_mega_sms = 0;

}  // namespace

// ---- host side ---------------------------------------------------------------------------------------
// Estimated consumer cost of a stage per row pair: (units per lane x 2 rows x ~27 instructions + the f64
// butterfly, ~35) spread over the G row workers -- the model behind the choice of units per lane.
static int mega_tile_target() {
    static const int v = getenv("B2_MEGA_TILE_KB") ? atoi(getenv("B2_MEGA_TILE_KB")) * 1024 : 60 * 1024;
    return v;
}
static double plan_cost(int nunits, int nbe, int ncw, int* S_out, int* G_out) {
    const int S = (nunits + 32 * nbe - 1) / (32 * nbe);
    if (S > ncw) return 1e30;
    const int G = ncw / S;
    *S_out = S; *G_out = G;
    return (54.0 * nbe + 35.0) / G;
}
double mega_plan_cost(const QMat& W, int nb_max) {
    const bool kq = W.type == T_Q4_K || W.type == T_Q6_K;
    const int nunits = (int)(kq ? W.k / 128 : W.k / 32);
    double best = 1e30;
    int S, G;
    for (int nbe = kq ? 1 : nb_max; nbe >= 1; nbe--) best = std::min(best, plan_cost(nunits, nbe, MEGA_NCW, &S, &G));
    return best * (double)(W.rows / 2);
}
int mega_plan_stage(MegaStage& st, int ncw, int nb_max_q, int tile_target) {
    const QMat& W = st.g.W;
    const bool kq = W.type == T_Q4_K || W.type == T_Q6_K;
    const int nb_max = kq ? 1 : nb_max_q;
    st.nunits = (int)(kq ? W.k / 128 : W.k / 32);
    // S warps cover a row (nbeff units per lane); G = ncw / S "row workers" take the rows of a tile
    // round-robin.  Pick the units-per-lane with the lowest estimated cost per row pair.
    double best = 1e30;
    int best_idle = 1 << 30;
    for (int nbe = nb_max; nbe >= 1; nbe--) {
        int S = 0, G = 0;
        const double c = plan_cost(st.nunits, nbe, ncw, &S, &G);
        if (c < best) { best = c; best_idle = 0; st.nbeff = nbe; st.S = S; st.G = G; }
    }
    if (best_idle == (1 << 30)) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: k=%lld does not map onto %d consumer warps", (long long)W.k, ncw);
    // tile = PT row pairs, up to 60 KB (one bulk copy per plane: a single producer thread issues about three
    // tiles per microsecond, so small tiles would bound the stream; 60 KB lets three tiles fit next to the attention
    // buffers of a short context -- 7B, ctx 512: 680 tok/s with 72 KB / two slots, 712 with 60 KB / three, 699 with
    // 46 KB / four); PT * S owner warps (<= 15) release a tile
    size_t rb = 0;
    for (int p = 0; p < W.n_planes; p++) rb += W.stride[p];
    int pt = (int)((size_t)(tile_target > 0 ? tile_target : mega_tile_target()) / (2 * rb));
    pt = std::max(1, std::min(pt, std::max(1, ncw / st.S)));
    st.PT = pt;
    st.RT = 2 * pt;
    st.P = 1; st.RPG = 0;
    return B2_OK;
}

// Per-stage constants that depend on the grid (known after mega_configure): the kernel reads them instead of dividing.
void mega_finish_stage(MegaStage& st, int grid) {
    if (st.kind != 0) return;
    const int pairs = (int)(st.g.W.rows >> 1);
    st.pairs_base = pairs / grid;
    st.pairs_rem = pairs % grid;
    st.inv_S = (65536 + st.S - 1) / st.S;
    st.inv_PT = (65536 + st.PT - 1) / st.PT;
    st.arrive_units = MEGA_EMPTY_BASE / (uint32_t)(st.G * st.S);
}

template <int TYPE, int NB, int NCW, int GRP, int HD, bool TP>
static int mega_launch_t(const MegaLaunch& L, cudaStream_t stream) {
    auto kern = mega_kernel<TYPE, NB, NCW, GRP, HD, TP, false>;
    if (!TP && L.trace) kern = mega_kernel<TYPE, NB, NCW, GRP, HD, false, !TP>;      // (no traced variant under tensor parallelism)
    static PerDeviceSmem configured[2];
    if (configured[L.trace ? 1 : 0].raise(L.smem_bytes)) {
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 0));
    }
    MegaSmem sm;
    sm.off_red = L.off_red; sm.off_xq = L.off_xq; sm.off_xd = L.off_xd; sm.off_xs = L.off_xs; sm.off_att = L.off_att; sm.off_ring = L.off_ring;
    sm.stage_bytes = L.stage_bytes; sm.ns = L.ns; sm.pf_bytes = L.pf_bytes; sm.debug = L.debug;
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
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, L.stages, L.s_begin, L.s_end, L.grid_count, sm, L.trace, L.pos, L.page_table, L.n_pages, L.tp));
    return B2_OK;
}

template <int TYPE, int NB, int NCW, int GRP, int HD, bool TP>
static int mega_occupancy_t(const MegaLaunch& L, int* max_clusters) {
    auto kern = mega_kernel<TYPE, NB, NCW, GRP, HD, TP>;
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

template <int TYPE, int NB, int NCW, int GRP, int HD, bool TP>
static int mega_static_smem_t(const MegaLaunch&, size_t* out) {
    cudaFuncAttributes fa;
    B2_CUDA(cudaFuncGetAttributes(&fa, mega_kernel<TYPE, NB, NCW, GRP, HD, TP>));
    *out = fa.sharedSizeBytes;
    return B2_OK;
}

#define B2_MEGA_CASE(FN, G, H, ...)                                                                        \
    if (L.type == T_Q4_0 && L.grp == G && L.head_dim == H && L.nb == 4)                                    \
        return L.tp.n > 1 ? FN<T_Q4_0, 4, MEGA_NCW, G, H, true>(__VA_ARGS__) : FN<T_Q4_0, 4, MEGA_NCW, G, H, false>(__VA_ARGS__); \
    if (L.type == T_Q4_0 && L.grp == G && L.head_dim == H && L.nb == 5)                                    \
        return L.tp.n > 1 ? FN<T_Q4_0, 5, MEGA_NCW, G, H, true>(__VA_ARGS__) : FN<T_Q4_0, 5, MEGA_NCW, G, H, false>(__VA_ARGS__)
#define B2_MEGA_CASE_1GPU(FN, G, H, ...)                                                                   \
    if (L.type == T_Q4_0 && L.grp == G && L.head_dim == H && L.tp.n <= 1) {                                \
        if (L.nb == 4) return FN<T_Q4_0, 4, MEGA_NCW, G, H, false>(__VA_ARGS__);                           \
        if (L.nb == 5) return FN<T_Q4_0, 5, MEGA_NCW, G, H, false>(__VA_ARGS__);                           \
    }
#define B2_MEGA_CASE_KQ(FN, G, H, ...)                                                                     \
    if (L.type == T_KQ && L.grp == G && L.head_dim == H && L.tp.n <= 1) return FN<T_KQ, 1, MEGA_NCW, G, H, false>(__VA_ARGS__)
#define B2_MEGA_DISPATCH(FN, ...)                                                                          \
    do {                                                                                                   \
        B2_MEGA_CASE_KQ(FN, 4, 64, __VA_ARGS__);                                                           \
        B2_MEGA_CASE_KQ(FN, 4, 128, __VA_ARGS__);                                                          \
        B2_MEGA_CASE_1GPU(FN, 1, 80, __VA_ARGS__);                                                         \
        B2_MEGA_CASE(FN, 1, 64, __VA_ARGS__);                                                              \
        B2_MEGA_CASE(FN, 1, 128, __VA_ARGS__);                                                             \
        B2_MEGA_CASE(FN, 8, 128, __VA_ARGS__);                                                             \
    } while (0)

bool mega_supported(int type, int grp, int head_dim) {
    if (type == T_KQ) return grp == 4 && (head_dim == 64 || head_dim == 128);
    return type == T_Q4_0 && ((grp == 1 && (head_dim == 64 || head_dim == 80 || head_dim == 128)) || (grp == 8 && head_dim == 128));
}

// Fill in the shared-memory layout and the grid for a model: k_max = longest activation vector,
// row_bytes_max = bytes of the largest stage tile row set (RT rows).
int mega_configure(MegaLaunch& L, int64_t k_max, size_t tile_bytes_max, int rt_s_max, int max_ctx) {
    if (!mega_supported(L.type, L.grp, L.head_dim)) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: no kernel for type %d, %d query heads per KV head, head_dim %d", L.type, L.grp, L.head_dim);
    const bool kq = L.type == T_Q4_K || L.type == T_Q6_K || L.type == T_KQ;
    size_t off = (2 * MEGA_NS_MAX + 2) * 8 + 32 * 8;
    L.off_red = (uint32_t)off; off += (size_t)2 * rt_s_max * 8; off = (off + 15) / 16 * 16;
    L.off_xq = (uint32_t)off; off += (size_t)k_max; off = (off + 15) / 16 * 16;
    L.off_xd = (uint32_t)off; off += (size_t)(kq ? k_max / 256 : k_max / 32) * 4; off = (off + 15) / 16 * 16;
    L.off_xs = (uint32_t)off; off += (size_t)(kq ? k_max / 16 : k_max / 32) * 4; off = (off + 15) / 16 * 16;
    L.off_att = (uint32_t)off;
    L.att_max_local = (max_ctx + 3) / 4 * 4;
    if (L.cluster != 1 && L.cluster != 2 && L.cluster != 4) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: cluster size %d (1, 2 or 4)", L.cluster);
    if (L.head_dim % (8 * L.cluster)) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: head_dim %d does not split over a cluster of %d", L.head_dim, L.cluster);
    off += attndev::attn_smem_bytes(L.grp, L.head_dim, L.att_max_local, L.cluster, MEGA_NCW * 32);
    off = (off + 127) / 128 * 128;
    L.off_ring = (uint32_t)off;
    L.stage_bytes = (uint32_t)((tile_bytes_max + 127) / 128 * 128);
    size_t static_smem = 0;                 // page table copy, stage descriptors
    {
        auto q = [&](size_t* o) -> int { B2_MEGA_DISPATCH(mega_static_smem_t, L, o); B2_FAIL(B2_ERR_UNSUPPORTED, "mega: dispatch"); };
        B2_TRY(q(&static_smem));
    }
    const size_t cap = (227 * 1024 - static_smem) / 128 * 128;          // 227 KB per CTA minus the kernel's static shared memory
    L.ring_room = cap > off ? (uint32_t)(cap - off) : 0u;
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
    if (L.grid < 64) B2_FAIL(B2_ERR_UNSUPPORTED, "mega: only %d co-resident CTAs; the row-sum buffer is sized for >= 64", L.grid);
    return B2_OK;
}

int mega_launch(const MegaLaunch& L, cudaStream_t stream) {
    B2_MEGA_DISPATCH(mega_launch_t, L, stream);
    B2_FAIL(B2_ERR_UNSUPPORTED, "mega: no kernel for type %d grp %d head_dim %d", L.type, L.grp, L.head_dim);
}

}  // namespace b2
