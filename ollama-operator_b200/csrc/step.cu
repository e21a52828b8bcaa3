// step.cu -- the two small bookkeeping kernels that bracket a decode step.  They keep the whole
// greedy loop on the device (token feedback, position, RoPE table), so a CUDA graph of one step
// can be replayed back to back with no host round trip.
#include <math.h>

#include "common.cuh"
#include "kernels.h"
#include "qmat_dev.cuh"

namespace b2 {

// step_begin: x = embedding row of the current token (ggml get_rows), position scalar for the
// layer kernels, arg-max accumulator cleared.
__global__ void step_begin_kernel(const StepBeginArgs a) {
    pdl_trigger();
    pdl_wait();
    // b2_decode of one sequence hands the state over in pinned host memory: every block reads the two words it needs from
    // there, block 0 also installs the state on the device for step_end
    const SeqState* src = a.st_host ? a.st_host : a.st;
    const int tok = src->cur_tok, pos = src->pos;
    if (a.st_host && blockIdx.x == 0 && threadIdx.x == 0) *a.st = *a.st_host;
    const int E = (int)a.embd.k;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) a.x[i] = qmat_elem(a.embd, tok, i);
    // hand-over buffers of the persistent kernel (GemvArgs::x_ll): the ones written first in a step start out poisoned; done
    // here, a whole kernel boundary ahead of the first value stored into them
#pragma unroll
    for (int b = 0; b < 6; b++)
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.ll_init.n[b]; i += gridDim.x * blockDim.x) a.ll_init.p[b][i] = __uint_as_float(LL_POISON);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *a.pos_out = pos;
        *a.argmax_key = 0ull;
        if (a.grid_count) *a.grid_count = 0u;
    }
}

int step_begin_launch(const StepBeginArgs& a, cudaStream_t st, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    const int E = (int)a.embd.k;
    cfg.gridDim = dim3((E + 255) / 256);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, step_begin_kernel, a));
    return B2_OK;
}

// A kernel that only waits: b2_profile_step queues it ahead of the step it times, so that the host has every launch of the
// step enqueued before the GPU reaches the first one and the events between the kernels measure kernels, not launch gaps.
__global__ void delay_kernel(unsigned long long ns) {
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns);
}
int delay_launch(cudaStream_t st, unsigned long long ns) {
    delay_kernel<<<1, 1, 0, st>>>(ns);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

// step_end: pick the next token (prompt token while any are left, else the arg-max of the
// logits), record it, advance the position.
__global__ void step_end_kernel(const StepEndArgs a) {
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    SeqState* s = a.st;
    int next;
    if (s->forced_idx < s->n_forced) {
        next = a.forced[s->forced_idx++];
    } else {
        const unsigned long long key = a.use_argmax ? *a.argmax_key : 0ull;
        next = (int)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
        if (a.out_toks && s->n_out < a.out_cap) a.out_toks[s->n_out] = next;
        s->n_out++;
    }
    s->cur_tok = next;
    s->pos++;
    if (a.st_host) { *a.st_host = *s; __threadfence_system(); }
}

int step_end_launch(const StepEndArgs& a, cudaStream_t st, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(32);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, step_end_kernel, a));
    return B2_OK;
}

}  // namespace b2
