// Persistent decode kernel: all transformer layers of one token in ONE launch.
//
// Why: at batch 1 a layer is five tiny kernels (9-66 MB of weights each, 1.5-10 us of HBM time) and
// the measured fixed cost per kernel -- launch, parameter/instruction fetch, barrier set-up, first
// TMA round trip, tail -- was ~6 us, more than the streaming itself (profiles/).
// Here one CTA per SM stays resident for the whole token:
//   * producer warps walk the weight matrices of all 4 x n_layer mat-vec phases in order and keep the
//     shared-memory ring full. They never look at activations, so while the consumers sit at a grid
//     barrier or in attention the next phase's weights are already landing (attn_out and the QKV
//     projection fit in the ring entirely).
//   * consumer warps run the phases: qkv -> [grid barrier] -> attention -> [gb] -> attn_out -> [gb] ->
//     gate_up -> [gb] -> ffn_down -> [gb]. A phase is exactly the mat-vec of matvec_phase.cuh (input
//     quantisation + dp4a tiles + fused epilogue) or the flash-decode body of attention_body.cuh run
//     by 8 of the consumer warps on CTA (kv_head, split) = blockIdx.
//   * the grid barrier is a monotonic counter in global memory (one atomic per CTA per phase, thread 0
//     polls it with acquire loads); data written by other SMs is read with ld.cg / TMA (L2).
// Replaces the per-op launch sequence of ggml-cuda's graph for the layer stack; the
// embedding gather, lm_head and sampler stay separate launches around it.
#include "attention_body.cuh"
#include "matvec_phase.cuh"
#include <stdlib.h>

namespace b2 {

constexpr int MG_HDR = 8192;        // barriers, parameter slots
constexpr int MG_BAR_CONS = 1;      // named barrier of all consumer warps (also used inside the mat-vec prologue)

template <int NCW>
struct __align__(16) MegaHeader {
    MvShared<NCW> mv;
    uint64_t attn_full[2];
    MvParams pc[2];                 // current / next phase, consumers (double buffered)
    MvParams pp[MV_MAX_PROD];       // current phase, each producer warp
    AttnParams pa;
};

__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// all consumer threads of all CTAs; `epoch` counts barriers passed by this CTA
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <int NCW>
__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int& epoch, unsigned int n_ctas,
                                             unsigned int* err, unsigned long long* ts = nullptr) {
    named_bar_sync(MG_BAR_CONS, NCW * 32);
    epoch += n_ctas;
    if (threadIdx.x == 0) {
        // release: everything this CTA wrote (ordered before us by the CTA barrier) is visible to whoever
        // acquires the counter afterwards; no return value needed, so a reduction, not an atomic
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
        unsigned int spins = 0;
        while (ld_acquire(counter) < epoch) {
            if (++spins > (1u << 24)) {
                volatile unsigned int* e = reinterpret_cast<volatile unsigned int*>(err);
                if (e && e[0] == 0) { e[1] = blockIdx.x; e[2] = 0; e[3] = epoch; e[4] = 0; e[0] = 0x30001u; }
                break;
            }
        }
        if (ts && blockIdx.x == 0) ts[epoch / n_ctas] = globaltimer_ns();     // B2_MEGA_TS: phase end times
    }
    named_bar_sync(MG_BAR_CONS, NCW * 32);
}

// copy a parameter block global -> shared with one warp
template <typename T>
__device__ __forceinline__ void warp_copy_params(T* dst, const T* src, int lane) {
    static_assert(sizeof(T) % 4 == 0, "parameter blocks are word multiples");
    const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
    uint32_t* d = reinterpret_cast<uint32_t*>(dst);
    for (int i = lane; i < (int)(sizeof(T) / 4); i += 32) d[i] = __ldg(s + i);
    __syncwarp();
}

template <int NCW, bool ACT2>
__global__ void __launch_bounds__((NCW + MV_MAX_PROD) * 32, 1) decode_layers_kernel(const MegaParams mp) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MegaHeader<NCW>& hd = *reinterpret_cast<MegaHeader<NCW>*>(smem_raw);
    static_assert(sizeof(MegaHeader<NCW>) <= MG_HDR, "header region too small");
    uint8_t* region = smem_raw + MG_HDR;                    // activation planes / attention tiles (never both)
    uint8_t* ring = smem_raw + mp.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x, n_ctas = gridDim.x;
    RingGeom geo; geo.slot_bytes = mp.slot_bytes; geo.stage_bytes = mp.stage_bytes; geo.n_stages = mp.n_stages;

    if (threadIdx.x == 0) {
        mv_ring_init<NCW>(hd.mv, geo.n_stages);
        mbar_init(&hd.attn_full[0], 1); mbar_init(&hd.attn_full[1], 1);
        fence_mbar_init();
    }
    __syncthreads();

    const int n_phase = mp.n_layer * 4;
    if (warp >= NCW) {
        // ------------------------------------------------------------ producers: every mat-vec phase, in order
        const int pi = warp - NCW;
        if (pi >= mp.n_prod) return;
        uint32_t qbase = 0;
        for (int ph = 0; ph < n_phase; ph++) {
            warp_copy_params(&hd.pp[pi], mp.mv + ph, lane);
            const MvParams& p = hd.pp[pi];
            const CtaWork work = cta_work(p, cta, n_ctas);
            mv_produce<NCW>(p, work, hd.mv, ring, geo, qbase, pi, mp.n_prod, lane);
            qbase += (uint32_t)work.total;
            __syncwarp();
        }
        return;
    }

    // ---------------------------------------------------------------- consumers: phase by phase
    unsigned int epoch = 0;
    if (mp.ts && cta == 0 && threadIdx.x == 0) mp.ts[0] = globaltimer_ns();
    uint32_t qbase = 0;
    uint32_t attn_use[2] = {0u, 0u};
    const int G_items = mp.attn_items;                          // n_kv * n_splits work items of the attention phase
    if (warp == 0) warp_copy_params(&hd.pc[0], mp.mv, lane);
    for (int l = 0; l < mp.n_layer; l++) {
        for (int k = 0; k < 4; k++) {
            const int ph = l * 4 + k;
            if (k == 1) {
                // ---- attention between the QKV projection and attn_out
                if (cta < G_items && warp < AT_THREADS / 32) {
                    const AttnParams& a = hd.pa;
                    const int sp = cta % a.n_splits, h = cta / a.n_splits;
                    if (a.head_dim == 64) attn_decode_body<64, false>(a, *reinterpret_cast<AttnShared<64>*>(region), hd.attn_full, attn_use, sp, h, threadIdx.x);
                    else attn_decode_body<128, false>(a, *reinterpret_cast<AttnShared<128>*>(region), hd.attn_full, attn_use, sp, h, threadIdx.x);
                }
                grid_barrier<NCW>(mp.bar, epoch, (unsigned)n_ctas, mp.err, mp.ts);
            }
            unsigned long long* t2 = (mp.ts && cta == 0 && threadIdx.x == 0) ? mp.ts + (5 * mp.n_layer + 2) + ph * 4 : nullptr;
            if (t2) t2[0] = globaltimer_ns();
            named_bar_sync(MG_BAR_CONS, NCW * 32);          // this phase's parameters (copied during the previous phase) are in place
            // the last consumer warp fetches what comes next while the others start on this phase
            if (warp == NCW - 1) {
                if (ph + 1 < n_phase) warp_copy_params(&hd.pc[(ph + 1) & 1], mp.mv + ph + 1, lane);
                if (k == 0) warp_copy_params(&hd.pa, mp.at + l, lane);
            }
            const MvParams& p = hd.pc[ph & 1];
            const CtaWork work = cta_work(p, cta, n_ctas);
            if (t2) t2[1] = globaltimer_ns();
            mv_consume_phase<NCW, ACT2>(p, work, hd.mv, region, ring, geo, qbase, warp, lane, t2 ? t2 + 2 : nullptr);
            qbase += (uint32_t)work.total;
            if (t2) t2[3] = globaltimer_ns();
            grid_barrier<NCW>(mp.bar, epoch, (unsigned)n_ctas, mp.err, mp.ts);
        }
    }
}

// ---------------------------------------------------------------- host side
static int mega_ncw() {
    int v = 16;
    if (const char* s = getenv("B2_MV_WARPS")) { int x = atoi(s); if (x == 16 || x == 20) v = x; }
    return v;
}

cudaError_t mega_init() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(decode_layers_kernel<16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MEGA_SMEM_MAX)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(decode_layers_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MEGA_SMEM_MAX)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(decode_layers_kernel<20, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MEGA_SMEM_MAX)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(decode_layers_kernel<20, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MEGA_SMEM_MAX);
}

// Fix the ring geometry for a whole layer stack: slot/stage sizes are the largest any phase needs, the
// region in front of the ring holds the larger of (activation planes of the widest input, attention tiles).
int mega_plan(MegaParams& mp, const MvParams* host_mv, int n_phase, int head_dim, int n_kv, int n_splits, int n_sms) {
    int slot = 0, act = 0;
    for (int i = 0; i < n_phase; i++) {
        const MvParams& p = host_mv[i];
        if (p.whole_rows) return -1;
        slot = p.slot_bytes > slot ? p.slot_bytes : slot;
        int a = (p.act_dual ? 4 : 2) * p.n_slices * B2_SLICE + (p.K / 32) * 4 + (p.K / 16) * 4;
        act = a > act ? a : act;
    }
    int attn = head_dim == 64 ? (int)sizeof(AttnShared<64>) : (int)sizeof(AttnShared<128>);
    int region = act > attn ? act : attn;
    mp.slot_bytes = slot;
    mp.stage_bytes = (MV_ROWS * slot + 127) / 128 * 128;
    mp.ring_off = (MG_HDR + region + 127) / 128 * 128;
    mp.n_prod = 2;
    if (const char* s = getenv("B2_MV_PROD")) { int v = atoi(s); if (v >= 1 && v <= MV_MAX_PROD) mp.n_prod = v; }
    int S = (MEGA_SMEM_MAX - mp.ring_off) / mp.stage_bytes;
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (S < 8 * mp.n_prod) mp.n_prod = 1;
    if (S < 8) return -2;
    S -= S % mp.n_prod;
    mp.n_stages = S;
    mp.attn_items = n_kv * n_splits;
    if (mp.attn_items > n_sms) return -3;
    return 0;
}

size_t mega_smem_bytes(const MegaParams& mp) { return (size_t)mp.ring_off + (size_t)mp.n_stages * mp.stage_bytes; }

cudaError_t mega_launch(const MegaParams& mp, int n_ctas, cudaStream_t stream, bool act2) {
    cudaError_t e = cudaMemsetAsync(mp.bar, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const int ncw = mega_ncw();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3((ncw + mp.n_prod) * 32);
    cfg.dynamicSmemBytes = mega_smem_bytes(mp); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;      // all CTAs co-resident: they wait on each other at the grid barrier
    at[0].val.cooperative = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (ncw == 20) return act2 ? cudaLaunchKernelEx(&cfg, decode_layers_kernel<20, true>, mp)
                               : cudaLaunchKernelEx(&cfg, decode_layers_kernel<20, false>, mp);
    return act2 ? cudaLaunchKernelEx(&cfg, decode_layers_kernel<16, true>, mp)
                : cudaLaunchKernelEx(&cfg, decode_layers_kernel<16, false>, mp);
}

}  // namespace b2
