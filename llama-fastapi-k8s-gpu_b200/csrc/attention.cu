// Split-K flash-decode attention over the paged f16 KV cache (one new token per sequence).
// Replaces ggml-cuda's mul_mat_vec_p021/nc (K.q), soft_max_f32 and the P.V mat-mul that the
// reference launches separately per layer, on a paged cache instead of
// llama.cpp's contiguous one (H7).
//
// Grid = (n_splits, n_kv_heads), 256 threads. A CTA owns one KV head and one slice of the context
// and serves all G = n_head / n_kv query heads of that KV head from a single read of K and V
// (GQA-aware). KV pages are [n_kv][block_tokens][head_dim] f16, so the tokens of one page for one
// KV head are one contiguous run: thread 0 brings 64-token tiles into shared memory with 1-D TMA
// bulk copies (double buffered, mbarrier complete_tx). Inside a tile every warp is a flash-decode of its
// own over 8 of the 64 positions, a lane owning head_dim / 32 dims (attention_body.cuh): scores of 4 heads x 8
// positions reduced over the lanes by one transposed butterfly, online softmax inside 8-lane groups, o += p . v
// with the probabilities passed by shuffle; running max / sum / output stay in registers across tiles and the
// eight warps are merged once at the end. No block-wide barrier between those steps.
// Merging the splits:
//   * n_splits = 2, 4 or 8 (default 8): the splits of one KV head form a thread-block CLUSTER; partials stay
//     in shared memory and are merged through distributed shared memory after one cluster barrier.
//     No global scratch, no atomics, no fence.
//   * other split counts: partials go to global scratch; the last CTA of a KV head to finish (atomic
//     ticket) merges them.
//
// Roofline: HBM; algorithmic bytes per launch = 2 * n_kv * head_dim * 2 B * (pos + 1). At chat
// context lengths the kernel is latency bound (three dependent memory round trips).
#include "attention_body.cuh"
#include <mutex>

namespace b2 {

size_t attn_scratch_floats(int n_head, int n_kv, int head_dim, int n_splits) {
    (void)n_kv;
    return (size_t)n_head * n_splits * (head_dim + 2);
}

template <int HD, bool CLUSTER, int NP>
__global__ void __launch_bounds__(at_threads(NP)) attn_decode_kernel(const AttnParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    AttnShared<HD, at_tt(NP)>& sh = *reinterpret_cast<AttnShared<HD, at_tt(NP)>*>(smem_raw);
    __shared__ uint64_t full[AT_NST];
    if (threadIdx.x == 0) { for (int i = 0; i < AT_NST; i++) mbar_init(&full[i], 1); fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) tl_stamp(p.tl, 0);
    griddep_launch();
    if (!p.early) {
        griddep_wait();     // q and this token's K/V come from the QKV kernel just before us
        if (threadIdx.x == 0 && (B2_ATTN_ONE_CTA < 0 || (int)(blockIdx.y * gridDim.x + blockIdx.x) == B2_ATTN_ONE_CTA)) tl_stamp(p.tl, 1);
    }
    uint32_t use[AT_NST] = {};
    attn_decode_body<HD, CLUSTER, NP>(p, sh, full, use, blockIdx.x, blockIdx.y, threadIdx.x, p.early != 0);
    if (p.tl) {
        __syncthreads();
        if (threadIdx.x == 0 && (B2_ATTN_ONE_CTA < 0 || (int)(blockIdx.y * gridDim.x + blockIdx.x) == B2_ATTN_ONE_CTA)) tl_stamp(p.tl, 3);
    }
}

// one launch for several sequences: blockIdx.z picks the sequence
template <int HD, bool CLUSTER, int NP>
__global__ void __launch_bounds__(at_threads(NP)) attn_decode_batch_kernel(const AttnParams p0, const SlotDev* tab, const BatchIdx b) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    AttnShared<HD, at_tt(NP)>& sh = *reinterpret_cast<AttnShared<HD, at_tt(NP)>*>(smem_raw);
    __shared__ uint64_t full[AT_NST];
    if (threadIdx.x == 0) { for (int i = 0; i < AT_NST; i++) mbar_init(&full[i], 1); fence_mbar_init(); }
    __syncthreads();
    griddep_launch();
    griddep_wait();
    const int z = blockIdx.z;
    const SlotDev sd = tab[b.slot[z]];
    AttnParams p = p0;
    const size_t qd = (size_t)p.n_head * p.head_dim;
    p.q = p0.q + z * qd; p.out = p0.out + z * qd;
    p.pos_ptr = sd.ints + 1; p.block_table = sd.block_table;
    p.scratch = p0.scratch + z * qd / p.head_dim * p.n_splits * (p.head_dim + 2);
    p.counters = p0.counters + z * p.n_kv;
    uint32_t use[AT_NST] = {};
    attn_decode_body<HD, CLUSTER, NP>(p, sh, full, use, blockIdx.x, blockIdx.y, threadIdx.x);
}

// kernel variants: head_dim x (cluster merge | scratch merge) x (one | two passes of 4 query heads)
template <int HD>
static cudaError_t set_smem_attrs() {
    cudaError_t e;
#define B2_ATTR(K, NP) if ((e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AttnShared<HD, at_tt(NP)>))) != cudaSuccess) return e
    B2_ATTR((attn_decode_kernel<HD, true, 1>), 1); B2_ATTR((attn_decode_kernel<HD, true, 2>), 2);
    B2_ATTR((attn_decode_kernel<HD, false, 1>), 1); B2_ATTR((attn_decode_kernel<HD, false, 2>), 2);
    B2_ATTR((attn_decode_batch_kernel<HD, true, 1>), 1); B2_ATTR((attn_decode_batch_kernel<HD, true, 2>), 2);
    B2_ATTR((attn_decode_batch_kernel<HD, false, 1>), 1); B2_ATTR((attn_decode_batch_kernel<HD, false, 2>), 2);
#undef B2_ATTR
    return cudaSuccess;
}
template <int HD>
static cudaError_t launch_one(cudaLaunchConfig_t& cfg, bool cluster, int G, const AttnParams& p) {
    const int np = G <= 4 ? 1 : 2;
    cfg.blockDim = dim3(at_threads(np));
    cfg.dynamicSmemBytes = np == 1 ? sizeof(AttnShared<HD, at_tt(1)>) : sizeof(AttnShared<HD, at_tt(2)>);
    if (G <= 4) return cluster ? cudaLaunchKernelEx(&cfg, attn_decode_kernel<HD, true, 1>, p) : cudaLaunchKernelEx(&cfg, attn_decode_kernel<HD, false, 1>, p);
    return cluster ? cudaLaunchKernelEx(&cfg, attn_decode_kernel<HD, true, 2>, p) : cudaLaunchKernelEx(&cfg, attn_decode_kernel<HD, false, 2>, p);
}
template <int HD>
static cudaError_t launch_batch(cudaLaunchConfig_t& cfg, bool cluster, int G, const AttnParams& p, const SlotDev* tab, const BatchIdx& b) {
    const int np = G <= 4 ? 1 : 2;
    cfg.blockDim = dim3(at_threads(np));
    cfg.dynamicSmemBytes = np == 1 ? sizeof(AttnShared<HD, at_tt(1)>) : sizeof(AttnShared<HD, at_tt(2)>);
    if (G <= 4) return cluster ? cudaLaunchKernelEx(&cfg, attn_decode_batch_kernel<HD, true, 1>, p, tab, b) : cudaLaunchKernelEx(&cfg, attn_decode_batch_kernel<HD, false, 1>, p, tab, b);
    return cluster ? cudaLaunchKernelEx(&cfg, attn_decode_batch_kernel<HD, true, 2>, p, tab, b) : cudaLaunchKernelEx(&cfg, attn_decode_batch_kernel<HD, false, 2>, p, tab, b);
}

cudaError_t attn_init() {
    cudaError_t e;
    if ((e = set_smem_attrs<64>()) != cudaSuccess) return e;
    return set_smem_attrs<128>();
}

cudaError_t attn_decode_batch_launch(const AttnParams& p, const SlotDev* tab, const BatchIdx& b, cudaStream_t stream) {
    const int G = p.n_kv > 0 ? p.n_head / p.n_kv : 0;
    if (p.n_kv <= 0 || p.n_head % p.n_kv || G > AT_GMAX || b.n < 1 || b.n > BATCH_MAX) return cudaErrorInvalidValue;
    if (p.n_splits < 1 || G * p.n_splits > AT_GMAX * AT_STAGE) return cudaErrorInvalidValue;
    const bool cluster = p.n_splits == 2 || p.n_splits == 4 || p.n_splits == 8;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.n_kv, b.n); cfg.stream = stream;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (g_chain_pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    if (cluster) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = p.n_splits; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
        na++;
    }
    cfg.attrs = at; cfg.numAttrs = na;
    if (p.head_dim == 64) {
        return launch_batch<64>(cfg, cluster, G, p, tab, b);
    } else if (p.head_dim == 128) {
        return launch_batch<128>(cfg, cluster, G, p, tab, b);
    }
    return cudaErrorInvalidValue;
}

cudaError_t attn_decode_launch(const AttnParams& p, cudaStream_t stream, bool pdl) {
    const int G = p.n_kv > 0 ? p.n_head / p.n_kv : 0;
    if (p.n_kv <= 0 || p.n_head % p.n_kv || G > AT_GMAX) return cudaErrorInvalidValue;
    if (p.n_splits < 1 || G * p.n_splits > AT_GMAX * AT_STAGE) return cudaErrorInvalidValue;   // merge staging fits sc[]
    const bool cluster = p.n_splits == 2 || p.n_splits == 4 || p.n_splits == 8;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.n_kv); cfg.stream = stream;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    if (cluster) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = p.n_splits; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
        na++;
    }
    cfg.attrs = at; cfg.numAttrs = na;
    if (p.head_dim == 64) {
        return launch_one<64>(cfg, cluster, G, p);
    } else if (p.head_dim == 128) {
        return launch_one<128>(cfg, cluster, G, p);
    }
    return cudaErrorInvalidValue;
}

}  // namespace b2

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of the global-scratch split merge (b2_attention_decode_static in the C ABI).
//
// Who needs it: a serving host that has already handed every byte of HBM it owns to weights and KV
// pages (the engine sizes its KV pool up front), or that runs this op on a latency-critical token
// loop where a driver allocation would serialise the device and may fail under memory pressure.
// Such a caller cannot cudaMalloc the split-merge scratch, so the scratch lives here, inside the
// library image: a fixed file-scope device array that every GPU loading the library carries until
// the process exits (STATIC_SCRATCH_FLOATS * 4 bytes + 256 bytes of tickets per GPU).
//
// `g_static_scratch` / `g_static_tickets` are device symbols. In host code those names are only
// stand-ins, not device addresses, so the launcher asks the runtime for the real address once per
// device (cudaGetSymbolAddress with that device current) and passes the pointer to the kernel.
namespace b2 {

constexpr size_t STATIC_SCRATCH_FLOATS = 256 * 1024;        // 1 MiB of split partials per GPU
constexpr int STATIC_MAX_KV_HEADS = 64;
constexpr int STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) float g_static_scratch[STATIC_SCRATCH_FLOATS];
static __device__ __align__(256) unsigned int g_static_tickets[STATIC_MAX_KV_HEADS];

static float* g_scratch_addr[STATIC_MAX_DEVICES];
static unsigned int* g_ticket_addr[STATIC_MAX_DEVICES];
static cudaError_t g_static_status[STATIC_MAX_DEVICES];
static std::once_flag g_static_once[STATIC_MAX_DEVICES];      // calls may come from any host thread

size_t attn_static_scratch_floats() { return STATIC_SCRATCH_FLOATS; }

// Once per device (that device must be current): resolve the device addresses of the two arrays
// and raise the kernel's dynamic shared-memory limit. Nothing here allocates.
cudaError_t attn_static_setup(int device) {
    if (device < 0 || device >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::call_once(g_static_once[device], [device]() {
        cudaError_t e;
        float* sp = nullptr;
        unsigned int* t = nullptr;
        if ((e = cudaGetSymbolAddress((void**)&sp, g_static_scratch)) == cudaSuccess &&
            (e = cudaGetSymbolAddress((void**)&t, g_static_tickets)) == cudaSuccess && (e = attn_init()) == cudaSuccess) {
            g_scratch_addr[device] = sp;
            g_ticket_addr[device] = t;
        }
        g_static_status[device] = e;
    });
    return g_static_status[device];
}

// Launch the scratch-merging kernel on the built-in arrays. All pointers in `p` except scratch and
// counters are the caller's device pointers. Every call shares the same arrays: the tickets are
// cleared here before the launch (they must start at zero); the partials keep whatever the last
// call left and are fully rewritten before they are read. Two calls that overlap on the GPU
// (different streams) would race on both arrays.
cudaError_t attn_decode_launch_static(AttnParams p, int device, cudaStream_t stream) {
    if (device < 0 || device >= STATIC_MAX_DEVICES || !g_scratch_addr[device]) return cudaErrorNotReady;
    p.scratch = g_scratch_addr[device];
    p.counters = g_ticket_addr[device];
    cudaError_t e = cudaMemsetAsync(p.counters, 0, sizeof(unsigned int) * STATIC_MAX_KV_HEADS, stream);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.n_kv); cfg.stream = stream;
    cfg.attrs = nullptr; cfg.numAttrs = 0;
    if (p.head_dim == 64) {
        return launch_one<64>(cfg, false, p.n_head / p.n_kv, p);
    }
    return launch_one<128>(cfg, false, p.n_head / p.n_kv, p);
}

}  // namespace b2
