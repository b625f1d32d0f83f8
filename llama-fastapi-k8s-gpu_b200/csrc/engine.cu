// Engine object + the C ABI declared in include/b200_gguf.h.
// Owns the device copies of the GGUF tensors (raw quant blocks, unchanged), the paged f16 KV pool,
// per-slot activation buffers and the captured CUDA graphs of one decode / prefill step.
// Stands in for libllama's llama_model + llama_context behind llama-cpp-python's Llama object
// (/root/reference/api.py:24-28,55-63).
#include "../../include/b200_gguf.h"
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

using namespace b2;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return fail(B2_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e));         \
    } while (0)

// kernel watchdogs (ptx.cuh mbar_wait) report through a host-mapped word; call after a sync
static int check_kernel_error() {
    unsigned int* w = kernel_error_word();
    if (w && *w) {
        unsigned int c = w[0];
        char buf[160];
        snprintf(buf, sizeof(buf), "kernel barrier wait timed out (watchdog code 0x%x, block %u, thread %u, tile %u, parity %u)",
                 c, w[1], w[2], w[3], w[4]);
        for (int i = 0; i < 8; i++) w[i] = 0;
        return fail(B2_ERR_CUDA, buf);
    }
    return B2_OK;
}

struct DevTensor {
    uint8_t* ptr = nullptr;
    int type = -1;
    int64_t ne0 = 0, ne1 = 0, nbytes = 0;
};

struct Layer {
    DevTensor attn_norm, wq, wk, wv, wo, ffn_norm, wgate, wup, wdown;
    MvParams p_qkv, p_o, p_gu, p_down;
    MvParams p_gate, p_up;         // un-fused pair when ffn_gate and ffn_up have different quant types
    bool gu_fused = true;
};

struct Slot {
    float *x = nullptr, *q = nullptr, *att = nullptr, *h = nullptr, *hg = nullptr, *logits = nullptr;
    float* tp_part = nullptr;      // tensor parallelism: this rank's partial sum of attn_output / ffn_down (rank 0: + the residual)
    int* ints = nullptr;           // [0]=token [1]=pos [2]=out_token [3]=step [4]=cursor [5]=pos_base
    int* block_table = nullptr;
    int* prompt = nullptr;         // [n_ctx]
    int* ring = nullptr;           // [n_ctx + 1] sampled token by position
    SamplerState* sst = nullptr;
    SampleParams* sp = nullptr;
    double* uniforms = nullptr;    // [8192]
    cudaGraphExec_t g_decode = nullptr, g_pre_step = nullptr, g_pre_last = nullptr, g_fwd = nullptr;
    int n_decode_nodes = 0, n_pre_step_nodes = 0, n_pre_last_nodes = 0, n_fwd_nodes = 0;
    const float* logits_view = nullptr;   // where the latest logits of this slot are (own buffer, or a row of the batch output)
    int n_past = 0;                // host mirror of ints[1]
    int n_sampled = 0;             // host mirror of ints[3]
};

struct b2_engine {
    b2_config cfg;
    cudaStream_t stream = nullptr;
    std::unordered_map<std::string, DevTensor> tensors;
    std::vector<Layer> layers;
    DevTensor tok_embd, out_norm, output;
    MvParams p_head;
    __half *k_pool = nullptr, *v_pool = nullptr;
    size_t layer_kv_elems = 0;
    float* rope_freq = nullptr;
    float* attn_scratch = nullptr;
    unsigned* attn_counters = nullptr;
    int n_splits = 1, max_blocks = 0, n_ctas = 148;
    std::vector<Slot> slots;
    int32_t* pinned = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int64_t launches = 0, weight_bytes = 0, kv_bytes = 0;
    bool use_graph = true, use_pdl = true, act2 = true, finalized = false, chain_pdl = true;
    int batch_attn_splits = 0;     // B2_BATCH_ATTN_SPLITS (tuning), read once at create
    int attn_early = 1;            // B2_ATTN_EARLY=0 (tuning): attention requests nothing before its dependency wait
    // tensor parallelism (b2_engine_tp_init): this engine holds 1 / tp_world of the heads, FFN rows and vocabulary rows
    void* tp_comm = nullptr;
    int tp_rank = 0, tp_world = 1;
    // all-reduce over peer memory (tp.cu): [2][n_embd] partial sums + flags + sequence number in one IPC-shared allocation
    float* tp_shared = nullptr;
    void* tp_peer[TP_MAX_WORLD] = {};
    bool tp_p2p = false;
    TpArParams tp_ar[2] = {};
    unsigned long long* tl = nullptr;   // B2_TIMELINE=1: 4 time stamps per kernel of a decode step (debugging aid)
    int tl_n = 0;                  // kernels recorded by the last forward_launches
    // prompt (GEMM) workspace, allocated on the first long prompt: PF_T tokens per chunk
    float *pf_x = nullptr, *pf_part = nullptr;     // residual stream [PF_T][n_embd]; K-split partial sums of the last GEMM
    size_t pf_part_floats = 0;
    __half *pf_xn = nullptr, *pf_qh = nullptr, *pf_att = nullptr, *pf_h = nullptr;
    int n_sms = 148;
    // batched decode (several sequences per step through the GEMM path)
    SlotDev* slot_tab = nullptr;   // device table, one entry per slot
    float *bd_q = nullptr, *bd_att = nullptr, *bd_scratch = nullptr;
    unsigned* bd_counters = nullptr;
    int* bd_out = nullptr;
    int attn_mma = -1;             // tensor-core flash-decode: -1 = batched steps only; 0 never; 1 always
    int batch_min = 3;             // fewer sequences than this step one after another through the mat-vec path (measured: 3 slots 4.1 vs 6.3 ms)
    bool gemm_ok = false;          // every layer matrix is a block format the GEMM reads
    int prefill_min = 8;           // prompts chunks shorter than this go token by token through the mat-vec path
    // GEMM launch plans (tile split, K splits, tensor maps) are built once per (first weight matrix, token count) and reused:
    // re-planning -- cuTensorMapEncodeTiled and a handful of getenv per call -- for each of the 129 GEMMs of every batched
    // step cost more host time than the step takes on the GPU
    std::unordered_map<uint64_t, GemmParams> gemm_plans;
    std::mutex mu;
};

// ------------------------------------------------------------------------------------------
// small device helpers
__global__ void prefill_next_kernel(const int* prompt, int* ints, SamplerState* st) {
    griddep_launch();
    griddep_wait();
    int cur = ints[4];
    int tok = prompt[cur];
    ints[0] = tok;
    ints[1] = ints[5] + cur;
    ints[4] = cur + 1;
    int n = st->n_hist;
    st->history[n & (HISTORY_LEN - 1)] = tok;
    st->n_hist = n + 1;
}

static cudaError_t launch_prefill_next(Slot& s, cudaStream_t st, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(1); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, prefill_next_kernel, (const int*)s.prompt, s.ints, s.sst);
}

// ------------------------------------------------------------------------------------------
extern "C" const char* b2_last_error(void) { return g_err.c_str(); }
extern "C" int b2_version(void) { return 100; }
extern "C" int b2_device_count(int* n) {
    CK(cudaGetDeviceCount(n));
    return B2_OK;
}

static int engine_create_impl(b2_engine* e, const b2_config* c);
extern "C" int b2_engine_create(const b2_config* c, b2_engine** out) {
    if (!c || !out) return fail(B2_ERR_ARG, "null argument");
    if (c->n_layer <= 0 || c->n_embd <= 0 || c->n_head <= 0 || c->n_head_kv <= 0 || c->n_vocab <= 0 ||
        c->n_ff <= 0 || c->n_ctx <= 0 || c->max_seqs <= 0 || c->kv_block_tokens <= 0 || c->kv_num_blocks <= 0)
        return fail(B2_ERR_ARG, "bad model/config dimensions");
    if (c->head_dim != 64 && c->head_dim != 128) return fail(B2_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    if (c->n_head % c->n_head_kv || c->n_head / c->n_head_kv > 8)
        return fail(B2_ERR_UNSUPPORTED, "n_head / n_head_kv must be an integer <= 8");
    if (c->n_embd % 256 || c->n_ff % 256 || (c->n_head * c->head_dim) % 256)
        return fail(B2_ERR_UNSUPPORTED, "n_embd, n_ff and n_head*head_dim must be multiples of 256");
    if (c->kv_block_tokens % 16) return fail(B2_ERR_ARG, "kv_block_tokens must be a multiple of 16");
    b2_engine* e = new b2_engine();
    e->cfg = *c;
    const int rc = engine_create_impl(e, c);
    if (rc) { b2_engine_destroy(e); return rc; }     // frees whatever was allocated before the failure (e.g. KV pool OOM)
    *out = e;
    return B2_OK;
}

static int engine_create_impl(b2_engine* e, const b2_config* c) {
    CK(cudaSetDevice(c->device));
    CK(mv_init());
    CK(attn_init());
    CK(attn_mma_init());
    e->use_graph = !(c->flags & B2_FLAG_NO_GRAPH);
    e->use_pdl = !(c->flags & B2_FLAG_NO_PDL);
    e->act2 = !(c->flags & B2_FLAG_ACT8);
    // tuning knobs are read here, once; the hot path never calls getenv
    e->chain_pdl = e->use_pdl && !getenv("B2_NO_CHAIN_PDL");
    if (const char* s = getenv("B2_BATCH_ATTN_SPLITS")) e->batch_attn_splits = atoi(s);
    if (const char* s = getenv("B2_ATTN_EARLY")) e->attn_early = atoi(s);
    if (getenv("B2_TIMELINE")) CK(cudaMalloc(&e->tl, sizeof(unsigned long long) * 4 * 1024));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    // one mat-vec CTA per SM: the second CTA slot of every SM is for the NEXT kernel's CTAs (matvec.cu)
    int per = mv_ctas_per_sm();
    if (const char* s = getenv("B2_CTAS_PER_SM")) per = atoi(s) > 0 ? atoi(s) : per;
    e->n_ctas = sms * per;
    e->n_sms = sms;
    CK(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&e->ev0));
    CK(cudaEventCreate(&e->ev1));
    CK(cudaMallocHost(&e->pinned, sizeof(int32_t) * (size_t)(c->max_seqs + c->n_ctx + 16)));
    // KV pool: [layer][block][kv_head][token][dim] f16
    e->layer_kv_elems = (size_t)c->kv_num_blocks * c->n_head_kv * c->kv_block_tokens * c->head_dim;
    size_t kvb = e->layer_kv_elems * c->n_layer * sizeof(__half);
    CK(cudaMalloc(&e->k_pool, kvb));
    CK(cudaMalloc(&e->v_pool, kvb));
    CK(cudaMemsetAsync(e->k_pool, 0, kvb, e->stream));
    CK(cudaMemsetAsync(e->v_pool, 0, kvb, e->stream));
    e->kv_bytes = 2 * (int64_t)kvb;
    // rope table, built like ggml: theta_scale = powf(base, -2/d), multiplied up in float32
    {
        std::vector<float> f(c->head_dim / 2);
        float scale = powf(c->rope_freq_base, -2.0f / (float)c->head_dim), v = 1.0f;
        for (int i = 0; i < c->head_dim / 2; i++) { f[i] = v; v *= scale; }
        CK(cudaMalloc(&e->rope_freq, f.size() * 4));
        CK(cudaMemcpy(e->rope_freq, f.data(), f.size() * 4, cudaMemcpyHostToDevice));
    }
    // 8 splits per KV head = one thread-block cluster merging through distributed shared memory
    // long contexts: 16 splits (global-scratch merge) put 128 CTAs on the 148 SMs: mistral-7b at an 8k context measured
    // 2.15 ms/token with 16 splits against 2.36 with 32 (two waves of CTAs)
    int ns = c->n_ctx > 2048 ? 16 : 8;
    while (ns > 1 && ns * 16 > c->n_ctx) ns >>= 1;
    if (const char* s = getenv("B2_ATTN_SPLITS")) if (atoi(s) > 0) ns = atoi(s);
    e->n_splits = ns;
    CK(cudaMalloc(&e->attn_scratch, attn_scratch_floats(c->n_head, c->n_head_kv, c->head_dim, ns) * 4));
    CK(cudaMalloc(&e->attn_counters, sizeof(unsigned) * c->n_head_kv));
    CK(cudaMemset(e->attn_counters, 0, sizeof(unsigned) * c->n_head_kv));
    e->max_blocks = (c->n_ctx + c->kv_block_tokens - 1) / c->kv_block_tokens;
    e->slots.resize(c->max_seqs);
    for (auto& s : e->slots) {
        CK(cudaMalloc(&s.x, sizeof(float) * c->n_embd));
        CK(cudaMalloc(&s.q, sizeof(float) * c->n_head * c->head_dim));
        CK(cudaMalloc(&s.att, sizeof(float) * c->n_head * c->head_dim));
        CK(cudaMalloc(&s.h, sizeof(float) * c->n_ff));
        CK(cudaMalloc(&s.hg, sizeof(float) * c->n_ff));
        CK(cudaMalloc(&s.tp_part, sizeof(float) * c->n_embd));
        CK(cudaMalloc(&s.logits, sizeof(float) * c->n_vocab));
        CK(cudaMalloc(&s.ints, sizeof(int) * 8));
        CK(cudaMemset(s.ints, 0, sizeof(int) * 8));
        CK(cudaMalloc(&s.block_table, sizeof(int) * e->max_blocks));
        CK(cudaMemset(s.block_table, 0, sizeof(int) * e->max_blocks));
        CK(cudaMalloc(&s.prompt, sizeof(int) * c->n_ctx));
        CK(cudaMalloc(&s.ring, sizeof(int) * (c->n_ctx + 1)));
        CK(cudaMemset(s.ring, 0, sizeof(int) * (c->n_ctx + 1)));
        CK(cudaMalloc(&s.sst, sizeof(SamplerState)));
        CK(cudaMemset(s.sst, 0, sizeof(SamplerState)));
        CK(cudaMalloc(&s.sp, sizeof(SampleParams)));
        CK(cudaMemset(s.sp, 0, sizeof(SampleParams)));
        CK(cudaMalloc(&s.uniforms, sizeof(double) * 8192));
        CK(cudaMemset(s.uniforms, 0, sizeof(double) * 8192));
    }
    {
        std::vector<SlotDev> tab(c->max_seqs);
        for (int i = 0; i < c->max_seqs; i++) {
            Slot& s = e->slots[i];
            tab[i].ints = s.ints; tab[i].block_table = s.block_table; tab[i].ring = s.ring;
            tab[i].sst = s.sst; tab[i].sp = s.sp; tab[i].uniforms = s.uniforms;
        }
        CK(cudaMalloc(&e->slot_tab, sizeof(SlotDev) * c->max_seqs));
        CK(cudaMemcpy(e->slot_tab, tab.data(), sizeof(SlotDev) * c->max_seqs, cudaMemcpyHostToDevice));
    }
    CK(cudaStreamSynchronize(e->stream));
    return B2_OK;
}

static void free_slot_graphs(Slot& s) {
    if (s.g_decode) cudaGraphExecDestroy(s.g_decode);
    if (s.g_pre_step) cudaGraphExecDestroy(s.g_pre_step);
    if (s.g_pre_last) cudaGraphExecDestroy(s.g_pre_last);
    if (s.g_fwd) cudaGraphExecDestroy(s.g_fwd);
    s.g_decode = s.g_pre_step = s.g_pre_last = s.g_fwd = nullptr;
}

extern "C" int b2_engine_destroy(b2_engine* e) {
    if (!e) return B2_OK;
    cudaSetDevice(e->cfg.device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (auto& kv : e->tensors) cudaFree(kv.second.ptr);
    for (auto& s : e->slots) {
        free_slot_graphs(s);
        cudaFree(s.x); cudaFree(s.q); cudaFree(s.att); cudaFree(s.h); cudaFree(s.hg); cudaFree(s.tp_part); cudaFree(s.logits);
        cudaFree(s.ints); cudaFree(s.block_table); cudaFree(s.prompt); cudaFree(s.ring); cudaFree(s.sst); cudaFree(s.sp);
        cudaFree(s.uniforms);
    }
    cudaFree(e->tl);
    for (int r = 0; r < e->tp_world; r++)
        if (e->tp_peer[r] && r != e->tp_rank) cudaIpcCloseMemHandle(e->tp_peer[r]);
    cudaFree(e->tp_shared);
    tp_comm_destroy(e->tp_comm);
    cudaFree(e->k_pool); cudaFree(e->v_pool); cudaFree(e->rope_freq);
    cudaFree(e->attn_scratch); cudaFree(e->attn_counters);
    cudaFree(e->pf_x); cudaFree(e->pf_part); cudaFree(e->pf_xn); cudaFree(e->pf_qh); cudaFree(e->pf_att); cudaFree(e->pf_h);
    cudaFree(e->slot_tab); cudaFree(e->bd_q); cudaFree(e->bd_att); cudaFree(e->bd_scratch); cudaFree(e->bd_counters); cudaFree(e->bd_out);
    if (e->pinned) cudaFreeHost(e->pinned);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    cudaGetLastError();
    delete e;
    return B2_OK;
}

extern "C" int b2_upload_tensor(b2_engine* e, const char* name, int type, int64_t ne0, int64_t ne1,
                                const void* host, int64_t nbytes) {
    if (!e || !name || !host) return fail(B2_ERR_ARG, "null argument");
    std::lock_guard<std::mutex> lk(e->mu);
    CK(cudaSetDevice(e->cfg.device));
    int be = block_elems(type), bb = block_bytes(type);
    if (!bb) return fail(B2_ERR_UNSUPPORTED, std::string("tensor ") + name + ": unsupported ggml type " + std::to_string(type));
    if (ne0 % be || ne0 / be * bb * ne1 != nbytes)
        return fail(B2_ERR_ARG, std::string("tensor ") + name + ": byte size does not match its shape");
    if (e->tensors.count(name)) return fail(B2_ERR_STATE, std::string("tensor ") + name + " uploaded twice");
    DevTensor t;
    t.type = type; t.ne0 = ne0; t.ne1 = ne1; t.nbytes = nbytes;
    CK(cudaMalloc(&t.ptr, (size_t)nbytes + 64));   // +64: aligned bulk copies may read a few bytes past a tile
    CK(cudaMemcpy(t.ptr, host, (size_t)nbytes, cudaMemcpyHostToDevice));
    CK(cudaMemset(t.ptr + nbytes, 0, 64));
    e->tensors[name] = t;
    e->weight_bytes += nbytes;
    return B2_OK;
}

static int get_tensor(b2_engine* e, const std::string& name, int64_t ne0, int64_t ne1, bool quant_ok, DevTensor& out) {
    auto it = e->tensors.find(name);
    if (it == e->tensors.end()) return fail(B2_ERR_STATE, "missing tensor " + name);
    const DevTensor& t = it->second;
    if (t.ne0 != ne0 || t.ne1 != ne1)
        return fail(B2_ERR_ARG, "tensor " + name + " has shape [" + std::to_string(t.ne0) + "," +
                                    std::to_string(t.ne1) + "], expected [" + std::to_string(ne0) + "," +
                                    std::to_string(ne1) + "]");
    if (!quant_ok && t.type != T_F32) return fail(B2_ERR_UNSUPPORTED, "tensor " + name + " must be F32");
    if (quant_ok && (t.type == T_F32 || t.type == T_F16))
        return fail(B2_ERR_UNSUPPORTED, "tensor " + name + ": F32/F16 weight matrices are not supported by the mat-vec yet");
    out = t;
    return B2_OK;
}

static MvSeg make_seg(const DevTensor& t, int out_off = 0) {
    MvSeg s = {};
    s.w = t.ptr; s.type = t.type; s.n_rows = (int)t.ne1; s.out_off = out_off;
    return s;
}

static int plan_or_fail(MvParams& p, int n_ctas, const char* what) {
    int rc = mv_plan(p, n_ctas);
    if (rc) return fail(B2_ERR_UNSUPPORTED, std::string("cannot plan mat-vec for ") + what + " (code " + std::to_string(rc) +
                                               (rc == -2 ? ": K > 32768 not supported)" : ")"));
    return B2_OK;
}

extern "C" int b2_tp_unique_id(void* out128) {
    if (!out128) return fail(B2_ERR_ARG, "null argument");
    std::string err;
    if (!tp_unique_id(out128, err)) return fail(B2_ERR_CUDA, err);
    return B2_OK;
}

extern "C" int b2_engine_tp_init(b2_engine* e, int rank, int world, const void* id128) {
    if (!e || !id128) return fail(B2_ERR_ARG, "null argument");
    if (e->finalized) return fail(B2_ERR_STATE, "b2_engine_tp_init must come before b2_engine_finalize");
    if (world < 2 || rank < 0 || rank >= world) return fail(B2_ERR_ARG, "bad rank / world");
    std::lock_guard<std::mutex> lk(e->mu);
    CK(cudaSetDevice(e->cfg.device));
    std::string err;
    if (!tp_comm_init(&e->tp_comm, rank, world, id128, err)) return fail(B2_ERR_CUDA, err);
    e->tp_rank = rank; e->tp_world = world;
    // first collective outside any stream capture: NCCL sets up its channels (and allocates) lazily
    Slot& s = e->slots[0];
    CK(cudaMemsetAsync(s.tp_part, 0, sizeof(float) * e->cfg.n_embd, e->stream));
    if (!tp_all_reduce_sum(e->tp_comm, s.tp_part, s.tp_part, e->cfg.n_embd, e->stream, err)) return fail(B2_ERR_CUDA, err);
    const int v_local = e->cfg.n_vocab / world;
    if (!tp_all_gather(e->tp_comm, s.logits + (size_t)rank * v_local, s.logits, v_local, e->stream, err)) return fail(B2_ERR_CUDA, err);
    CK(cudaStreamSynchronize(e->stream));
    // map every rank's reduction buffers into every other rank (CUDA IPC); when that is not possible the NCCL all-reduce stays
    if (world <= TP_MAX_WORLD && !getenv("B2_TP_NCCL")) {
        const size_t n = (size_t)e->cfg.n_embd;
        const size_t bytes = 2 * n * sizeof(float) + 256;
        CK(cudaMalloc(&e->tp_shared, bytes));
        CK(cudaMemset(e->tp_shared, 0, bytes));
        cudaIpcMemHandle_t mine;
        bool ok = cudaIpcGetMemHandle(&mine, e->tp_shared) == cudaSuccess;
        unsigned char* dh = nullptr;
        CK(cudaMalloc(&dh, (size_t)world * (sizeof(mine) + 8)));
        std::vector<unsigned char> rec(sizeof(mine) + 8, 0), all((size_t)world * (sizeof(mine) + 8));
        memcpy(rec.data(), &mine, sizeof(mine)); rec[sizeof(mine)] = ok ? 1 : 0;
        CK(cudaMemcpy(dh + (size_t)rank * rec.size(), rec.data(), rec.size(), cudaMemcpyHostToDevice));
        if (!tp_all_gather_bytes(e->tp_comm, dh + (size_t)rank * rec.size(), dh, rec.size(), e->stream, err)) return fail(B2_ERR_CUDA, err);
        CK(cudaStreamSynchronize(e->stream));
        CK(cudaMemcpy(all.data(), dh, all.size(), cudaMemcpyDeviceToHost));
        cudaFree(dh);
        for (int r = 0; r < world && ok; r++) ok = all[(size_t)r * rec.size() + sizeof(mine)] == 1;
        for (int r = 0; r < world && ok; r++) {
            if (r == rank) { e->tp_peer[r] = e->tp_shared; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, all.data() + (size_t)r * rec.size(), sizeof(h));
            if (cudaIpcOpenMemHandle(&e->tp_peer[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { e->tp_peer[r] = nullptr; ok = false; }
        }
        cudaGetLastError();
        // everybody must agree on the path: one more (tiny) collective
        float* flag = s.tp_part;
        const float mine_ok = ok ? 0.f : 1.f;
        CK(cudaMemcpy(flag, &mine_ok, 4, cudaMemcpyHostToDevice));
        if (!tp_all_reduce_sum(e->tp_comm, flag, flag, 1, e->stream, err)) return fail(B2_ERR_CUDA, err);
        float bad = 0.f;
        CK(cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        CK(cudaMemset(flag, 0, 4));
        e->tp_p2p = bad == 0.f;
        if (e->tp_p2p) {
            for (int b = 0; b < 2; b++) {
                TpArParams& a = e->tp_ar[b];
                a.x = nullptr;      // per call: the slot's hidden vector
                for (int r = 0; r < world; r++) {
                    a.part[r] = reinterpret_cast<const float*>(e->tp_peer[r]) + (size_t)b * n;
                    a.flag_out[r] = reinterpret_cast<unsigned int*>(reinterpret_cast<float*>(e->tp_peer[r]) + 2 * n);
                }
                a.flags_in = reinterpret_cast<const unsigned int*>(e->tp_shared + 2 * n);
                a.seq = reinterpret_cast<unsigned int*>(e->tp_shared + 2 * n) + 32;
                a.rank = rank; a.world = world; a.n = (int)n; a.err = kernel_error_word();
            }
        }
        if (getenv("B2_TP_VERBOSE")) fprintf(stderr, "[b2] tensor parallel rank %d/%d: hidden-vector all-reduce over %s\n", rank, world, e->tp_p2p ? "peer memory" : "NCCL");
    }
    return B2_OK;
}

extern "C" int b2_engine_tp_mode(b2_engine* e, int* rank, int* world, int* peer_memory) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    if (rank) *rank = e->tp_rank; if (world) *world = e->tp_world; if (peer_memory) *peer_memory = e->tp_p2p ? 1 : 0;
    return B2_OK;
}

// block formats the tensor-core GEMM dequantises (gemm.cu): all of them -- Q4_0, Q8_0, Q4_K, Q5_K, Q6_K with hand-written
// expansions, Q4_1, Q5_0, Q5_1, Q2_K, Q3_K element by element
static bool gemm_type_ok(int t) { return block_elems(t) != 1 && block_bytes(t) != 0; }

extern "C" int b2_engine_finalize(b2_engine* e) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    std::lock_guard<std::mutex> lk(e->mu);
    const b2_config& c = e->cfg;
    int rc;
    int qd = c.n_head * c.head_dim, kvd = c.n_head_kv * c.head_dim;
    if ((rc = get_tensor(e, "token_embd.weight", c.n_embd, c.n_vocab, true, e->tok_embd))) {
        // F16/F32 embeddings are fine for the gather
        auto it = e->tensors.find("token_embd.weight");
        if (it == e->tensors.end() || it->second.ne0 != c.n_embd || it->second.ne1 != c.n_vocab) return rc;
        e->tok_embd = it->second;
    }
    if ((rc = get_tensor(e, "output_norm.weight", c.n_embd, 1, false, e->out_norm))) return rc;
    const int tpw = e->tp_world;
    if (c.n_vocab % (2 * tpw)) return fail(B2_ERR_UNSUPPORTED, "n_vocab must be a multiple of 2 * tp_world");
    if (tpw > 1) {      // the output matrix is split by rows: this rank holds vocabulary rows [rank, rank + 1) * n_vocab / tp_world
        if ((rc = get_tensor(e, "output.weight", c.n_embd, c.n_vocab / tpw, true, e->output))) return rc;
    } else if (e->tensors.count("output.weight")) {
        if ((rc = get_tensor(e, "output.weight", c.n_embd, c.n_vocab, true, e->output))) return rc;
    } else {
        if ((rc = get_tensor(e, "token_embd.weight", c.n_embd, c.n_vocab, true, e->output))) return rc;  // tied
    }
    e->layers.resize(c.n_layer);
    for (int l = 0; l < c.n_layer; l++) {
        Layer& L = e->layers[l];
        std::string p = "blk." + std::to_string(l) + ".";
        if ((rc = get_tensor(e, p + "attn_norm.weight", c.n_embd, 1, false, L.attn_norm))) return rc;
        if ((rc = get_tensor(e, p + "ffn_norm.weight", c.n_embd, 1, false, L.ffn_norm))) return rc;
        if ((rc = get_tensor(e, p + "attn_q.weight", c.n_embd, qd, true, L.wq))) return rc;
        if ((rc = get_tensor(e, p + "attn_k.weight", c.n_embd, kvd, true, L.wk))) return rc;
        if ((rc = get_tensor(e, p + "attn_v.weight", c.n_embd, kvd, true, L.wv))) return rc;
        if ((rc = get_tensor(e, p + "attn_output.weight", qd, c.n_embd, true, L.wo))) return rc;
        if ((rc = get_tensor(e, p + "ffn_gate.weight", c.n_embd, c.n_ff, true, L.wgate))) return rc;
        if ((rc = get_tensor(e, p + "ffn_up.weight", c.n_embd, c.n_ff, true, L.wup))) return rc;
        if ((rc = get_tensor(e, p + "ffn_down.weight", c.n_ff, c.n_embd, true, L.wdown))) return rc;
        MvParams q = {};
        q.nseg = 3; q.seg[0] = make_seg(L.wq); q.seg[1] = make_seg(L.wk); q.seg[2] = make_seg(L.wv);
        q.K = c.n_embd; q.mode = MV_QKV; q.norm_w = (const float*)L.attn_norm.ptr; q.eps = c.rms_eps;
        q.k_cache = e->k_pool + (size_t)l * e->layer_kv_elems; q.v_cache = e->v_pool + (size_t)l * e->layer_kv_elems;
        q.rope_freq = e->rope_freq; q.head_dim = c.head_dim; q.n_kv = c.n_head_kv; q.block_tokens = c.kv_block_tokens;
        q.err = kernel_error_word();
        if ((rc = plan_or_fail(q, e->n_ctas, "attn_q/k/v"))) return rc;
        L.p_qkv = q;
        MvParams o = {};
        // tensor parallelism: every rank produces a partial sum over its K range; rank 0's carries the residual, then all-reduce
        const int part_mode = (tpw > 1 && e->tp_rank != 0) ? MV_STORE : MV_RESID;
        o.nseg = 1; o.seg[0] = make_seg(L.wo); o.K = qd; o.mode = part_mode; o.err = kernel_error_word();
        if ((rc = plan_or_fail(o, e->n_ctas, "attn_output"))) return rc;
        L.p_o = o;
        MvParams g = {};
        g.nseg = 2; g.seg[0] = make_seg(L.wgate); g.seg[1] = make_seg(L.wup); g.K = c.n_embd; g.mode = MV_GATEUP;
        g.norm_w = (const float*)L.ffn_norm.ptr; g.eps = c.rms_eps; g.err = kernel_error_word();
        if (L.wgate.type == L.wup.type) {
            if ((rc = plan_or_fail(g, e->n_ctas, "ffn_gate/up"))) return rc;
            L.p_gu = g;
        } else {          // mixed pair: two plain mat-vecs (each with the fused norm) and a SiLU-multiply
            L.gu_fused = false;
            MvParams a = {}, b = {};
            a.nseg = 1; a.seg[0] = make_seg(L.wgate); a.K = c.n_embd; a.mode = MV_STORE;
            a.norm_w = (const float*)L.ffn_norm.ptr; a.eps = c.rms_eps; a.err = kernel_error_word();
            b = a; b.seg[0] = make_seg(L.wup);
            if ((rc = plan_or_fail(a, e->n_ctas, "ffn_gate"))) return rc;
            if ((rc = plan_or_fail(b, e->n_ctas, "ffn_up"))) return rc;
            L.p_gate = a; L.p_up = b;
        }
        MvParams d = {};
        d.nseg = 1; d.seg[0] = make_seg(L.wdown); d.K = c.n_ff; d.mode = part_mode; d.err = kernel_error_word();
        if ((rc = plan_or_fail(d, e->n_ctas, "ffn_down"))) return rc;
        L.p_down = d;
    }
    MvParams h = {};
    h.nseg = 1; h.seg[0] = make_seg(e->output); h.K = c.n_embd; h.mode = MV_STORE;
    h.norm_w = (const float*)e->out_norm.ptr; h.eps = c.rms_eps; h.err = kernel_error_word();
    if ((rc = plan_or_fail(h, e->n_ctas, "output"))) return rc;
    e->p_head = h;
    e->gemm_ok = !getenv("B2_NO_GEMM_PREFILL") && tpw == 1;     // tensor parallelism: prompts and batches go through the mat-vec path
    for (auto& L : e->layers)
        for (const DevTensor* t : {&L.wq, &L.wk, &L.wv, &L.wo, &L.wgate, &L.wup, &L.wdown})
            if (!gemm_type_ok(t->type)) e->gemm_ok = false;
    if (!gemm_type_ok(e->output.type) && block_elems(e->output.type) != 1) e->gemm_ok = false;   // (an F16/F32 output matrix only rules out the batched lm_head)
    if (const char* s = getenv("B2_PREFILL_MIN")) if (atoi(s) > 0) e->prefill_min = atoi(s);
    if (const char* s = getenv("B2_BATCH_MIN")) if (atoi(s) > 0) e->batch_min = atoi(s);
    if (const char* s = getenv("B2_ATTN_MMA")) e->attn_mma = atoi(s) ? 1 : 0;
    if (e->gemm_ok) CK(gemm_init());
    e->finalized = true;
    return B2_OK;
}

extern "C" int b2_engine_device_bytes(b2_engine* e, int64_t* w, int64_t* kv) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    if (w) *w = e->weight_bytes;
    if (kv) *kv = e->kv_bytes;
    return B2_OK;
}

// ------------------------------------------------------------------------------------------
// forward of one slot's pending token; `timing` (optional) receives per-kind event pairs
struct KindTimer {
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[8];
};

static int forward_launches(b2_engine* e, Slot& s, bool with_head, bool with_sample, bool pdl, KindTimer* kt) {
    const b2_config& c = e->cfg;
    cudaStream_t st = e->stream;
    auto tick = [&](int kind, cudaEvent_t* a) {
        if (!kt) return;
        cudaEvent_t x, y;
        cudaEventCreate(&x); cudaEventCreate(&y);
        cudaEventRecord(x, st);
        kt->ev[kind].push_back({x, y});
        *a = y;
    };
    auto tock = [&](cudaEvent_t y) { if (kt) cudaEventRecord(y, st); };
    cudaEvent_t ye = nullptr;
    int* tok = s.ints + 0; int* pos = s.ints + 1;
    const bool tp = e->tp_world > 1;
    std::string tp_err;
    int tl_k = 0;
    auto next_tl = [&]() -> unsigned long long* { return (e->tl && !kt && tl_k < 1024) ? e->tl + 4 * (tl_k++) : nullptr; };
    tick(0, &ye);
    CK(embed_row_launch(e->tok_embd.type, e->tok_embd.ptr, c.n_embd, tok, s.x, st, pdl));
    tock(ye);
    e->launches++;
    for (int l = 0; l < c.n_layer; l++) {
        Layer& L = e->layers[l];
        MvParams q = L.p_qkv;
        q.x = s.x; q.q_out = s.q; q.pos_ptr = pos; q.block_table = s.block_table;
        // the tensor-core kernel is for batched steps; at batch 1 the per-warp scalar kernel is faster at every context
        // length measured (8k: 2.36 against 2.42 ms/token at equal splits) and keeps q and the probabilities in float32
        const bool attn_plain = e->attn_mma != 1;
        q.trigger_late = pdl && attn_plain && e->attn_early;      // attention may start once we are past our own wait: it prefetches old K/V rows
        q.tl = next_tl();
        tick(1, &ye);
        CK(mv_launch(q, e->n_ctas, st, e->act2, pdl));
        tock(ye);
        AttnParams a = {};
        a.q = s.q; a.k_cache = q.k_cache; a.v_cache = q.v_cache; a.block_table = s.block_table; a.pos_ptr = pos;
        a.out = s.att; a.scratch = e->attn_scratch; a.counters = e->attn_counters;
        a.n_head = c.n_head; a.n_kv = c.n_head_kv; a.head_dim = c.head_dim; a.block_tokens = c.kv_block_tokens;
        a.n_splits = e->n_splits; a.scale = 1.0f / sqrtf((float)c.head_dim); a.err = kernel_error_word();
        a.max_pages = e->max_blocks;
        a.early = q.trigger_late;
        a.tl = next_tl();
        tick(2, &ye);
        if (!attn_plain) CK(attn_decode_mma_launch(a, nullptr, nullptr, st, pdl));
        else CK(attn_decode_launch(a, st, pdl));
        tock(ye);
        MvParams o = L.p_o;
        float* part0 = !tp ? s.x : e->tp_p2p ? e->tp_shared : s.tp_part;
        float* part1 = !tp ? s.x : e->tp_p2p ? e->tp_shared + c.n_embd : s.tp_part;
        auto reduce = [&](int buf, float* part) -> bool {      // sum of the ranks' partials -> s.x on every rank
            if (e->tp_p2p) {
                TpArParams ar = e->tp_ar[buf];
                ar.x = s.x;
                e->launches++;
                return tp_allreduce_launch(ar, st, pdl) == cudaSuccess;
            }
            return tp_all_reduce_sum(e->tp_comm, part, s.x, c.n_embd, st, tp_err);
        };
        o.x = s.att; o.y = part0; o.resid = s.x; o.tl = next_tl();
        tick(3, &ye);
        CK(mv_launch(o, e->n_ctas, st, e->act2, pdl));
        if (tp && !reduce(0, part0)) return fail(B2_ERR_CUDA, tp_err.empty() ? "all-reduce launch failed" : tp_err);
        tock(ye);
        tick(4, &ye);
        if (L.gu_fused) {
            MvParams g = L.p_gu;
            g.x = s.x; g.y = s.h; g.tl = next_tl();
            CK(mv_launch(g, e->n_ctas, st, e->act2, pdl));
        } else {
            MvParams a = L.p_gate, b = L.p_up;
            a.x = s.x; a.y = s.hg; b.x = s.x; b.y = s.h;
            CK(mv_launch(a, e->n_ctas, st, e->act2, pdl));
            CK(mv_launch(b, e->n_ctas, st, e->act2, pdl));
            CK(silu_mul_launch(s.hg, s.h, c.n_ff, st, pdl));
            e->launches += 2;
        }
        tock(ye);
        MvParams d = L.p_down;
        d.x = s.h; d.y = part1; d.resid = s.x; d.tl = next_tl();
        tick(5, &ye);
        CK(mv_launch(d, e->n_ctas, st, e->act2, pdl));
        if (tp && !reduce(1, part1)) return fail(B2_ERR_CUDA, tp_err.empty() ? "all-reduce launch failed" : tp_err);
        tock(ye);
        e->launches += 5;
    }
    if (with_head) {
        MvParams h = e->p_head;
        const int v_local = c.n_vocab / e->tp_world;
        h.x = s.x; h.y = s.logits + (size_t)e->tp_rank * v_local; h.tl = next_tl();
        tick(6, &ye);
        CK(mv_launch(h, e->n_ctas, st, e->act2, pdl));
        // every rank ends up with the full logit vector (in place) and samples the same token from the same uniforms
        if (tp && !tp_all_gather(e->tp_comm, s.logits + (size_t)e->tp_rank * v_local, s.logits, v_local, st, tp_err)) return fail(B2_ERR_CUDA, tp_err);
        tock(ye);
        e->launches++;
    }
    if (!kt) e->tl_n = tl_k;
    if (with_sample) {
        tick(7, &ye);
        CK(sample_launch(s.logits, c.n_vocab, s.sp, s.sst, s.uniforms, s.ints + 3, s.ints + 2, tok, pos, s.ring, c.n_ctx + 1, st, pdl));
        tock(ye);
        e->launches++;
    }
    return B2_OK;
}

enum GraphKind { GK_DECODE, GK_PRE_STEP, GK_PRE_LAST, GK_FWD };

static int record_step(b2_engine* e, Slot& s, GraphKind k, bool pdl) {
    if (k == GK_PRE_STEP || k == GK_PRE_LAST) {
        CK(launch_prefill_next(s, e->stream, pdl));
        e->launches++;
    }
    bool head = (k != GK_PRE_STEP), samp = (k == GK_DECODE);
    return forward_launches(e, s, head, samp, pdl, nullptr);
}

// run one step of kind k for slot s, through a captured graph when enabled
static int run_step(b2_engine* e, Slot& s, GraphKind k) {
    s.logits_view = s.logits;
    cudaGraphExec_t* ge = k == GK_DECODE ? &s.g_decode : k == GK_PRE_STEP ? &s.g_pre_step : k == GK_PRE_LAST ? &s.g_pre_last : &s.g_fwd;
    int* nn = k == GK_DECODE ? &s.n_decode_nodes : k == GK_PRE_STEP ? &s.n_pre_step_nodes : k == GK_PRE_LAST ? &s.n_pre_last_nodes : &s.n_fwd_nodes;
    if (!e->use_graph) return record_step(e, s, k, e->use_pdl);
    if (!*ge) {
        for (int attempt = 0; attempt < 2 && !*ge; attempt++) {
            bool pdl = e->use_pdl && attempt == 0;
            int64_t before = e->launches;
            cudaGraph_t g = nullptr;
            CK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
            int rc = record_step(e, s, k, pdl);
            cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
            *nn = (int)(e->launches - before);
            e->launches = before;
            if (rc == B2_OK && ce == cudaSuccess && g) {
                cudaGraphExec_t x = nullptr;
                ce = cudaGraphInstantiate(&x, g, 0);
                if (ce == cudaSuccess) *ge = x;
            }
            if (g) cudaGraphDestroy(g);
            if (!*ge) {
                cudaGetLastError();
                if (!pdl) {            // graphs unusable here: fall back to direct launches for good
                    e->use_graph = false;
                    return record_step(e, s, k, e->use_pdl);
                }
                e->use_pdl = false;    // PDL edges rejected by capture/instantiate: retry without
            }
        }
    }
    CK(cudaGraphLaunch(*ge, e->stream));
    e->launches += *nn;
    return B2_OK;
}

static int check_slot(b2_engine* e, int slot) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    if (!e->finalized) return fail(B2_ERR_STATE, "engine not finalized");
    if (slot < 0 || slot >= (int)e->slots.size()) return fail(B2_ERR_ARG, "slot out of range");
    return B2_OK;
}

extern "C" int b2_seq_set_blocks(b2_engine* e, int slot, const int32_t* blocks, int n) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    if (n < 0 || n > e->max_blocks) return fail(B2_ERR_ARG, "too many KV pages for n_ctx");
    for (int i = 0; i < n; i++)
        if (blocks[i] < 0 || blocks[i] >= e->cfg.kv_num_blocks) return fail(B2_ERR_ARG, "KV page id out of range");
    CK(cudaSetDevice(e->cfg.device));
    if (n) CK(cudaMemcpyAsync(e->slots[slot].block_table, blocks, sizeof(int) * n, cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));   // `blocks` is caller memory
    return B2_OK;
}

extern "C" int b2_seq_begin(b2_engine* e, int slot, int n_past, const int32_t* hist, int n_hist,
                            const b2_sampling* sp, const double* uni, int n_uni) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    if (!sp) return fail(B2_ERR_ARG, "null sampling");
    if (n_past < 0 || n_past >= e->cfg.n_ctx) return fail(B2_ERR_ARG, "n_past out of range");
    if (n_hist < 0 || n_hist > HISTORY_LEN) return fail(B2_ERR_ARG, "history longer than 64");
    if (n_uni < 0 || n_uni > 8192) return fail(B2_ERR_ARG, "at most 8192 uniforms");
    if (sp->temperature > 0.f && (sp->top_k < 1 || sp->top_k > SAMPLER_MAX_K))
        return fail(B2_ERR_UNSUPPORTED, "top_k must be in 1..1024 on the device sampler");
    CK(cudaSetDevice(e->cfg.device));
    Slot& s = e->slots[slot];
    SamplerState h = {};
    for (int i = 0; i < n_hist; i++) h.history[i] = hist[i];
    h.n_hist = n_hist;
    SampleParams p = {};
    p.temperature = sp->temperature; p.top_p = sp->top_p; p.min_p = sp->min_p;
    p.repeat_penalty = sp->repeat_penalty; p.freq_penalty = sp->frequency_penalty;
    p.presence_penalty = sp->presence_penalty; p.top_k = sp->top_k;
    p.penalty_last_n = sp->penalty_last_n;
    int ints[8] = {0, n_past, 0, 0, 0, n_past, 0, 0};
    CK(cudaMemcpyAsync(s.sst, &h, sizeof(h), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(s.sp, &p, sizeof(p), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(s.ints, ints, sizeof(ints), cudaMemcpyHostToDevice, e->stream));
    if (n_uni) CK(cudaMemcpyAsync(s.uniforms, uni, sizeof(double) * n_uni, cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    s.n_past = n_past;
    s.n_sampled = 0;
    return B2_OK;
}

constexpr int PF_T = 512;    // tokens per GEMM chunk (TMEM columns of one CTA)

static int ensure_prefill_workspace(b2_engine* e) {
    if (e->pf_x) return B2_OK;
    const b2_config& c = e->cfg;
    const size_t qd = (size_t)c.n_head * c.head_dim, kvd = (size_t)c.n_head_kv * c.head_dim;
    // partial sums: up to 8 K splits of the widest-but-one outputs, 2 of gate|up
    size_t wide = std::max<size_t>({(size_t)2 * 2 * c.n_ff, (size_t)8 * (qd + 2 * kvd), (size_t)8 * c.n_embd});
    e->pf_part_floats = PF_T * wide;
    CK(cudaMalloc(&e->pf_x, PF_T * (size_t)c.n_embd * 4));
    CK(cudaMalloc(&e->pf_part, e->pf_part_floats * 4));
    CK(cudaMalloc(&e->pf_xn, PF_T * (size_t)c.n_embd * 2));
    CK(cudaMalloc(&e->pf_qh, PF_T * qd * 2));
    CK(cudaMalloc(&e->pf_att, PF_T * qd * 2));
    CK(cudaMalloc(&e->pf_h, PF_T * (size_t)c.n_ff * 2));
    // rows [T, T_pad) of the f16 operands are read by the MMA (their columns are dropped): keep them finite
    CK(cudaMemsetAsync(e->pf_xn, 0, PF_T * (size_t)c.n_embd * 2, e->stream));
    CK(cudaMemsetAsync(e->pf_att, 0, PF_T * qd * 2, e->stream));
    CK(cudaMemsetAsync(e->pf_h, 0, PF_T * (size_t)c.n_ff * 2, e->stream));
    return B2_OK;
}

// One GEMM over up to three matrices that share the input; returns the number of K-split partial sums in pf_part.
static int gemm_call(b2_engine* e, const DevTensor* const* ws, int nw, const __half* x, int T, int* n_parts, long* part_stride) {
    const uint64_t key = ((uint64_t)(uintptr_t)ws[0]->ptr << 10) ^ (uint64_t)T;     // the first matrix identifies the call site
    auto it = e->gemm_plans.find(key);
    if (it == e->gemm_plans.end()) {
        GemmParams p = {};
        p.nseg = nw; p.K = (int)ws[0]->ne0;
        int off = 0;
        for (int i = 0; i < nw; i++) {
            p.seg[i].w = ws[i]->ptr; p.seg[i].type = ws[i]->type; p.seg[i].n_rows = (int)ws[i]->ne1; p.seg[i].out_off = off;
            off += (int)ws[i]->ne1;
        }
        p.x = x; p.T = T; p.y = e->pf_part; p.ldy = off; p.split_stride = (long)T * off;
        p.err = kernel_error_word();
        const int max_splits = (int)std::min<size_t>(16, e->pf_part_floats / ((size_t)T * off));
        if (gemm_plan(p, e->n_sms, max_splits)) return fail(B2_ERR_UNSUPPORTED, "GEMM shape not supported");
        it = e->gemm_plans.emplace(key, p).first;
    }
    GemmParams& p = it->second;
    if (p.x != x || p.y != e->pf_part) return fail(B2_ERR_STATE, "GEMM plan cache: workspace moved");
    CK(gemm_launch(p, e->stream));
    e->launches++;
    *n_parts = p.n_splits; *part_stride = p.split_stride;
    return B2_OK;
}

// One chunk of T <= PF_T prompt tokens (device ids at `tokens`) at positions pos0.. through every layer:
// tensor-core GEMMs over all T tokens at once, so each weight is read and dequantised once per chunk.
struct ChainPdl {      // programmatic dependent launch for every kernel launched while this is alive
    bool prev;
    explicit ChainPdl(bool on) : prev(g_chain_pdl) { g_chain_pdl = on; }
    ~ChainPdl() { g_chain_pdl = prev; }
};

static int prefill_chunk(b2_engine* e, Slot& s, const int* tokens, int T, int pos0) {
    ChainPdl chain(e->chain_pdl && e->use_pdl);
    const b2_config& c = e->cfg;
    cudaStream_t st = e->stream;
    const int E = c.n_embd, F = c.n_ff, hd = c.head_dim, T_pad = (T + 15) / 16 * 16;
    int rc, np = 0;
    long ps = 0;
    CK(embed_rows_launch(e->tok_embd.type, e->tok_embd.ptr, E, tokens, T, e->pf_x, st));
    e->launches++;
    for (int l = 0; l < c.n_layer; l++) {
        Layer& L = e->layers[l];
        __half* kc = e->k_pool + (size_t)l * e->layer_kv_elems;
        __half* vc = e->v_pool + (size_t)l * e->layer_kv_elems;
        // x += ffn_down partials of the previous layer; xn = attn_norm(x)
        CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, (const float*)L.attn_norm.ptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));
        const DevTensor* qkv[3] = {&L.wq, &L.wk, &L.wv};
        if ((rc = gemm_call(e, qkv, 3, e->pf_xn, T, &np, &ps))) return rc;
        CK(rope_store_launch(e->pf_part, np, ps, T, pos0, c.n_head, c.n_head_kv, hd, c.kv_block_tokens, s.block_table, e->rope_freq,
                             e->pf_qh, kc, vc, st));
        CK(attn_prefill_launch(e->pf_qh, kc, vc, s.block_table, T, T_pad, pos0, c.n_head, c.n_head_kv, hd, c.kv_block_tokens, e->pf_att, st));
        const DevTensor* wo[1] = {&L.wo};
        if ((rc = gemm_call(e, wo, 1, e->pf_att, T, &np, &ps))) return rc;
        CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, (const float*)L.ffn_norm.ptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));
        const DevTensor* gu[2] = {&L.wgate, &L.wup};
        if ((rc = gemm_call(e, gu, 2, e->pf_xn, T, &np, &ps))) return rc;
        CK(silu_mul_half_launch(e->pf_part, np, ps, F, T, T_pad, e->pf_h, st));
        const DevTensor* wd[1] = {&L.wdown};
        if ((rc = gemm_call(e, wd, 1, e->pf_h, T, &np, &ps))) return rc;
        e->launches += 5;
    }
    CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, nullptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));   // last residual add
    e->launches++;
    return B2_OK;
}

extern "C" int b2_prefill(b2_engine* e, int slot, const int32_t* tokens, int n) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    if (n <= 0) return fail(B2_ERR_ARG, "empty prompt chunk");
    if (s.n_past + n > e->cfg.n_ctx) return fail(B2_ERR_ARG, "prompt does not fit n_ctx");
    for (int i = 0; i < n; i++)
        if (tokens[i] < 0 || tokens[i] >= e->cfg.n_vocab) return fail(B2_ERR_ARG, "token id out of range");
    CK(cudaSetDevice(e->cfg.device));
    int hdr[2] = {0, s.n_past};   // cursor, pos_base
    CK(cudaMemcpyAsync(s.prompt, tokens, sizeof(int) * n, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(s.ints + 4, hdr, sizeof(hdr), cudaMemcpyHostToDevice, e->stream));
    if (e->gemm_ok && n >= e->prefill_min) {
        // tensor-core path: all layers for up to PF_T tokens at a time, then the output head for the last token only
        if ((rc = ensure_prefill_workspace(e))) return rc;
        for (int c0 = 0; c0 < n; c0 += PF_T) {
            int T = n - c0 < PF_T ? n - c0 : PF_T;
            if ((rc = prefill_chunk(e, s, s.prompt + c0, T, s.n_past + c0))) return rc;
            if (c0 + T == n)
                CK(cudaMemcpyAsync(s.x, e->pf_x + (size_t)(T - 1) * e->cfg.n_embd, sizeof(float) * e->cfg.n_embd,
                                   cudaMemcpyDeviceToDevice, e->stream));
        }
        CK(prefill_finish_launch(s.prompt, n, s.n_past + n - 1, s.ints, s.sst, e->stream));
        MvParams h = e->p_head;
        h.x = s.x; h.y = s.logits;
        CK(mv_launch(h, e->n_ctas, e->stream, e->act2, false));
        s.logits_view = s.logits;
        e->launches += 2;
    } else {
        for (int i = 0; i < n; i++) {
            rc = run_step(e, s, i == n - 1 ? GK_PRE_LAST : GK_PRE_STEP);
            if (rc) return rc;
        }
    }
    CK(cudaStreamSynchronize(e->stream));
    if ((rc = check_kernel_error())) return rc;
    s.n_past += n;     // device pos stays at the last token; b2_sample / set_token move it on
    return B2_OK;
}

__global__ void set_ints_kernel(int* ints, int tok, int pos) { if (tok >= 0) ints[0] = tok; ints[1] = pos; }

extern "C" int b2_sample(b2_engine* e, int slot, int32_t* out) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    CK(cudaSetDevice(e->cfg.device));
    // sampler bumps pos: make it point at the last evaluated token first
    set_ints_kernel<<<1, 1, 0, e->stream>>>(s.ints, -1, s.n_past - 1);
    CK(sample_launch(s.logits, e->cfg.n_vocab, s.sp, s.sst, s.uniforms, s.ints + 3, s.ints + 2, s.ints + 0,
                     s.ints + 1, s.ring, e->cfg.n_ctx + 1, e->stream, false));
    e->launches += 2;
    CK(cudaMemcpyAsync(e->pinned, s.ints + 2, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    if (out) *out = e->pinned[0];
    s.n_sampled += 1;
    return B2_OK;
}

static int ensure_batch_workspace(b2_engine* e) {
    int rc = ensure_prefill_workspace(e); if (rc) return rc;
    if (e->bd_q) return B2_OK;
    const b2_config& c = e->cfg;
    const size_t qd = (size_t)c.n_head * c.head_dim;
    CK(cudaMalloc(&e->bd_q, BATCH_MAX * qd * 4));
    CK(cudaMalloc(&e->bd_att, BATCH_MAX * qd * 4));
    CK(cudaMalloc(&e->bd_scratch, BATCH_MAX * attn_scratch_floats(c.n_head, c.n_head_kv, c.head_dim, e->n_splits) * 4));
    CK(cudaMalloc(&e->bd_counters, sizeof(unsigned) * BATCH_MAX * c.n_head_kv));
    CK(cudaMemsetAsync(e->bd_counters, 0, sizeof(unsigned) * BATCH_MAX * c.n_head_kv, e->stream));
    CK(cudaMalloc(&e->bd_out, sizeof(int) * BATCH_MAX));
    return B2_OK;
}

// One decode step for b.n sequences at once: the weights are read (and dequantised) once for all of them by the
// tensor-core GEMM; attention and sampling run one CTA group per sequence from the per-slot device tables.
static int batch_step(b2_engine* e, const BatchIdx& b) {
    ChainPdl chain(e->chain_pdl && e->use_pdl);
    const b2_config& c = e->cfg;
    cudaStream_t st = e->stream;
    const int T = b.n, T_pad = (T + 15) / 16 * 16, E = c.n_embd, F = c.n_ff, hd = c.head_dim, qd = c.n_head * hd;
    int rc, np = 0;
    long ps = 0;
    CK(embed_batch_launch(e->tok_embd.type, e->tok_embd.ptr, E, e->slot_tab, b, e->pf_x, st));
    e->launches++;
    for (int l = 0; l < c.n_layer; l++) {
        Layer& L = e->layers[l];
        __half* kc = e->k_pool + (size_t)l * e->layer_kv_elems;
        __half* vc = e->v_pool + (size_t)l * e->layer_kv_elems;
        CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, (const float*)L.attn_norm.ptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));
        const DevTensor* qkv[3] = {&L.wq, &L.wk, &L.wv};
        if ((rc = gemm_call(e, qkv, 3, e->pf_xn, T, &np, &ps))) return rc;
        CK(rope_store_batch_launch(e->pf_part, np, ps, e->slot_tab, b, c.n_head, c.n_head_kv, hd, c.kv_block_tokens, e->rope_freq,
                                   e->bd_q, kc, vc, st));
        AttnParams a = {};
        a.q = e->bd_q; a.k_cache = kc; a.v_cache = vc; a.out = e->bd_att; a.scratch = e->bd_scratch; a.counters = e->bd_counters;
        a.n_head = c.n_head; a.n_kv = c.n_head_kv; a.head_dim = hd; a.block_tokens = c.kv_block_tokens;
        // enough CTAs to fill the GPU ~2x over, no more (8 sequences: 4 splits 4.36 ms/step, 8 splits 4.81): every extra split of a sequence re-pays the kernel's fixed latency
        int ns = e->n_splits;
        while (ns > 1 && (long)ns * c.n_head_kv * T > 2L * e->n_sms) ns >>= 1;
        if (e->batch_attn_splits > 0 && e->batch_attn_splits <= e->n_splits) ns = e->batch_attn_splits;
        a.n_splits = ns; a.scale = 1.0f / sqrtf((float)hd); a.err = kernel_error_word(); a.max_pages = e->max_blocks;
        if (e->attn_mma != 0) {      // the tensor-core kernel writes f16 in the o-projection GEMM's operand order itself
            a.out_h = e->pf_att; a.out_T_pad = T_pad;
            CK(attn_decode_mma_launch(a, e->slot_tab, &b, st, false));
        } else {
            CK(attn_decode_batch_launch(a, e->slot_tab, b, st));
            CK(to_half_blocked_launch(e->bd_att, T, T_pad, qd, e->pf_att, st));
        }
        const DevTensor* wo[1] = {&L.wo};
        if ((rc = gemm_call(e, wo, 1, e->pf_att, T, &np, &ps))) return rc;
        CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, (const float*)L.ffn_norm.ptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));
        const DevTensor* gu[2] = {&L.wgate, &L.wup};
        if ((rc = gemm_call(e, gu, 2, e->pf_xn, T, &np, &ps))) return rc;
        CK(silu_mul_half_launch(e->pf_part, np, ps, F, T, T_pad, e->pf_h, st));
        const DevTensor* wd[1] = {&L.wdown};
        if ((rc = gemm_call(e, wd, 1, e->pf_h, T, &np, &ps))) return rc;
        e->launches += e->attn_mma != 0 ? 5 : 6;
    }
    // x += last ffn_down; output norm; lm_head (one K range: the logits are complete rows); sample every sequence
    CK(resid_norm_half_launch(e->pf_x, e->pf_part, np, ps, (const float*)e->out_norm.ptr, c.rms_eps, E, T, T_pad, e->pf_xn, st));
    {
        const uint64_t key = (((uint64_t)(uintptr_t)e->output.ptr << 10) ^ (uint64_t)T) + 1;    // + 1: not a gemm_call site (one K range)
        auto it = e->gemm_plans.find(key);
        if (it == e->gemm_plans.end()) {
            GemmParams p = {};
            p.nseg = 1; p.K = E; p.seg[0].w = e->output.ptr; p.seg[0].type = e->output.type; p.seg[0].n_rows = c.n_vocab;
            p.x = e->pf_xn; p.T = T; p.y = e->pf_part; p.ldy = c.n_vocab; p.split_stride = 0; p.err = kernel_error_word();
            if (gemm_plan(p, e->n_sms, 1)) return fail(B2_ERR_UNSUPPORTED, "GEMM shape not supported");
            it = e->gemm_plans.emplace(key, p).first;
        }
        CK(gemm_launch(it->second, st));
    }
    CK(sample_batch_launch(e->pf_part, c.n_vocab, c.n_vocab, e->slot_tab, b, c.n_ctx + 1, e->bd_out, st));
    e->launches += 3;
    return B2_OK;
}

extern "C" int b2_decode(b2_engine* e, const int32_t* slots, int n, int32_t* out) {
    if (!e || !slots || n <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (!e->finalized) return fail(B2_ERR_STATE, "engine not finalized");
    std::lock_guard<std::mutex> lk(e->mu);
    if (n > (int)e->slots.size()) return fail(B2_ERR_ARG, "more slots than max_seqs");
    CK(cudaSetDevice(e->cfg.device));
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || slots[i] >= (int)e->slots.size()) return fail(B2_ERR_ARG, "slot out of range");
        if (e->slots[slots[i]].n_past + 1 > e->cfg.n_ctx) return fail(B2_ERR_ARG, "context window full");
        for (int j = 0; j < i; j++) if (slots[j] == slots[i]) return fail(B2_ERR_ARG, "slot listed twice");
    }
    const bool batched = e->gemm_ok && block_elems(e->output.type) != 1 && n >= e->batch_min &&
                         (size_t)n * e->cfg.n_vocab <= (e->pf_part_floats ? e->pf_part_floats : (size_t)-1);
    if (batched) {
        int rc = ensure_batch_workspace(e); if (rc) return rc;
        for (int i0 = 0; i0 < n; i0 += BATCH_MAX) {
            BatchIdx b = {};
            b.n = n - i0 < BATCH_MAX ? n - i0 : BATCH_MAX;
            for (int i = 0; i < b.n; i++) b.slot[i] = slots[i0 + i];
            if ((size_t)b.n * e->cfg.n_vocab > e->pf_part_floats) return fail(B2_ERR_UNSUPPORTED, "vocabulary too large for the batch workspace");
            if ((rc = batch_step(e, b))) return rc;
            // logits of this step stay in the batch output until the next GEMM call reuses the buffer
            for (int i = 0; i < b.n; i++) e->slots[b.slot[i]].logits_view = e->pf_part + (size_t)i * e->cfg.n_vocab;
            CK(cudaMemcpyAsync(e->pinned + i0, e->bd_out, sizeof(int) * b.n, cudaMemcpyDeviceToHost, e->stream));
        }
        for (int i = 0; i < n; i++) { Slot& s = e->slots[slots[i]]; s.n_past += 1; s.n_sampled += 1; }
    } else {
        for (int i = 0; i < n; i++) {
            Slot& s = e->slots[slots[i]];
            int rc = run_step(e, s, GK_DECODE);
            if (rc) return rc;
            CK(cudaMemcpyAsync(e->pinned + i, s.ints + 2, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
            s.n_past += 1;
            s.n_sampled += 1;
        }
    }
    CK(cudaStreamSynchronize(e->stream));
    { int rc2 = check_kernel_error(); if (rc2) return rc2; }
    if (out) for (int i = 0; i < n; i++) out[i] = e->pinned[i];
    return B2_OK;
}

extern "C" int b2_decode_n(b2_engine* e, int slot, int n, int32_t* out) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    if (n <= 0) return fail(B2_ERR_ARG, "n must be positive");
    if (s.n_past + n > e->cfg.n_ctx) return fail(B2_ERR_ARG, "context window full");
    CK(cudaSetDevice(e->cfg.device));
    for (int i = 0; i < n; i++) {
        if (e->tl && i == n - 1) {      // debugging aid: record the last step of the call
            std::vector<unsigned long long> init(4 * 1024, 0ull);
            for (int k = 0; k < 1024; k++) init[4 * k] = ~0ull;
            CK(cudaMemcpyAsync(e->tl, init.data(), init.size() * 8, cudaMemcpyHostToDevice, e->stream));
            CK(cudaStreamSynchronize(e->stream));
        }
        rc = run_step(e, s, GK_DECODE);
        if (rc) return rc;
    }
    if (e->tl) {
        CK(cudaStreamSynchronize(e->stream));
        std::vector<unsigned long long> t(4 * 1024);
        CK(cudaMemcpy(t.data(), e->tl, t.size() * 8, cudaMemcpyDeviceToHost));
        const int nk = e->tl_n;
        const char* names[5] = {"qkv", "attn", "o", "gate_up", "down"};
        double sum[6][4] = {};
        for (int k = 0; k < nk; k++) {
            const unsigned long long* r = &t[4 * k];
            const int kind = k < nk - 1 ? k % 5 : 5;
            const double prev_end = k ? (double)t[4 * (k - 1) + 3] : (double)r[0];
            sum[kind][0] += ((double)r[0] - prev_end) * 1e-3;                 // first CTA in, relative to the previous kernel's end
            sum[kind][1] += r[1] ? ((double)r[1] - (double)r[0]) * 1e-3 : 0;  // in -> past the dependency wait
            sum[kind][2] += r[2] ? ((double)r[2] - (double)r[1]) * 1e-3 : 0;  // wait -> prologue done
            sum[kind][3] += ((double)r[3] - (double)r[0]) * 1e-3;             // in -> out
        }
        const int L = e->cfg.n_layer;
        fprintf(stderr, "[b2 timeline] one decode step, us per launch (mean over layers): start-prev_end | in->wait | wait->prologue | in->out\n");
        for (int kind = 0; kind < 5; kind++)
            fprintf(stderr, "[b2 timeline] %-8s %7.2f %7.2f %7.2f %7.2f\n", names[kind], sum[kind][0] / L, sum[kind][1] / L, sum[kind][2] / L, sum[kind][3] / L);
        fprintf(stderr, "[b2 timeline] %-8s %7.2f %7.2f %7.2f %7.2f\n", "lm_head", sum[5][0], sum[5][1], sum[5][2], sum[5][3]);
        fprintf(stderr, "[b2 timeline] step span %.1f us\n", ((double)t[4 * (nk - 1) + 3] - (double)t[0]) * 1e-3);
    }
    // step i forwards the token at position n_past+i and its sample lands at ring[n_past+i+1]
    CK(cudaMemcpyAsync(e->pinned, s.ring + s.n_past + 1, sizeof(int) * n, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    if ((rc = check_kernel_error())) return rc;
    if (out) memcpy(out, e->pinned, sizeof(int) * n);
    s.n_past += n;
    s.n_sampled += n;
    return B2_OK;
}

extern "C" int b2_seq_rewind(b2_engine* e, int slot, int n_keep, int32_t pending, const int32_t* hist, int n_hist) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    if (n_keep < 0 || n_keep > s.n_past) return fail(B2_ERR_ARG, "n_keep out of range");
    if (pending < 0 || pending >= e->cfg.n_vocab) return fail(B2_ERR_ARG, "token id out of range");
    if (n_hist < 0 || n_hist > HISTORY_LEN) return fail(B2_ERR_ARG, "history longer than 64");
    CK(cudaSetDevice(e->cfg.device));
    int dropped = s.n_past - n_keep;
    SamplerState h = {};
    for (int i = 0; i < n_hist; i++) h.history[i] = hist[i];
    h.n_hist = n_hist;
    s.n_sampled = s.n_sampled > dropped ? s.n_sampled - dropped : 0;
    int ints[4] = {pending, n_keep, pending, s.n_sampled};
    CK(cudaMemcpyAsync(s.sst, &h, sizeof(h), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(s.ints, ints, sizeof(ints), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    s.n_past = n_keep;
    return B2_OK;
}

extern "C" int b2_seq_set_token(b2_engine* e, int slot, int32_t token) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    if (token < 0 || token >= e->cfg.n_vocab) return fail(B2_ERR_ARG, "token id out of range");
    CK(cudaSetDevice(e->cfg.device));
    Slot& s = e->slots[slot];
    set_ints_kernel<<<1, 1, 0, e->stream>>>(s.ints, token, s.n_past);
    e->launches++;
    CK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_forward_pending(b2_engine* e, int slot) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    if (s.n_past + 1 > e->cfg.n_ctx) return fail(B2_ERR_ARG, "context window full");
    CK(cudaSetDevice(e->cfg.device));
    set_ints_kernel<<<1, 1, 0, e->stream>>>(s.ints, -1, s.n_past);
    e->launches++;
    rc = run_step(e, s, GK_FWD);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    if ((rc = check_kernel_error())) return rc;
    s.n_past += 1;
    return B2_OK;
}

extern "C" int b2_get_logits(b2_engine* e, int slot, float* out) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    CK(cudaSetDevice(e->cfg.device));
    const Slot& s = e->slots[slot];
    CK(cudaMemcpyAsync(out, s.logits_view ? s.logits_view : s.logits, sizeof(float) * e->cfg.n_vocab, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return B2_OK;
}

extern "C" int b2_synchronize(b2_engine* e) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaStreamSynchronize(e->stream));
    return B2_OK;
}

extern "C" int b2_timer_start(b2_engine* e) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaStreamSynchronize(e->stream));
    CK(cudaEventRecord(e->ev0, e->stream));
    return B2_OK;
}
extern "C" int b2_timer_stop(b2_engine* e, float* ms) {
    if (!e || !ms) return fail(B2_ERR_ARG, "null argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaEventRecord(e->ev1, e->stream));
    CK(cudaEventSynchronize(e->ev1));
    CK(cudaEventElapsedTime(ms, e->ev0, e->ev1));
    return B2_OK;
}
extern "C" int b2_kernel_launch_count(b2_engine* e, int64_t* n) {
    if (!e || !n) return fail(B2_ERR_ARG, "null argument");
    *n = e->launches;
    return B2_OK;
}
extern "C" int b2_engine_mode(b2_engine* e, int* graph, int* pdl, int* act_planes, int* n_ctas, int* n_splits) {
    if (!e) return fail(B2_ERR_ARG, "null engine");
    if (graph) *graph = e->use_graph; if (pdl) *pdl = e->use_pdl; if (act_planes) *act_planes = e->act2 ? 2 : 1;
    if (n_ctas) *n_ctas = e->n_ctas; if (n_splits) *n_splits = e->n_splits;
    return B2_OK;
}

static int64_t seg_bytes(const MvParams& p) {
    int64_t b = 0;
    for (int i = 0; i < p.nseg; i++) b += (int64_t)p.seg[i].n_rows * p.seg[i].row_bytes;
    return b;
}

extern "C" int b2_profile_decode_step(b2_engine* e, int slot, float* ms, int32_t* nl, int64_t* bytes) {
    int rc = check_slot(e, slot); if (rc) return rc;
    std::lock_guard<std::mutex> lk(e->mu);
    Slot& s = e->slots[slot];
    if (s.n_past + 1 > e->cfg.n_ctx) return fail(B2_ERR_ARG, "context window full");
    CK(cudaSetDevice(e->cfg.device));
    KindTimer kt;
    rc = forward_launches(e, s, true, true, false, &kt);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    const b2_config& c = e->cfg;
    for (int k = 0; k < 8; k++) {
        float tot = 0.f;
        for (auto& pr : kt.ev[k]) {
            float t = 0.f;
            cudaEventElapsedTime(&t, pr.first, pr.second);
            tot += t;
            cudaEventDestroy(pr.first); cudaEventDestroy(pr.second);
        }
        ms[k] = tot; nl[k] = (int)kt.ev[k].size(); bytes[k] = 0;
    }
    int be = block_elems(e->tok_embd.type), bb = block_bytes(e->tok_embd.type);
    bytes[0] = (int64_t)c.n_embd / be * bb;
    for (auto& L : e->layers) {
        bytes[1] += seg_bytes(L.p_qkv); bytes[3] += seg_bytes(L.p_o);
        bytes[4] += L.gu_fused ? seg_bytes(L.p_gu) : seg_bytes(L.p_gate) + seg_bytes(L.p_up); bytes[5] += seg_bytes(L.p_down);
    }
    bytes[2] = (int64_t)c.n_layer * 2 * c.n_head_kv * c.head_dim * 2 * (s.n_past + 1);
    bytes[6] = seg_bytes(e->p_head);
    bytes[7] = (int64_t)c.n_vocab * 4;
    s.n_past += 1;
    return B2_OK;
}

// ------------------------------------------------------------------------------------------
// single-operator entry points (host buffers in, host buffers out)
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
    template <typename T> T* as() { return (T*)p; }
};

extern "C" int b2_op_dequantize(int device, int type, const void* blocks, int64_t n, float* out) {
    if (!blocks || !out || n < 0) return fail(B2_ERR_ARG, "bad arguments");
    int be = block_elems(type), bb = block_bytes(type);
    if (!bb) return fail(B2_ERR_UNSUPPORTED, "unsupported ggml type");
    if (n % be) return fail(B2_ERR_ARG, "n_elems is not a whole number of blocks");
    if (n == 0) return B2_OK;
    CK(cudaSetDevice(device));
    DevBuf in, o;
    CK(in.alloc((size_t)(n / be * bb))); CK(o.alloc((size_t)n * 4));
    CK(cudaMemcpy(in.p, blocks, (size_t)(n / be * bb), cudaMemcpyHostToDevice));
    CK(dequantize_launch(type, in.as<uint8_t>(), n, o.as<float>(), 0));
    CK(cudaMemcpy(out, o.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

static int op_matvec_common(int device, MvParams& p, const void* w0, const void* w1, int n_rows, int K,
                            const float* x, const float* nw, const float* resid, float* y, int planes, int n_ctas) {
    if (n_ctas <= 0) { int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device); n_ctas = sms; }
    CK(mv_init());
    int rc = mv_plan(p, n_ctas);
    if (rc) return fail(B2_ERR_UNSUPPORTED, "mat-vec shape/type not supported (plan code " + std::to_string(rc) + ")");
    DevBuf dw0, dw1, dx, dn, dy;
    size_t wb0 = (size_t)n_rows * p.seg[0].row_bytes;
    CK(dw0.alloc(wb0 + 64)); CK(cudaMemcpy(dw0.p, w0, wb0, cudaMemcpyHostToDevice));
    p.seg[0].w = dw0.as<uint8_t>();
    if (w1) {
        size_t wb1 = (size_t)n_rows * p.seg[1].row_bytes;
        CK(dw1.alloc(wb1 + 64)); CK(cudaMemcpy(dw1.p, w1, wb1, cudaMemcpyHostToDevice));
        p.seg[1].w = dw1.as<uint8_t>();
    }
    CK(dx.alloc((size_t)K * 4)); CK(cudaMemcpy(dx.p, x, (size_t)K * 4, cudaMemcpyHostToDevice));
    p.x = dx.as<float>();
    if (nw) { CK(dn.alloc((size_t)K * 4)); CK(cudaMemcpy(dn.p, nw, (size_t)K * 4, cudaMemcpyHostToDevice)); p.norm_w = dn.as<float>(); }
    CK(dy.alloc((size_t)n_rows * 4));
    if (resid) CK(cudaMemcpy(dy.p, resid, (size_t)n_rows * 4, cudaMemcpyHostToDevice));
    p.y = dy.as<float>(); p.resid = dy.as<float>();
    p.err = kernel_error_word();
    if ((rc = mv_plan(p, n_ctas))) return fail(B2_ERR_UNSUPPORTED, "mat-vec plan failed");     // again, now that the weights have their addresses (tensor maps)
    CK(mv_launch(p, n_ctas, 0, planes != 1, false));
    CK(cudaDeviceSynchronize());
    { int rc2 = check_kernel_error(); if (rc2) return rc2; }

    CK(cudaMemcpy(y, dy.p, (size_t)n_rows * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

extern "C" int b2_op_matvec(int device, int type, const void* w, int n_rows, int K, const float* x,
                            const float* nw, float eps, const float* resid, float* y, int planes, int n_ctas) {
    if (!w || !x || !y || n_rows <= 0 || K <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (n_rows % 2) return fail(B2_ERR_ARG, "n_rows must be even");
    CK(cudaSetDevice(device));
    MvParams p = {};
    p.nseg = 1; p.seg[0].type = type; p.seg[0].n_rows = n_rows; p.K = K;
    p.mode = resid ? MV_RESID : MV_STORE; p.eps = eps;
    return op_matvec_common(device, p, w, nullptr, n_rows, K, x, nw, resid, y, planes, n_ctas);
}

extern "C" int b2_op_gate_up(int device, int tg, const void* wg, int tu, const void* wu, int n_rows, int K,
                             const float* x, const float* nw, float eps, float* y, int n_ctas) {
    if (!wg || !wu || !x || !y || n_rows <= 0 || K <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (n_rows % 2) return fail(B2_ERR_ARG, "n_rows must be even");
    CK(cudaSetDevice(device));
    if (tg != tu) {     // what the engine does for a mixed pair: two plain mat-vecs with the fused norm, then SiLU-multiply
        std::vector<float> g(n_rows);
        int rc = b2_op_matvec(device, tg, wg, n_rows, K, x, nw, eps, nullptr, g.data(), 2, n_ctas);
        if (rc) return rc;
        if ((rc = b2_op_matvec(device, tu, wu, n_rows, K, x, nw, eps, nullptr, y, 2, n_ctas))) return rc;
        DevBuf dg, dh;
        CK(dg.alloc((size_t)n_rows * 4)); CK(dh.alloc((size_t)n_rows * 4));
        CK(cudaMemcpy(dg.p, g.data(), (size_t)n_rows * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dh.p, y, (size_t)n_rows * 4, cudaMemcpyHostToDevice));
        CK(silu_mul_launch(dg.as<float>(), dh.as<float>(), n_rows, 0, false));
        CK(cudaMemcpy(y, dh.p, (size_t)n_rows * 4, cudaMemcpyDeviceToHost));
        return B2_OK;
    }
    MvParams p = {};
    p.nseg = 2; p.seg[0].type = tg; p.seg[0].n_rows = n_rows; p.seg[1].type = tu; p.seg[1].n_rows = n_rows;
    p.K = K; p.mode = MV_GATEUP; p.eps = eps;
    return op_matvec_common(device, p, wg, wu, n_rows, K, x, nw, nullptr, y, 2, n_ctas);
}

// ------------------------------------------------------------------------------------------
// measurement aid: streaming rate of the mat-vec kernel alone, on synthetic weight bytes
__global__ void fill_hash_kernel(uint32_t* p, size_t n_words, uint32_t seed) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t v = (uint32_t)i * 2654435761u + seed;
        v ^= v >> 15; v *= 2246822519u; v ^= v >> 13; v *= 3266489917u; v ^= v >> 16;
        p[i] = v & 0x3BFF3BFFu;        // clears the top exponent bit of every f16 that may sit here: scale fields stay finite
    }
}

extern "C" int b2_bench_matvec(int device, int mode, int type, int n_rows, int K, int n_copies, int iters, int n_ctas,
                               int planes, int pdl, float* us_out) {
    if (!us_out || n_rows <= 0 || K <= 0 || n_copies <= 0 || iters <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (mode != MV_STORE && mode != MV_RESID && mode != MV_GATEUP) return fail(B2_ERR_ARG, "mode must be 0, 1 or 3");
    CK(cudaSetDevice(device));
    CK(mv_init());
    if (n_ctas <= 0) { int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device); n_ctas = sms * mv_ctas_per_sm(); }
    MvParams p = {};
    const int nmat = mode == MV_GATEUP ? 2 : 1;
    p.nseg = nmat; p.K = K; p.mode = mode;
    for (int m = 0; m < nmat; m++) { p.seg[m].type = type; p.seg[m].n_rows = n_rows; }
    int rc = mv_plan(p, n_ctas);
    if (rc) return fail(B2_ERR_UNSUPPORTED, "mat-vec shape/type not supported (plan code " + std::to_string(rc) + ")");
    const size_t wb = (size_t)n_rows * p.seg[0].row_bytes, wpad = (wb + 64 + 255) / 256 * 256;
    DevBuf dw, dx, dy;
    CK(dw.alloc(wpad * nmat * n_copies)); CK(dx.alloc((size_t)K * 4)); CK(dy.alloc((size_t)n_rows * 4));
    fill_hash_kernel<<<1184, 256>>>(dw.as<uint32_t>(), wpad * nmat * n_copies / 4, 12345u);
    fill_hash_kernel<<<64, 256>>>(dx.as<uint32_t>(), (size_t)K, 777u);       // finite floats of mixed magnitude
    CK(cudaMemset(dy.p, 0, (size_t)n_rows * 4));
    CK(cudaGetLastError());
    p.x = dx.as<float>(); p.y = dy.as<float>(); p.resid = dy.as<float>(); p.err = kernel_error_word();
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaDeviceSynchronize());
    float ms = 0.f;
    std::vector<MvParams> pv((size_t)n_copies, p);   // one planned parameter block per weight copy (the tensor maps hold the addresses)
    for (int c2 = 0; c2 < n_copies; c2++) {
        const uint8_t* base = dw.as<uint8_t>() + (size_t)c2 * wpad * nmat;
        for (int m = 0; m < nmat; m++) pv[c2].seg[m].w = base + (size_t)m * wpad;
        if (mv_plan(pv[c2], n_ctas)) return fail(B2_ERR_UNSUPPORTED, "mat-vec plan failed");
    }
    for (int pass = 0; pass < 2; pass++) {          // pass 0 warms up
        if (pass) CK(cudaEventRecord(e0, st));
        for (int i = 0; i < iters; i++) CK(mv_launch(pv[i % n_copies], n_ctas, st, planes != 1, pdl != 0));
        if (pass) CK(cudaEventRecord(e1, st));
        CK(cudaStreamSynchronize(st));
    }
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    if (int rc2 = check_kernel_error()) return rc2;
    *us_out = ms * 1000.f / (float)iters;
    return B2_OK;
}

__global__ void sum_parts_kernel(const float* parts, int n_parts, long stride, const float* y_in, long n, float* y) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = y_in ? y_in[i] : 0.f;
    for (int s = 0; s < n_parts; s++) v += parts[(size_t)s * stride + i];
    y[i] = v;
}

extern "C" int b2_op_gemm(int device, int type, const void* w, int n_rows, int K, const float* x, int T,
                          const float* y_in, float* y, int n_splits) {
    if (!w || !x || !y || n_rows <= 0 || K <= 0 || T <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (T > 512) return fail(B2_ERR_ARG, "at most 512 tokens per call");
    int be = block_elems(type), bb = block_bytes(type);
    if (be == 1 || !bb) return fail(B2_ERR_UNSUPPORTED, "weight type must be a quantised block format");
    if (K % 256) return fail(B2_ERR_ARG, "K must be a multiple of 256");
    if (n_splits < 0 || n_splits > 16) return fail(B2_ERR_ARG, "bad n_splits");
    CK(cudaSetDevice(device));
    CK(gemm_init());
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t wbytes = (size_t)n_rows * (K / be) * bb, out_n = (size_t)T * n_rows;
    DevBuf dw, dx, dxh, dpart, dyin, dy;
    CK(dw.alloc(wbytes + 64));
    CK(cudaMemcpy(dw.p, w, wbytes, cudaMemcpyHostToDevice));
    GemmParams p = {};
    p.nseg = 1; p.K = K; p.seg[0].w = dw.as<uint8_t>(); p.seg[0].type = type; p.seg[0].n_rows = n_rows; p.seg[0].out_off = 0;
    p.T = T; p.ldy = n_rows; p.split_stride = (long)out_n; p.err = kernel_error_word();
    if (gemm_plan(p, sms, 16)) return fail(B2_ERR_UNSUPPORTED, "GEMM shape not supported");
    if (n_splits > 0) p.n_splits = n_splits;
    if (p.n_splits > K / 256) p.n_splits = K / 256;
    CK(dx.alloc((size_t)T * K * 4)); CK(dxh.alloc((size_t)p.T_pad * K * 2));
    CK(dpart.alloc(out_n * 4 * p.n_splits)); CK(dy.alloc(out_n * 4));
    CK(cudaMemcpy(dx.p, x, (size_t)T * K * 4, cudaMemcpyHostToDevice));
    CK(to_half_blocked_launch(dx.as<float>(), T, p.T_pad, K, dxh.as<__half>(), 0));
    if (y_in) { CK(dyin.alloc(out_n * 4)); CK(cudaMemcpy(dyin.p, y_in, out_n * 4, cudaMemcpyHostToDevice)); }
    p.x = dxh.as<__half>(); p.y = dpart.as<float>();
    DevBuf ddbg;
    if (getenv("B2_GEMM_DBG")) { CK(ddbg.alloc(64 * 8 * 8)); CK(cudaMemset(ddbg.p, 0, 64 * 8 * 8)); p.dbg = ddbg.as<long long>(); }
    CK(gemm_launch(p, 0));
    sum_parts_kernel<<<(unsigned)((out_n + 255) / 256), 256>>>(dpart.as<float>(), p.n_splits, (long)out_n,
                                                               y_in ? dyin.as<float>() : nullptr, (long)out_n, dy.as<float>());
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    if (int rc = check_kernel_error()) return rc;
    CK(cudaMemcpy(y, dy.p, out_n * 4, cudaMemcpyDeviceToHost));
    if (p.dbg) {      // debugging aid: pipeline timeline of CTA 0, cycles relative to the first record
        std::vector<long long> t(64 * 8);
        cudaMemcpy(t.data(), ddbg.p, t.size() * 8, cudaMemcpyDeviceToHost);
        long long t0 = t[0];
        fprintf(stderr, "[b2 gemm] stages A %d B %d raw %d splits %d; per K block: deq(wait_empty, got, arrived) mma(b_ok, a_ok, committed) raw(wait, got)\n",
                p.n_stages, p.n_bstages, p.n_raw, p.n_splits);
        for (int i = 0; i < 24; i++)
            fprintf(stderr, "[b2 gemm] kb %2d  deq %7lld %7lld %7lld   mma %7lld %7lld %7lld   raw %7lld %7lld\n", i, t[i * 8] - t0, t[i * 8 + 1] - t0,
                    t[i * 8 + 2] - t0, t[i * 8 + 3] - t0, t[i * 8 + 4] - t0, t[i * 8 + 5] - t0, t[i * 8 + 6] ? t[i * 8 + 6] - t0 : 0, t[i * 8 + 7] ? t[i * 8 + 7] - t0 : 0);
    }
    return B2_OK;
}

extern "C" int b2_op_rmsnorm(int device, const float* x, const float* w, float eps, int K, int rows, float* y) {
    if (!x || !w || !y || K <= 0 || rows < 0) return fail(B2_ERR_ARG, "bad arguments");
    if (rows == 0) return B2_OK;
    CK(cudaSetDevice(device));
    DevBuf dx, dw, dy;
    size_t n = (size_t)K * rows * 4;
    CK(dx.alloc(n)); CK(dw.alloc((size_t)K * 4)); CK(dy.alloc(n));
    CK(cudaMemcpy(dx.p, x, n, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw.p, w, (size_t)K * 4, cudaMemcpyHostToDevice));
    CK(rmsnorm_launch(dx.as<float>(), dw.as<float>(), eps, K, rows, dy.as<float>(), 0));
    CK(cudaMemcpy(y, dy.p, n, cudaMemcpyDeviceToHost));
    return B2_OK;
}

extern "C" int b2_op_attention_decode(int device, const float* q, const uint16_t* k, const uint16_t* v, int T,
                                      int n_head, int n_kv, int hd, int bt, const int32_t* blocks, int nblk_total,
                                      int n_splits, float* out) {
    if (!q || !k || !v || !blocks || !out || T <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (hd != 64 && hd != 128) return fail(B2_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    if (n_head % n_kv || n_head / n_kv > 8) return fail(B2_ERR_UNSUPPORTED, "n_head / n_kv must be an integer <= 8");
    if (bt <= 0 || bt % 16) return fail(B2_ERR_ARG, "block_tokens must be a multiple of 16");
    int need = (T + bt - 1) / bt;
    const bool use_mma = n_splits < 0;      // negative: the tensor-core kernel with -n_splits splits
    if (use_mma) n_splits = -n_splits;
    if (n_splits < 1) n_splits = 1;
    for (int i = 0; i < need; i++) if (blocks[i] < 0 || blocks[i] >= nblk_total) return fail(B2_ERR_ARG, "page id out of range");
    CK(cudaSetDevice(device));
    CK(attn_init());
    size_t pool_elems = (size_t)nblk_total * n_kv * bt * hd;
    std::vector<uint16_t> hk(pool_elems, 0), hv(pool_elems, 0);
    for (int t = 0; t < T; t++) {
        int page = blocks[t / bt], slot = t % bt;
        for (int h = 0; h < n_kv; h++) {
            size_t dst = (((size_t)page * n_kv + h) * bt + slot) * hd, src = ((size_t)t * n_kv + h) * hd;
            memcpy(&hk[dst], &k[src], hd * 2); memcpy(&hv[dst], &v[src], hd * 2);
        }
    }
    DevBuf dk, dv, dq, dout, dbt, dpos, dscr, dcnt;
    CK(dk.alloc(pool_elems * 2)); CK(dv.alloc(pool_elems * 2));
    CK(cudaMemcpy(dk.p, hk.data(), pool_elems * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dv.p, hv.data(), pool_elems * 2, cudaMemcpyHostToDevice));
    CK(dq.alloc((size_t)n_head * hd * 4)); CK(cudaMemcpy(dq.p, q, (size_t)n_head * hd * 4, cudaMemcpyHostToDevice));
    CK(dout.alloc((size_t)n_head * hd * 4));
    CK(dbt.alloc((size_t)need * 4)); CK(cudaMemcpy(dbt.p, blocks, (size_t)need * 4, cudaMemcpyHostToDevice));
    int pos = T - 1;
    CK(dpos.alloc(4)); CK(cudaMemcpy(dpos.p, &pos, 4, cudaMemcpyHostToDevice));
    CK(dscr.alloc(attn_scratch_floats(n_head, n_kv, hd, n_splits) * 4));
    CK(dcnt.alloc((size_t)n_kv * 4)); CK(cudaMemset(dcnt.p, 0, (size_t)n_kv * 4));
    AttnParams a = {};
    a.q = dq.as<float>(); a.k_cache = dk.as<__half>(); a.v_cache = dv.as<__half>(); a.block_table = dbt.as<int>();
    a.pos_ptr = dpos.as<int>(); a.out = dout.as<float>(); a.scratch = dscr.as<float>(); a.counters = dcnt.as<unsigned>();
    a.n_head = n_head; a.n_kv = n_kv; a.head_dim = hd; a.block_tokens = bt; a.n_splits = n_splits;
    a.scale = 1.0f / sqrtf((float)hd); a.err = kernel_error_word(); a.max_pages = need;
    if (use_mma) { CK(attn_mma_init()); CK(attn_decode_mma_launch(a, nullptr, nullptr, 0, false)); }
    else CK(attn_decode_launch(a, 0, false));
    CK(cudaDeviceSynchronize());
    { int rc2 = check_kernel_error(); if (rc2) return rc2; }
    CK(cudaMemcpy(out, dout.p, (size_t)n_head * hd * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

extern "C" int b2_op_sample(int device, const float* logits, int V, const b2_sampling* sp, const int32_t* hist,
                            int n_hist, double u, int32_t* tok) {
    if (!logits || !sp || !tok || V <= 0) return fail(B2_ERR_ARG, "bad arguments");
    if (n_hist < 0 || n_hist > HISTORY_LEN) return fail(B2_ERR_ARG, "history longer than 64");
    if (sp->temperature > 0.f && (sp->top_k < 1 || sp->top_k > SAMPLER_MAX_K))
        return fail(B2_ERR_UNSUPPORTED, "top_k must be in 1..1024 on the device sampler");
    CK(cudaSetDevice(device));
    DevBuf dl, dsp, dst, du, di;
    CK(dl.alloc((size_t)V * 4)); CK(cudaMemcpy(dl.p, logits, (size_t)V * 4, cudaMemcpyHostToDevice));
    SampleParams p = {};
    p.temperature = sp->temperature; p.top_p = sp->top_p; p.min_p = sp->min_p; p.repeat_penalty = sp->repeat_penalty;
    p.freq_penalty = sp->frequency_penalty; p.presence_penalty = sp->presence_penalty; p.top_k = sp->top_k;
    p.penalty_last_n = sp->penalty_last_n;
    CK(dsp.alloc(sizeof(p))); CK(cudaMemcpy(dsp.p, &p, sizeof(p), cudaMemcpyHostToDevice));
    SamplerState h = {};
    for (int i = 0; i < n_hist; i++) h.history[i] = hist[i];
    h.n_hist = n_hist;
    CK(dst.alloc(sizeof(h))); CK(cudaMemcpy(dst.p, &h, sizeof(h), cudaMemcpyHostToDevice));
    std::vector<double> uu(8192, u);
    CK(du.alloc(sizeof(double) * 8192)); CK(cudaMemcpy(du.p, uu.data(), sizeof(double) * 8192, cudaMemcpyHostToDevice));
    CK(di.alloc(sizeof(int) * 8)); CK(cudaMemset(di.p, 0, sizeof(int) * 8));
    int* ints = di.as<int>();
    CK(sample_launch(dl.as<float>(), V, dsp.as<SampleParams>(), dst.as<SamplerState>(), du.as<double>(), ints + 3,
                     ints + 2, ints + 0, ints + 1, nullptr, 0, 0, false));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(tok, ints + 2, sizeof(int), cudaMemcpyDeviceToHost));
    return B2_OK;
}

// ------------------------------------------------------------------------------------------
// Allocation-free decode attention (see the block comment at the end of attention.cu).
extern "C" int b2_attention_decode_static(int device, const float* dev_q, const uint16_t* dev_k_pool,
                                          const uint16_t* dev_v_pool, const int32_t* dev_block_table, int max_pages,
                                          const int32_t* dev_pos, int n_head, int n_kv, int head_dim, int block_tokens,
                                          int n_splits, float* dev_out, void* stream) {
    if (!dev_q || !dev_k_pool || !dev_v_pool || !dev_block_table || !dev_pos || !dev_out)
        return fail(B2_ERR_ARG, "null device pointer");
    if (head_dim != 64 && head_dim != 128) return fail(B2_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    if (n_kv <= 0 || n_head <= 0 || n_head % n_kv || n_head / n_kv > 8)
        return fail(B2_ERR_UNSUPPORTED, "n_head / n_kv must be an integer <= 8");
    if (block_tokens <= 0 || block_tokens % 16) return fail(B2_ERR_ARG, "block_tokens must be a multiple of 16");
    if (max_pages <= 0 || max_pages > 512) return fail(B2_ERR_ARG, "max_pages must be in 1..512");
    if (n_splits < 2 || (n_head / n_kv) * n_splits > 512)
        return fail(B2_ERR_ARG, "n_splits must be >= 2 (one split needs no scratch) and G * n_splits <= 512");
    if (n_kv > 64) return fail(B2_ERR_ARG, "built-in ticket array holds at most 64 KV heads");
    size_t need = attn_scratch_floats(n_head, n_kv, head_dim, n_splits);
    if (need > attn_static_scratch_floats())
        return fail(B2_ERR_ARG, "split partials need " + std::to_string(need * 4) + " bytes, the built-in scratch holds " +
                                    std::to_string(attn_static_scratch_floats() * 4) + " (n_head * n_splits * (head_dim + 2) floats)");
    if (((uintptr_t)dev_k_pool | (uintptr_t)dev_v_pool) & 15)
        return fail(B2_ERR_ARG, "the K and V pools must be 16-byte aligned (bulk copies read them)");
    if (((uintptr_t)dev_q | (uintptr_t)dev_out | (uintptr_t)dev_block_table | (uintptr_t)dev_pos) & 3)
        return fail(B2_ERR_ARG, "q, out, the page table and the position must be 4-byte aligned");
    int prev_device = -1;
    CK(cudaGetDevice(&prev_device));
    if (prev_device != device) CK(cudaSetDevice(device));
    // first call on a device: resolve the symbol addresses (no allocation). Never touches the host-mapped watchdog word
    // the engine paths use (creating it allocates pinned memory): a barrier time-out in this call is not reported.
    cudaError_t ce = attn_static_setup(device);
    if (ce == cudaSuccess) {
        AttnParams a = {};
        a.q = dev_q; a.k_cache = (const __half*)dev_k_pool; a.v_cache = (const __half*)dev_v_pool;
        a.block_table = dev_block_table; a.pos_ptr = dev_pos; a.out = dev_out;
        a.n_head = n_head; a.n_kv = n_kv; a.head_dim = head_dim; a.block_tokens = block_tokens; a.n_splits = n_splits;
        a.max_pages = max_pages; a.scale = 1.0f / sqrtf((float)head_dim); a.err = nullptr;
        ce = attn_decode_launch_static(a, device, (cudaStream_t)stream);
    }
    if (prev_device != device) cudaSetDevice(prev_device);      // leave the caller's current device as it was
    if (ce != cudaSuccess) return fail(B2_ERR_CUDA, std::string("b2_attention_decode_static: ") + cudaGetErrorString(ce));
    return B2_OK;
}
