// Internal launch interfaces between the engine (engine.cu) and the kernels. Not part of the
// C ABI -- that is include/b200_gguf.h.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <utility>

namespace b2 {

// Kernels of the prompt / batched-decode chains are launched with programmatic dependent launch while this is set
// (engine.cu sets it around a chain): a kernel may start while its predecessor still runs and must execute
// griddepcontrol.wait before it touches anything the predecessor writes (or reads, if it overwrites it).
extern thread_local bool g_chain_pdl;
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*k)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_chain_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k, std::forward<Args>(args)...);
}

// ---------------------------------------------------------------- mat-vec (matvec.cu)
enum MvMode { MV_STORE = 0, MV_RESID = 1, MV_QKV = 2, MV_GATEUP = 3 };
#ifndef B2_MV_NCW
#define B2_MV_NCW 19
#endif
constexpr int MV_NCW = B2_MV_NCW;            // consumer warps per CTA
constexpr int MV_MAX_PROD = 4;               // TMA producer warps per CTA (upper bound; 2 are launched)
constexpr int MV_ROWS = 4;                   // rows per row group (two RoPE pairs; one warp, one ring stage)
constexpr int MV_MAX_STAGES = 48;            // ring stages (upper bound)
constexpr int MV_SHARED_HDR = 2048;          // barriers + reduction scratch at the start of dynamic smem

struct MvSeg {
    alignas(64) unsigned char tmap[128];   // CUtensorMap over the raw rows (8-byte elements); valid when MvParams::use_tma
    const uint8_t* w;   // quantised rows, unchanged GGUF bytes
    int type;           // ggml type id
    int n_rows;
    int row_bytes;      // filled by mv_plan
    int slice_bytes;    // bytes of one full 2048-element K-slice of a row (filled by mv_plan)
    int cta0, ncta;     // CTA range that works on this segment (filled by mv_plan)
    int out_off;        // element offset into y / resid
};

struct MvParams {
    MvSeg seg[3];
    int nseg;
    int K;
    int mode;
    const float* x;         // [K] float32 input
    const float* norm_w;    // optional fused RMSNorm weight [K]
    float eps;
    float* y;
    const float* resid;
    // MV_QKV epilogue
    float* q_out;
    __half* k_cache;        // this layer's K pool
    __half* v_cache;
    const int* pos_ptr;     // position of the token being decoded (device)
    const int* block_table; // this sequence's page table (device)
    const float* rope_freq; // [head_dim / 2]
    int head_dim, n_kv, block_tokens;
    unsigned int* err;      // host-mapped error word (kernel watchdogs)
    unsigned long long* tl; // optional timeline record of this launch (B2_TIMELINE=1): [0] first CTA in, [1] last CTA past its
                            // dependency wait, [2] last CTA done with its prologue, [3] last CTA out (%globaltimer, ns)
    int trigger_late;       // 1: let the next kernel launch only after this one's own dependency wait (its prologue
                            //    may then read what this kernel's predecessors wrote, e.g. the token position)
    // plan (mv_plan)
    int n_slices;           // K-slices of 2048 elements per row
    int slot_bytes;         // one row-slice slot of a stage
    int stage_bytes, n_stages; // ring: tile q of a CTA lives in stage q % n_stages
    int n_prod;             // producer warps
    int use_tma;            // 1: a tile (R row-slices) is ONE 2-D tensor copy, rows land densely (pitch = slice bytes);
                            // 2: the same as a 3-D copy of two half-slices per row (Q8_0: a slice is wider than one box);
                            // 0: R 1-D bulk copies into slots of slot_bytes (rows or slices off a 16-byte boundary)
    int act_plane;          // bytes of one int8 activation plane (K padded to whole slices)
    int act_off;            // byte offset of the activation block in dynamic shared memory
    int part_off;           // byte offset of the partial-sum slots (groups spanning several tiles)
    int ring_off;           // byte offset of the ring in dynamic shared memory (after the activations)
};

cudaError_t mv_init();   // once per device, outside stream capture
int mv_ctas_per_sm();    // 1 (default build)
int mv_plan(MvParams& p, int n_ctas);
size_t mv_smem_bytes(const MvParams& p);
cudaError_t mv_launch(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl);

// ---------------------------------------------------------------- batched steps: per-slot device state
struct SamplerState;
struct SampleParams;
struct SlotDev {              // one entry per sequence slot, device-resident table
    int* ints;                // [0]=token [1]=pos [2]=out_token [3]=step ...
    int* block_table;
    int* ring;
    SamplerState* sst;
    SampleParams* sp;
    double* uniforms;
};
constexpr int BATCH_MAX = 64;
struct BatchIdx { int n; int slot[BATCH_MAX]; };   // passed by value: which slots take part in this step

// ---------------------------------------------------------------- flash-decode (attention.cu)
struct AttnParams {
    const float* q;          // [n_head * head_dim] float32 (already rotated)
    const __half* k_cache;   // layer pool [num_blocks][n_kv][block_tokens][head_dim]
    const __half* v_cache;
    const int* block_table;  // [max_blocks] pages of this sequence
    const int* pos_ptr;      // current position; attends to positions [0, pos]
    float* out;              // [n_head * head_dim] float32
    float* scratch;          // split-K partials: [n_kv][n_splits][G][head_dim + 2]
    unsigned int* counters;  // [n_kv] arrival tickets (zero between launches)
    int n_head, n_kv, head_dim, block_tokens, n_splits;
    int max_pages;           // entries of block_table that may be read
    float scale;             // 1 / sqrt(head_dim)
    unsigned int* err;       // host-mapped error word (kernel watchdogs)
    unsigned long long* tl;  // optional timeline record (see MvParams::tl)
    int early;               // 1: launched with programmatic dependent launch behind a kernel that triggers late (attention_body.cuh)
    __half* out_h;           // batched tensor-core kernel only: write the output as f16 in the GEMM's operand order
    int out_T_pad;           //   [n_head * head_dim / 8][out_T_pad][8] (row = sequence index) instead of float `out`
};
cudaError_t attn_init();
size_t attn_scratch_floats(int n_head, int n_kv, int head_dim, int n_splits);
cudaError_t attn_decode_launch(const AttnParams& p, cudaStream_t stream, bool pdl);
// one launch for b.n sequences: q / out are [b.n][n_head*head_dim]; position and page table come from tab[b.slot[z]];
// scratch / counters need b.n times the single-sequence size
cudaError_t attn_decode_batch_launch(const AttnParams& p, const SlotDev* tab, const BatchIdx& b, cudaStream_t stream);
// tensor-core flash-decode (attention_mma.cu): same scratch / counters protocol; tab == nullptr: the one sequence p describes
cudaError_t attn_mma_init();
cudaError_t attn_decode_mma_launch(const AttnParams& p, const SlotDev* tab, const BatchIdx* b, cudaStream_t stream, bool pdl);
// allocation-free variant: split partials in a fixed device array inside the library (attention.cu)
size_t attn_static_scratch_floats();
cudaError_t attn_static_setup(int device);
cudaError_t attn_decode_launch_static(AttnParams p, int device, cudaStream_t stream);

// ---------------------------------------------------------------- tensor-core GEMM (gemm.cu)
struct GemmSeg {
    const uint8_t* w;       // quantised rows [n_rows][K], unchanged GGUF bytes
    int type, n_rows;
    int out_off;            // column offset of this matrix in the output row
    int row_bytes, tile0, n_tiles;   // filled by gemm_plan
    int span_bytes, box_bytes;       // one 256-weight span of a row; its (padded) row in the shared-memory box
    int use_tma;                     // 1: spans are staged by 2-D TMA; 0: direct loads (rows not 16-byte multiples)
    alignas(64) unsigned char tmap[128];   // CUtensorMap over the raw rows (u16 elements)
};
struct GemmParams {
    GemmSeg seg[3];         // matrices that share the input (q|k|v, gate|up)
    int nseg, K;
    const __half* x;        // activations in MMA operand order [K/8][T_pad][8] f16 (rows >= T: anything)
    int T, T_pad;           // T_pad: multiple of 16, <= 512 (gemm_plan)
    float* y;               // partial sums of K split s at y + s * split_stride, element (t, n) at [t * ldy + n]
    int ldy;
    long split_stride;
    unsigned int* err;
    long long* dbg;         // optional: CTA 0 records clock64() per K block (8 slots each), scripts/gemm_micro.py
    int n_tiles, n_splits, n_stages, n_bstages, n_raw, swap;   // filled by gemm_plan (swap: 256-row variant, tokens as the MMA's M)
};
cudaError_t gemm_init();
int gemm_plan(GemmParams& p, int n_sms, int max_splits);
size_t gemm_smem_bytes(const GemmParams& p);
cudaError_t gemm_launch(const GemmParams& p, cudaStream_t stream);
cudaError_t to_half_blocked_launch(const float* x, int T, int T_pad, int K, __half* y, cudaStream_t s);

// ---------------------------------------------------------------- prompt / batch kernels (prefill.cu)
// f16 activation matrices are in the GEMM operand order: element (t, k) at ((k / 8) * T_pad + t) * 8 + k % 8.
// "parts": n_parts partial sums (K splits of the GEMM that produced the input) part_stride floats apart.
struct SamplerState;
cudaError_t embed_rows_launch(int type, const uint8_t* table, int K, const int* tokens, int T, float* x, cudaStream_t s);
// x[t] += sum of parts (if n_parts > 0); then, if w: xn[t] = f16(rmsnorm(x[t]) * w) in operand order
cudaError_t resid_norm_half_launch(float* x, const float* parts, int n_parts, long part_stride, const float* w, float eps,
                                   int K, int T, int T_pad, __half* xn, cudaStream_t s);
// qkv parts [T][n_head*hd | n_kv*hd | n_kv*hd]: RoPE on q,k at pos0 + t; q -> qh [T][n_head*hd] f16 (token-major),
// k, v -> paged cache
cudaError_t rope_store_launch(const float* parts, int n_parts, long part_stride, int T, int pos0, int n_head, int n_kv, int hd,
                              int bt, const int* block_table, const float* rope_freq, __half* qh, __half* k_cache,
                              __half* v_cache, cudaStream_t s);
// causal attention of T queries at pos0.. over the paged cache; out in operand order (T_pad)
cudaError_t attn_prefill_launch(const __half* qh, const __half* k_cache, const __half* v_cache, const int* block_table,
                                int T, int T_pad, int pos0, int n_head, int n_kv, int hd, int bt, __half* out, cudaStream_t s);
// gate|up parts [T][2F] -> h = f16(silu(gate) * up) in operand order
cudaError_t silu_mul_half_launch(const float* parts, int n_parts, long part_stride, int F, int T, int T_pad, __half* h, cudaStream_t s);
// batched decode step (one token per slot of b)
cudaError_t embed_batch_launch(int type, const uint8_t* table, int K, const SlotDev* tab, const BatchIdx& b, float* x, cudaStream_t s);
cudaError_t rope_store_batch_launch(const float* parts, int n_parts, long part_stride, const SlotDev* tab, const BatchIdx& b,
                                    int n_head, int n_kv, int hd, int bt, const float* rope_freq, float* q_out,
                                    __half* k_cache, __half* v_cache, cudaStream_t s);
cudaError_t prefill_finish_launch(const int* prompt, int n, int pos_last, int* ints, SamplerState* st, cudaStream_t s);

// ---------------------------------------------------------------- tensor-parallel collectives (tp.cu; NCCL loaded at run time)
bool tp_unique_id(void* out128, std::string& err);
bool tp_comm_init(void** comm, int rank, int world, const void* id128, std::string& err);
void tp_comm_destroy(void* comm);
bool tp_all_reduce_sum(void* comm, const float* send, float* recv, size_t n, cudaStream_t s, std::string& err);
bool tp_all_gather(void* comm, const float* send, float* recv, size_t n, cudaStream_t s, std::string& err);
bool tp_all_gather_bytes(void* comm, const void* send, void* recv, size_t bytes, cudaStream_t s, std::string& err);
constexpr int TP_MAX_WORLD = 8;
struct TpArParams {
    float* x;                                   // out: the reduced vector (local)
    const float* part[TP_MAX_WORLD];            // every rank's partial: [rank] local, the others mapped peer memory
    unsigned int* flag_out[TP_MAX_WORLD];       // every peer's flag array (we write slot `rank` of it)
    const unsigned int* flags_in;               // our flag array (peers write their slot)
    unsigned int* seq;                          // sequence number of this reduction stream (local)
    int rank, world, n;                         // n floats, a multiple of 4
    unsigned int* err;
};
cudaError_t tp_allreduce_launch(const TpArParams& p, cudaStream_t stream, bool pdl);

// ---------------------------------------------------------------- small kernels (misc.cu)
// One pinned, device-mapped word per process that kernel watchdogs write to; 0 = no error.
unsigned int* kernel_error_word();
cudaError_t dequantize_launch(int type, const uint8_t* raw, long n_elems, float* out, cudaStream_t s);
// x[0..K) = dequantised row `*token_ptr` of the embedding table
cudaError_t embed_row_launch(int type, const uint8_t* table, int K, const int* token_ptr, float* x,
                             cudaStream_t s, bool pdl);
cudaError_t silu_mul_launch(const float* g, float* h, int n, cudaStream_t s, bool pdl);   // h = silu(g) * h
cudaError_t rmsnorm_launch(const float* x, const float* w, float eps, int K, int rows, float* y, cudaStream_t s);

// ---------------------------------------------------------------- sampler (sampler.cu)
struct SampleParams {
    float temperature, top_p, min_p, repeat_penalty, freq_penalty, presence_penalty;
    int top_k;          // 1..SAMPLER_MAX_K
    int penalty_last_n; // <= 64... window over the history ring
};
constexpr int SAMPLER_MAX_K = 1024;
constexpr int HISTORY_LEN = 64;
struct SamplerState {            // per sequence slot, device resident
    int history[HISTORY_LEN];    // ring of the last tokens (prompt + generated)
    int n_hist;                  // total tokens ever pushed
};
// logits[V] (modified in place by the penalties) -> *out_token; also pushes the token into the
// history ring, stores it in *next_token_ptr, bumps *pos_ptr and records the token at
// out_ring[new pos] (device-resident decode loop: the host can read many steps back at once).
cudaError_t sample_launch(float* logits, int n_vocab, const SampleParams* params_dev, SamplerState* st,
                          const double* uniforms, int* step_ptr, int* out_token, int* next_token_ptr,
                          int* pos_ptr, int* out_ring, int ring_len, cudaStream_t s, bool pdl);

// logits row z (row_stride floats apart) belongs to slot b.slot[z]; also writes the token to out_tokens[z]
cudaError_t sample_batch_launch(float* logits, long row_stride, int n_vocab, const SlotDev* tab, const BatchIdx& b,
                                int ring_len, int* out_tokens, cudaStream_t s);

}  // namespace b2
