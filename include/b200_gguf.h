/* b200_gguf.h -- C ABI of the B200-native GGUF inference engine (libb200gguf.so).
 *
 * This is the boundary the Python host (`llama_cpp.Llama`, the drop-in for the object that
 * /root/reference/api.py:7,24-28,55-63 drives) sits on, through ctypes. The reference reaches
 * its engine through llama-cpp-python 0.2.77's ctypes binding of libllama (un-vendored
 * upstream, pinned at /root/reference/docker/Dockerfile.base:30); each entry point below names
 * the libllama call(s) of that binding it stands in for. Plain pointers and sizes only; no
 * torch or C++ types cross this line.
 *
 * Conventions: every function returns 0 on success and a negative code on failure;
 * b2_last_error() returns a thread-local message for the most recent failure on the calling
 * thread. Functions may be called from any host thread (each call selects the engine's device);
 * calls on ONE engine must not overlap in time (the Python host serialises them).
 * "host" pointers are ordinary CPU memory, "dev" pointers are CUDA device memory.
 */
#ifndef B200_GGUF_H
#define B200_GGUF_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK 0
#define B2_ERR_ARG -1
#define B2_ERR_CUDA -2
#define B2_ERR_UNSUPPORTED -3
#define B2_ERR_STATE -4

/* ggml tensor type ids accepted for weights (gguf/constants.py:4059-4094) */
#define B2_TYPE_F32 0
#define B2_TYPE_F16 1
#define B2_TYPE_Q4_0 2
#define B2_TYPE_Q8_0 8
#define B2_TYPE_Q4_K 12
#define B2_TYPE_Q5_K 13
#define B2_TYPE_Q6_K 14

typedef struct b2_engine b2_engine;

typedef struct b2_config {
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab;
    float rms_eps;
    float rope_freq_base;
    int32_t n_ctx;            /* max tokens per sequence (api.py:17,27 passes 1024) */
    int32_t max_seqs;         /* sequence slots that can be live at once */
    int32_t kv_block_tokens;  /* tokens per KV page */
    int32_t kv_num_blocks;    /* pages in the pool shared by all slots */
    int32_t device;           /* CUDA device ordinal */
    int32_t flags;            /* B2_FLAG_* */
} b2_config;

#define B2_FLAG_NO_GRAPH 1      /* launch kernels directly instead of replaying CUDA graphs */
#define B2_FLAG_NO_PDL 2        /* no programmatic dependent launch between kernels */
#define B2_FLAG_ACT8 4          /* single int8 activation plane (faster, ~0.5% noise) */

typedef struct b2_sampling {
    float temperature;        /* <= 0: greedy */
    float top_p;
    float min_p;
    float repeat_penalty;
    float frequency_penalty;
    float presence_penalty;
    int32_t top_k;            /* 1..1024 */
    int32_t penalty_last_n;   /* <= 64 */
} b2_sampling;

/* ---- library ------------------------------------------------------------------------- */
const char* b2_last_error(void);
int b2_version(void);
int b2_device_count(int* n);

/* ---- engine lifecycle: replaces llama_load_model_from_file + llama_new_context_with_model
 *      behind Llama.__init__ (api.py:24-28) and llama_free / llama_free_model. ------------ */
int b2_engine_create(const b2_config* cfg, b2_engine** out);
int b2_engine_destroy(b2_engine* e);
/* Copy one GGUF tensor's raw blocks, unchanged, into HBM (the upload half of llama.cpp's
 * llama_model_loader::load_all_data). ne0 = contiguous length K, ne1 = rows (1 for vectors). */
int b2_upload_tensor(b2_engine* e, const char* name, int ggml_type, int64_t ne0, int64_t ne1,
                     const void* host_data, int64_t nbytes);
/* Tensor parallelism for models that do not fit (or should not occupy) one GPU: one process per GPU, each creating an
 * engine whose b2_config carries the LOCAL head counts and n_ff (1 / world of the model's) and the FULL n_embd and n_vocab,
 * and uploading its shard of every matrix (llama_cpp/tp_shard.py: whole rows of attn_q/k/v, ffn_gate/up and output, K
 * ranges of attn_output and ffn_down cut at 256-element block boundaries). The engine then all-reduces (sum, NCCL) the
 * hidden vector after attn_output and ffn_down and all-gathers the logits. No counterpart in the reference, which only
 * runs independent replicas (/root/reference/helm/values.yaml:17). b2_tp_unique_id: rank 0 makes the 128-byte NCCL id,
 * the host distributes it; b2_engine_tp_init: collective over all ranks, after b2_engine_create and before
 * b2_engine_finalize. In this mode prompts and multi-slot steps run through the mat-vec path. */
int b2_tp_unique_id(void* host_out_128_bytes);
int b2_engine_tp_init(b2_engine* e, int rank, int world, const void* host_id_128_bytes);
/* All tensors uploaded: validate the set, build the per-layer launch plans. */
int b2_engine_finalize(b2_engine* e);
int b2_engine_device_bytes(b2_engine* e, int64_t* weight_bytes, int64_t* kv_bytes);

/* ---- sequences (paged KV): replaces llama_kv_cache_* bookkeeping (llama.cpp llama_kv_cache).
 *      The page allocator lives in the host (llama_cpp/kv_pages.py); the engine only holds
 *      each slot's page table. -------------------------------------------------------------- */
int b2_seq_set_blocks(b2_engine* e, int slot, const int32_t* host_blocks, int n_blocks);
/* Start (or restart) a slot: n_past tokens are already in its KV pages; `host_history` are the
 * last <= 64 tokens of the sequence so far (for the penalties); `host_uniforms` are n_uniform
 * (<= 8192) doubles in [0,1) for the multinomial draws of this sequence. */
int b2_seq_begin(b2_engine* e, int slot, int n_past, const int32_t* host_history, int n_history,
                 const b2_sampling* sampling, const double* host_uniforms, int n_uniform);

/* ---- forward: replaces llama_decode (llama.cpp llama_decode_internal) + llama_get_logits +
 *      llama_sample_* under Llama.eval / Llama.sample (api.py:55-63). ------------------------ */
/* Append n_tokens prompt tokens at positions n_past.. to the slot, filling its KV pages; the
 * logits of the LAST token are left in the slot's logit buffer. */
int b2_prefill(b2_engine* e, int slot, const int32_t* host_tokens, int n_tokens);
/* Sample from the slot's logit buffer with the slot's sampling settings; the sampled token
 * becomes the slot's next input. */
int b2_sample(b2_engine* e, int slot, int32_t* host_token_out);
/* One decode step for each listed slot: forward the slot's pending token, sample the next.
 * Writes n sampled tokens. */
int b2_decode(b2_engine* e, const int32_t* host_slots, int n, int32_t* host_tokens_out);
/* n decode steps of ONE slot queued back to back (the sampled token never leaves the device
 * between steps); the n sampled tokens come back in one copy at the end. Used for run-ahead
 * decoding: the host checks stop conditions afterwards and rewinds with b2_seq_rewind. */
int b2_decode_n(b2_engine* e, int slot, int n, int32_t* host_tokens_out);
/* Forget everything after the first n_keep tokens of the slot (KV pages beyond are simply
 * overwritten later); the token at position n_keep becomes the pending input. Needs the
 * recent-token history as it stood at that point (for the penalties). */
int b2_seq_rewind(b2_engine* e, int slot, int n_keep, int32_t pending_token,
                  const int32_t* host_history, int n_history);
/* Teacher forcing / tests: overwrite the slot's pending input token. */
int b2_seq_set_token(b2_engine* e, int slot, int32_t token);
/* Forward the slot's pending token without sampling (logits left in the slot's buffer). */
int b2_forward_pending(b2_engine* e, int slot);
/* Copy the slot's current logits [n_vocab] to the host (llama_get_logits). */
int b2_get_logits(b2_engine* e, int slot, float* host_out);
int b2_synchronize(b2_engine* e);

/* ---- measurement hooks (bench.py): CUDA events on the engine's own stream --------------- */
int b2_timer_start(b2_engine* e);
int b2_timer_stop(b2_engine* e, float* ms_out);          /* synchronises */
int b2_kernel_launch_count(b2_engine* e, int64_t* n);   /* kernels launched since create */
/* How the engine is running: CUDA graphs on/off, PDL on/off, int8 activation planes (1|2),
 * CTAs per mat-vec launch, attention split count. Any pointer may be NULL. */
int b2_engine_mode(b2_engine* e, int* graph, int* pdl, int* act_planes, int* n_ctas, int* n_splits);
/* Tensor-parallel state: rank, world size (1 = not tensor parallel) and whether the hidden-vector all-reduce runs over
 * mapped peer memory (1) or through NCCL (0). Any pointer may be NULL. */
int b2_engine_tp_mode(b2_engine* e, int* rank, int* world, int* peer_memory);
/* Per-kernel-kind device time of ONE un-graphed, un-overlapped decode step of `slot`
 * (events around every launch). kinds: 0 embed, 1 qkv, 2 attention, 3 attn_out, 4 gate_up,
 * 5 ffn_down, 6 lm_head, 7 sampler. ms_out[8], launches_out[8], bytes_out[8] (algorithmic). */
int b2_profile_decode_step(b2_engine* e, int slot, float* ms_out, int32_t* launches_out,
                           int64_t* bytes_out);

/* Streaming rate of the decode mat-vec kernel alone (scripts/mv_micro.py): synthetic weight bytes made on the
 * device, `iters` launches back to back on one stream cycling through n_copies weight sets (size them past the
 * L2), chained by programmatic dependent launch when pdl != 0. mode: 0 store, 1 residual add, 3 fused gate|up
 * (two matrices of n_rows). us_per_launch_out = mean device time per launch (CUDA events). */
int b2_bench_matvec(int device, int mode, int ggml_type, int n_rows, int K, int n_copies, int iters, int n_ctas,
                    int act_planes, int pdl, float* us_per_launch_out);

/* ---- single-operator entry points (parity tests call the kernels through these).
 *      All pointers are HOST memory; the call stages data through device scratch. ---------- */
/* ggml dequantize_row_* / gguf.quants.dequantize */
int b2_op_dequantize(int device, int ggml_type, const void* host_blocks, int64_t n_elems, float* host_out);
/* y[n_rows] = W[n_rows, K] . x  (+ optional fused RMSNorm of x, + optional residual);
 * ggml-cuda quantize_q8_1 + mul_mat_vec_q (+ rms_norm_f32, add). act_planes = 1 or 2. */
int b2_op_matvec(int device, int ggml_type, const void* host_w, int n_rows, int K, const float* host_x,
                 const float* host_norm_w, float eps, const float* host_resid, float* host_y,
                 int act_planes, int n_ctas);
/* fused gate/up: y[n_rows] = silu(Wg . x) * (Wu . x), x optionally RMS-normed first */
int b2_op_gate_up(int device, int type_gate, const void* host_wg, int type_up, const void* host_wu,
                  int n_rows, int K, const float* host_x, const float* host_norm_w, float eps,
                  float* host_y, int n_ctas);
/* Prefill / batched GEMM on the tensor cores (tcgen05, in-shared-memory dequantisation):
 * y[t][n] = (y_in ? y_in[t][n] : 0) + sum_k f16(x[t][k]) * W[n][k]; x [T][K] f32 (rounded to f16 on the device),
 * W [n_rows][K] raw GGUF blocks, T <= 512, K % 256 == 0. Replaces ggml-cuda's dequantise-to-f16 + cublasGemmEx
 * prompt path that create_chat_completion reaches (/root/reference/api.py:55-63).
 * n_splits: K ranges computed by separate CTAs and added in order (0 = the engine's own choice, else 1..16). */
int b2_op_gemm(int device, int type, const void* host_w, int n_rows, int K, const float* host_x, int T,
               const float* host_y_in, float* host_y, int n_splits);
/* ggml rms_norm_f32 * weight */
int b2_op_rmsnorm(int device, const float* host_x, const float* host_w, float eps, int K, int rows, float* host_y);
/* Decode attention for one sequence: q [n_head*head_dim] f32 (already rotated), K/V as f16
 * [n_ctx_tokens][n_kv][head_dim] (token-major; the op scatters them into pages in the order
 * host_blocks gives), attends to all n_ctx_tokens. out [n_head*head_dim] f32. * n_splits < 0 selects the tensor-core kernel (attention_mma.cu, q rounded to f16) with -n_splits splits. */
int b2_op_attention_decode(int device, const float* host_q, const uint16_t* host_k, const uint16_t* host_v,
                           int n_ctx_tokens, int n_head, int n_kv, int head_dim, int block_tokens,
                           const int32_t* host_blocks, int n_blocks_total, int n_splits, float* host_out);
/* penalties + greedy/top-k/top-p/min-p/temperature draw over logits[n_vocab] with the given
 * recent-token history and uniform u in [0,1). */
int b2_op_sample(int device, const float* host_logits, int n_vocab, const b2_sampling* sampling,
                 const int32_t* host_history, int n_history, double u, int32_t* host_token_out);

/* ---- decode attention that allocates nothing (no counterpart in the reference: llama.cpp
 *      allocates its compute buffers itself; this exists for hosts of THIS library).
 *
 * Same kernel and same result as the split path of b2_op_attention_decode, for callers that
 * cannot ask the driver for memory at call time: a host that has already committed all of its
 * HBM to weights and KV pages, or one whose token loop must not hit the allocator (an
 * allocation serialises the device and can fail under memory pressure). The split-merge
 * scratch is a fixed 1 MiB array (plus 256 B of tickets) built into the library; every GPU that
 * loads the library carries it until the process exits.
 *
 * All pointers are DEVICE pointers owned by the caller: q [n_head*head_dim] f32, the paged K and
 * V pools [page][n_kv][block_tokens][head_dim] f16, the page table (max_pages entries), *dev_pos
 * = index of the newest token (attends to 0..pos), out [n_head*head_dim] f32. `stream` is a
 * cudaStream_t (NULL = default stream); the call only enqueues work.
 * Limits: n_splits >= 2, (n_head / n_kv) * n_splits <= 512, n_head * n_splits * (head_dim + 2) <= 262144 floats,
 * n_kv <= 64, max_pages <= 512; anything larger is rejected with B2_ERR_ARG and a message naming the limit.
 * The K and V pools must be 16-byte aligned (the kernel reads them with bulk copies), the other pointers 4-byte
 * aligned; misaligned pointers are rejected. *dev_pos must be < max_pages * block_tokens: the kernel reads the
 * position on the device and attends to at most max_pages * block_tokens positions whatever it says.
 * The caller's current CUDA device is left as it was.
 * The first call on each device also looks up the array's device address (cudaGetSymbolAddress)
 * and sets a kernel attribute; make that first call before starting any stream capture.
 * Every call uses the same array: it clears the tickets, the partials are overwritten before they
 * are read, and two calls that overlap on the GPU (different streams) race with each other.
 * No cudaMalloc, cudaFree or allocating library routine is called, also when this is the very first call
 * into the library. The engine paths report a kernel that timed out on a barrier through a pinned host word;
 * creating that word allocates, so this entry point does without it and cannot report such a time-out. */
int b2_attention_decode_static(int device, const float* dev_q, const uint16_t* dev_k_pool,
                               const uint16_t* dev_v_pool, const int32_t* dev_block_table, int max_pages,
                               const int32_t* dev_pos, int n_head, int n_kv, int head_dim, int block_tokens,
                               int n_splits, float* dev_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_GGUF_H */
