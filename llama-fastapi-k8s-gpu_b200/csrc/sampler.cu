// On-device sampler: repetition/frequency/presence penalties over the last 64 tokens, then either
// greedy arg-max or top-k -> top-p -> min-p -> temperature -> multinomial draw. Restates the order
// llama-cpp-python 0.2.77 applies under create_chat_completion(temperature=1.2, top_p=0.9,
// frequency_penalty=0.7, presence_penalty=0.8) (/root/reference/api.py:55-63; SURVEY.md 2.2 H8,
// 2.3 K12): upstream copies all n_vocab logits to the host every token and sorts on the CPU; here
// one CTA selects the k largest logits with a 3-pass radix select, sorts only those, and 4 bytes
// (the token id) go back to the host. The uniform random numbers come from the host (MT19937
// stream restated in llama_cpp/sampling.py), one double per token.
//
// Roofline: HBM/L2 (reads the logit vector a handful of times); launch-latency bound in practice.
#include "kernels.h"
#include "ptx.cuh"
#include <limits.h>
#include <math.h>

namespace b2 {

constexpr int SM_THREADS = 1024;

struct SmShared {
    unsigned hist[2048];
    unsigned wsum[32];
    unsigned sel, k_rem, n_gt, n_eq;
    float cval[SAMPLER_MAX_K];
    int cidx[SAMPLER_MAX_K];
    float redv[32];
    int redi[32];
};

__device__ __forceinline__ unsigned okey(float x) {      // larger float <-> larger key
    unsigned u = __float_as_uint(x);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Find the first bin i (bins ordered from the largest digit down) whose inclusive prefix count
// reaches k. Writes sh.sel = digit, sh.k_rem = k - (count of strictly larger digits).
__device__ void select_bin(SmShared& sh, int nbins, unsigned k) {
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    // thread t owns reversed bins 2t, 2t+1  (reversed index r <-> digit nbins-1-r)
    unsigned h0 = 0, h1 = 0;
    if (2 * t < nbins) h0 = sh.hist[nbins - 1 - 2 * t];
    if (2 * t + 1 < nbins) h1 = sh.hist[nbins - 2 - 2 * t];
    unsigned s = h0 + h1, inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) sh.wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        unsigned v = sh.wsum[lane], iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned x = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += x; }
        sh.wsum[lane] = iv - v;      // exclusive warp offsets
    }
    __syncthreads();
    unsigned excl = sh.wsum[w] + inc - s;          // count of digits larger than bin 2t
    if (excl < k && k <= excl + h0) { sh.sel = nbins - 1 - 2 * t; sh.k_rem = k - excl; }
    else if (excl + h0 < k && k <= excl + s) { sh.sel = nbins - 2 - 2 * t; sh.k_rem = k - excl - h0; }
    __syncthreads();
}

__device__ __forceinline__ void sample_body(float* logits, int V, const SampleParams* prm,
                                            SamplerState* st, const double* uniforms,
                                            int n_uniform, int* step_ptr, int* out_token,
                                            int* next_token_ptr, int* pos_ptr, int* out_ring,
                                            int ring_len, int* out_token2) {
    __shared__ SmShared sh;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const SampleParams P = *prm;
    const int n_hist = st->n_hist;

    // ---- penalties over the last tokens (llama_sample_repetition_penalties)
    int last_n = min(min(P.penalty_last_n, HISTORY_LEN), n_hist);
    if (last_n > 0 && (P.repeat_penalty != 1.0f || P.freq_penalty != 0.0f || P.presence_penalty != 0.0f)) {
        if (t < last_n) {
            int tok = st->history[(n_hist - 1 - t) & (HISTORY_LEN - 1)];
            bool first = true; int count = 0;
            for (int j = 0; j < last_n; j++) {
                int o = st->history[(n_hist - 1 - j) & (HISTORY_LEN - 1)];
                if (o == tok) { count++; if (j < t) first = false; }
            }
            if (first && tok >= 0 && tok < V) {
                float l = logits[tok];
                l = (l <= 0.f) ? l * P.repeat_penalty : l / P.repeat_penalty;
                l -= (float)count * P.freq_penalty + P.presence_penalty;
                logits[tok] = l;
            }
        }
        __syncthreads();
    }

    int result = 0;
    if (P.temperature <= 0.f || P.top_k == 1) {
        // ---- greedy: first maximum (ties -> lowest id)
        float bv = -INFINITY; int bi = INT_MAX;
        for (int i = t; i < V; i += SM_THREADS) { float v = logits[i]; if (v > bv) { bv = v; bi = i; } }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, o); int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { sh.redv[w] = bv; sh.redi[w] = bi; }
        __syncthreads();
        if (t == 0) {
            for (int i = 1; i < SM_THREADS / 32; i++)
                if (sh.redv[i] > bv || (sh.redv[i] == bv && sh.redi[i] < bi)) { bv = sh.redv[i]; bi = sh.redi[i]; }
            result = bi;
        }
    } else {
        // ---- top-k by radix select on the order-preserving key, 11 + 11 + 10 bits
        const unsigned k = (unsigned)min(min(P.top_k, SAMPLER_MAX_K), V);
        unsigned prefix = 0;
        for (int pass = 0; pass < 3; pass++) {
            const int nb = pass < 2 ? 2048 : 1024;
            for (int i = t; i < 2048; i += SM_THREADS) sh.hist[i] = 0;
            __syncthreads();
            for (int i = t; i < V; i += SM_THREADS) {
                unsigned key = okey(logits[i]);
                if (pass == 0) atomicAdd(&sh.hist[key >> 21], 1u);
                else if (pass == 1) { if ((key >> 21) == prefix) atomicAdd(&sh.hist[(key >> 10) & 2047u], 1u); }
                else { if ((key >> 10) == prefix) atomicAdd(&sh.hist[key & 1023u], 1u); }
            }
            __syncthreads();
            select_bin(sh, nb, pass == 0 ? k : sh.k_rem);
            prefix = (pass == 0) ? sh.sel : (pass == 1 ? ((prefix << 11) | sh.sel) : ((prefix << 10) | sh.sel));
            __syncthreads();
        }
        const unsigned T = prefix, n_eq_take = sh.k_rem, n_eq_total = sh.hist[T & 1023u];
        if (t == 0) { sh.n_gt = 0; sh.n_eq = 0; }
        __syncthreads();
        const unsigned n_gt_total = k - n_eq_take;
        if (n_eq_total == n_eq_take) {
            for (int i = t; i < V; i += SM_THREADS) {
                float v = logits[i];
                if (okey(v) >= T) { unsigned s = atomicAdd(&sh.n_gt, 1u); if (s < k) { sh.cval[s] = v; sh.cidx[s] = i; } }
            }
        } else {
            // ties at the k-th value: keep the lowest ids (contiguous ranges + ordered ranks)
            for (int i = t; i < V; i += SM_THREADS) {
                float v = logits[i];
                if (okey(v) > T) { unsigned s = atomicAdd(&sh.n_gt, 1u); sh.cval[s] = v; sh.cidx[s] = i; }
            }
            int per = (V + SM_THREADS - 1) / SM_THREADS, a = t * per, b = min(V, a + per);
            unsigned mine = 0;
            for (int i = a; i < b; i++) mine += (okey(logits[i]) == T);
            unsigned inc = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { unsigned x = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += x; }
            if (lane == 31) sh.wsum[w] = inc;
            __syncthreads();
            if (w == 0) {
                unsigned v = sh.wsum[lane], iv = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { unsigned x = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += x; }
                sh.wsum[lane] = iv - v;
            }
            __syncthreads();
            unsigned rank = sh.wsum[w] + inc - mine;
            for (int i = a; i < b && rank < n_eq_take; i++) {
                float v = logits[i];
                if (okey(v) == T) { sh.cval[n_gt_total + rank] = v; sh.cidx[n_gt_total + rank] = i; rank++; }
            }
        }
        __syncthreads();
        // ---- bitonic sort of the k candidates: value descending, id ascending
        unsigned n2 = 1; while (n2 < k) n2 <<= 1;
        for (unsigned i = k + t; i < n2; i += SM_THREADS) { sh.cval[i] = -INFINITY; sh.cidx[i] = INT_MAX; }
        __syncthreads();
        for (unsigned size = 2; size <= n2; size <<= 1) {
            for (unsigned stride = size >> 1; stride > 0; stride >>= 1) {
                unsigned i = t, j = i ^ stride;
                if (i < n2 && j > i) {
                    float vi = sh.cval[i], vj = sh.cval[j]; int ii = sh.cidx[i], ij = sh.cidx[j];
                    bool j_first = (vj > vi) || (vj == vi && ij < ii);     // j should come before i
                    bool desc = (i & size) == 0;
                    if (desc ? j_first : !j_first) { sh.cval[i] = vj; sh.cval[j] = vi; sh.cidx[i] = ij; sh.cidx[j] = ii; }
                }
                __syncthreads();
            }
        }
        if (t == 0) {
            // the rest runs over <= k sorted candidates (llama_sample_top_p/min_p/temp/token)
            int n = (int)k;
            float* l = sh.cval;
            if (P.top_p < 1.0f) {
                float mx = l[0], sum = 0.f;
                for (int i = 0; i < n; i++) sum += expf(l[i] - mx);
                float cum = 0.f; int last = n;
                for (int i = 0; i < n; i++) {
                    cum += expf(l[i] - mx) / sum;
                    if (cum >= P.top_p) { last = i + 1; break; }
                }
                n = last;
            }
            if (P.min_p > 0.0f && n > 1) {
                float min_logit = l[0] + logf(P.min_p);
                int i = 1;
                for (; i < n; i++) if (l[i] < min_logit) break;
                n = i;
            }
            float inv_t = 1.0f / P.temperature;
            float mx = l[0] * inv_t, sumf = 0.f;
            for (int i = 0; i < n; i++) { float e = expf(l[i] * inv_t - mx); sh.hist[i] = __float_as_uint(e); sumf += e; }
            // std::discrete_distribution: normalise in double, cumulative sums, lower_bound(u)
            double tot = 0.0;
            for (int i = 0; i < n; i++) tot += (double)(__uint_as_float(sh.hist[i]) / sumf);
            double u = uniforms[*step_ptr % n_uniform];
            double c = 0.0; int pick = n - 1;
            for (int i = 0; i < n - 1; i++) {
                c += (double)(__uint_as_float(sh.hist[i]) / sumf) / tot;
                if (c >= u) { pick = i; break; }
            }
            result = sh.cidx[pick];
        }
    }
    if (t == 0) {
        *out_token = result;
        if (out_token2) *out_token2 = result;
        *next_token_ptr = result;
        *step_ptr = *step_ptr + 1;      // one uniform consumed per sampled token (greedy too)
        st->history[n_hist & (HISTORY_LEN - 1)] = result;
        st->n_hist = n_hist + 1;
        int np = *pos_ptr + 1;          // position the sampled token will occupy
        *pos_ptr = np;
        if (out_ring && np < ring_len) out_ring[np] = result;
    }
}

__global__ void __launch_bounds__(SM_THREADS) sample_kernel(float* logits, int V, const SampleParams* prm,
                                                            SamplerState* st, const double* uniforms,
                                                            int n_uniform, int* step_ptr, int* out_token,
                                                            int* next_token_ptr, int* pos_ptr, int* out_ring,
                                                            int ring_len) {
    griddep_launch();
    griddep_wait();
    sample_body(logits, V, prm, st, uniforms, n_uniform, step_ptr, out_token, next_token_ptr, pos_ptr, out_ring, ring_len, nullptr);
}

// one CTA per sequence of the batch
__global__ void __launch_bounds__(SM_THREADS) sample_batch_kernel(float* logits, long row_stride, int V, const SlotDev* tab,
                                                                  const BatchIdx b, int n_uniform, int ring_len, int* out_tokens) {
    griddep_launch();
    griddep_wait();
    const int z = blockIdx.x;
    const SlotDev sd = tab[b.slot[z]];
    sample_body(logits + (size_t)z * row_stride, V, sd.sp, sd.sst, sd.uniforms, n_uniform, sd.ints + 3, sd.ints + 2, sd.ints + 0,
                sd.ints + 1, sd.ring, ring_len, out_tokens + z);
}

cudaError_t sample_batch_launch(float* logits, long row_stride, int n_vocab, const SlotDev* tab, const BatchIdx& b,
                                int ring_len, int* out_tokens, cudaStream_t s) {
    return launch_chain(sample_batch_kernel, dim3(b.n), dim3(SM_THREADS), 0, s, logits, row_stride, n_vocab, tab, b, 8192, ring_len, out_tokens);
}

cudaError_t sample_launch(float* logits, int n_vocab, const SampleParams* params_dev, SamplerState* st,
                          const double* uniforms, int* step_ptr, int* out_token, int* next_token_ptr,
                          int* pos_ptr, int* out_ring, int ring_len, cudaStream_t s, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(SM_THREADS); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    int n_uniform = 8192;
    return cudaLaunchKernelEx(&cfg, sample_kernel, logits, n_vocab, params_dev, st, uniforms, n_uniform,
                              step_ptr, out_token, next_token_ptr, pos_ptr, out_ring, ring_len);
}

}  // namespace b2
