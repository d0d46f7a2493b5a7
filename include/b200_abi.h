/*
 * b200_abi.h — C-ABI of libb200ops.so: the B200 (sm_100a) replacement for the
 * GPU operators on micro-vllm's paged-KV attention hot path.
 *
 * Every entry point below is what the reference's Python operator modules bind
 * instead of flash-attn / Triton.  Each declaration cites the reference
 * interface it replaces (paths relative to the reference repo root).
 *
 * Conventions:
 *   - plain pointers and sizes only; device pointers come from tensor.data_ptr()
 *   - strides are in ELEMENTS, innermost (head_dim / hidden) stride must be 1
 *   - activations / caches / rotary tables are bf16, indices are int32,
 *     workspaces are fp32
 *   - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream)
 *   - return value: 0 = OK, <0 = argument error (B200_E_*), >0 = cudaError_t
 *   - no entry point allocates, synchronises or touches host memory after
 *     argument validation, so all of them are CUDA-graph capturable
 */
#ifndef B200_ABI_H
#define B200_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 17

/* argument errors */
#define B200_E_NULL        (-1)  /* required pointer is NULL            */
#define B200_E_SHAPE       (-2)  /* unsupported shape (head_dim, group) */
#define B200_E_ALIGN       (-3)  /* pointer / stride not 16-byte aligned */
#define B200_E_RANGE       (-4)  /* value out of range                  */
#define B200_E_UNSUPPORTED (-5)  /* feature not implemented             */

/* weight dtype selector for the norm kernels */
#define B200_DT_BF16 0
#define B200_DT_F32  1

int b200_version(void);

/* Human-readable description of the most recent non-zero return on this thread. */
const char* b200_last_error(void);

/* Number of SMs of the current device (used by the host-side split heuristic). */
int b200_sm_count(void);

/* ------------------------------------------------------------------------ *
 * SwiGLU  — replaces kernel/swiglu.py:6-46 (_swiglu_kernel), launched from
 * swiglu_fused (:49-93) and swiglu_gemm (:100-148).
 *   out[r, c] = bf16( silu(f32(gate_up[r, d + c])) * f32(gate_up[r, c]) )
 * gate_up is [rows, 2*d] (up = first half, gate = second half), out [rows, d].
 * ------------------------------------------------------------------------ */
int b200_swiglu(const void* gate_up, void* out,
                int64_t rows, int64_t d,
                int64_t stride_in, int64_t stride_out,
                void* stream);

/* ------------------------------------------------------------------------ *
 * RMSNorm — replaces kernel/rmsnorm_add.py:9-34 (_rmsnorm_kernel) and
 * :133-167 (_rmsnorm_gemm_kernel); wrappers rmsnorm (:37-53), rmsnorm_ (:170-206).
 *   y = bf16( f32(x) * rsqrt(mean(f32(x)^2) + eps) * f32(w) )
 * ------------------------------------------------------------------------ */
int b200_rmsnorm(const void* x, const void* w, void* y,
                 int64_t rows, int64_t n,
                 int64_t stride_x, int64_t stride_y,
                 float eps, int w_dtype,
                 void* stream);

/* ------------------------------------------------------------------------ *
 * Residual-add + RMSNorm — replaces kernel/rmsnorm_add.py:56-86,
 * kernel/rmsnorm_residual.py:11-41 and :74-124; wrappers
 * rmsnorm_residual_fused / _gemm / _hybrid.
 *   s = f32(x) + f32(r); res_out = bf16(s); rrms = rsqrt(mean(s^2) + eps)
 *   y = bf16( f32(res_out) * rrms * f32(w) )      (second pass re-reads the
 *   ROUNDED residual, rmsnorm_residual.py:119)
 * ------------------------------------------------------------------------ */
int b200_rmsnorm_residual(const void* x, const void* r, const void* w,
                          void* y, void* res_out,
                          int64_t rows, int64_t n,
                          int64_t stride_x, int64_t stride_r,
                          int64_t stride_y, int64_t stride_res,
                          float eps, int w_dtype,
                          void* stream);

/* ------------------------------------------------------------------------ *
 * store_kvcache — replaces core/cache_manager.py:74-140 (store_kvcache_kernel),
 * launched from store_kvcache (:143-194) / store_kvcache_batch (:196-255).
 *   cache[slot / block_size, slot % block_size, h, :] = key/value[t, h, :]
 *   for every token t with slot_mapping[t] != -1   (bit-exact copy)
 * key/value [T, H, D] with element strides (kt, kh) / (vt, vh), D contiguous;
 * caches [n_blocks, block_size, H, D] with element strides (cb, cs, ch).
 * ------------------------------------------------------------------------ */
int b200_store_kvcache(const void* key, const void* value,
                       void* k_cache, void* v_cache,
                       const int32_t* slot_mapping,
                       int64_t T, int64_t H, int64_t D,
                       int64_t kt, int64_t kh, int64_t vt, int64_t vh,
                       int64_t n_blocks, int64_t block_size,
                       int64_t cb, int64_t cs, int64_t ch,
                       void* stream);

/* ------------------------------------------------------------------------ *
 * Greedy token selection — replaces core/layer/sampler.py:21-43 evaluated at
 * top_k = 1 (torch.topk(logits / temp, 1): the index of the row maximum; the
 * softmax / top-p / multinomial over one candidate are the identity), called
 * from core/engine.py:276-280.  Runs as the last launch of the captured decode
 * step so the ids never leave the device.
 *   out_ids[r] = FIRST index of the maximum of logits[r, 0..vocab)   (int64)
 * logits [rows, vocab] bf16 (B200_DT_BF16) or fp32 (B200_DT_F32), row stride in
 * elements, rows <= 65535.  NaN (sign clear) orders above +inf like torch.topk.
 * workspace: b200_greedy_argmax_workspace_bytes(rows) bytes, 8-byte aligned,
 * ZERO-FILLED ONCE by the caller; the kernel leaves it zeroed (replay safe).
 * ------------------------------------------------------------------------ */
int64_t b200_greedy_argmax_workspace_bytes(int64_t rows);
int b200_greedy_argmax(const void* logits, int logits_dtype,
                       int64_t rows, int64_t vocab, int64_t stride,
                       int64_t* out_ids,
                       void* workspace, int64_t workspace_bytes,
                       void* stream);

/* ------------------------------------------------------------------------ *
 * Arg-max without a caller workspace.  NO COUNTERPART IN THE REFERENCE: this
 * entry exists for hosts that cannot allocate device memory themselves, not to
 * replace a reference call.
 *
 * Who needs it: an engine that sized its KV pool to fill the GPU and has
 * nothing left for a late cudaMalloc; a latency-critical decode loop that must
 * not pay cudaMalloc's device-wide synchronisation; a cgo / JNI / ctypes host
 * that owns only the logits and the id buffer it passes in.  For them the
 * scratch that b200_greedy_argmax asks the caller for (16 bytes per row) is a
 * fixed array inside the library: B200_ARGMAX_NOALLOC_MAX_ROWS * 16 bytes =
 * 64 KiB, resident on every GPU that loads libb200ops.so until the process
 * exits.  Nothing is allocated or freed by this call, on any path.
 *
 * Same inputs, outputs and result as b200_greedy_argmax (FIRST index of the
 * row maximum, int64), minus the workspace arguments.
 *   rows > B200_ARGMAX_NOALLOC_MAX_ROWS -> B200_E_RANGE (use the workspace form).
 *   The first call on each device looks up the array's device address (that
 *     device must be current, as for every entry here); make that first call
 *     before starting any stream capture on the device.
 *   All calls on one device share the one array.  Each call clears the part it
 *     uses in stream order, so back-to-back calls on ONE stream are fine; calls
 *     that can run concurrently on the GPU (different streams, or different
 *     host threads feeding different streams) race on it and must be serialised
 *     by the caller.
 * ------------------------------------------------------------------------ */
#define B200_ARGMAX_NOALLOC_MAX_ROWS 4096
int b200_greedy_argmax_noalloc(const void* logits, int logits_dtype,
                               int64_t rows, int64_t vocab, int64_t stride,
                               int64_t* out_ids, void* stream);

/* Top-k / top-p sampling: replaces the reference's Inductor-compiled Sampler._sample_impl
 * (/root/reference/core/layer/sampler.py:21-43; called with top_k = 50 for decode, 1000 for prefill,
 * core/engine.py:233, :278) in ONE launch, one CTA per row:
 *   x = logits / temperatures[row];  the top_k largest (ties: lowest token id), sorted descending;
 *   probs = softmax(x) over them (fp32);  cum = cumsum(probs);  kept = cum < top_ps[row], first always kept;
 *   renormalise;  draw = first kept candidate whose cumulative probability exceeds uniforms[row].
 * `uniforms` f32 [rows] in [0, 1) comes from the caller's generator (the reference draws with
 * torch.multinomial; the DISTRIBUTION is the contract, the RNG stream is not reproducible across
 * implementations).  temperatures / top_ps / uniforms may be NULL (1.0 / 1.0 / 0.0: the arg-max).
 * top_k in [1, 1024] (clamped to vocab).  Optional debug outputs for parity tests: dbg_idx i32 [rows, top_k]
 * = candidate token ids in sorted order, dbg_prob f32 [rows, top_k] = renormalised probabilities (0 for the
 * candidates the top-p mask removed); both NULL or both set.  out_ids i64 [rows].  No scratch; capturable. */
int b200_topk_topp_sample(const void* logits, int logits_dtype, int64_t rows, int64_t vocab, int64_t stride,
                          int top_k, const float* temperatures, const float* top_ps, const float* uniforms,
                          int64_t* out_ids, int32_t* dbg_idx, float* dbg_prob, void* stream);

/* ------------------------------------------------------------------------ *
 * Device-resident decode metadata — replaces, between decode steps, the per-step rebuild of
 * core/cache_manager.py:509-546 (cache_batch_data: B slice-assign kernels + 2 H2D copies) driven by
 * core/engine.py:259-262:   cache_seqlens[0..B) += delta   and every pending block-table entry
 * (KVCacheManager.append handing out a new page = one int32) is written from the device-side list
 *   updates = int32 [1 + 3 * max_updates] = { count, (row, col, block_id) x count },
 * whose count the kernel resets to 0.  The list is read at run time, so the launch may sit inside a captured
 * CUDA graph.  updates may be NULL (lengths only).  One CTA.
 * ------------------------------------------------------------------------ */
int b200_table_advance(int32_t* cache_seqlens, int B, int delta, int32_t* block_table, int64_t bt_stride,
                       int rows, int cols, int32_t* updates, int max_updates, void* stream);

/* ------------------------------------------------------------------------ *
 * L2 prefetch of `rows` slabs of `row_bytes` bytes, `stride_bytes` apart, starting at ptr (rows = 1: one contiguous
 * range) (cp.async.bulk.prefetch.L2; reads nothing into the SM, writes nothing).  The
 * decode runner launches it on a side stream beside the NVLink reduction after o_proj / down_proj
 * (core/layer/model_graph.py:210, :219), a latency-bound kernel that leaves HBM idle, with the weights of the GEMM
 * that follows the reduction.  ptr and stride 16-byte aligned; at most 64 CTAs.
 * ------------------------------------------------------------------------ */
int b200_l2_prefetch(const void* ptr, int64_t row_bytes, int64_t rows, int64_t stride_bytes, void* stream);

/* ------------------------------------------------------------------------ *
 * Weight-streaming GEMM on tcgen05 tensor cores (TMA-fed, TMEM accumulators, swap-AB:
 * the weights are the 128-row UMMA M operand, the batch is the UMMA N) — replaces the
 * torch.matmul (cuBLAS) calls of a decode layer AND the elementwise launches behind them:
 * core/layer/model_graph.py:145-147 (QKV projection + bias add, B200_GEMM_EPI_BIAS), :171
 * (o_proj), :183-184 (gate_up projection + kernel/swiglu.py::swiglu_gemm,
 * B200_GEMM_EPI_SWIGLU), :185 (down_proj).  SURVEY.md §8f N4.
 *   out[m, n] = bf16( sum_k f32(x[m, k]) * f32(w[n, k]) )                 (fp32 accumulate)
 *   EPI_BIAS:   out = bf16( f32(bf16(sum)) + f32(bias[n]) )               (two roundings, like the reference)
 *   EPI_SWIGLU: w rows are interleaved (row 2j = up_j, row 2j+1 = gate_j);
 *               out[m, j] = bf16( silu(f32(bf16(gate))) * f32(bf16(up)) ), out is [M, N/2]
 * x [M, K] bf16, 1 <= M <= 128; w [N, K] bf16 (nn.Linear layout, K contiguous); K % 64 == 0,
 * N % 8 == 0; strides in elements (ldx, ldw multiples of 8; ldo multiple of 4).  HBM-bound on w.
 * One persistent CTA per SM, the (128-row tile, 64-element k-block) units dealt out in equal
 * contiguous shares (stream-K); tiles shared between CTAs are reduced by all of them in fixed
 * order (deterministic).  workspace: b200_stream_gemm_workspace_bytes(M, N, grid_hint) bytes,
 * 16-byte aligned, ZERO-FILLED ONCE by the caller; the kernel leaves its counters zeroed
 * (replay safe).  grid_hint: 0 = one CTA per SM (never more: the CTAs wait on one another).
 * flags: B200_GEMM_FLAG_PDL launches with programmatic stream serialization: the first weight
 * boxes are requested before griddepcontrol.wait (precondition: the preceding kernel in the
 * stream writes neither w nor bias).  With PDL, B200_GEMM_FLAG_L2_AHEAD(n) also prefetches the
 * next n weight boxes (16 KB each) of every CTA's share into L2 before the wait: when the
 * preceding kernel idles in a cross-GPU reduction, HBM already streams this product's weights.
 * ------------------------------------------------------------------------ */
#define B200_GEMM_EPI_NONE   0
#define B200_GEMM_EPI_BIAS   1
#define B200_GEMM_EPI_SWIGLU 2
#define B200_GEMM_FLAG_PDL   1
#define B200_GEMM_FLAG_L2_AHEAD(n) (((n) & 0xffff) << 8)
int64_t b200_stream_gemm_workspace_bytes(int M, int64_t N, int grid_hint);
/* debug only: uint64 [CTAs of the split-KV grid][8] device buffer the next b200_paged_attn_decode launches stamp with
 * %globaltimer (entry, first loads issued, prologue done, first stage landed, loop done, partial written, exit,
 * merge done); NULL = off.  tools/trace_decode.py */
void b200_debug_decode_trace(void* dev_ptr);
/* debug only: uint64 [grid][8] device buffer the next launches stamp with %globaltimer per CTA (NULL = off) */
void b200_debug_stream_gemm_trace(void* dev_ptr);
int b200_stream_gemm(const void* x, const void* w, const void* bias, void* out,
                     void* workspace, int64_t workspace_bytes,
                     int M, int64_t K, int64_t N, int64_t ldx, int64_t ldw, int64_t ldo,
                     int epilogue, int flags, int grid_hint, void* stream);

/* ------------------------------------------------------------------------ *
 * Weight-streaming GEMM for decode-sized batches (legacy mma.sync form, kept for the
 * [in, out] weight layout) — replaces the torch.matmul
 * (cuBLAS) calls of a decode layer: core/layer/model_graph.py:145-147 (QKV
 * projection; `bias` reproduces the separate bf16 add), :171 (o_proj), :183
 * (gate_up), :185 (down_proj), on the [in, out]-contiguous weights of
 * prepare_weights (:95-116).  SURVEY.md §8f N4.
 *   out[m, n] = bf16( sum_k f32(x[m, k]) * f32(w[k, n]) )        (fp32 accumulate)
 *   with bias: out = bf16( f32(bf16(sum)) + f32(bias[n]) )        (two roundings, like the reference)
 * x [M, K] bf16, 1 <= M <= 64; w [K, N] bf16 row-major; K % 64 == 0, N % 64 == 0;
 * strides in elements (x, w multiples of 8; out multiple of 4).  HBM-bound on w.
 * ksplit: number of k-ranges (0 = b200_skinny_gemm_splits); variant: 0 = 128-column
 * slices, 1 = 64, 2 = 256.  workspace (needed when ksplit > 1):
 * b200_skinny_gemm_workspace_bytes(M, N, ksplit) bytes, 16-byte aligned, ZERO-FILLED
 * ONCE by the caller; the kernel leaves its counters zeroed (graph-replay safe).
 * Summation order is fixed (k ascending inside a range, ranges ascending): results
 * are deterministic for a given (ksplit, variant).
 * ------------------------------------------------------------------------ */
int b200_skinny_gemm_splits(int M, int64_t K, int64_t N, int variant);
int64_t b200_skinny_gemm_workspace_bytes(int M, int64_t N, int ksplit);
int b200_skinny_gemm(const void* x, const void* w, const void* bias, void* out,
                     void* workspace, int64_t workspace_bytes,
                     int M, int64_t K, int64_t N,
                     int64_t stride_x, int64_t stride_w, int64_t stride_out,
                     int ksplit, int variant, void* stream);

/* ------------------------------------------------------------------------ *
 * Paged decode attention — replaces the flash_attn_with_kvcache call at
 * core/layer/model_graph.py:161-168 and core/paged_attention.py:281-293
 * (seqlen_q == 1).  One launch does, for every batch row b:
 *   pos = cache_seqlens[b]
 *   q,k_new rotated at cos/sin[pos] (NeoX halves or interleaved pairs),
 *   rotated k_new and raw v_new written to page block_table[b, pos/bs],
 *   slot pos % bs,
 *   out[b] = softmax(scale * q . K[0..pos]^T) . V[0..pos]     (split-KV,
 *   fp32 online softmax, partials merged by the last-arriving CTA).
 * k_new/v_new may be NULL (no append, no rotary, attends over [0, pos)).
 * cos/sin may be NULL (no rotary).
 *
 * Workspace (caller-owned, sized by b200_decode_workspace_bytes, zeroed ONCE
 * by the caller before first use; the kernel leaves it reusable, also by
 * launches of other shapes):
 *   [ 65536 arrival counters (fixed) | lse partials | O partials ]
 * ------------------------------------------------------------------------ */
typedef struct b200_decode_args {
    const void* q;            /* [B, Hq, D] bf16                                  */
    const void* k_new;        /* [B, Hkv, D] bf16 or NULL                         */
    const void* v_new;        /* [B, Hkv, D] bf16 or NULL                         */
    void*       k_cache;      /* [n_blocks, block_size, Hkv, D] bf16              */
    void*       v_cache;
    const void* rotary_cos;   /* [rot_rows, rotary_dim/2] bf16 or NULL            */
    const void* rotary_sin;
    const int32_t* cache_seqlens; /* [B]                                          */
    const int32_t* block_table;   /* [B, max_blocks_per_seq]                      */
    void*       out;          /* [B, Hq, D] bf16                                  */
    float*      softmax_lse;  /* optional [B, Hq] fp32 (natural log) or NULL      */
    void*       workspace;    /* see b200_decode_workspace_bytes                  */
    int64_t     workspace_bytes;
    int32_t B, Hq, Hkv, D;
    int32_t rotary_dim;       /* 0 = no rotary; multiple of 16, <= D              */
    int32_t rotary_interleaved; /* 0 = NeoX halves (model_graph.py:166), 1 = GPT-J pairs (flash-attn default, paged_attention.py:281) */
    int32_t rot_rows;         /* rows in the cos/sin tables                       */
    int32_t n_blocks, block_size, max_blocks_per_seq;
    int64_t q_stride_b, q_stride_h;
    int64_t k_stride_b, k_stride_h;
    int64_t v_stride_b, v_stride_h;
    int64_t o_stride_b, o_stride_h;
    int64_t kc_stride_blk, kc_stride_tok, kc_stride_h;   /* k_cache strides */
    int64_t vc_stride_blk, vc_stride_tok, vc_stride_h;   /* v_cache strides */
    int64_t bt_stride_b;      /* block_table row stride                           */
    int32_t num_splits;       /* >= 1; grid is (num_splits, Hkv*groups, B)        */
    int32_t variant;          /* tuning variant, 0 = default                      */
    float   softmax_scale;
    int32_t flags;            /* B200_DECODE_FLAG_*                               */
} b200_decode_args;

/* Launch with programmatic stream serialization (PDL): the kernel may start, read
 * cache_seqlens / block_table and prefetch cached K/V pages while the PRECEDING kernel in
 * the stream is still draining; q, k_new, v_new, the workspace and all writes wait for it.
 * Only set it when the preceding kernel does not produce cache_seqlens, block_table or
 * cache pages this call reads (true inside the decode layer loop: the predecessor is the
 * QKV projection, model_graph.py:141-148). */
#define B200_DECODE_FLAG_PDL 1
/* with PDL: release the dependent kernel after the streaming loop instead of at kernel start */
#define B200_DECODE_FLAG_PDL_LATE 2

/* Balanced (stream-K) scheduling: a fixed grid of #SMs x resident CTAs, independent of the
 * lengths; the 64-token chunks of all (sequence, kv head) pairs are laid end to end and cut into
 * equal shares, one per CTA, computed on the device from cache_seqlens, so ragged batches
 * (continuous batching) and small batches keep every SM equally busy.  num_splits is ignored;
 * the workspace must hold b200_decode_balanced_workspace_bytes(D) bytes (same zero-once rule);
 * B <= 512. */
#define B200_DECODE_FLAG_BALANCED 4
/* Split counts 2..8 launch the splits of a (sequence, kv head, group) as ONE thread-block cluster and merge the
 * partials through distributed shared memory (no workspace needed, none touched).  This flag selects the older form:
 * fp32 partials in the workspace, merged by the last CTA to arrive on a counter (the only form for > 8 splits). */
#define B200_DECODE_FLAG_NO_CLUSTER 8

int64_t b200_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                                    int32_t num_splits);
int64_t b200_decode_balanced_workspace_bytes(int32_t D);

int b200_paged_attn_decode(const b200_decode_args* args, void* stream);

/* Stand-alone LSE combine of split partials laid out as in the decode
 * workspace: o_part [B, Hq, S, D] fp32 (already normalised per split),
 * lse_part [B, Hq, S] fp32 (natural log, -inf = empty split).  The decode
 * entry point fuses this step; it is exported for callers that produce
 * partials themselves (flash-attn's combine kernel equivalent). */
int b200_lse_combine(const float* o_part, const float* lse_part,
                     void* out, float* lse_out,
                     int32_t B, int32_t Hq, int32_t S, int32_t D,
                     int64_t o_stride_b, int64_t o_stride_h,
                     void* stream);

/* ------------------------------------------------------------------------ *
 * Paged prefill attention — replaces RoPE.forward (core/layer/rope.py:10-39)
 * + the flash_attn_with_kvcache call at core/layer/model_prefill.py:49-52.
 * For every batch row b with S new tokens on top of P_b = cache_seqlens[b]
 * cached tokens:
 *   1. q, k rotated at positions P_b + i (or raw when cos == NULL: the
 *      reference pre-rotates with RoPE.forward),
 *   2. K (rotated) and V scattered into pages through block_table,
 *   3. causal (bottom-right aligned) attention of the S queries over
 *      positions [0, P_b + S) read back from the pages, on tcgen05/TMEM.
 * ------------------------------------------------------------------------ */
typedef struct b200_prefill_args {
    const void* q;            /* [B, S, Hq, D] bf16  */
    const void* k;            /* [B, S, Hkv, D] bf16 */
    const void* v;            /* [B, S, Hkv, D] bf16 */
    void*       k_cache;
    void*       v_cache;
    const void* rotary_cos;   /* optional */
    const void* rotary_sin;
    const int32_t* cache_seqlens; /* [B] prefix lengths P_b                 */
    const int32_t* block_table;   /* [B, max_blocks_per_seq]                */
    void*       out;          /* [B, S, Hq, D] bf16 */
    void*       workspace;    /* b200_prefill_workspace_bytes: rotated Q + the attention kernel's work-item counter */
    int64_t     workspace_bytes;
    int32_t B, S, Hq, Hkv, D;
    int32_t rotary_dim, rotary_interleaved, rot_rows;
    int32_t n_blocks, block_size, max_blocks_per_seq;
    int32_t max_prefix;       /* upper bound on cache_seqlens[b] (grid sizing; table capacity if unknown) */
    int32_t causal;
    int64_t q_stride_b, q_stride_s, q_stride_h;
    int64_t k_stride_b, k_stride_s, k_stride_h;
    int64_t v_stride_b, v_stride_s, v_stride_h;
    int64_t o_stride_b, o_stride_s, o_stride_h;
    int64_t kc_stride_blk, kc_stride_tok, kc_stride_h;
    int64_t vc_stride_blk, vc_stride_tok, vc_stride_h;
    int64_t bt_stride_b;
    int32_t variant;
    float   softmax_scale;
} b200_prefill_args;

int64_t b200_prefill_workspace_bytes(int32_t B, int32_t S, int32_t Hq, int32_t D);

int b200_paged_attn_prefill(const b200_prefill_args* args, void* stream);

/* Debug / profiling hook, not part of the operator surface: while `dev_buf` (device memory,
 * 16 * 64 int64) is non-NULL every prefill launch records clock64 stamps of the pipeline
 * hand-offs of CTA (0,0,0) into it (tools/trace_prefill.py); NULL switches it off. */
void b200_debug_prefill_trace(void* dev_buf);

/* ------------------------------------------------------------------------ *
 * Peer-memory all-reduce fused with the residual add and the RMSNorm that
 * follow it — replaces, in ONE launch per call site,
 *   all_reduce(attn_out)  + rmsnorm_residual_gemm(attn_out, h, ln_2.w, ...)
 *       (core/layer/model_graph.py:210 + :213-216, core/parallel_config.py:65-69)
 *   all_reduce(mlp_out).add_(residual)  + next layer's rmsnorm_(h, ln_1.w, ...)
 *       (core/layer/model_graph.py:219 + :143 / final ln_f :222)
 * Every rank's partial sums x_r [rows, n] bf16 live in a buffer the other
 * ranks of the node can load from directly (NVLink 5 / NVSwitch peer memory,
 * exported with b200_ipc_export and mapped with b200_ipc_import).  Each rank
 * runs the same kernel: a per-CTA flag barrier over peer memory (monotonic
 * epoch, no reset, CUDA-graph replay safe), then
 *   s   = bf16( sum_{r=0..world-1} f32(x_r) )          fixed rank order: all
 *                                                      ranks get identical bits
 *   res_out = bf16( f32(s) + f32(residual) )           (residual may be NULL)
 *   ss  = mean( t^2 ),  t = f32(s)+f32(residual)  (ss_rounded = 0, the
 *         rmsnorm_residual.py:104-108 order) or t = f32(res_out) (ss_rounded
 *         = 1, the add_ -> rmsnorm_ order of model_graph.py:219,:143)
 *   y   = bf16( f32(res_out) * rsqrt(ss + eps) * f32(w) )   (w may be NULL:
 *         no norm, y untouched)
 * world == 1 runs the same arithmetic on the local x without any barrier.
 *
 * Protocol requirements (the Python PeerComm enforces them):
 *   - all ranks issue the same sequence of calls with the same rows / n;
 *   - CONSECUTIVE calls must read x from DIFFERENT buffers (two alternating
 *     slots): the barrier of call k+1 is what tells a rank that its peers
 *     finished reading the buffer of call k;
 *   - flags: uint32 [B200_AR_MAX_CTAS][B200_AR_MAX_WORLD] (+ [B200_AR_MAX_CTAS], row-sharded form) per rank, zeroed
 *     once; epoch: LOCAL uint32 [B200_AR_MAX_CTAS + 1], zeroed once; word
 *     [B200_AR_MAX_CTAS] is set to 1 if a wait timed out (peer missing).
 * ------------------------------------------------------------------------ */
#define B200_AR_MAX_WORLD 8
#define B200_AR_MAX_CTAS  1024

/* OR-ed into `mode`: launch with programmatic stream serialization (the kernel's prologue overlaps the tail of the
 * previous kernel of the stream; it waits for that kernel before reading x / residual).  Every form also releases a
 * PDL-launched SUCCESSOR at its start, so a weight-streaming GEMM behind it fetches weights during the exchange. */
#define B200_AR_FLAG_PDL 0x100
/* OR-ed into `mode` with B200_AR_MODE_NVLS: the owner sends y as 16-byte flag-in-data packets (plain peer stores into
 * every rank's y staging, ystage_peers[], the LL form's region: rows * n * 4 bytes) and every rank unpacks its local copy into `y` (any local
 * tensor) — one one-way hop instead of a release fence + a counter round trip.  epoch must hold MAX_CTAS + 11 words. */
#define B200_AR_FLAG_PACKET_TAIL 0x200

typedef struct b200_ar_args {
    const void* x_peers[B200_AR_MAX_WORLD];  /* rank r's partial sums, bf16 [rows, n] (entry `rank` is local) */
    void*       flag_peers[B200_AR_MAX_WORLD]; /* rank r's flag array (entry `rank` is local)                 */
    void*       epoch;        /* local epoch counters + error word                 */
    const void* residual;     /* local bf16 [rows, n] or NULL                      */
    const void* weight;       /* norm weight [n] (w_dtype) or NULL = no norm       */
    void*       y;            /* bf16 [rows, n] normed output (NULL iff weight NULL) */
    void*       res_out;      /* bf16 [rows, n] reduced (+ residual) output        */
    int64_t rows, n;
    int64_t stride_x, stride_r, stride_y, stride_res;   /* elements */
    float   eps;
    int32_t w_dtype;          /* B200_DT_*                                         */
    int32_t rank, world;      /* world in {1, 2, 4, 8}                             */
    int32_t ss_rounded;       /* see above                                         */
    int32_t timeout_ms;       /* flag wait bound; 0 = default 2000                 */
    int32_t mode;             /* 0 = one-shot, or B200_AR_MODE_ROW_SHARDED         */
    void*   y_peers[B200_AR_MAX_WORLD];  /* row-sharded: rank r's y buffer (entry `rank` == y) */
    void*   xstage_peers[B200_AR_MAX_WORLD];  /* LL form: rank r's staging for pushed partial rows  */
    void*   ystage_peers[B200_AR_MAX_WORLD];  /* LL form: rank r's staging for pushed normalised rows */
    void*   trace;            /* optional uint64 [min(rows, MAX_CTAS)][4]: %globaltimer (ns) at kernel start, barrier
                                 passed, rows reduced, end of every CTA (tools/bench_allreduce.py); NULL = off        */
    const void* x_mc;         /* NVLS form: multicast address of the x slot (same offset as x_peers[rank])            */
    void*       y_mc;         /* NVLS form: multicast address of the y buffer (same offset as y)                      */
    void*       flags_mc;     /* NVLS form: multicast address of the flag region (same offset as flag_peers[rank])    */
    void*       ystage_mc;    /* reserved (multicast address of the y staging; unused: unicast packets are faster)    */
} b200_ar_args;

/* Row-sharded ("two-shot") form for world > 1, rows <= B200_AR_MAX_CTAS: rank r OWNS rows
 * [r*ceil(rows/world), ...).  The owner of a row waits for every peer's per-row "ready" flag, loads the
 * row from all ranks, reduces, adds ITS residual row, normalises, and stores y into EVERY rank's y
 * buffer (peer stores over NVLink), then raises the row's "done" flag on every rank.  Each rank moves
 * (world-1)/world of one activation in each direction instead of (world-1) activations inbound.
 * Contract differences: `residual` is read and `res_out` written ONLY for owned rows (the residual
 * stream stays row-sharded across calls — what the decode loop needs); y must live in peer-visible
 * memory; flags: a second array u32 [B200_AR_MAX_CTAS] follows the first in every rank's flag region
 * (total (B200_AR_MAX_WORLD + 1) * B200_AR_MAX_CTAS words); consecutive calls need NOT alternate
 * buffers (the done flag is the proof that the owner finished reading). */
#define B200_AR_MODE_ROW_SHARDED 1

/* LL push form (flag-in-data): row ownership and the row-sharded residual contract of ROW_SHARDED, but no
 * remote loads, no flags, no fences.  Non-owners store their partial row into the owner's x staging as
 * 16-byte packets {d0, epoch, d1, epoch}; the owner polls its own staging, reduces in rank order,
 * normalises and stores y as packets into every rank's y staging; each rank unpacks its copy into the
 * plain local `y`.  y / x need not alias peer-visible buffers except x_peers[rank] (the local partial sums).
 * Staging (peer visible, zeroed once): x staging world * ceil(rows/world) * n * 4 bytes, y staging
 * rows * n * 4 bytes.  epoch words [B200_AR_MAX_CTAS + 1] (call counter) and [+2] (arrival counter) are
 * used: `epoch` must hold B200_AR_MAX_CTAS + 4 words. */
#define B200_AR_MODE_LL 2

/* NVLS form (in-switch reduction; the form SURVEY.md §8f N1 names): every rank's slab is bound to ONE multicast
 * object (cuMulticastCreate / cuMulticastBindMem; `x_mc`, `y_mc`, `flags_mc` are addresses inside its mapping).
 * Row ownership and the row-sharded residual contract are those of ROW_SHARDED.  The owner of a row reads the SUM
 * of the row over all ranks with multimem.ld_reduce (fp32 accumulation inside the NVSwitch, one rounding to bf16
 * by the switch's own rule: measured within 1 bf16 ulp of the exact sum, not always round-to-nearest-even),
 * adds its residual row, normalises, and multimem.st's y into EVERY rank's y buffer; the two flag hops are
 * multimem.red on monotonic counters (u32 words B200_AR_MAX_CTAS * (B200_AR_MAX_WORLD + 1) and + 32 of every
 * rank's flag region; local targets in epoch words [B200_AR_MAX_CTAS + 4 .. + 6], so `epoch` must hold
 * B200_AR_MAX_CTAS + 8 words).  grid = ceil(rows / world) CTAs per rank.  When the call returns on any rank every
 * owner has finished reading x: the x slot may be rewritten immediately (no slot alternation). */
#define B200_AR_MODE_NVLS 4

/* One-hop NVLS form: EVERY rank asks the switch for the sum of EVERY row (multimem.ld_reduce of the whole slot; at
 * decode sizes that is ~1 MB per GPU, far below NVLink bandwidth), adds its residual, normalises and stores y
 * LOCALLY.  One flag hop (the `ready` counter), no second hop, no trailing barrier: latency of one switch round
 * trip at any world size.  Contract = the one-shot form's: residual / res_out are full replicated tensors, y is any
 * local tensor, and CONSECUTIVE calls must read x from different slots.  The switch rounds the fp32 sum to bf16 with
 * its own rule (measured: within 1 bf16 ulp of the exact sum, not round-to-nearest-even; tools/bench_allreduce.py),
 * identically for every requester, so ranks stay bit-identical.  Needs x_mc and flags_mc. */
#define B200_AR_MODE_NVLS_ONEHOP 8

int b200_allreduce_residual_rmsnorm(const b200_ar_args* args, void* stream);

/* ------------------------------------------------------------------------
 * Column all-gather over peer memory — the collective of a VOCAB-PARALLEL lm_head.  The reference keeps lm_head
 * replicated under tensor parallelism (/root/reference/models/adapter.py:25, core/layer/model_graph.py:223): every
 * rank streams the whole [vocab, hidden] matrix.  With this call every rank multiplies by its vocab / world rows and
 * the [rows, vocab] logits are assembled on EVERY rank: rank r's block src [rows, cols_local] lands at columns
 * [r * cols_local, (r + 1) * cols_local) of every rank's dst (one multimem.st per 16 bytes when dst_mc / flags_mc are
 * given, else `world` peer stores).  The call returns on a rank (stream order) only after every rank's block has
 * landed in that rank's dst.  flag_peers / epoch: the regions of b200_ar_args (one more counter word each).
 * Caller's duty (WAR): a rank's dst may only be overwritten after that rank consumed it; any barrier-carrying
 * collective between two gathers that every rank enters after its consumer kernels (the decode step's reductions)
 * provides that.  flags: bit 0 = launch with programmatic stream serialization.
 * ------------------------------------------------------------------------ */
typedef struct b200_gather_args {
    const void* src;                          /* local bf16 [rows, cols_local]                                 */
    void*       dst_peers[B200_AR_MAX_WORLD]; /* rank r's gather buffer bf16 [rows, world * cols_local]        */
    void*       dst_mc;                       /* multicast address of the gather buffer, or NULL               */
    void*       flag_peers[B200_AR_MAX_WORLD];/* rank r's flag region (as in b200_ar_args)                     */
    void*       flags_mc;                     /* multicast address of the flag region, or NULL (with dst_mc)   */
    void*       epoch;                        /* local epoch words (as in b200_ar_args; needs MAX_CTAS + 10)   */
    int64_t rows, cols_local;
    int64_t stride_src, stride_dst;           /* elements                                                      */
    int32_t rank, world;
    int32_t timeout_ms;                       /* 0 = default 2000                                              */
    int32_t flags;
} b200_gather_args;
int b200_allgather_cols(const b200_gather_args* args, void* stream);

/* Peer-memory plumbing (init time only; these DO allocate / synchronise).
 * b200_comm_alloc: cudaMalloc + zero fill of an IPC-exportable buffer on the current device.
 * b200_ipc_export: 64-byte handle another process of the node can map.
 * b200_ipc_import: map a peer's buffer (enables peer access lazily); b200_ipc_close unmaps. */
#define B200_IPC_HANDLE_BYTES 64
int b200_comm_alloc(int64_t bytes, void** dev_ptr);
int b200_comm_free(void* dev_ptr);
int b200_ipc_export(const void* dev_ptr, void* handle_out);
int b200_ipc_import(const void* handle, void** peer_ptr);
int b200_ipc_close(void* peer_ptr);

#ifdef __cplusplus
}
#endif
#endif /* B200_ABI_H */
