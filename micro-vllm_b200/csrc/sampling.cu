// Greedy token selection inside the decode graph.
//   reference: core/layer/sampler.py:21-43 evaluated at top_k = 1 (torch.topk(logits / temp, 1) ->
//   the index of the row maximum; softmax / top-p / multinomial over ONE candidate are the identity),
//   driven from core/engine.py:276-280 after ModelGraphRunner.forward returned the logits.
// The reference leaves the graph, runs an Inductor-compiled sampler and broadcasts the ids; here the
// row arg-max is one more launch at the tail of the captured decode step, so the ids are produced on
// every rank (logits are replicated) with no host round trip.
//
// HBM-bound: rows * vocab * sizeof(logit) bytes read once (Qwen2.5 vocab 152 064 x 32 rows bf16 =
// 9.7 MB).  grid = (column chunks, rows) so that decode-sized batches still cover every SM; each
// thread keeps a 64-bit key (order-preserving image of the value << 32 | ~index) so the maximum of
// keys is the FIRST index of the largest value; chunks merge with one atomicMax per CTA and the last
// CTA of a row (arrival counter) publishes the id and re-zeroes the scratch -> graph-replay safe.
#include "common.cuh"
#include <math_constants.h>

namespace b200 {

constexpr int AM_THREADS = 256;
constexpr int AM_VPT = 4;                                  // 16-byte vectors in flight per thread
constexpr int AM_CHUNK_VEC = AM_THREADS * AM_VPT;          // vectors per CTA

// order-preserving map of fp32 bits to u32 (larger float -> larger unsigned); NaNs with the sign bit
// clear order above +inf, like torch.topk which treats NaN as the largest value
__device__ __forceinline__ uint32_t orderable(uint32_t bits) {
    return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}

template <typename T> struct LogitVec;
template <> struct LogitVec<__nv_bfloat16> {
    static constexpr int N = 8;
    __device__ static void keys(const uint4& v, uint32_t (&o)[8]) {
        const uint32_t* w = &v.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) { o[2 * i] = orderable(w[i] << 16); o[2 * i + 1] = orderable(w[i] & 0xffff0000u); }
    }
    __device__ static uint32_t key1(const __nv_bfloat16* p) {
        return orderable((uint32_t)(*reinterpret_cast<const uint16_t*>(p)) << 16);
    }
};
template <> struct LogitVec<float> {
    static constexpr int N = 4;
    __device__ static void keys(const uint4& v, uint32_t (&o)[4]) {
        o[0] = orderable(v.x); o[1] = orderable(v.y); o[2] = orderable(v.z); o[3] = orderable(v.w);
    }
    __device__ static uint32_t key1(const float* p) { return orderable(__float_as_uint(*p)); }
};

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
        k = other > k ? other : k;
    }
    return k;
}

// VEC: 16-byte loads (base pointer and row stride 16-byte aligned); the ragged tail of a row (vocab not
// a multiple of the vector width) is handled element-wise by the last chunk.
template <typename T, bool VEC>
__global__ void __launch_bounds__(AM_THREADS)
greedy_argmax_kernel(const T* __restrict__ logits, int64_t stride, int vocab, int nchunks,
                     unsigned long long* __restrict__ keys, unsigned int* __restrict__ arrivals,
                     int64_t* __restrict__ out_ids) {
    constexpr int N = LogitVec<T>::N;
    const int row = blockIdx.y;
    const T* __restrict__ src = logits + (int64_t)row * stride;
    uint32_t best = 0, bidx = 0;                            // key 0 sorts below every finite value
    if (VEC) {
        const int nvec = vocab / N;
        const int v0 = blockIdx.x * AM_CHUNK_VEC + threadIdx.x;
        uint4 v[AM_VPT];
#pragma unroll
        for (int u = 0; u < AM_VPT; ++u) {
            const int c = v0 + u * AM_THREADS;
            if (c < nvec) v[u] = ld_nc_v4(src + (int64_t)c * N);
        }
#pragma unroll
        for (int u = 0; u < AM_VPT; ++u) {                  // increasing index order: strict > keeps the first
            const int c = v0 + u * AM_THREADS;
            if (c < nvec) {
                uint32_t k[N];
                LogitVec<T>::keys(v[u], k);
#pragma unroll
                for (int i = 0; i < N; ++i)
                    if (k[i] > best) { best = k[i]; bidx = (uint32_t)(c * N + i); }
            }
        }
        if (blockIdx.x == nchunks - 1) {
            for (int i = nvec * N + threadIdx.x; i < vocab; i += AM_THREADS) {
                const uint32_t k = LogitVec<T>::key1(src + i);
                if (k > best) { best = k; bidx = (uint32_t)i; }
            }
        }
    } else {
        const int lo = blockIdx.x * (AM_CHUNK_VEC * N);
        const int hi = min(vocab, lo + AM_CHUNK_VEC * N);
        for (int i = lo + threadIdx.x; i < hi; i += AM_THREADS) {
            const uint32_t k = LogitVec<T>::key1(src + i);
            if (k > best) { best = k; bidx = (uint32_t)i; }
        }
    }
    unsigned long long key = ((unsigned long long)best << 32) | (0xffffffffu - bidx);
    key = warp_max_u64(key);
    __shared__ unsigned long long s_key[AM_THREADS / 32];
    if ((threadIdx.x & 31) == 0) s_key[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 1; w < AM_THREADS / 32; ++w) key = s_key[w] > key ? s_key[w] : key;
        if (nchunks == 1) {
            out_ids[row] = (int64_t)(0xffffffffu - (uint32_t)key);
            return;
        }
        atomicMax(keys + row, key);
        __threadfence();
        const unsigned int prev = atomicAdd(arrivals + row, 1u);
        if (prev == (unsigned int)nchunks - 1) {            // every other CTA's atomicMax is visible
            __threadfence();
            const unsigned long long fin = atomicExch(keys + row, 0ull);
            arrivals[row] = 0;
            out_ids[row] = (int64_t)(0xffffffffu - (uint32_t)fin);
        }
    }
}


// ------------------------------------------------------------------ top-k / top-p sampling
// The reference sampler (core/layer/sampler.py:21-43, called with top_k = 50 for decode and 1000 for prefill,
// core/engine.py:233, :278):  x = logits / temp;  vals, idxs = topk(x, k);  probs = softmax(vals);  sorted descending;
// cum = cumsum(probs);  valid = cum < top_p, valid[0] = True;  renormalise;  multinomial(1);  map back to token ids.
// One CTA per row does all of it in ONE launch (the reference leaves the graph for an Inductor-compiled function):
//   1. radix select on the order-preserving integer image of the logits (8-bit digits, 2 passes for bf16, 4 for
//      fp32; the row is re-read from L2) -> the k-th largest key;  2. ordered collection of the k candidates (keys
//      above the threshold, then the lowest-index ties: deterministic);  3. bitonic sort in shared memory by
//      (value descending, index ascending);  4. fp32 softmax of logit / temp, inclusive scan, top-p prefix,
//      renormalisation;  5. inverse-CDF draw with the caller's uniform number u[row] in [0, 1) — the distribution of
//      torch.multinomial over the same probabilities (the reference's RNG stream itself is torch-internal).
constexpr int TK_THREADS = 1024;
constexpr int TK_MAX_K = 1024;
constexpr int TK_EPT = 8;                                  // consecutive elements per thread in the ordered pass

template <typename T> struct TkKey;
template <> struct TkKey<__nv_bfloat16> {
    static constexpr int BITS = 16;
    __device__ static uint32_t key(const __nv_bfloat16* p) {
        return orderable((uint32_t)(*reinterpret_cast<const uint16_t*>(p)) << 16) >> 16;
    }
    __device__ static float val(const __nv_bfloat16* p) { return __bfloat162float(*p); }
};
template <> struct TkKey<float> {
    static constexpr int BITS = 32;
    __device__ static uint32_t key(const float* p) { return orderable(__float_as_uint(*p)); }
    __device__ static float val(const float* p) { return *p; }
};

__device__ __forceinline__ float block_sum_1024(float v, float* s_red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = (threadIdx.x < 32) ? s_red[threadIdx.x] : 0.0f;
    if (threadIdx.x < 32) t = warp_sum(t);
    if (threadIdx.x == 0) s_red[32] = t;
    __syncthreads();
    return s_red[32];
}

// inclusive scan over the 1024 threads of the CTA (value of thread i = sum of threads 0..i)
template <typename V>
__device__ __forceinline__ V block_scan_incl_1024(V v, V* s_w) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const V t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    __syncthreads();
    if (lane == 31) s_w[w] = v;
    __syncthreads();
    if (w == 0) {
        V t = s_w[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const V u = __shfl_up_sync(0xffffffffu, t, o);
            if (lane >= o) t += u;
        }
        s_w[lane] = t;
    }
    __syncthreads();
    return v + (w > 0 ? s_w[w - 1] : (V)0);
}

template <typename T>
__global__ void __launch_bounds__(TK_THREADS)
topk_topp_sample_kernel(const T* __restrict__ logits, int64_t stride, int vocab, int k,
                        const float* __restrict__ temps, const float* __restrict__ top_ps,
                        const float* __restrict__ uniforms, int64_t* __restrict__ out_ids,
                        int32_t* __restrict__ dbg_idx, float* __restrict__ dbg_prob) {
    constexpr int BITS = TkKey<T>::BITS;
    __shared__ unsigned int s_hist[256];
    __shared__ unsigned int s_u32[40];                      // scan scratch [0..31], misc [32..]
    __shared__ float s_f32[40];
    __shared__ unsigned long long s_cand[TK_MAX_K];
    const int row = blockIdx.x, tid = threadIdx.x;
    const T* __restrict__ src = logits + (int64_t)row * stride;
    if (k > vocab) k = vocab;

    // ---- 1. radix select: the k-th largest key T (count(key > T) < k <= count(key >= T))
    uint32_t prefix = 0;
    int kk = k;                                             // rank still to find inside the current prefix bucket
#pragma unroll 1
    for (int shift = BITS - 8; shift >= 0; shift -= 8) {
        if (tid < 256) s_hist[tid] = 0;
        __syncthreads();
        const int hs = shift + 8;                           // bits above the current digit
        for (int i = tid; i < vocab; i += TK_THREADS) {
            const uint32_t key = TkKey<T>::key(src + i);
            if (hs >= BITS || (key >> hs) == (prefix >> hs)) atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            unsigned int above = 0;
            int d = 255;
            for (; d > 0; --d) {
                if (above + s_hist[d] >= (unsigned int)kk) break;
                above += s_hist[d];
            }
            s_u32[32] = (uint32_t)d;
            s_u32[33] = above;
        }
        __syncthreads();
        prefix |= s_u32[32] << shift;
        kk -= (int)s_u32[33];
        __syncthreads();
    }
    const uint32_t thr = prefix;
    const int n_gt = k - kk, need_eq = kk;                  // candidates above the threshold / ties to take (lowest indices)

    // ---- 2. ordered collection: keys > thr anywhere (their order is fixed by the sort), keys == thr in index order
    if (tid == 0) { s_u32[34] = 0; s_u32[35] = 0; }         // [34] = #greater collected, [35] = ties seen so far
    for (int i = tid; i < TK_MAX_K; i += TK_THREADS) s_cand[i] = 0ull;
    __syncthreads();
    for (int base = 0; base < vocab; base += TK_THREADS * TK_EPT) {
        const int i0 = base + tid * TK_EPT;
        uint32_t keys[TK_EPT];
        unsigned int n_eq = 0;
#pragma unroll
        for (int j = 0; j < TK_EPT; ++j) {
            const int i = i0 + j;
            keys[j] = 0;
            if (i < vocab) {
                const uint32_t key = TkKey<T>::key(src + i);
                keys[j] = key;
                if (key > thr) {
                    const unsigned int pos = atomicAdd(&s_u32[34], 1u);
                    if (pos < (unsigned int)TK_MAX_K)
                        s_cand[pos] = ((unsigned long long)key << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i);
                } else if (key == thr) {
                    ++n_eq;
                }
            }
        }
        const unsigned int seen = s_u32[35];                // ties before this tile (written behind the scan's barriers)
        if (seen < (unsigned int)need_eq) {                 // uniform across the CTA
            const unsigned int incl = block_scan_incl_1024<unsigned int>(n_eq, s_u32);
            unsigned int rank = seen + incl - n_eq;
#pragma unroll
            for (int j = 0; j < TK_EPT; ++j) {
                const int i = i0 + j;
                if (i < vocab && keys[j] == thr) {
                    if (rank < (unsigned int)need_eq)
                        s_cand[n_gt + rank] = ((unsigned long long)thr << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i);
                    ++rank;
                }
            }
            __syncthreads();
            if (tid == TK_THREADS - 1) s_u32[35] = seen + incl;
            __syncthreads();
        }
    }
    __syncthreads();

    // ---- 3. bitonic sort, descending by (key, ~index): largest value first, ties by lowest index
    int n = 32;
    while (n < k) n <<= 1;
    for (int size = 2; size <= n; size <<= 1) {
        for (int st = size >> 1; st > 0; st >>= 1) {
            __syncthreads();
            const int i = tid, j = tid ^ st;
            if (i < n && j > i) {
                const unsigned long long a = s_cand[i], b = s_cand[j];
                const bool desc = (i & size) == 0;
                if (desc ? (a < b) : (a > b)) { s_cand[i] = b; s_cand[j] = a; }
            }
        }
    }
    __syncthreads();

    // ---- 4. softmax of logit / temp over the k candidates (fp32), top-p prefix, renormalisation
    const float temp = temps ? temps[row] : 1.0f;
    const float top_p = top_ps ? top_ps[row] : 1.0f;
    const bool live = tid < k;
    const uint32_t my_idx = live ? (0xffffffffu - (uint32_t)(s_cand[tid] & 0xffffffffull)) : 0u;
    const float x = live ? TkKey<T>::val(src + my_idx) / temp : -CUDART_INF_F;
    if (tid == 0) s_f32[33] = x;                            // sorted: candidate 0 holds the maximum
    __syncthreads();
    const float e = live ? expf(x - s_f32[33]) : 0.0f;
    const float denom = block_sum_1024(e, s_f32);
    const float p = e / denom;
    const float cum = block_scan_incl_1024<float>(live ? p : 0.0f, s_f32);
    const bool valid = live && (tid == 0 || cum < top_p);   // reference :33-35
    const float kept = block_sum_1024(valid ? p : 0.0f, s_f32);
    const float pn = valid ? p / kept : 0.0f;
    if (dbg_idx) {
        if (tid < k) { dbg_idx[(int64_t)row * k + tid] = (int32_t)my_idx; dbg_prob[(int64_t)row * k + tid] = pn; }
    }
    // ---- 5. inverse CDF with the caller's uniform number: first candidate whose cumulative probability exceeds u
    const float cn = block_scan_incl_1024<float>(pn, s_f32);
    const float u = uniforms ? uniforms[row] : 0.0f;
    const unsigned int below = block_scan_incl_1024<unsigned int>((valid && cn <= u) ? 1u : 0u, s_u32);
    const unsigned int n_valid = block_scan_incl_1024<unsigned int>(valid ? 1u : 0u, s_u32);
    if (tid == TK_THREADS - 1) {
        unsigned int pick = below;                          // number of valid candidates entirely below u
        if (pick >= n_valid) pick = n_valid - 1;            // u beyond the rounded total: the last kept candidate
        s_u32[36] = pick;
    }
    __syncthreads();
    if (tid == (int)s_u32[36]) out_ids[row] = (int64_t)my_idx;
}

}  // namespace b200
using namespace b200;

extern "C" int64_t b200_greedy_argmax_workspace_bytes(int64_t rows) {
    return rows < 0 ? -1 : rows * 16;
}

extern "C" int b200_greedy_argmax(const void* logits, int logits_dtype, int64_t rows, int64_t vocab,
                                  int64_t stride, int64_t* out_ids, void* workspace, int64_t workspace_bytes,
                                  void* stream) {
    if (!logits || !out_ids) B200_FAIL(B200_E_NULL, "greedy_argmax: null pointer");
    if (rows < 0 || rows > 65535 || vocab <= 0 || vocab > 0x7fffffffLL || stride < vocab)
        B200_FAIL(B200_E_RANGE, "greedy_argmax: rows=%lld vocab=%lld stride=%lld", (long long)rows, (long long)vocab,
                  (long long)stride);
    if (logits_dtype != B200_DT_BF16 && logits_dtype != B200_DT_F32)
        B200_FAIL(B200_E_UNSUPPORTED, "greedy_argmax: logits dtype %d", logits_dtype);
    if (rows == 0) return 0;
    const int n = logits_dtype == B200_DT_BF16 ? 8 : 4;
    const int esz = logits_dtype == B200_DT_BF16 ? 2 : 4;
    const int64_t per_cta = (int64_t)AM_CHUNK_VEC * n;
    const int nchunks = (int)((vocab + per_cta - 1) / per_cta);
    if (nchunks > 1) {
        if (!workspace) B200_FAIL(B200_E_NULL, "greedy_argmax: workspace required for vocab > %lld", (long long)per_cta);
        if (workspace_bytes < rows * 16 || (reinterpret_cast<uintptr_t>(workspace) & 7))
            B200_FAIL(B200_E_RANGE, "greedy_argmax: workspace of %lld bytes (8-byte aligned) needed, got %lld",
                      (long long)(rows * 16), (long long)workspace_bytes);
    }
    auto keys = (unsigned long long*)workspace;
    auto arrivals = (unsigned int*)((char*)workspace + rows * 8);
    const bool vec = aligned16(logits) && ((stride * esz) % 16 == 0) && vocab >= n;
    cudaStream_t st = (cudaStream_t)stream;
    const dim3 grid((unsigned)nchunks, (unsigned)rows);
    if (logits_dtype == B200_DT_BF16) {
        auto L = (const __nv_bfloat16*)logits;
        if (vec) greedy_argmax_kernel<__nv_bfloat16, true><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
        else     greedy_argmax_kernel<__nv_bfloat16, false><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
    } else {
        auto L = (const float*)logits;
        if (vec) greedy_argmax_kernel<float, true><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
        else     greedy_argmax_kernel<float, false><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

// ------------------------------------------------------------------ arg-max for callers that cannot allocate
// Same kernel as b200_greedy_argmax, but the per-row merge scratch (one 64-bit key + one arrival counter per row)
// lives in the library image instead of a caller buffer.  Meant for hosts that have no way to hand over device
// scratch: a serving process whose KV pool already took every spare byte of HBM (a late cudaMalloc would fail),
// a decode loop that cannot afford cudaMalloc's device-wide synchronisation, or a cgo / JNI binding that only
// owns its input and output buffers.  Every GPU that loads libb200ops.so carries these 64 KiB until exit.
static __device__ __align__(256) unsigned char g_argmax_static_scratch[B200_ARGMAX_NOALLOC_MAX_ROWS * 16];

namespace {
constexpr int kMaxDevices = 64;
// Device address of g_argmax_static_scratch per device ordinal.  In host code the symbol name above is only a
// handle for the runtime, not something a kernel can dereference, so the real address is looked up once per device
// (with that device current) and cached here.  Racing first calls resolve the same value: the store is benign.
void* g_argmax_static_ptr[kMaxDevices] = {};

int argmax_static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { ::b200::set_error("greedy_argmax_noalloc: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
    if (dev < 0 || dev >= kMaxDevices) B200_FAIL(B200_E_RANGE, "greedy_argmax_noalloc: device ordinal %d (0..%d)", dev, kMaxDevices - 1);
    void* p = __atomic_load_n(&g_argmax_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress(&p, g_argmax_static_scratch);
        if (e != cudaSuccess || !p) {
            ::b200::set_error("greedy_argmax_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
        }
        __atomic_store_n(&g_argmax_static_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}
}  // namespace

extern "C" int b200_greedy_argmax_noalloc(const void* logits, int logits_dtype, int64_t rows, int64_t vocab,
                                          int64_t stride, int64_t* out_ids, void* stream) {
    if (!logits || !out_ids) B200_FAIL(B200_E_NULL, "greedy_argmax_noalloc: null pointer");
    if (rows < 0 || vocab <= 0 || vocab > 0x7fffffffLL || stride < vocab)
        B200_FAIL(B200_E_RANGE, "greedy_argmax_noalloc: rows=%lld vocab=%lld stride=%lld", (long long)rows,
                  (long long)vocab, (long long)stride);
    if (rows > B200_ARGMAX_NOALLOC_MAX_ROWS)
        B200_FAIL(B200_E_RANGE, "greedy_argmax_noalloc: rows=%lld exceeds the built-in scratch (%d rows); use "
                  "b200_greedy_argmax with a caller workspace", (long long)rows, B200_ARGMAX_NOALLOC_MAX_ROWS);
    if (logits_dtype != B200_DT_BF16 && logits_dtype != B200_DT_F32)
        B200_FAIL(B200_E_UNSUPPORTED, "greedy_argmax_noalloc: logits dtype %d", logits_dtype);
    if (rows == 0) return 0;
    void* scratch = nullptr;
    if (int rc = argmax_static_scratch(&scratch)) return rc;     // first call per device: resolves the address
    cudaStream_t st = (cudaStream_t)stream;
    // The array is shared by every call and keeps whatever the previous one left (an aborted launch could leave
    // keys or counters behind), so the part this call uses is zeroed in stream order before the kernel reads it.
    cudaError_t e = cudaMemsetAsync(scratch, 0, (size_t)rows * 16, st);
    if (e != cudaSuccess) { ::b200::set_error("greedy_argmax_noalloc: clearing scratch: %s", cudaGetErrorString(e)); return (int)e; }
    const int n = logits_dtype == B200_DT_BF16 ? 8 : 4;
    const int esz = logits_dtype == B200_DT_BF16 ? 2 : 4;
    const int64_t per_cta = (int64_t)AM_CHUNK_VEC * n;
    const int nchunks = (int)((vocab + per_cta - 1) / per_cta);
    auto keys = (unsigned long long*)scratch;
    auto arrivals = (unsigned int*)((char*)scratch + rows * 8);
    const bool vec = aligned16(logits) && ((stride * esz) % 16 == 0) && vocab >= n;
    const dim3 grid((unsigned)nchunks, (unsigned)rows);
    if (logits_dtype == B200_DT_BF16) {
        auto L = (const __nv_bfloat16*)logits;
        if (vec) greedy_argmax_kernel<__nv_bfloat16, true><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
        else     greedy_argmax_kernel<__nv_bfloat16, false><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
    } else {
        auto L = (const float*)logits;
        if (vec) greedy_argmax_kernel<float, true><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
        else     greedy_argmax_kernel<float, false><<<grid, AM_THREADS, 0, st>>>(L, stride, (int)vocab, nchunks, keys, arrivals, out_ids);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_topk_topp_sample(const void* logits, int logits_dtype, int64_t rows, int64_t vocab, int64_t stride,
                                     int top_k, const float* temperatures, const float* top_ps, const float* uniforms,
                                     int64_t* out_ids, int32_t* dbg_idx, float* dbg_prob, void* stream) {
    if (!logits || !out_ids) B200_FAIL(B200_E_NULL, "topk_topp_sample: null pointer");
    if (rows < 0 || vocab <= 0 || vocab > 0x7fffffffLL || stride < vocab)
        B200_FAIL(B200_E_RANGE, "topk_topp_sample: rows=%lld vocab=%lld stride=%lld", (long long)rows, (long long)vocab, (long long)stride);
    if (top_k < 1 || top_k > TK_MAX_K) B200_FAIL(B200_E_RANGE, "topk_topp_sample: top_k=%d (1..%d)", top_k, TK_MAX_K);
    if ((dbg_idx == nullptr) != (dbg_prob == nullptr)) B200_FAIL(B200_E_NULL, "topk_topp_sample: dbg_idx and dbg_prob go together");
    if (logits_dtype != B200_DT_BF16 && logits_dtype != B200_DT_F32)
        B200_FAIL(B200_E_UNSUPPORTED, "topk_topp_sample: logits dtype %d", logits_dtype);
    if (rows == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (logits_dtype == B200_DT_BF16)
        topk_topp_sample_kernel<__nv_bfloat16><<<(unsigned)rows, TK_THREADS, 0, st>>>(
            (const __nv_bfloat16*)logits, stride, (int)vocab, top_k, temperatures, top_ps, uniforms, out_ids, dbg_idx, dbg_prob);
    else
        topk_topp_sample_kernel<float><<<(unsigned)rows, TK_THREADS, 0, st>>>(
            (const float*)logits, stride, (int)vocab, top_k, temperatures, top_ps, uniforms, out_ids, dbg_idx, dbg_prob);
    B200_CHECK_LAUNCH();
    return 0;
}
