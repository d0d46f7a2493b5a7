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
