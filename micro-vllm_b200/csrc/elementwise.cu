// Bandwidth-bound elementwise / row-reduction kernels of the hot path:
//   SwiGLU               (reference: kernel/swiglu.py:6-46)
//   RMSNorm              (reference: kernel/rmsnorm_add.py:9-34, :133-167)
//   residual + RMSNorm   (reference: kernel/rmsnorm_residual.py:74-124)
//   store_kvcache        (reference: core/cache_manager.py:74-140)
// All of them are HBM-bound: 128-bit loads/stores, rows kept in registers so
// every input byte is read from HBM exactly once, fp32 math, one RNE rounding
// per output element.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>

namespace b200 {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ============================================================== SwiGLU
// grid = (column chunks, rows): no integer division on the hot path.  One thread owns VPT
// 16-byte (8 x bf16) output vectors of one row, strided by the block so that every warp access
// is a contiguous 512 B; all 2*VPT loads are issued before the first use.
template <int VPT>
__global__ void __launch_bounds__(256)
swiglu_vec_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ out,
                  int dv /* d/8 */, int64_t stride_in, int64_t stride_out) {
    const int64_t row = blockIdx.y;
    const int c0 = blockIdx.x * (VPT * 256) + threadIdx.x;
    const __nv_bfloat16* up_p = gu + row * stride_in;
    const __nv_bfloat16* gate_p = up_p + (int64_t)dv * 8;
    uint4 up[VPT], gate[VPT];
#pragma unroll
    for (int u = 0; u < VPT; ++u) {
        const int c = c0 + u * 256;
        if (c < dv) {
            up[u] = ld_nc_v4(up_p + (int64_t)c * 8);
            gate[u] = ld_nc_v4(gate_p + (int64_t)c * 8);
        }
    }
#pragma unroll
    for (int u = 0; u < VPT; ++u) {
        const int c = c0 + u * 256;
        if (c < dv) {
            const uint32_t* pu = &up[u].x;
            const uint32_t* pg = &gate[u].x;
            uint4 o;
            uint32_t* po = &o.x;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float g0 = bf16lo(pg[i]), g1 = bf16hi(pg[i]);
                const float u0 = bf16lo(pu[i]), u1 = bf16hi(pu[i]);
                // silu(g) = g * sigmoid(g) = g / (1 + exp(-g))   (swiglu.py:40-42)
                const float s0 = __fdividef(g0, 1.0f + __expf(-g0));
                const float s1 = __fdividef(g1, 1.0f + __expf(-g1));
                po[i] = pack_bf16(s0 * u0, s1 * u1);
            }
            st_na_v4(out + row * stride_out + (int64_t)c * 8, o);
        }
    }
}

__global__ void __launch_bounds__(256)
swiglu_scalar_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ out,
                     int64_t rows, int64_t d, int64_t stride_in, int64_t stride_out) {
    const int64_t total = rows * d;
    for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * 256) {
        const int64_t r = idx / d, c = idx - r * d;
        const float u = __bfloat162float(gu[r * stride_in + c]);
        const float g = __bfloat162float(gu[r * stride_in + d + c]);
        const float s = g * (1.0f / (1.0f + __expf(-g)));
        out[r * stride_out + c] = __float2bfloat16_rn(s * u);
    }
}

// ============================================================== RMSNorm family
template <typename WT> struct WLoad;
template <> struct WLoad<__nv_bfloat16> {
    // 8 weights starting at element offset e (multiple of 8)
    static __device__ __forceinline__ void load8(const __nv_bfloat16* w, int64_t e, float (&o)[8]) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(w + e));
        const uint32_t* p = &v.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) { o[2 * i] = bf16lo(p[i]); o[2 * i + 1] = bf16hi(p[i]); }
    }
    static __device__ __forceinline__ float load1(const __nv_bfloat16* w, int64_t e) { return __bfloat162float(w[e]); }
};
template <> struct WLoad<float> {
    static __device__ __forceinline__ void load8(const float* w, int64_t e, float (&o)[8]) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(w + e));
        const float4 b = __ldg(reinterpret_cast<const float4*>(w + e + 4));
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    }
    static __device__ __forceinline__ float load1(const float* w, int64_t e) { return w[e]; }
};

__device__ __forceinline__ float block_sum(float v, float* red /* [32] smem */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    if (nw == 1) return v;
    if (lane == 0) red[wid] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : 0.0f;
    t = warp_sum(t);
    return t;
}

// One CTA per row; thread t owns vectors t, t+T, ... (VPT of them) and keeps
// them in registers between the sum-of-squares pass and the scale pass.
template <int VPT, bool RESIDUAL, typename WT>
__global__ void __launch_bounds__(VPT == 4 ? 512 : 1024)
rmsnorm_row_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ r,
                   const WT* __restrict__ w, __nv_bfloat16* __restrict__ y,
                   __nv_bfloat16* __restrict__ res_out,
                   int64_t n, int64_t sx, int64_t sr, int64_t sy, int64_t sres, float eps) {
    __shared__ float red[32];
    const int64_t row = blockIdx.x;
    const int nvec = (int)(n >> 3);
    const int T = blockDim.x;
    const __nv_bfloat16* xr = x + row * sx;
    uint4 xv[VPT], rv[VPT];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * T;
        if (v < nvec) {
            xv[i] = ld_v4(xr + (int64_t)v * 8);
            if (RESIDUAL) rv[i] = ld_v4(r + row * sr + (int64_t)v * 8);
        }
    }
    float ss = 0.0f;
    float f[VPT][8];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * T;
        if (v < nvec) {
            const uint32_t* px = &xv[i].x;
            const uint32_t* pr = &rv[i].x;
            uint4 ro;
            uint32_t* pro = &ro.x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float a = bf16lo(px[k]), b = bf16hi(px[k]);
                if (RESIDUAL) {
                    a += bf16lo(pr[k]);
                    b += bf16hi(pr[k]);
                    // mean of squares uses the UNROUNDED sum (rmsnorm_residual.py:104-108)
                    ss += a * a;
                    ss += b * b;
                    const uint32_t pk = pack_bf16(a, b);
                    pro[k] = pk;
                    // second pass re-reads the bf16-ROUNDED residual (rmsnorm_residual.py:119)
                    a = bf16lo(pk); b = bf16hi(pk);
                } else {
                    ss += a * a;
                    ss += b * b;
                }
                f[i][2 * k] = a; f[i][2 * k + 1] = b;
            }
            if (RESIDUAL) st_na_v4(res_out + row * sres + (int64_t)v * 8, ro);
        }
    }
    const float tot = block_sum(ss, red);
    const float rrms = rsqrtf(tot / (float)n + eps);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * T;
        if (v < nvec) {
            float wf[8];
            WLoad<WT>::load8(w, (int64_t)v * 8, wf);
            uint4 o;
            uint32_t* po = &o.x;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                po[k] = pack_bf16(f[i][2 * k] * rrms * wf[2 * k], f[i][2 * k + 1] * rrms * wf[2 * k + 1]);
            st_na_v4(y + row * sy + (int64_t)v * 8, o);
        }
    }
}

// Generic fallback (n % 8 != 0, unaligned rows, or n too large for registers):
// two passes over global memory exactly like the reference kernels.
template <bool RESIDUAL, typename WT>
__global__ void __launch_bounds__(256)
rmsnorm_generic_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ r,
                       const WT* __restrict__ w, __nv_bfloat16* __restrict__ y,
                       __nv_bfloat16* __restrict__ res_out,
                       int64_t n, int64_t sx, int64_t sr, int64_t sy, int64_t sres, float eps) {
    __shared__ float red[32];
    const int64_t row = blockIdx.x;
    float ss = 0.0f;
    for (int64_t c = threadIdx.x; c < n; c += blockDim.x) {
        float a = __bfloat162float(x[row * sx + c]);
        if (RESIDUAL) {
            a += __bfloat162float(r[row * sr + c]);
            res_out[row * sres + c] = __float2bfloat16_rn(a);
        }
        ss += a * a;
    }
    const float tot = block_sum(ss, red);
    const float rrms = rsqrtf(tot / (float)n + eps);
    for (int64_t c = threadIdx.x; c < n; c += blockDim.x) {
        // each thread re-reads only elements it wrote itself -> no barrier needed
        const float a = RESIDUAL ? __bfloat162float(res_out[row * sres + c])
                                 : __bfloat162float(x[row * sx + c]);
        y[row * sy + c] = __float2bfloat16_rn(a * rrms * WLoad<WT>::load1(w, c));
    }
}

template <bool RESIDUAL, typename WT>
static int launch_rmsnorm(const void* x, const void* r, const void* w, void* y, void* res_out,
                          int64_t rows, int64_t n, int64_t sx, int64_t sr, int64_t sy, int64_t sres,
                          float eps, cudaStream_t st) {
    auto X = (const __nv_bfloat16*)x; auto R = (const __nv_bfloat16*)r; auto W = (const WT*)w;
    auto Y = (__nv_bfloat16*)y; auto RO = (__nv_bfloat16*)res_out;
    bool vec = (n % 8 == 0) && aligned16(x) && aligned16(y) && aligned16(w) &&
               (sx % 8 == 0) && (sy % 8 == 0);
    if (RESIDUAL) vec = vec && aligned16(r) && aligned16(res_out) && (sr % 8 == 0) && (sres % 8 == 0);
    const int64_t nvec = n / 8;
    if (!vec || nvec > 1024 * 8) {
        rmsnorm_generic_kernel<RESIDUAL, WT><<<(unsigned)rows, 256, 0, st>>>(X, R, W, Y, RO, n, sx, sr, sy, sres, eps);
        B200_CHECK_LAUNCH();
        return 0;
    }
    // Few rows (decode): maximise threads per row (latency).  Many rows (prefill):
    // 4-8 vectors per thread so each SM keeps >=100 KB of loads in flight.
    int vpt = 1;
    if (rows >= 512) vpt = nvec >= 512 ? 4 : (nvec >= 128 ? 2 : 1);
    while ((nvec + vpt - 1) / vpt > (vpt == 4 ? 512 : 1024)) vpt *= 2;
    int threads = (int)((nvec + vpt - 1) / vpt);
    threads = (threads + 31) & ~31;
#define RMS_LAUNCH(V) rmsnorm_row_kernel<V, RESIDUAL, WT><<<(unsigned)rows, threads, 0, st>>>(X, R, W, Y, RO, n, sx, sr, sy, sres, eps)
    switch (vpt) {
        case 1: RMS_LAUNCH(1); break;
        case 2: RMS_LAUNCH(2); break;
        case 4: RMS_LAUNCH(4); break;
        default: RMS_LAUNCH(8); break;
    }
#undef RMS_LAUNCH
    B200_CHECK_LAUNCH();
    return 0;
}

// ============================================================== store_kvcache
// thread = one 16-byte vector of one (token, head); K and V handled together.
__global__ void __launch_bounds__(256)
store_kvcache_vec_kernel(const __nv_bfloat16* __restrict__ key, const __nv_bfloat16* __restrict__ value,
                         __nv_bfloat16* __restrict__ kc, __nv_bfloat16* __restrict__ vc,
                         const int32_t* __restrict__ slots, int64_t T, int H, int dv,
                         int64_t kt, int64_t kh, int64_t vt, int64_t vh,
                         int64_t n_blocks, int block_size, int64_t cb, int64_t cs, int64_t ch) {
    const int64_t total = T * H * dv;
    for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
        const int c = (int)(idx % dv);
        const int64_t th = idx / dv;
        const int h = (int)(th % H);
        const int64_t t = th / H;
        const int slot = __ldg(slots + t);
        if (slot < 0) continue;                       // -1 = skip (cache_manager.py:113-115)
        const int64_t blk = slot / block_size;
        if (blk >= n_blocks) continue;                // memory safety; the reference would fault
        const int off = slot - (int)blk * block_size;
        const int64_t dst = blk * cb + (int64_t)off * cs + (int64_t)h * ch + (int64_t)c * 8;
        const uint4 k = ld_nc_v4(key + t * kt + (int64_t)h * kh + (int64_t)c * 8);
        const uint4 v = ld_nc_v4(value + t * vt + (int64_t)h * vh + (int64_t)c * 8);
        *reinterpret_cast<uint4*>(kc + dst) = k;
        *reinterpret_cast<uint4*>(vc + dst) = v;
    }
}

__global__ void __launch_bounds__(256)
store_kvcache_scalar_kernel(const __nv_bfloat16* __restrict__ key, const __nv_bfloat16* __restrict__ value,
                            __nv_bfloat16* __restrict__ kc, __nv_bfloat16* __restrict__ vc,
                            const int32_t* __restrict__ slots, int64_t T, int H, int D,
                            int64_t kt, int64_t kh, int64_t vt, int64_t vh,
                            int64_t n_blocks, int block_size, int64_t cb, int64_t cs, int64_t ch) {
    const int64_t total = T * H * D;
    for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
        const int c = (int)(idx % D);
        const int64_t th = idx / D;
        const int h = (int)(th % H);
        const int64_t t = th / H;
        const int slot = __ldg(slots + t);
        if (slot < 0) continue;
        const int64_t blk = slot / block_size;
        if (blk >= n_blocks) continue;
        const int off = slot - (int)blk * block_size;
        const int64_t dst = blk * cb + (int64_t)off * cs + (int64_t)h * ch + c;
        kc[dst] = key[t * kt + (int64_t)h * kh + c];
        vc[dst] = value[t * vt + (int64_t)h * vh + c];
    }
}

// ============================================================== device-resident block table
// One tiny launch keeps the decode metadata on the device between steps:  cache_seqlens[:B] += delta  and the
// pending block-table entries (a page handed out by KVCacheManager.append = ONE int32 write) are applied from a
// device-side update list  updates = [count, (row, col, block) x count]  whose count the kernel zeroes again.
// The list is read from device memory at run time, so the launch can sit inside a captured CUDA graph (the first
// node of a multi-step decode graph) while the host only uploads the few entries that changed.
__global__ void __launch_bounds__(256)
table_advance_kernel(int32_t* __restrict__ seqlens, int B, int delta, int32_t* __restrict__ table, int64_t bt_stride,
                     int rows, int cols, int32_t* __restrict__ updates, int max_updates) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (updates != nullptr) {
        int n = updates[0];
        if (n > max_updates) n = max_updates;
        for (int i = tid; i < n; i += gridDim.x * blockDim.x) {
            const int r = updates[1 + 3 * i], c = updates[2 + 3 * i], blk = updates[3 + 3 * i];
            if (r >= 0 && r < rows && c >= 0 && c < cols) table[(int64_t)r * bt_stride + c] = blk;
        }
    }
    if (delta != 0)
        for (int i = tid; i < B; i += gridDim.x * blockDim.x) seqlens[i] += delta;
    if (updates != nullptr) {
        __syncthreads();                       // single CTA (see the launch): every thread has read the count
        if (tid == 0) updates[0] = 0;
    }
}

// ============================================================== L2 prefetch
// Pulls `rows` slabs of `row_bytes` bytes, `stride` bytes apart, into L2 with bulk prefetches (no registers, no shared
// memory, nothing written): the decode runner launches it on a side stream next to the NVLink reduction — a
// latency-bound kernel that leaves HBM idle — so that the weights of the GEMM behind the reduction are already on
// their way (core/layer/model_graph.py).
__global__ void __launch_bounds__(256)
l2_prefetch_kernel(const uint8_t* __restrict__ base, int64_t row_bytes, int64_t rows, int64_t stride, int chunk) {
    const int64_t cpr = (row_bytes + chunk - 1) / chunk;
    const int64_t n_chunks = cpr * rows;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cpr, off = (i - r * cpr) * chunk;
        const uint32_t sz = (uint32_t)((row_bytes - off < chunk) ? (row_bytes - off) & ~15LL : chunk);
        if (sz) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(base + r * stride + off), "r"(sz) : "memory");
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_table_advance(int32_t* cache_seqlens, int B, int delta, int32_t* block_table, int64_t bt_stride,
                                  int rows, int cols, int32_t* updates, int max_updates, void* stream) {
    if (B < 0 || (B > 0 && !cache_seqlens)) B200_FAIL(B200_E_NULL, "table_advance: cache_seqlens is NULL");
    if (updates && (!block_table || max_updates < 0 || rows < 0 || cols < 0 || bt_stride < cols))
        B200_FAIL(B200_E_RANGE, "table_advance: update list without a block table (rows=%d cols=%d stride=%lld)", rows, cols, (long long)bt_stride);
    if (B == 0 && !updates) return 0;
    table_advance_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(cache_seqlens, B, delta, block_table, bt_stride, rows, cols,
                                                              updates, max_updates);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_l2_prefetch(const void* ptr, int64_t row_bytes, int64_t rows, int64_t stride_bytes, void* stream) {
    if (!ptr) B200_FAIL(B200_E_NULL, "l2_prefetch: null pointer");
    if (row_bytes < 0 || rows < 0 || (rows > 1 && stride_bytes < row_bytes))
        B200_FAIL(B200_E_RANGE, "l2_prefetch: row_bytes=%lld rows=%lld stride=%lld", (long long)row_bytes, (long long)rows, (long long)stride_bytes);
    if (!aligned16(ptr) || (rows > 1 && stride_bytes % 16)) B200_FAIL(B200_E_ALIGN, "l2_prefetch: pointer / stride must be 16-byte aligned");
    if (row_bytes < 16 || rows == 0) return 0;
    const int chunk = 2048;
    const int64_t n_chunks = ((row_bytes + chunk - 1) / chunk) * rows;
    int64_t blocks = (n_chunks + 255) / 256;
    if (blocks > 64) blocks = 64;                        // a sliver of the chip: the kernels it runs beside need the rest
    l2_prefetch_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)ptr, row_bytes, rows, stride_bytes, chunk);
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_version(void) { return B200_ABI_VERSION; }
extern "C" const char* b200_last_error(void) { return g_err; }

extern "C" int b200_sm_count(void) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}

extern "C" int b200_swiglu(const void* gate_up, void* out, int64_t rows, int64_t d,
                           int64_t stride_in, int64_t stride_out, void* stream) {
    if (!gate_up || !out) B200_FAIL(B200_E_NULL, "swiglu: null pointer");
    if (rows < 0 || d <= 0) B200_FAIL(B200_E_RANGE, "swiglu: rows=%lld d=%lld", (long long)rows, (long long)d);
    if (rows == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    auto G = (const __nv_bfloat16*)gate_up; auto O = (__nv_bfloat16*)out;
    const bool vec = (d % 8 == 0) && (stride_in % 8 == 0) && (stride_out % 8 == 0) && aligned16(gate_up) && aligned16(out);
    if (vec && rows <= 65535 * 32768LL && d / 8 < (1 << 30)) {
        const int dv = (int)(d / 8);
        // rows go to grid.y (<= 65535 per launch); few rows (decode) -> 1 vector per thread so the
        // whole chip is used, many rows (prefill) -> 4 vectors per thread for bytes in flight
        for (int64_t r0 = 0; r0 < rows; r0 += 65535) {
            const unsigned nr = (unsigned)((rows - r0 < 65535) ? rows - r0 : 65535);
            const __nv_bfloat16* g = G + r0 * stride_in;
            __nv_bfloat16* o = O + r0 * stride_out;
            if (rows >= 256 && dv >= 1024) {
                swiglu_vec_kernel<4><<<dim3((dv + 1023) / 1024, nr), 256, 0, st>>>(g, o, dv, stride_in, stride_out);
            } else if (rows >= 256 && dv >= 512) {
                swiglu_vec_kernel<2><<<dim3((dv + 511) / 512, nr), 256, 0, st>>>(g, o, dv, stride_in, stride_out);
            } else {
                swiglu_vec_kernel<1><<<dim3((dv + 255) / 256, nr), 256, 0, st>>>(g, o, dv, stride_in, stride_out);
            }
        }
    } else {
        int64_t blocks = (rows * d + 255) / 256;
        if (blocks > 148 * 32) blocks = 148 * 32;
        swiglu_scalar_kernel<<<(unsigned)blocks, 256, 0, st>>>(G, O, rows, d, stride_in, stride_out);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200_rmsnorm(const void* x, const void* w, void* y, int64_t rows, int64_t n,
                            int64_t stride_x, int64_t stride_y, float eps, int w_dtype, void* stream) {
    if (!x || !w || !y) B200_FAIL(B200_E_NULL, "rmsnorm: null pointer");
    if (rows < 0 || n <= 0) B200_FAIL(B200_E_RANGE, "rmsnorm: rows=%lld n=%lld", (long long)rows, (long long)n);
    if (rows == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (w_dtype == B200_DT_BF16)
        return launch_rmsnorm<false, __nv_bfloat16>(x, nullptr, w, y, nullptr, rows, n, stride_x, 0, stride_y, 0, eps, st);
    if (w_dtype == B200_DT_F32)
        return launch_rmsnorm<false, float>(x, nullptr, w, y, nullptr, rows, n, stride_x, 0, stride_y, 0, eps, st);
    B200_FAIL(B200_E_UNSUPPORTED, "rmsnorm: weight dtype %d", w_dtype);
}

extern "C" int b200_rmsnorm_residual(const void* x, const void* r, const void* w, void* y, void* res_out,
                                     int64_t rows, int64_t n, int64_t stride_x, int64_t stride_r,
                                     int64_t stride_y, int64_t stride_res, float eps, int w_dtype, void* stream) {
    if (!x || !r || !w || !y || !res_out) B200_FAIL(B200_E_NULL, "rmsnorm_residual: null pointer");
    if (rows < 0 || n <= 0) B200_FAIL(B200_E_RANGE, "rmsnorm_residual: rows=%lld n=%lld", (long long)rows, (long long)n);
    if (rows == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (w_dtype == B200_DT_BF16)
        return launch_rmsnorm<true, __nv_bfloat16>(x, r, w, y, res_out, rows, n, stride_x, stride_r, stride_y, stride_res, eps, st);
    if (w_dtype == B200_DT_F32)
        return launch_rmsnorm<true, float>(x, r, w, y, res_out, rows, n, stride_x, stride_r, stride_y, stride_res, eps, st);
    B200_FAIL(B200_E_UNSUPPORTED, "rmsnorm_residual: weight dtype %d", w_dtype);
}

extern "C" int b200_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                                  const int32_t* slot_mapping, int64_t T, int64_t H, int64_t D,
                                  int64_t kt, int64_t kh, int64_t vt, int64_t vh,
                                  int64_t n_blocks, int64_t block_size,
                                  int64_t cb, int64_t cs, int64_t ch, void* stream) {
    if (!key || !value || !k_cache || !v_cache || !slot_mapping) B200_FAIL(B200_E_NULL, "store_kvcache: null pointer");
    if (T < 0 || H <= 0 || D <= 0 || block_size <= 0 || n_blocks <= 0) B200_FAIL(B200_E_RANGE, "store_kvcache: bad sizes");
    if (T == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    auto K = (const __nv_bfloat16*)key; auto V = (const __nv_bfloat16*)value;
    auto KC = (__nv_bfloat16*)k_cache; auto VC = (__nv_bfloat16*)v_cache;
    const bool vec = (D % 8 == 0) && aligned16(key) && aligned16(value) && aligned16(k_cache) && aligned16(v_cache) &&
                     (kt % 8 == 0) && (kh % 8 == 0) && (vt % 8 == 0) && (vh % 8 == 0) &&
                     (cb % 8 == 0) && (cs % 8 == 0) && (ch % 8 == 0);
    if (vec) {
        const int64_t total = T * H * (D / 8);
        int64_t blocks = (total + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        store_kvcache_vec_kernel<<<(unsigned)blocks, 256, 0, st>>>(K, V, KC, VC, slot_mapping, T, (int)H, (int)(D / 8),
                                                                  kt, kh, vt, vh, n_blocks, (int)block_size, cb, cs, ch);
    } else {
        const int64_t total = T * H * D;
        int64_t blocks = (total + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        store_kvcache_scalar_kernel<<<(unsigned)blocks, 256, 0, st>>>(K, V, KC, VC, slot_mapping, T, (int)H, (int)D,
                                                                     kt, kh, vt, vh, n_blocks, (int)block_size, cb, cs, ch);
    }
    B200_CHECK_LAUNCH();
    return 0;
}
