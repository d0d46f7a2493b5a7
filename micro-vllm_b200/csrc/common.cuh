// Shared device helpers for libb200ops (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/b200_abi.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb200ops is written for sm_100a (B200) only"
#endif

namespace b200 {

void set_error(const char* fmt, ...);

#define B200_FAIL(code, ...) do { ::b200::set_error(__VA_ARGS__); return (code); } while (0)
#define B200_CHECK_LAUNCH() do { cudaError_t e__ = cudaGetLastError(); \
    if (e__ != cudaSuccess) { ::b200::set_error("launch failed: %s", cudaGetErrorString(e__)); return (int)e__; } } while (0)

// Kernel launch with (optionally) programmatic stream serialization: the kernel may begin while the previous kernel of
// the stream is still running and must execute griddepcontrol.wait before it touches anything that kernel produces.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_maybe_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                           bool pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
__device__ __forceinline__ void pdl_release_next() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait_prev() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---------------------------------------------------------------- bf16 packing
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// round-to-nearest-even pack of two floats into a bf16x2 word (lo = a, hi = b)
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}

__device__ __forceinline__ float bf16_round(float a) {
    return __bfloat162float(__float2bfloat16_rn(a));
}

// ---------------------------------------------------------------- global ld/st
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ld_v4(const void* p) {
    return *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ void st_na_v4(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// Rotate (or copy) one 16-element block `j` of a head row.  NeoX: elements
// [8j, 8j+8) pair with [rd/2 + 8j, rd/2 + 8j + 8); interleaved: pairs (2i, 2i+1)
// inside [16j, 16j+16).  fp32 math, one RNE rounding (flash-attn rotary contract).
// Results are written as 16-byte chunks through `store(chunk_index, uint4)`.
template <typename Store>
__device__ __forceinline__ void rope_block(const __nv_bfloat16* __restrict__ src, int j, int rd, bool interleaved,
                                           const __nv_bfloat16* __restrict__ cosr, const __nv_bfloat16* __restrict__ sinr,
                                           Store&& store) {
    if (16 * j >= rd) {             // pass-through part of the head (d >= rotary_dim)
        store(2 * j, ld_v4(src + 16 * j));
        store(2 * j + 1, ld_v4(src + 16 * j + 8));
        return;
    }
    const uint4 cv = ld_v4(cosr + 8 * j), sv = ld_v4(sinr + 8 * j);
    const uint32_t* pc = &cv.x; const uint32_t* ps = &sv.x;
    if (!interleaved) {
        const int half = rd >> 1;
        const uint4 a = ld_v4(src + 8 * j), b = ld_v4(src + half + 8 * j);
        const uint32_t* pa = &a.x; const uint32_t* pb = &b.x;
        uint4 oa, ob; uint32_t* poa = &oa.x; uint32_t* pob = &ob.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float x0l = bf16lo(pa[i]), x0h = bf16hi(pa[i]);
            const float x1l = bf16lo(pb[i]), x1h = bf16hi(pb[i]);
            const float cl = bf16lo(pc[i]), chh = bf16hi(pc[i]);
            const float sl = bf16lo(ps[i]), shh = bf16hi(ps[i]);
            poa[i] = pack_bf16(x0l * cl - x1l * sl, x0h * chh - x1h * shh);
            pob[i] = pack_bf16(x0l * sl + x1l * cl, x0h * shh + x1h * chh);
        }
        store(j, oa);
        store((half >> 3) + j, ob);
    } else {
        const uint4 a = ld_v4(src + 16 * j), b = ld_v4(src + 16 * j + 8);
        const uint32_t* pa = &a.x; const uint32_t* pb = &b.x;
        uint4 oa, ob; uint32_t* poa = &oa.x; uint32_t* pob = &ob.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // pair index within the block: words 0..3 of `a` are pairs 0..3, `b` pairs 4..7
            const float ca = (i & 1) ? bf16hi(pc[i >> 1]) : bf16lo(pc[i >> 1]);
            const float sa = (i & 1) ? bf16hi(ps[i >> 1]) : bf16lo(ps[i >> 1]);
            const float cb = (i & 1) ? bf16hi(pc[2 + (i >> 1)]) : bf16lo(pc[2 + (i >> 1)]);
            const float sb = (i & 1) ? bf16hi(ps[2 + (i >> 1)]) : bf16lo(ps[2 + (i >> 1)]);
            const float a0 = bf16lo(pa[i]), a1 = bf16hi(pa[i]);
            const float b0 = bf16lo(pb[i]), b1 = bf16hi(pb[i]);
            poa[i] = pack_bf16(a0 * ca - a1 * sa, a0 * sa + a1 * ca);
            pob[i] = pack_bf16(b0 * cb - b1 * sb, b0 * sb + b1 * cb);
        }
        store(2 * j, oa);
        store(2 * j + 1, ob);
    }
}

}  // namespace b200
