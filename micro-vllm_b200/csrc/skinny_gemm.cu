// Weight-streaming GEMM for decode-sized batches (M <= 64 rows): out[M, N] = x[M, K] . W[K, N] (+ bias).
//   reference: the four torch.matmul calls of a decode layer, core/layer/model_graph.py:145-147 (QKV
//   projection + bias), :171 (o_proj), :183 (gate_up), :185 (down_proj), on the weights prepare_weights
//   (:95-116) transposed to [in, out] contiguous — SURVEY.md §8f N4 (the GEMMs either side of the path).
//
// At M <= 64 these products are HBM-bound on the weight matrix (K*N*2 bytes read once; 2*M flops per
// weight): the job is to stream W at the HBM roofline, not to feed the tensor cores.  cuBLAS picks
// output-tile grids without split-K for deep-K narrow-N shapes, so a [32, 14336] x [14336, 4096] product runs
// on 64 CTAs at 0.5 of the roofline (profiles/r01/skinny_gemm_*.json).  Here:
//   * grid = (N / BN column slices, KS k-ranges) sized to the resident CTAs of all SMs;
//   * 256 threads = 4 column groups x 2 halves of a 64-row stage; every CTA streams its [k-range, BN] panel of
//     W through a 3-8 stage cp.async ring (~100 KB, two CTAs per SM; 128-bit copies, XOR-swizzled for
//     conflict-free ldmatrix) together with the matching [M, 64] slab of x (L2 resident: x is tiny and shared
//     by all column slices).  The per-thread copy plan (pointers, swizzled offsets) lives in registers: the
//     first version recomputed it from threadIdx every stage and was instruction-issue bound (ncu);
//   * mma.sync m16n8k16 bf16 -> fp32 with x as the row operand (ldmatrix) and W as the column operand
//     (ldmatrix.trans straight from the [k][n] tile); all fragments of a stage are requested before the first
//     mma.  The legacy tensor path sustains one HMMA per 8 clk per sub-partition on B200 (tools/micro), the
//     roofline needs one per 22 (M = 32) / 11 (M = 64);
//   * the two k-halves of a CTA merge through shared memory; split-K partials go to an fp32 workspace and the
//     LAST CTA to arrive on a column slice (one acq_rel atomic per CTA) sums them in k-range order —
//     deterministic, one launch, CUDA-graph replay safe (the counter is reset by that CTA);
//   * epilogue rounds once to bf16; a bias is added the way the reference does it (bf16 matmul result, then
//     a bf16 add: two roundings).
#include "common.cuh"

namespace b200 {

constexpr int SG_THREADS = 256;          // 8 warps = 4 column groups (BN / 4 columns each) x 2 halves of a stage's rows
constexpr int SG_KC = 64;                // W rows per stage; warps 0-3 take rows 0-31, warps 4-7 rows 32-63
// ring depth: as many stages as fit in ~110 KB (two CTAs per SM), between 3 and 8 — the bytes in flight per SM
// ((stages - 1) x stage x 2 CTAs = 130-170 KB) are what keeps HBM busy at its loaded latency
__host__ __device__ constexpr int sg_stage_bytes(int mt, int bn) { return SG_KC * bn * 2 + 16 * mt * SG_KC * 2; }
__host__ __device__ constexpr int sg_stages(int mt, int bn) {
    return 110 * 1024 / sg_stage_bytes(mt, bn) > 8 ? 8 : (110 * 1024 / sg_stage_bytes(mt, bn) < 3 ? 3 : 110 * 1024 / sg_stage_bytes(mt, bn));
}
constexpr int SG_COUNTERS = 4096;        // arrival counters at the head of the workspace (one per column slice)

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct SkinnyParams {
    const __nv_bfloat16* x; const __nv_bfloat16* w; const __nv_bfloat16* bias; __nv_bfloat16* out;
    float* partial; unsigned int* counters;
    int M, KS;
    int64_t K, N, ldx, ldw, ldo;
};

// MT = number of 16-row tiles of x (M <= 16 * MT); BN = columns per CTA (64 / 128 / 256).
template <int MT, int BN>
__global__ void __launch_bounds__(SG_THREADS, (MT == 4 && BN == 256) ? 1 : 2)
skinny_gemm_kernel(const SkinnyParams p) {
    constexpr int WCH = BN / 8;                         // 16-byte chunks per W tile row
    constexpr int W_STAGE = SG_KC * BN * 2;             // bytes
    constexpr int X_STAGE = 16 * MT * SG_KC * 2;
    constexpr int NT2 = BN / 64;                        // ldmatrix.x4.trans per k16 per warp (2 n8 tiles each)
    constexpr int SG_STAGES = sg_stages(MT, BN);
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t sW = smem_u32(smem);
    const uint32_t sX = sW + SG_STAGES * W_STAGE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n0 = (int64_t)blockIdx.x * BN;
    const int kz = blockIdx.y;
    const int units = (int)(p.K / SG_KC);
    const int u0 = (int)((int64_t)kz * units / p.KS), u1 = (int)((int64_t)(kz + 1) * units / p.KS);
    const int n_it = u1 - u0;

    // rows of x beyond M stay zero in every stage (never written by cp.async)
    for (int i = tid; i < SG_STAGES * 16 * MT * 8; i += SG_THREADS) {
        const int row = (i >> 3) % (16 * MT);
        if (row >= p.M) *reinterpret_cast<uint4*>(smem + SG_STAGES * W_STAGE + (size_t)i * 16) = make_uint4(0, 0, 0, 0);
    }

    // Per-thread copy plan, hoisted out of the loop (the ncu source view of the first version showed ~150 issued
    // instructions per warp and stage, two thirds of them address arithmetic: the kernel was issue bound, not
    // memory bound).  Thread t copies chunk c = t % WCH of rows t / WCH + i * (256 / WCH): the swizzle term
    // (row & 7) is the same for every i, so destinations are dst0 + i * 4096 and sources advance by a constant.
    constexpr int NW = SG_KC * WCH / SG_THREADS;        // W chunks per thread and stage (2 / 4 / 8)
    constexpr int NX = (MT + 1) / 2;                    // x chunks per thread and stage
    const int wrow0 = tid / WCH, wc = tid % WCH;
    uint32_t w_dst0 = sW + wrow0 * (BN * 2) + ((wc ^ (wrow0 & 7)) << 4);
    asm volatile("" : "+r"(w_dst0));
    const int64_t w_rstep = (int64_t)(SG_THREADS / WCH) * p.ldw;             // elements between a thread's rows
    const __nv_bfloat16* w_src[NW];                     // one live pointer per copy: 2 integer adds per stage each
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        w_src[i] = p.w + ((int64_t)u0 * SG_KC + wrow0) * p.ldw + n0 + wc * 8 + i * w_rstep;
        asm volatile("" : "+l"(w_src[i]));              // opaque: keep it in a register instead of recomputing from tid
    }
    int64_t w_kstep = (int64_t)SG_KC * p.ldw * 2;       // bytes
    asm volatile("" : "+l"(w_kstep));
    const int xrow0 = tid >> 3, xc = tid & 7;
    uint32_t x_dst0 = sX + xrow0 * 128 + ((xc ^ (xrow0 & 7)) << 4);
    const __nv_bfloat16* x_src = p.x + (int64_t)xrow0 * p.ldx + (int64_t)u0 * SG_KC + xc * 8;
    asm volatile("" : "+r"(x_dst0), "+l"(x_src));
    const int64_t x_rstep = 32 * p.ldx;
    bool x_ok[NX];
#pragma unroll
    for (int i = 0; i < NX; ++i) x_ok[i] = (xrow0 + 32 * i < p.M) && (xrow0 + 32 * i < 16 * MT);

    auto load_stage = [&](int slot) {                   // loads the NEXT stage of this CTA's k-range into `slot`
        const uint32_t wdst = slot * W_STAGE + w_dst0;
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            cp_async16(wdst + i * 4096, w_src[i]);
            w_src[i] = reinterpret_cast<const __nv_bfloat16*>(reinterpret_cast<const char*>(w_src[i]) + w_kstep);
        }
        const uint32_t xdst = slot * X_STAGE + x_dst0;
#pragma unroll
        for (int i = 0; i < NX; ++i)
            if (x_ok[i]) cp_async16(xdst + i * 4096, x_src + i * x_rstep);
        x_src += SG_KC;
    };

    float acc[MT][2 * NT2][4];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int n = 0; n < 2 * NT2; ++n)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[m][n][e] = 0.0f;

#pragma unroll
    for (int s = 0; s < SG_STAGES - 1; ++s) {
        if (s < n_it) load_stage(s);
        cp_async_commit();
    }
    const int lrow = lane & 7, lmat = lane >> 3;
    const int cgrp = warp & 3, khalf = warp >> 2;       // column group, half of the stage's 64 rows
    // ldmatrix offsets inside a stage: x rows m*16 + ... differ by m * 2048 bytes (same swizzle term)
    uint32_t a_off[2], b_off[2][NT2];
#pragma unroll
    for (int k2 = 0; k2 < 2; ++k2) {
        const int kk = khalf * 2 + k2;
        const int row = (lmat & 1) * 8 + lrow;
        const int ca = kk * 2 + (lmat >> 1);
        a_off[k2] = row * 128 + ((ca ^ (row & 7)) << 4);
        const int krow = kk * 16 + (lmat & 1) * 8 + lrow;
#pragma unroll
        for (int t = 0; t < NT2; ++t) {
            const int c = cgrp * (BN / 32) + t * 2 + (lmat >> 1);
            b_off[k2][t] = sW + krow * (BN * 2) + ((c ^ (krow & 7)) << 4);
            asm volatile("" : "+r"(b_off[k2][t]));
        }
        a_off[k2] += sX;
        asm volatile("" : "+r"(a_off[k2]));
    }
    int slot = 0, lslot = SG_STAGES - 1;
    for (int it = 0; it < n_it; ++it) {
        cp_async_wait<SG_STAGES - 2>();
        __syncthreads();                                // stage `it` visible to all; slot of stage it-1 is free
        if (it + SG_STAGES - 1 < n_it) load_stage(lslot);
        cp_async_commit();
        const uint32_t wbase = slot * W_STAGE, xbase = slot * X_STAGE;
        // every fragment of this warp's two k16 steps is requested before the first mma
        uint32_t a[2][MT][4], b[2][NT2][4];
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
#pragma unroll
            for (int m = 0; m < MT; ++m) ldsm_x4(xbase + a_off[k2] + m * 2048, a[k2][m]);
#pragma unroll
            for (int t = 0; t < NT2; ++t) ldsm_x4_trans(wbase + b_off[k2][t], b[k2][t]);
        }
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
            for (int t = 0; t < NT2; ++t)
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    mma_bf16(acc[m][2 * t], a[k2][m], b[k2][t][0], b[k2][t][1]);
                    mma_bf16(acc[m][2 * t + 1], a[k2][m], b[k2][t][2], b[k2][t][3]);
                }
        slot = (slot + 1 == SG_STAGES) ? 0 : slot + 1;
        lslot = (lslot + 1 == SG_STAGES) ? 0 : lslot + 1;
    }
    cp_async_wait<0>();
    __syncthreads();                                    // pipeline buffers are free: reuse them for the k-half merge
    {
        constexpr int ACC = MT * 2 * NT2 * 4;
        float* xch = reinterpret_cast<float*>(smem) + ((size_t)cgrp * 32 + lane) * ACC;
        if (khalf == 1) {
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int n = 0; n < 2 * NT2; ++n)
                    *reinterpret_cast<float4*>(xch + (m * 2 * NT2 + n) * 4) =
                        make_float4(acc[m][n][0], acc[m][n][1], acc[m][n][2], acc[m][n][3]);
        }
        __syncthreads();
        if (khalf == 1 && p.KS == 1) return;
        if (khalf == 0) {
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int n = 0; n < 2 * NT2; ++n) {
                    const float4 v = *reinterpret_cast<const float4*>(xch + (m * 2 * NT2 + n) * 4);
                    acc[m][n][0] += v.x; acc[m][n][1] += v.y; acc[m][n][2] += v.z; acc[m][n][3] += v.w;
                }
        }
    }

    // accumulator element (m, n, e): row = m*16 + lane/4 + (e >= 2 ? 8 : 0), col = cgrp*(BN/4) + n*8 + (lane%4)*2 + (e&1)
    const int g = lane >> 2, q = lane & 3;
    const int wcol = cgrp * (BN / 4);                   // only the khalf == 0 warps hold the merged sums
    if (p.KS == 1) {
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int n = 0; n < 2 * NT2; ++n)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = m * 16 + g + h * 8;
                    if (row >= p.M) continue;
                    const int64_t col = n0 + wcol + n * 8 + q * 2;
                    uint32_t pk = pack_bf16(acc[m][n][2 * h], acc[m][n][2 * h + 1]);
                    if (p.bias) {
                        const uint32_t bb = *reinterpret_cast<const uint32_t*>(p.bias + col);
                        pk = pack_bf16(bf16lo(pk) + bf16lo(bb), bf16hi(pk) + bf16hi(bb));
                    }
                    *reinterpret_cast<uint32_t*>(p.out + (int64_t)row * p.ldo + col) = pk;
                }
        return;
    }
    // split-K: fp32 partials [KS][M][N], last CTA of the column slice reduces in k-range order
    float* part = p.partial + (int64_t)kz * p.M * p.N;
    if (khalf == 0)
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int n = 0; n < 2 * NT2; ++n)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int row = m * 16 + g + h * 8;
                if (row >= p.M) continue;
                const int64_t col = n0 + wcol + n * 8 + q * 2;
                *reinterpret_cast<float2*>(part + (int64_t)row * p.N + col) = make_float2(acc[m][n][2 * h], acc[m][n][2 * h + 1]);
            }
    __syncthreads();                                    // every partial store of this CTA happens-before thread 0's release
    __shared__ unsigned int s_last;
    if (tid == 0) {
        unsigned int prev;                              // one acq_rel atomic per CTA (cumulative over the barrier)
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(prev) : "l"(p.counters + blockIdx.x) : "memory");
        s_last = (prev == (unsigned int)p.KS - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    constexpr int V4 = BN / 4;                          // float4 per row of the slice
    for (int i = tid; i < p.M * V4; i += SG_THREADS) {
        const int row = i / V4, c4 = i % V4;
        const int64_t col = n0 + c4 * 4;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        const float* src = p.partial + (int64_t)row * p.N + col;
        const int64_t zs = (int64_t)p.M * p.N;
        for (int z0 = 0; z0 < p.KS; z0 += 8) {          // 8 independent L2 loads in flight, summed in k-range order
            float4 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (z0 + j < p.KS) v[j] = __ldcg(reinterpret_cast<const float4*>(src + (z0 + j) * zs));
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (z0 + j < p.KS) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
        }
        uint32_t lo = pack_bf16(s.x, s.y), hi = pack_bf16(s.z, s.w);
        if (p.bias) {
            const uint2 bb = *reinterpret_cast<const uint2*>(p.bias + col);
            lo = pack_bf16(bf16lo(lo) + bf16lo(bb.x), bf16hi(lo) + bf16hi(bb.x));
            hi = pack_bf16(bf16lo(hi) + bf16lo(bb.y), bf16hi(hi) + bf16hi(bb.y));
        }
        *reinterpret_cast<uint2*>(p.out + (int64_t)row * p.ldo + col) = make_uint2(lo, hi);
    }
    if (tid == 0) p.counters[blockIdx.x] = 0;           // replay safe
}

template <int MT, int BN>
static int launch_skinny(const SkinnyParams& p, cudaStream_t st) {
    constexpr int smem = sg_stages(MT, BN) * sg_stage_bytes(MT, BN);
    // not a stream operation (legal during capture); per device, so it is not cached in a static
    cudaError_t e = cudaFuncSetAttribute(skinny_gemm_kernel<MT, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("skinny_gemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    skinny_gemm_kernel<MT, BN><<<dim3((unsigned)(p.N / BN), (unsigned)p.KS), SG_THREADS, smem, st>>>(p);
    B200_CHECK_LAUNCH();
    return 0;
}

static int skinny_bn(int variant, int64_t N) {
    int bn = variant == 1 ? 64 : (variant == 2 ? 256 : 128);
    while (bn > 64 && N % bn != 0) bn >>= 1;
    return bn;
}

// split count from launch constants only: fill `ctas_per_sm` resident CTAs on every SM, at least 8 stages per range
static int skinny_auto_ks(int M, int64_t K, int64_t N, int bn, int n_sm) {
    const int mt = M <= 16 ? 1 : (M <= 32 ? 2 : 4);
    const int smem = sg_stages(mt, bn) * sg_stage_bytes(mt, bn);
    int per_sm = (int)(220 * 1024 / smem);
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    const int64_t slices = N / bn, units = K / SG_KC;
    int64_t ks = ((int64_t)n_sm * per_sm) / slices;
    if (ks > units / 8) ks = units / 8;
    if (ks < 1) ks = 1;
    if (ks > 64) ks = 64;
    return (int)ks;
}

}  // namespace b200
using namespace b200;

extern "C" int b200_skinny_gemm_splits(int M, int64_t K, int64_t N, int variant) {
    if (M <= 0 || K <= 0 || N <= 0 || K % SG_KC != 0 || N % 64 != 0) return -1;
    int dev = 0, n_sm = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) n_sm = n;
    }
    return skinny_auto_ks(M, K, N, skinny_bn(variant, N), n_sm);
}

extern "C" int64_t b200_skinny_gemm_workspace_bytes(int M, int64_t N, int ksplit) {
    if (M <= 0 || N <= 0 || ksplit < 1) return -1;
    return (int64_t)SG_COUNTERS * 4 + (ksplit > 1 ? (int64_t)ksplit * M * N * 4 : 0);
}

extern "C" int b200_skinny_gemm(const void* x, const void* w, const void* bias, void* out,
                                void* workspace, int64_t workspace_bytes,
                                int M, int64_t K, int64_t N,
                                int64_t stride_x, int64_t stride_w, int64_t stride_out,
                                int ksplit, int variant, void* stream) {
    if (!x || !w || !out) B200_FAIL(B200_E_NULL, "skinny_gemm: null pointer");
    if (M < 0 || M > 64) B200_FAIL(B200_E_SHAPE, "skinny_gemm: M=%d (1..64 rows; larger batches are compute bound, use a dense GEMM)", M);
    if (K <= 0 || N <= 0 || K % SG_KC != 0 || N % 64 != 0)
        B200_FAIL(B200_E_SHAPE, "skinny_gemm: K=%lld must be a multiple of %d and N=%lld of 64", (long long)K, SG_KC, (long long)N);
    if (stride_x < K || stride_w < N || stride_out < N || stride_x % 8 || stride_w % 8 || stride_out % 4)
        B200_FAIL(B200_E_ALIGN, "skinny_gemm: strides x=%lld w=%lld out=%lld", (long long)stride_x, (long long)stride_w, (long long)stride_out);
    if (!aligned16(x) || !aligned16(w) || (reinterpret_cast<uintptr_t>(out) & 7) || (bias && (reinterpret_cast<uintptr_t>(bias) & 7)))
        B200_FAIL(B200_E_ALIGN, "skinny_gemm: x / w must be 16-byte aligned, out / bias 8-byte aligned");
    if (M == 0) return 0;
    const int bn = skinny_bn(variant, N);
    if (N / bn > SG_COUNTERS) B200_FAIL(B200_E_RANGE, "skinny_gemm: N=%lld gives more than %d column slices", (long long)N, SG_COUNTERS);
    int ks = ksplit;
    if (ks <= 0) ks = b200_skinny_gemm_splits(M, K, N, variant);
    if (ks > K / SG_KC) ks = (int)(K / SG_KC);
    if (ks < 1 || ks > 65535) B200_FAIL(B200_E_RANGE, "skinny_gemm: ksplit=%d", ks);
    SkinnyParams p;
    p.x = (const __nv_bfloat16*)x; p.w = (const __nv_bfloat16*)w; p.bias = (const __nv_bfloat16*)bias;
    p.out = (__nv_bfloat16*)out; p.M = M; p.KS = ks; p.K = K; p.N = N;
    p.ldx = stride_x; p.ldw = stride_w; p.ldo = stride_out;
    p.partial = nullptr; p.counters = nullptr;
    if (ks > 1) {
        const int64_t need = b200_skinny_gemm_workspace_bytes(M, N, ks);
        if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 15))
            B200_FAIL(B200_E_RANGE, "skinny_gemm: workspace of %lld bytes (16-byte aligned, zero-filled once) needed, got %lld",
                      (long long)need, (long long)workspace_bytes);
        p.counters = (unsigned int*)workspace;
        p.partial = (float*)((char*)workspace + (int64_t)SG_COUNTERS * 4);
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int mt = M <= 16 ? 1 : (M <= 32 ? 2 : 4);
#define SG_CASE(MT_, BN_) if (mt == MT_ && bn == BN_) return launch_skinny<MT_, BN_>(p, st)
    SG_CASE(1, 64); SG_CASE(1, 128); SG_CASE(1, 256);
    SG_CASE(2, 64); SG_CASE(2, 128); SG_CASE(2, 256);
    SG_CASE(4, 64); SG_CASE(4, 128); SG_CASE(4, 256);
#undef SG_CASE
    B200_FAIL(B200_E_UNSUPPORTED, "skinny_gemm: no kernel for mt=%d bn=%d", mt, bn);
}
