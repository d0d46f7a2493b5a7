// Paged prefill attention for sm_100a on tcgen05 tensor cores — replaces
// RoPE.forward (core/layer/rope.py:10-39) + flash_attn_with_kvcache at
// core/layer/model_prefill.py:49-52.
//
// Two launches per layer call:
//   1. prefill_rope_append_kernel: rotate q and k at positions P_b + i (fp32, one
//      rounding), scatter rotated K and raw V into the pages named by block_table,
//      stage rotated Q in a workspace.                       (HBM-bound, elementwise)
//   2. prefill_attn_kernel: one CTA = 128 query rows of one head.  A TMA producer
//      warp streams Q and 128-token K/V tiles straight out of the paged cache
//      (4-D tensor map: d, head, slot, page -> the page id is a TMA coordinate,
//      128B-swizzled so the tiles are UMMA-canonical).  One thread issues
//      tcgen05.mma: S = Q.K^T into a double-buffered TMEM accumulator and
//      O += P.V into a second TMEM accumulator (V is consumed MN-major, no
//      transpose).  Eight softmax warps put two threads on every query row (TMEM lane =
//      row, 64 key columns each): tcgen05.ld the scores, online softmax in fp32 with
//      lazy rescaling, write P as a K-major bf16 operand to shared memory, rescale O in
//      TMEM only when the running max moved by more than 2^8.   (tensor-core bound)
#include "common.cuh"
#include "tc05.cuh"
#include <cuda.h>
#include <math_constants.h>
#include <type_traits>
#include <cstdlib>

#ifndef B200_PREFILL_DEFAULT_SM
#define B200_PREFILL_DEFAULT_SM 2   // softmax arithmetic of variant 0: packed fp32x2 + 25 % polynomial exp2 (fastest, profiles/r01/prefill_bench.json)
#endif

namespace b200 {

// ---- packed fp32x2 arithmetic (FFMA2 / FADD2 on sm_100): halves the issue slots of the softmax
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
    uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
    uint64_t d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d;
}
__device__ __forceinline__ uint64_t fsub2(uint64_t a, uint64_t b) {
    uint64_t d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d;
}
__device__ __forceinline__ uint64_t fadd2_rm(uint64_t a, uint64_t b) {   // round toward -inf
    uint64_t d; asm("add.rm.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d;
}
// 2^x for a pair on the FMA pipe instead of the 16-lane/clk MUFU: Cody-Waite split with the
// magic-number floor (x + 1.5*2^23 rounded down leaves floor(x) in the low mantissa bits), a
// degree-3 minimax polynomial for 2^frac on [0,1) (rel. error ~1e-4, below the 2^-9 step of the
// bf16 probabilities it feeds) and the integer part shifted into the exponent field.  x <= 127.
__device__ __forceinline__ void exp2_poly2(uint64_t x, float& e0, float& e1) {
    float x0, x1; f2_unpack(x, x0, x1);
    const uint64_t xc = f2_pack(fmaxf(x0, -127.0f), fmaxf(x1, -127.0f));
    const uint64_t magic = f2_pack(12582912.0f, 12582912.0f);
    const uint64_t t = fadd2_rm(xc, magic);
    const uint64_t fr = fsub2(xc, fsub2(t, magic));
    uint64_t pl = ffma2(f2_pack(0.077119089663028717f, 0.077119089663028717f), fr,
                        f2_pack(0.227564394474029541f, 0.227564394474029541f));
    pl = ffma2(pl, fr, f2_pack(0.695146143436431885f, 0.695146143436431885f));
    pl = ffma2(pl, fr, f2_pack(1.0f, 1.0f));
    float t0, t1, p0, p1; f2_unpack(t, t0, t1); f2_unpack(pl, p0, p1);
    e0 = __uint_as_float((__float_as_uint(t0) << 23) + __float_as_uint(p0));
    e1 = __uint_as_float((__float_as_uint(t1) << 23) + __float_as_uint(p1));
}
// exp2 of 64 scores (scaled, reference max subtracted) -> packed bf16x2 probabilities + row sum.
// NPOLY of every 4 pairs take the polynomial, the rest MUFU.EX2; two independent sum chains.
template <int NPOLY, int HC, int OFF = 0, int TOTAL = HC>
__device__ __forceinline__ float softmax_exp_tile(const uint32_t (&su)[TOTAL], float sl2, float mr, uint32_t (&pw)[HC / 2]) {
    const uint64_t s2 = f2_pack(sl2, sl2), m2 = f2_pack(-mr, -mr);
    uint64_t acc0 = 0ull, acc1 = 0ull;
#pragma unroll
    for (int i = 0; i < HC / 2; ++i) {
        const uint64_t x = ffma2(f2_pack(__uint_as_float(su[OFF + 2 * i]), __uint_as_float(su[OFF + 2 * i + 1])), s2, m2);
        float e0, e1;
        if ((i & 3) < NPOLY) {
            exp2_poly2(x, e0, e1);
        } else {
            float x0, x1; f2_unpack(x, x0, x1);
            e0 = fast_exp2(x0); e1 = fast_exp2(x1);
        }
        if (i & 1) acc1 = fadd2(acc1, f2_pack(e0, e1)); else acc0 = fadd2(acc0, f2_pack(e0, e1));
        pw[i] = pack_bf16(e0, e1);
    }
    float a, b; f2_unpack(fadd2(acc0, acc1), a, b);
    return a + b;
}

// Zero the V rows of tokens >= `valid` in a [128 tokens][D] shared-memory tile (DSLABS slabs of [128][64] bf16, one
// 128-byte swizzle row per token: zeroing whole rows is swizzle-invariant).  A masked token's probability is exactly
// 0, but the TMA box brings in whatever bytes sit in never-written cache slots, and 0 * NaN / Inf would poison every
// output row of the tile (flash-attn clears out-of-bounds V; the decode kernel zero-fills these rows).
template <int DSLABS>
__device__ __forceinline__ void zero_v_tail(uint32_t v_tile, int valid, int lane, int nlanes) {
    const int n = (128 - valid) * DSLABS * 8;
    for (int i = lane; i < n; i += nlanes) {
        const int chunk = i & 7, t = i >> 3;
        const int sl = t % DSLABS, row = valid + t / DSLABS;
        asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};"
                     :: "r"(v_tile + (uint32_t)(sl * (128 * 128) + row * 128 + chunk * 16)), "r"(0u) : "memory");
    }
    fence_async_smem();                 // generic-proxy writes -> visible to the tensor core's async proxy
}

// ------------------------------------------------------------------ kernel 1: RoPE + append
struct AppendParams {
    const __nv_bfloat16* q; const __nv_bfloat16* k; const __nv_bfloat16* v;
    __nv_bfloat16* k_cache; __nv_bfloat16* v_cache; __nv_bfloat16* q_rot;
    const __nv_bfloat16* cos; const __nv_bfloat16* sin;
    const int32_t* seqlens; const int32_t* block_table;
    int B, S, Hq, Hkv, D, rotary_dim, interleaved, rot_rows;
    int n_blocks, block_size, max_blocks_per_seq;
    int64_t q_sb, q_ss, q_sh, k_sb, k_ss, k_sh, v_sb, v_ss, v_sh;
    int64_t kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh, bt_sb;
    int* work_counter;     // work-item counter of the persistent attention kernel that follows: zeroed here
};

__global__ void __launch_bounds__(128)
prefill_rope_append_kernel(const __grid_constant__ AppendParams p) {
    const int s = blockIdx.x, b = blockIdx.y;
    // the persistent attention kernel is launched with programmatic stream serialization: its CTAs may take their SMs,
    // set up barriers / TMEM / descriptors and then sit in griddepcontrol.wait while this grid is still running
    pdl_release_next();
    if (s == 0 && b == 0 && threadIdx.x == 0 && p.work_counter != nullptr) *p.work_counter = 0;
    const int D = p.D, JB = D / 16, CPR = D / 8;
    const int pos = __ldg(p.seqlens + b) + s;
    const int rd = p.cos ? p.rotary_dim : 0;
    const int rpos = min(pos, p.rot_rows - 1);
    const __nv_bfloat16* cosr = p.cos ? p.cos + (int64_t)rpos * (p.rotary_dim >> 1) : nullptr;
    const __nv_bfloat16* sinr = p.sin ? p.sin + (int64_t)rpos * (p.rotary_dim >> 1) : nullptr;
    const int pe = pos / p.block_size;
    const int page = (pe < p.max_blocks_per_seq) ? __ldg(p.block_table + (int64_t)b * p.bt_sb + pe) : -1;
    const bool page_ok = page >= 0 && page < p.n_blocks;
    const int slot = pos - pe * p.block_size;

    const int q_tasks = p.Hq * JB, k_tasks = p.Hkv * JB, v_tasks = p.Hkv * CPR;
    for (int t = threadIdx.x; t < q_tasks + k_tasks + v_tasks; t += blockDim.x) {
        if (t < q_tasks) {
            const int h = t / JB, j = t - h * JB;
            const __nv_bfloat16* src = p.q + (int64_t)b * p.q_sb + (int64_t)s * p.q_ss + (int64_t)h * p.q_sh;
            __nv_bfloat16* dst = p.q_rot + (((int64_t)b * p.S + s) * p.Hq + h) * D;
            rope_block(src, j, rd, p.interleaved != 0, cosr, sinr,
                       [&](int c, const uint4& v) { *reinterpret_cast<uint4*>(dst + c * 8) = v; });
        } else if (t < q_tasks + k_tasks) {
            if (!page_ok) continue;
            const int u = t - q_tasks;
            const int h = u / JB, j = u - h * JB;
            const __nv_bfloat16* src = p.k + (int64_t)b * p.k_sb + (int64_t)s * p.k_ss + (int64_t)h * p.k_sh;
            __nv_bfloat16* dst = p.k_cache + (int64_t)page * p.kc_sblk + (int64_t)slot * p.kc_stok + (int64_t)h * p.kc_sh;
            rope_block(src, j, rd, p.interleaved != 0, cosr, sinr,
                       [&](int c, const uint4& v) { *reinterpret_cast<uint4*>(dst + c * 8) = v; });
        } else {
            if (!page_ok) continue;
            const int u = t - q_tasks - k_tasks;
            const int h = u / CPR, c = u - h * CPR;
            const __nv_bfloat16* src = p.v + (int64_t)b * p.v_sb + (int64_t)s * p.v_ss + (int64_t)h * p.v_sh;
            __nv_bfloat16* dst = p.v_cache + (int64_t)page * p.vc_sblk + (int64_t)slot * p.vc_stok + (int64_t)h * p.vc_sh;
            *reinterpret_cast<uint4*>(dst + c * 8) = ld_v4(src + c * 8);
        }
    }
}

// ------------------------------------------------------------------ kernel 2: attention
struct PrefillParams {
    const int32_t* seqlens; const int32_t* block_table; __nv_bfloat16* out;
    int B, S, Hq, Hkv, G;
    int n_blocks, block_size, max_blocks_per_seq, causal;
    int64_t o_sb, o_ss, o_sh, bt_sb;
    float scale_log2;
    int nx, ny;           // 1-D grids: blockIdx.x = (x * B + b) * ny + y, x = causal weight rank (0 = heaviest row tile)
    long long* trace;     // debug only (b200_debug_prefill_trace): [16 events][64 tiles] clock64 stamps of CTA (0, 0, 0)
};

// Pipeline trace of ONE CTA (the heaviest: blockIdx 0,0,0): which side of each mbarrier hand-off waits.
__device__ __forceinline__ void pf_stamp(const PrefillParams& p, bool who, int ev, int j) {
    if (p.trace != nullptr && who && j < 64 && blockIdx.x == 0)
        p.trace[ev * 64 + j] = clock64();
}

// ------------------------------------------------------------------ kernel 2b: attention, operands in TMEM
// First-round kernel (one query tile per CTA; its predecessor with Q and P staged in shared memory is in the
// git history, retired in round 2).  The two A operands never touch shared memory: the
// rotated Q tile is written ONCE into tensor memory (tcgen05.st by the softmax warps) and the
// probabilities P go registers -> TMEM as packed bf16x2.  With Q and P in shared memory every KV
// tile moves 224 KB through the 128 B/clk shared-memory port (MMA operand reads 128 KB, P 32 KB,
// TMA fill 64 KB) against 1024 tensor-pipe cycles -> shared-memory bound at ~55 %; with both in
// TMEM only the K and V tiles remain (128 KB per tile) and the freed 64 KB buy a third K/V stage.
template <int D_>
struct PCfg2 {
    static constexpr int D = D_;
    static constexpr int BM = 128, BN = 128;
    static constexpr int SLAB_B = 128 * 128;
    static constexpr int DSLABS = D / 64;
    static constexpr int TILE_B = DSLABS * SLAB_B;
    static constexpr int KV_STAGES = 3;
    static constexpr int OFF_K = 0;
    static constexpr int OFF_V = OFF_K + KV_STAGES * TILE_B;
    static constexpr int OFF_BAR = OFF_V + KV_STAGES * TILE_B;
    static constexpr int NBARS = 20;
    static constexpr int OFF_XCH = OFF_BAR + NBARS * 8 + 16;
    static constexpr int SMEM = OFF_XCH + 2 * 2 * 128 * 4 + 1024;
    static constexpr int SOFTMAX_THREADS = 256;
    static constexpr int THREADS = 64 + SOFTMAX_THREADS;
    static constexpr int TMEM_COLS = 512;
    static constexpr int TM_S0 = 0, TM_S1 = 128, TM_O = 256, TM_Q = 384, TM_P = 448;   // Q: D/2 columns, P: 64 columns
};

enum { B2_Q = 0, B2_KF = 1, B2_KE = 4, B2_VF = 7, B2_VE = 10, B2_SF = 13, B2_SE = 15, B2_PF = 17, B2_PE = 18 };

// SM = softmax arithmetic: 0 = scalar FFMA/MUFU/FADD with a 256-thread row-max exchange (first version);
// k > 0 = packed fp32x2 arithmetic, 4 independent max chains, per-lane-quarter 64-thread exchange
// barriers, and (k-1) of every 4 score pairs exponentiated by the FMA-pipe polynomial (MUFU relief).
template <class C, int SM>
__global__ void __launch_bounds__(C::THREADS, 1)
prefill_attn_tmem_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                         const __nv_bfloat16* __restrict__ q_rot, const __grid_constant__ PrefillParams p) {
    constexpr int D = C::D;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_u32 = smem_u32(smem_raw);
    const uint32_t base = (raw_u32 + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw_u32);
    const uint32_t sK = base + C::OFF_K, sV = base + C::OFF_V;
    const uint32_t bars = base + C::OFF_BAR;
    auto bar = [&](int i) { return bars + 8u * (uint32_t)i; };
    volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + C::NBARS * 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // longest-processing-time-first order over the WHOLE grid: every (head, batch) CTA of the heaviest causal
    // row tile is dispatched before any CTA of the next one, so the tail of the launch is made of short CTAs
    const int hq = (int)blockIdx.x % p.ny;
    const int xb = (int)blockIdx.x / p.ny;
    const int b = xb % p.B;
    const int qt = p.nx - 1 - xb / p.B;
    const int q0 = qt * C::BM;
    const int kvh = hq / p.G;
    const int prefix = __ldg(p.seqlens + b);
    const int seqlen_k = prefix + p.S;
    const int kv_end = p.causal ? min(seqlen_k, prefix + q0 + C::BM) : seqlen_k;
    const int n_tiles = (kv_end + C::BN - 1) / C::BN;

    if (tid == 0) {
        mbar_init(bar(B2_Q), C::SOFTMAX_THREADS);
        for (int s = 0; s < C::KV_STAGES; ++s) {
            mbar_init(bar(B2_KF + s), 1); mbar_init(bar(B2_KE + s), 1);
            mbar_init(bar(B2_VF + s), 1); mbar_init(bar(B2_VE + s), 1);
        }
        for (int s = 0; s < 2; ++s) { mbar_init(bar(B2_SF + s), 1); mbar_init(bar(B2_SE + s), C::SOFTMAX_THREADS); }
        mbar_init(bar(B2_PF), C::SOFTMAX_THREADS); mbar_init(bar(B2_PE), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32((const void*)s_tmem)), "r"((uint32_t)C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp == 0) {
        // ===================================================== TMA producer (K/V ring, 3 stages)
        if (lane == 0) {
            const int32_t* bt = p.block_table + (int64_t)b * p.bt_sb;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % C::KV_STAGES;
                const uint32_t ph = (uint32_t)(j / C::KV_STAGES) & 1u;
                const int n0 = j * C::BN;
                const int pe = n0 / p.block_size;
                int page = (pe < p.max_blocks_per_seq) ? __ldg(bt + pe) : 0;
                if (page < 0 || page >= p.n_blocks) page = 0;
                const int tok = n0 - pe * p.block_size;
                mbar_wait(bar(B2_KE + st), ph ^ 1u);
                pf_stamp(p, true, 9, j);
                mbar_expect_tx(bar(B2_KF + st), C::TILE_B);
#pragma unroll
                for (int sl = 0; sl < C::DSLABS; ++sl)
                    tma_load_4d(sK + st * C::TILE_B + sl * C::SLAB_B, &tm_k, bar(B2_KF + st), sl * 64, kvh, tok, page);
                mbar_wait(bar(B2_VE + st), ph ^ 1u);
                pf_stamp(p, true, 10, j);
                mbar_expect_tx(bar(B2_VF + st), C::TILE_B);
#pragma unroll
                for (int sl = 0; sl < C::DSLABS; ++sl)
                    tma_load_4d(sV + st * C::TILE_B + sl * C::SLAB_B, &tm_v, bar(B2_VF + st), sl * 64, kvh, tok, page);
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp waits, one elected lane issues)
        constexpr uint32_t IDESC_QK = umma_idesc_bf16(C::BM, C::BN, 0, 0);   // A=Q (TMEM), B=K K-major
        constexpr uint32_t IDESC_PV = umma_idesc_bf16(C::BM, D, 0, 1);       // A=P (TMEM), B=V MN-major
        auto issue_pv = [&](int t) {
            const int st = t % C::KV_STAGES;
            pf_stamp(p, lane == 0, 7, t);
            mbar_wait(bar(B2_VF + st), (uint32_t)(t / C::KV_STAGES) & 1u);
            if (t == n_tiles - 1 && seqlen_k - t * C::BN < C::BN) {
                zero_v_tail<C::DSLABS>(sV + st * C::TILE_B, seqlen_k - t * C::BN, lane, 32);
                __syncwarp();
            }
            mbar_wait(bar(B2_PF), (uint32_t)t & 1u);
            pf_stamp(p, lane == 0, 8, t);
            tc_fence_after();
            const uint32_t v_base = sV + st * C::TILE_B;
            const uint32_t first = (t > 0) ? 1u : 0u;
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < C::BN / 16; ++k) {
                    const uint64_t db = umma_desc_sw128(v_base + k * 2048, C::SLAB_B, 1024);
                    umma_bf16_ts(tmem + C::TM_O, tmem + C::TM_P + k * 8, db, IDESC_PV, k > 0 ? 1u : first);
                }
                tc_commit(bar(B2_VE + st));
                tc_commit(bar(B2_PE));
            }
            __syncwarp();
        };
        mbar_wait(bar(B2_Q), 0);
        tc_fence_after();
        for (int j = 0; j < n_tiles; ++j) {
            const int st = j % C::KV_STAGES;
            const int sb = j & 1;
            mbar_wait(bar(B2_KF + st), (uint32_t)(j / C::KV_STAGES) & 1u);
            mbar_wait(bar(B2_SE + sb), ((uint32_t)(j >> 1) & 1u) ^ 1u);
            pf_stamp(p, lane == 0, 6, j);
            tc_fence_after();
            const uint32_t k_base = sK + st * C::TILE_B;
            const uint32_t s_tm = tmem + (sb ? C::TM_S1 : C::TM_S0);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < D / 16; ++k) {
                    const uint64_t db = umma_desc_sw128(k_base + (k >> 2) * C::SLAB_B + (k & 3) * 32, 16, 1024);
                    umma_bf16_ts(s_tm, tmem + C::TM_Q + k * 8, db, IDESC_QK, k > 0 ? 1u : 0u);
                }
                tc_commit(bar(B2_KE + st));
                tc_commit(bar(B2_SF + sb));
            }
            __syncwarp();
            if (j > 0) issue_pv(j - 1);
        }
        issue_pv(n_tiles - 1);
    } else {
        // ===================================================== softmax / correction / epilogue (8 warps)
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const int qpos = p.causal ? (prefix + q0 + row) : 0x7fffffff;
        const float sl2 = p.scale_log2;
        constexpr int HC = C::BN / 2;
        constexpr int HD = D / 2;
        const bool live = (q0 + row) < p.S;

        // ---- Q tile -> tensor memory, once: thread (row, half) stores D/2 bf16 = D/4 packed columns
        {
            constexpr int QW = HD / 2;                         // 32-bit words per thread: 32 (D=128) or 16 (D=64)
            uint32_t qw[QW];
            if (live) {
                const __nv_bfloat16* src = q_rot + (((int64_t)b * p.S + (q0 + row)) * p.Hq + hq) * D + half * HD;
#pragma unroll
                for (int i = 0; i < QW / 4; ++i) {
                    const uint4 v = ld_nc_v4(src + i * 8);
                    qw[4 * i] = v.x; qw[4 * i + 1] = v.y; qw[4 * i + 2] = v.z; qw[4 * i + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < QW; ++i) qw[i] = 0u;
            }
            const uint32_t tq = tmem + lane_sel + C::TM_Q + half * QW;
            if constexpr (QW == 32) tmem_st32(tq, *reinterpret_cast<const uint32_t(*)[32]>(qw));
            else                    tmem_st16(tq, *reinterpret_cast<const uint32_t(*)[16]>(qw));
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(bar(B2_Q));
        }

        float m_ref = -CUDART_INF_F, l = 0.0f;
        volatile float* xch = reinterpret_cast<volatile float*>(smem + C::OFF_XCH);

        for (int j = 0; j < n_tiles; ++j) {
            const int sb = j & 1;
            const int n0 = j * C::BN + half * HC;
            const bool tr = (tid == 64);                       // first softmax thread
            pf_stamp(p, tr, 11, j);
            mbar_wait(bar(B2_SF + sb), (uint32_t)(j >> 1) & 1u);
            pf_stamp(p, tr, 0, j);
            tc_fence_after();
            uint32_t su[HC];
#define SF(c) __uint_as_float(su[(c)])
            {
                const uint32_t ts = tmem + lane_sel + (sb ? C::TM_S1 : C::TM_S0) + half * HC;
#pragma unroll
                for (int c = 0; c < HC / 32; ++c) {
                    uint32_t r[32];
                    tmem_ld32(ts + c * 32, r);
#pragma unroll
                    for (int i = 0; i < 32; ++i) su[c * 32 + i] = r[i];
                }
                tmem_wait_ld();
            }
            tc_fence_before();
            mbar_arrive(bar(B2_SE + sb));
            pf_stamp(p, tr, 1, j);

            const bool masked = n0 + HC - 1 > prefix + q0 || n0 + HC > seqlen_k;   // warp-uniform
            if (masked) {
#pragma unroll
                for (int c = 0; c < HC; ++c) {
                    const int tok = n0 + c;
                    if (tok > qpos || tok >= seqlen_k) su[c] = 0xff800000u;
                }
            }
            float mx;
            if constexpr (SM == 0) {
                mx = SF(0);
#pragma unroll
                for (int c = 1; c < HC; ++c) mx = fmaxf(mx, SF(c));
            } else {                                           // 4 independent chains (FMNMX3-friendly)
                float m4[4] = {SF(0), SF(1), SF(2), SF(3)};
#pragma unroll
                for (int c = 4; c < HC; c += 8) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) m4[u] = fmaxf(fmaxf(m4[u], SF(c + u)), (c + 4 + u < HC) ? SF(c + 4 + u) : SF(c + u));
                }
                mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
            }
            xch[(j & 1) * 256 + half * 128 + row] = mx;
            if constexpr (SM == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
            else asm volatile("bar.sync %0, 64;" :: "r"(1 + quarter) : "memory");   // only the 2 warps sharing these rows
            mx = fmaxf(mx, xch[(j & 1) * 256 + (half ^ 1) * 128 + row]);
            pf_stamp(p, tr, 2, j);
            float alpha = 1.0f;
            bool grow = false;
            if (j == 0) {
                m_ref = mx;
            } else if ((mx - m_ref) * sl2 > 8.0f) {
                alpha = fast_exp2((m_ref - mx) * sl2);
                m_ref = mx;
                l *= alpha;
                grow = true;
            }
            const float mr = (m_ref == -CUDART_INF_F) ? 0.0f : m_ref * sl2;
            float rs = 0.0f;
            uint32_t pw[HC / 2];                               // probabilities, packed bf16x2 (A operand of P.V)
            if constexpr (SM == 0) {
#pragma unroll
                for (int c = 0; c < HC; c += 2) {
                    const float e0 = fast_exp2(SF(c) * sl2 - mr), e1 = fast_exp2(SF(c + 1) * sl2 - mr);
                    rs += e0 + e1;
                    pw[c >> 1] = pack_bf16(e0, e1);
                }
            } else if (masked) {                               // masked scores must give exact zeros: MUFU only
                rs = softmax_exp_tile<0, HC>(su, sl2, mr, pw);
            } else {
                rs = softmax_exp_tile<SM - 1, HC>(su, sl2, mr, pw);
            }
            l += rs;

            pf_stamp(p, tr, 3, j);
            mbar_wait(bar(B2_PE), ((uint32_t)j & 1u) ^ 1u);    // PV_{j-1} retired: P columns free, O stable
            pf_stamp(p, tr, 4, j);
            tc_fence_after();
            tmem_st32(tmem + lane_sel + C::TM_P + half * (HC / 2), pw);
            if (j > 0 && __any_sync(0xffffffffu, grow)) {
                const uint32_t to = tmem + lane_sel + C::TM_O + half * HD;
#pragma unroll
                for (int c = 0; c < HD / 32; ++c) {
                    uint32_t r[32];
                    tmem_ld32(to + c * 32, r);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                    tmem_st32(to + c * 32, r);
                }
            }
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(bar(B2_PF));
            pf_stamp(p, tr, 5, j);
#undef SF
        }

        asm volatile("bar.sync 1, 256;" ::: "memory");
        xch[half * 128 + row] = l;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        l += xch[(half ^ 1) * 128 + row];
        mbar_wait(bar(B2_PE), (uint32_t)(n_tiles - 1) & 1u);
        tc_fence_after();
        const float inv = (l > 0.0f) ? 1.0f / l : 0.0f;
        __nv_bfloat16* orow = p.out + (int64_t)b * p.o_sb + (int64_t)(q0 + row) * p.o_ss + (int64_t)hq * p.o_sh + half * HD;
        const uint32_t to = tmem + lane_sel + C::TM_O + half * HD;
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
            uint32_t r[32];
            tmem_ld32(to + c * 32, r);
            tmem_wait_ld();
            if (live) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 v;
                    v.x = pack_bf16(__uint_as_float(r[i * 8 + 0]) * inv, __uint_as_float(r[i * 8 + 1]) * inv);
                    v.y = pack_bf16(__uint_as_float(r[i * 8 + 2]) * inv, __uint_as_float(r[i * 8 + 3]) * inv);
                    v.z = pack_bf16(__uint_as_float(r[i * 8 + 4]) * inv, __uint_as_float(r[i * 8 + 5]) * inv);
                    v.w = pack_bf16(__uint_as_float(r[i * 8 + 6]) * inv, __uint_as_float(r[i * 8 + 7]) * inv);
                    *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"((uint32_t)C::TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ kernel 2c: two Q tiles per CTA
// One CTA owns TWO 128-row query tiles that share every K/V tile it streams (two query heads of one KV
// group at the same rows when G is even, else two consecutive row tiles of one head): K/V traffic from L2
// halves, and the tensor pipe works on one tile's GEMMs while the other tile's softmax runs (ping-pong).
//   TMEM   : S_A | S_B | O_A | O_B  (4 x 128 columns).  The probabilities overwrite the scores in place:
//            P_x = packed bf16x2 in score columns 0..31 (keys 0..63) and 64..95 (keys 64..127) of S_x,
//            consumed as the TMEM A operand of P.V.
//            No "S free" barrier is needed — Q.K^T of tile j+1 is issued after P.V of tile j by the same
//            thread, and tcgen05.mma executes in issue order.
//   smem   : Q_A, Q_B (TMA, once), 2-stage K ring, 2-stage V ring (128B-swizzled UMMA tiles).
//   warps  : 0 = TMA producer, 1 = MMA issuer, 2-9 = softmax of tile A, 10-17 = softmax of tile B; two
//            threads per query row (64 key columns each, row max exchanged through shared memory behind a
//            64-thread named barrier), two passes over the columns in 32-column chunks of tcgen05.ld (max,
//            then exp / pack / tcgen05.st into the thread's OWN score columns): <= 96 registers at 18 warps.
template <int D_>
struct PCfg3 {
    static constexpr int D = D_;
    static constexpr int BM = 128, BN = 128;
    static constexpr int SLAB_B = 128 * 128;
    static constexpr int DSLABS = D / 64;
    static constexpr int TILE_B = DSLABS * SLAB_B;
    static constexpr int KV_STAGES = 2;
    static constexpr int OFF_Q = 0;                                  // Q_A, Q_B
    static constexpr int OFF_K = OFF_Q + 2 * TILE_B;
    static constexpr int OFF_V = OFF_K + KV_STAGES * TILE_B;
    static constexpr int OFF_BAR = OFF_V + KV_STAGES * TILE_B;
    static constexpr int NBARS = 16;
    static constexpr int OFF_XCH = OFF_BAR + NBARS * 8 + 16;         // row max / row sum exchange [2 tiles][2][2][128] f32
    static constexpr int SMEM = OFF_XCH + 2 * 2 * 2 * 128 * 4 + 1024;
    static constexpr int GROUP_THREADS = 256;                        // two threads per query row
    static constexpr int THREADS = 64 + 2 * GROUP_THREADS;
    static constexpr int TMEM_COLS = 512;
    static constexpr int TM_S = 0, TM_O = 256;                       // + 128 * tile
};

enum { B3_Q = 0, B3_KF = 1, B3_KE = 3, B3_VF = 5, B3_VE = 7, B3_SF = 9, B3_PF = 11, B3_PD = 13 };

struct PairParams {
    PrefillParams p;
    int pair_heads;        // 1: tiles = heads (2*y, 2*y+1) at rows x*128;  0: tiles = rows (2*x, 2*x+1)*128 of head y
};

template <class C, int SM>
__global__ void __launch_bounds__(C::THREADS, 1)
prefill_attn_pair_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ PairParams pp) {
    constexpr int D = C::D;
    const PrefillParams& p = pp.p;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_u32 = smem_u32(smem_raw);
    const uint32_t base = (raw_u32 + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw_u32);
    const uint32_t sQ = base + C::OFF_Q, sK = base + C::OFF_K, sV = base + C::OFF_V;
    const uint32_t bars = base + C::OFF_BAR;
    auto bar = [&](int i) { return bars + 8u * (uint32_t)i; };
    volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + C::NBARS * 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // longest-processing-time-first order over the whole grid (see prefill_attn_tmem_kernel)
    const int by = (int)blockIdx.x % p.ny;
    const int xb = (int)blockIdx.x / p.ny;
    const int b = xb % p.B;
    const int xt = p.nx - 1 - xb / p.B;
    // tile t in {0,1}: head hq_t, first row q0_t
    const int hq0 = pp.pair_heads ? 2 * by : by;
    const int hq1 = pp.pair_heads ? hq0 + 1 : hq0;
    const int q00 = pp.pair_heads ? xt * C::BM : 2 * xt * C::BM;
    const int q01 = pp.pair_heads ? q00 : q00 + C::BM;
    const int q01_load = (q01 < p.S) ? q01 : q00;                    // a dead second tile re-reads tile 0 (never stored)
    const int kvh = hq0 / p.G;                                        // == hq1 / G (G even when heads are paired)
    const int prefix = __ldg(p.seqlens + b);
    const int seqlen_k = prefix + p.S;
    const int kv_end = p.causal ? min(seqlen_k, prefix + q01 + C::BM) : seqlen_k;
    const int n_tiles = (kv_end + C::BN - 1) / C::BN;

    if (tid == 0) {
        mbar_init(bar(B3_Q), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar(B3_KF + s), 1); mbar_init(bar(B3_KE + s), 1);
            mbar_init(bar(B3_VF + s), 1); mbar_init(bar(B3_VE + s), 1);
            mbar_init(bar(B3_SF + s), 1); mbar_init(bar(B3_PF + s), C::GROUP_THREADS); mbar_init(bar(B3_PD + s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32((const void*)s_tmem)), "r"((uint32_t)C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp == 0) {
        // ===================================================== TMA producer (warp-uniform loop, elected lane issues)
        if (elect_one()) {
            mbar_expect_tx(bar(B3_Q), 2 * C::TILE_B);
#pragma unroll
            for (int sl = 0; sl < C::DSLABS; ++sl) {
                tma_load_4d(sQ + sl * C::SLAB_B, &tm_q, bar(B3_Q), sl * 64, hq0, q00, b);
                tma_load_4d(sQ + C::TILE_B + sl * C::SLAB_B, &tm_q, bar(B3_Q), sl * 64, hq1, q01_load, b);
            }
        }
        __syncwarp();
        const int32_t* bt = p.block_table + (int64_t)b * p.bt_sb;
        for (int j = 0; j < n_tiles; ++j) {
            const int st = j & 1;
            const uint32_t ph = (uint32_t)(j >> 1) & 1u;
            const int n0 = j * C::BN;
            const int pe = n0 / p.block_size;
            int page = (pe < p.max_blocks_per_seq) ? __ldg(bt + pe) : 0;
            if (page < 0 || page >= p.n_blocks) page = 0;          // memory safety; tokens are masked or undefined
            const int tok = n0 - pe * p.block_size;
            mbar_wait(bar(B3_KE + st), ph ^ 1u);
            pf_stamp(p, lane == 0, 9, j);
            if (elect_one()) {
                mbar_expect_tx(bar(B3_KF + st), C::TILE_B);
#pragma unroll
                for (int sl = 0; sl < C::DSLABS; ++sl)
                    tma_load_4d(sK + st * C::TILE_B + sl * C::SLAB_B, &tm_k, bar(B3_KF + st), sl * 64, kvh, tok, page);
            }
            __syncwarp();
            mbar_wait(bar(B3_VE + st), ph ^ 1u);
            pf_stamp(p, lane == 0, 10, j);
            if (elect_one()) {
                mbar_expect_tx(bar(B3_VF + st), C::TILE_B);
#pragma unroll
                for (int sl = 0; sl < C::DSLABS; ++sl)
                    tma_load_4d(sV + st * C::TILE_B + sl * C::SLAB_B, &tm_v, bar(B3_VF + st), sl * 64, kvh, tok, page);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp waits, one elected lane issues)
        constexpr uint32_t IDESC_QK = umma_idesc_bf16(C::BM, C::BN, 0, 0);   // A=Q K-major (smem), B=K K-major
        constexpr uint32_t IDESC_PV = umma_idesc_bf16(C::BM, D, 0, 1);       // A=P (TMEM), B=V MN-major
        auto issue_qk = [&](int t, int j) {                 // S_t = Q_t . K_j^T ; K_j must have landed
            const uint32_t k_base = sK + (j & 1) * C::TILE_B;
            const uint32_t q_base = sQ + t * C::TILE_B;
            const uint32_t s_tm = tmem + C::TM_S + t * 128;
            pf_stamp(p, lane == 0 && t == 0, 6, j);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < D / 16; ++k) {
                    const uint64_t da = umma_desc_sw128(q_base + (k >> 2) * C::SLAB_B + (k & 3) * 32, 16, 1024);
                    const uint64_t db = umma_desc_sw128(k_base + (k >> 2) * C::SLAB_B + (k & 3) * 32, 16, 1024);
                    umma_bf16(s_tm, da, db, IDESC_QK, k > 0 ? 1u : 0u);
                }
                tc_commit(bar(B3_SF + t));
                if (t == 1) tc_commit(bar(B3_KE + (j & 1)));          // both tiles have consumed K_j
            }
            __syncwarp();
        };
        auto issue_pv = [&](int t, int j) {                 // O_t += P_t . V_j ; waits for tile t's softmax
            pf_stamp(p, lane == 0 && t == 0, 7, j);
            mbar_wait(bar(B3_PF + t), (uint32_t)j & 1u);
            pf_stamp(p, lane == 0 && t == 0, 8, j);
            tc_fence_after();
            const uint32_t v_base = sV + (j & 1) * C::TILE_B;
            const uint32_t p_tm = tmem + C::TM_S + t * 128;            // P overwrote S in place
            const uint32_t o_tm = tmem + C::TM_O + t * 128;
            const uint32_t first = (j > 0) ? 1u : 0u;
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < C::BN / 16; ++k) {
                    const uint64_t db = umma_desc_sw128(v_base + k * 2048, C::SLAB_B, 1024);
                    // P words of tokens 0..63 sit in score columns 0..31, tokens 64..127 in columns 64..95
                    umma_bf16_ts(o_tm, p_tm + (k < 4 ? k * 8 : 64 + (k - 4) * 8), db, IDESC_PV, k > 0 ? 1u : first);
                }
                tc_commit(bar(B3_PD + t));
                if (t == 1) tc_commit(bar(B3_VE + (j & 1)));          // both tiles have consumed V_j
            }
            __syncwarp();
        };
        mbar_wait(bar(B3_Q), 0);
        mbar_wait(bar(B3_KF), 0);
        tc_fence_after();
        issue_qk(0, 0);
        issue_qk(1, 0);
        for (int j = 0; j < n_tiles; ++j) {
            const bool more = (j + 1 < n_tiles);
            mbar_wait(bar(B3_VF + (j & 1)), (uint32_t)(j >> 1) & 1u);
            if (!more && seqlen_k - j * C::BN < C::BN) {
                zero_v_tail<C::DSLABS>(sV + (j & 1) * C::TILE_B, seqlen_k - j * C::BN, lane, 32);
                __syncwarp();
            }
            if (more) mbar_wait(bar(B3_KF + ((j + 1) & 1)), (uint32_t)((j + 1) >> 1) & 1u);
            issue_pv(0, j);
            if (more) issue_qk(0, j + 1);
            issue_pv(1, j);
            if (more) issue_qk(1, j + 1);
        }
    } else {
        // ===================================================== softmax / correction / epilogue
        // 16 warps: tile t = (w >> 3), column half = (w >> 2) & 1, TMEM lane quarter = warp & 3.  TWO threads per
        // query row (64 key columns each) keep four warps on every scheduler, so one warp's MUFU / FMA / ALU
        // stalls are filled by the others (a lone warp per row serialises the three pipes: measured 1780 clk
        // for the exp pass of 128 columns against 1100 clk for two lock-stepped 64-column warps).
        const int w = warp - 2;
        const int t = w >> 3;                                  // which Q tile
        const int half = (w >> 2) & 1;                         // which 64 key columns / which half of D
        const int quarter = warp & 3;                          // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const int hq = t ? hq1 : hq0;
        const int q0 = t ? q01 : q00;
        const int qpos = p.causal ? (prefix + q0 + row) : 0x7fffffff;
        const float sl2 = p.scale_log2;
        constexpr int HC = C::BN / 2, HD = D / 2;
        const uint32_t ts = tmem + lane_sel + C::TM_S + t * 128 + half * HC;     // my 64 score columns; my P words
        const uint32_t to = tmem + lane_sel + C::TM_O + t * 128 + half * HD;     // overwrite their first 32
        constexpr int NPOLY = SM > 0 ? SM - 1 : 0;
        volatile float* xch = reinterpret_cast<volatile float*>(smem + C::OFF_XCH) + t * 512;   // [parity][half][row]
        const int pair_bar = 1 + t * 4 + quarter;              // named barrier of the 2 warps that share these rows
        float m_ref = -CUDART_INF_F, l = 0.0f;

        for (int j = 0; j < n_tiles; ++j) {
            const int n0 = j * C::BN + half * HC;
            const bool tr = (tid == 64);
            pf_stamp(p, tr, 11, j);
            mbar_wait(bar(B3_SF + t), (uint32_t)j & 1u);
            pf_stamp(p, tr, 0, j);
            pf_stamp(p, tr, 1, j);
            tc_fence_after();
            const bool masked = (n0 + HC - 1 > prefix + q0) || (n0 + HC > seqlen_k);   // warp-uniform
            // ---- pass 1: maximum of my 64 columns, then across the two halves of the row
            float m4[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
#pragma unroll
            for (int c = 0; c < HC / 32; ++c) {
                uint32_t r[32];
                tmem_ld32(ts + c * 32, r);
                tmem_wait_ld();
                if (masked) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int tok = n0 + c * 32 + i;
                        if (tok > qpos || tok >= seqlen_k) r[i] = 0xff800000u;
                    }
                }
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        m4[u] = fmaxf(fmaxf(m4[u], __uint_as_float(r[i + u])), __uint_as_float(r[i + 4 + u]));
                }
            }
            float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
            xch[(j & 1) * 256 + half * 128 + row] = mx;
            asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
            mx = fmaxf(mx, xch[(j & 1) * 256 + (half ^ 1) * 128 + row]);
            pf_stamp(p, tr, 2, j);
            // lazy rescale: keep the old reference max unless it is off by more than 2^8
            // (both threads of a row take the same decision from the same numbers)
            float alpha = 1.0f;
            bool grow = false;
            if (j == 0) {
                m_ref = mx;
            } else if ((mx - m_ref) * sl2 > 8.0f) {
                alpha = fast_exp2((m_ref - mx) * sl2);
                m_ref = mx;
                l *= alpha;
                grow = true;
            }
            if (j > 0 && __any_sync(0xffffffffu, grow)) {      // warp-uniform: tcgen05.ld/st are warp-collective
                mbar_wait(bar(B3_PD + t), (uint32_t)(j - 1) & 1u);   // P.V of tile j-1 retired: O stable
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < HD / 32; ++c) {
                    uint32_t r[32];
                    tmem_ld32(to + c * 32, r);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                    tmem_st32(to + c * 32, r);
                }
            }
            const float mr = (m_ref == -CUDART_INF_F) ? 0.0f : m_ref * sl2;
            // ---- pass 2: probabilities over my own score columns (16 packed words per 32 scores land behind my
            //      own read frontier: no other thread reads or writes these columns)
            float rs = 0.0f;
#pragma unroll
            for (int c = 0; c < HC / 32; ++c) {
                uint32_t r[32];
                uint32_t pw[16];
                tmem_ld32(ts + c * 32, r);
                tmem_wait_ld();
                if (masked) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int tok = n0 + c * 32 + i;
                        if (tok > qpos || tok >= seqlen_k) r[i] = 0xff800000u;
                    }
                    rs += softmax_exp_tile<0, 32>(r, sl2, mr, pw);       // exact zeros for masked keys
                } else {
                    rs += softmax_exp_tile<NPOLY, 32>(r, sl2, mr, pw);
                }
                tmem_st16(ts + c * 16, pw);
            }
            l += rs;
            pf_stamp(p, tr, 3, j);
            pf_stamp(p, tr, 4, j);
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(bar(B3_PF + t));
            pf_stamp(p, tr, 5, j);
        }

        // epilogue: row sum across the two halves, O / l -> bf16 -> global (each thread half of D)
        asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
        xch[half * 128 + row] = l;
        asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
        l += xch[(half ^ 1) * 128 + row];
        mbar_wait(bar(B3_PD + t), (uint32_t)(n_tiles - 1) & 1u);
        tc_fence_after();
        const float inv = (l > 0.0f) ? 1.0f / l : 0.0f;
        const bool live = (q0 + row) < p.S;
        __nv_bfloat16* orow = p.out + (int64_t)b * p.o_sb + (int64_t)(q0 + row) * p.o_ss + (int64_t)hq * p.o_sh + half * HD;
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
            uint32_t r[32];
            tmem_ld32(to + c * 32, r);
            tmem_wait_ld();
            if (live) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 v;
                    v.x = pack_bf16(__uint_as_float(r[i * 8 + 0]) * inv, __uint_as_float(r[i * 8 + 1]) * inv);
                    v.y = pack_bf16(__uint_as_float(r[i * 8 + 2]) * inv, __uint_as_float(r[i * 8 + 3]) * inv);
                    v.z = pack_bf16(__uint_as_float(r[i * 8 + 4]) * inv, __uint_as_float(r[i * 8 + 5]) * inv);
                    v.w = pack_bf16(__uint_as_float(r[i * 8 + 6]) * inv, __uint_as_float(r[i * 8 + 7]) * inv);
                    *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"((uint32_t)C::TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ kernel 2d: persistent pair kernel
// The pair kernel above pays ~16 000 clk per CTA outside its steady-state loop (measured: ncu sm__cycles_active x 148
// minus tiles x 3356 clk, the same figure at S = 512, 2048 and 8192 — profiles/r02/prefill_small_r02.md): CTA launch
// behind the previous CTA's exit, barrier init, tcgen05.alloc behind the previous dealloc, descriptor fetch, Q load
// latency, pipeline fill, drain and the epilogue with the tensor pipe idle.  At the reference's prompt sizes (<= 2048
// tokens, engine.py:45) a CTA has 1..16 KV tiles, so that fixed cost was 40-60 % of the kernel.
// Here ONE CTA PER SM stays resident and pulls work items (the same pair of 128-row query tiles, in the same
// longest-first order) from a device counter that kernel 1 zeroes:
//   * TMEM, mbarriers and descriptors are set up once per SM; every mbarrier phase runs on a tile counter that keeps
//     counting across items, so the K/V ring never drains between items: the producer is already loading the next
//     item's Q (single buffer, released by a tcgen05.commit behind the item's last Q.K^T) and K/V tiles while the
//     softmax warps finish the previous item's last tile and epilogue;
//   * the next item's first Q.K^T is issued right behind the last P.V of the current item (in-order tensor pipe:
//     it may overwrite S/P), and its first P.V — the first write to O — is ordered behind the epilogue's tcgen05.ld
//     of O by the softmax warps themselves (same threads: epilogue, then the next item's first tile);
//   * dynamic pull (one atomicAdd per item by the producer behind the item's last tile loads, published with the prefix length
//     through an 8-slot shared-memory ring + mbarrier) keeps the longest-first list balanced like the hardware CTA
//     scheduler did.
template <int D_>
struct PCfg4 : PCfg3<D_> {
    using Base = PCfg3<D_>;
    static constexpr int NBARS = 32;
    static constexpr int NSLOT = 8;
    static constexpr int OFF_XCH = Base::OFF_BAR + NBARS * 8 + 16;           // row max exchange [2 tiles][2][2][128] f32
    static constexpr int OFF_LX = OFF_XCH + 2 * 2 * 2 * 128 * 4;             // row sum exchange [2 tiles][2][128] f32
    static constexpr int OFF_ITEM = OFF_LX + 2 * 2 * 128 * 4;                // item ring [NSLOT] {id, prefix}
    static constexpr int OFF_STG = OFF_ITEM + NSLOT * 8;                     // epilogue staging: 16 warps x [16 rows][64 B]
    static constexpr int SMEM = OFF_STG + 16 * 1024 + 1024;
};
enum { B4_QE = 15, B4_IF = 16 };

struct PersistParams {
    PrefillParams p;
    int pair_heads;
    int total;             // work items = nx * ny * B
    int* counter;          // zeroed by prefill_rope_append_kernel of the same call
};

template <class C, int SM>
__global__ void __launch_bounds__(C::THREADS, 1)
prefill_attn_persist_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                            const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ PersistParams pp) {
    constexpr int D = C::D;
    const PrefillParams& p = pp.p;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_u32 = smem_u32(smem_raw);
    const uint32_t base = (raw_u32 + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw_u32);
    const uint32_t sQ = base + C::OFF_Q, sK = base + C::OFF_K, sV = base + C::OFF_V;
    const uint32_t bars = base + C::OFF_BAR;
    auto bar = [&](int i) { return bars + 8u * (uint32_t)i; };
    volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + C::NBARS * 8);
    volatile int* s_item = reinterpret_cast<volatile int*>(smem + C::OFF_ITEM);
    const int total = pp.total;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        mbar_init(bar(B3_Q), 1);
        mbar_init(bar(B4_QE), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar(B3_KF + s), 1); mbar_init(bar(B3_KE + s), 1);
            mbar_init(bar(B3_VF + s), 1); mbar_init(bar(B3_VE + s), 1);
            mbar_init(bar(B3_SF + s), 1); mbar_init(bar(B3_PF + s), C::GROUP_THREADS); mbar_init(bar(B3_PD + s), 1);
        }
        for (int s = 0; s < C::NSLOT; ++s) mbar_init(bar(B4_IF + s), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32((const void*)s_tmem)), "r"((uint32_t)C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;
    // everything above overlapped the tail of prefill_rope_append_kernel (PDL); everything below reads what it wrote
    // (work counter, rotated Q, K/V pages) or writes `out`
    pdl_wait_prev();

    // one work item: the pair of query tiles of prefill_attn_pair_kernel, same longest-first numbering
    struct Item { int b, hq0, hq1, q00, q01, q01_load, kvh, prefix, seqlen_k, n_tiles; };
    auto decode = [&](int id, int prefix) {
        Item I;
        const int by = id % p.ny;
        const int xb = id / p.ny;
        I.b = xb % p.B;
        const int xt = p.nx - 1 - xb / p.B;
        I.hq0 = pp.pair_heads ? 2 * by : by;
        I.hq1 = pp.pair_heads ? I.hq0 + 1 : I.hq0;
        I.q00 = pp.pair_heads ? xt * C::BM : 2 * xt * C::BM;
        I.q01 = pp.pair_heads ? I.q00 : I.q00 + C::BM;
        I.q01_load = (I.q01 < p.S) ? I.q01 : I.q00;
        I.kvh = I.hq0 / p.G;
        I.prefix = prefix;
        I.seqlen_k = prefix + p.S;
        const int kv_end = p.causal ? min(I.seqlen_k, prefix + I.q01 + C::BM) : I.seqlen_k;
        I.n_tiles = (kv_end + C::BN - 1) / C::BN;
        return I;
    };
    // consumers: wait for item number `it` of this CTA; false = no more work
    auto next_item = [&](int it, Item& I) {
        const int slot = it % C::NSLOT;
        mbar_wait(bar(B4_IF + slot), (uint32_t)(it / C::NSLOT) & 1u);
        const int id = s_item[2 * slot], prefix = s_item[2 * slot + 1];
        if (id >= total) return false;
        I = decode(id, prefix);
        return true;
    };

    if (warp == 0) {
        // ===================================================== scheduler + TMA producer
        int nid = 0, npf = 0;
        auto fetch = [&]() {
            int v = 0, pf = 0;
            if (lane == 0) {
                v = atomicAdd(pp.counter, 1);
                if (v < total) pf = __ldg(p.seqlens + (v / p.ny) % p.B);
            }
            nid = __shfl_sync(0xffffffffu, v, 0);
            npf = __shfl_sync(0xffffffffu, pf, 0);
        };
        auto publish = [&](int it) {
            if (lane == 0) {
                const int slot = it % C::NSLOT;
                s_item[2 * slot] = nid; s_item[2 * slot + 1] = npf;
                mbar_arrive(bar(B4_IF + slot));                       // release: the slot is visible to the waiters
            }
            __syncwarp();
        };
        fetch();
        publish(0);
        uint32_t g = 0;                                               // K/V tiles issued so far (all items)
        for (int it = 0; nid < total; ++it) {
            const Item I = decode(nid, npf);
            mbar_wait(bar(B4_QE), (uint32_t)(it & 1) ^ 1u);           // previous item's last Q.K^T has read Q
            if (elect_one()) {
                mbar_expect_tx(bar(B3_Q), 2 * C::TILE_B);
#pragma unroll
                for (int sl = 0; sl < C::DSLABS; ++sl) {
                    tma_load_4d(sQ + sl * C::SLAB_B, &tm_q, bar(B3_Q), sl * 64, I.hq0, I.q00, I.b);
                    tma_load_4d(sQ + C::TILE_B + sl * C::SLAB_B, &tm_q, bar(B3_Q), sl * 64, I.hq1, I.q01_load, I.b);
                }
            }
            __syncwarp();
            const int32_t* bt = p.block_table + (int64_t)I.b * p.bt_sb;
            const int j_pull = max(0, I.n_tiles - 3);
            for (int j = 0; j < I.n_tiles; ++j, ++g) {
                const int st = (int)(g & 1u);
                const uint32_t ph = (g >> 1) & 1u;
                const int n0 = j * C::BN;
                const int pe = n0 / p.block_size;
                int page = (pe < p.max_blocks_per_seq) ? __ldg(bt + pe) : 0;
                if (page < 0 || page >= p.n_blocks) page = 0;          // memory safety; tokens are masked or undefined
                const int tok = n0 - pe * p.block_size;
                mbar_wait(bar(B3_KE + st), ph ^ 1u);
                pf_stamp(p, lane == 0 && it == 0, 9, j);
                if (elect_one()) {
                    mbar_expect_tx(bar(B3_KF + st), C::TILE_B);
#pragma unroll
                    for (int sl = 0; sl < C::DSLABS; ++sl)
                        tma_load_4d(sK + st * C::TILE_B + sl * C::SLAB_B, &tm_k, bar(B3_KF + st), sl * 64, I.kvh, tok, page);
                }
                __syncwarp();
                mbar_wait(bar(B3_VE + st), ph ^ 1u);
                pf_stamp(p, lane == 0 && it == 0, 10, j);
                if (elect_one()) {
                    mbar_expect_tx(bar(B3_VF + st), C::TILE_B);
#pragma unroll
                    for (int sl = 0; sl < C::DSLABS; ++sl)
                        tma_load_4d(sV + st * C::TILE_B + sl * C::SLAB_B, &tm_v, bar(B3_VF + st), sl * 64, I.kvh, tok, page);
                }
                __syncwarp();
                if (j == j_pull) {
                    // Pull the next item (behind this tile's loads) three tiles before the end of this one: late enough that the list stays a
                    // DYNAMIC longest-first queue (pulling at the start of every item hands each CTA its first two items
                    // at launch — a static round-robin split, measured 63 us instead of 48 us at S = 2048 with 256 items
                    // on 148 SMs), early enough that the atomic, the next Q load and its first Q.K^T hide behind this
                    // item's remaining tiles; items of <= 3 tiles (the end of the list) pull at once.
                    fetch();
                    publish(it + 1);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp waits, one elected lane issues)
        constexpr uint32_t IDESC_QK = umma_idesc_bf16(C::BM, C::BN, 0, 0);   // A=Q K-major (smem), B=K K-major
        constexpr uint32_t IDESC_PV = umma_idesc_bf16(C::BM, D, 0, 1);       // A=P (TMEM), B=V MN-major
        bool trace_it = true;
        // gj = tile number over all items (barrier stages / phases); `last` = the item's last tile: Q is free behind it
        auto issue_qk = [&](int t, uint32_t gj, int j, bool last) {
            const uint32_t k_base = sK + (gj & 1u) * C::TILE_B;
            const uint32_t q_base = sQ + t * C::TILE_B;
            const uint32_t s_tm = tmem + C::TM_S + t * 128;
            pf_stamp(p, lane == 0 && t == 0 && trace_it, 6, j);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < D / 16; ++k) {
                    const uint64_t da = umma_desc_sw128(q_base + (k >> 2) * C::SLAB_B + (k & 3) * 32, 16, 1024);
                    const uint64_t db = umma_desc_sw128(k_base + (k >> 2) * C::SLAB_B + (k & 3) * 32, 16, 1024);
                    umma_bf16(s_tm, da, db, IDESC_QK, k > 0 ? 1u : 0u);
                }
                tc_commit(bar(B3_SF + t));
                if (t == 1) {
                    tc_commit(bar(B3_KE + (gj & 1u)));                // both tiles have consumed K_j
                    if (last) tc_commit(bar(B4_QE));                  // ... and nobody reads Q of this item any more
                }
            }
            __syncwarp();
        };
        auto issue_pv = [&](int t, uint32_t gj, int j) {              // O_t += P_t . V_j ; waits for tile t's softmax
            pf_stamp(p, lane == 0 && t == 0 && trace_it, 7, j);
            mbar_wait(bar(B3_PF + t), gj & 1u);
            pf_stamp(p, lane == 0 && t == 0 && trace_it, 8, j);
            tc_fence_after();
            const uint32_t v_base = sV + (gj & 1u) * C::TILE_B;
            const uint32_t p_tm = tmem + C::TM_S + t * 128;            // P overwrote S in place
            const uint32_t o_tm = tmem + C::TM_O + t * 128;
            const uint32_t first = (j > 0) ? 1u : 0u;
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < C::BN / 16; ++k) {
                    const uint64_t db = umma_desc_sw128(v_base + k * 2048, C::SLAB_B, 1024);
                    umma_bf16_ts(o_tm, p_tm + (k < 4 ? k * 8 : 64 + (k - 4) * 8), db, IDESC_PV, k > 0 ? 1u : first);
                }
                tc_commit(bar(B3_PD + t));
                if (t == 1) tc_commit(bar(B3_VE + (gj & 1u)));         // both tiles have consumed V_j
            }
            __syncwarp();
        };
        uint32_t g = 0;
        Item I;
        for (int it = 0; next_item(it, I); ++it) {
            trace_it = (it == 0);
            const int n_tiles = I.n_tiles;
            mbar_wait(bar(B3_Q), (uint32_t)it & 1u);
            mbar_wait(bar(B3_KF + (g & 1u)), (g >> 1) & 1u);
            tc_fence_after();
            issue_qk(0, g, 0, n_tiles == 1);
            issue_qk(1, g, 0, n_tiles == 1);
            for (int j = 0; j < n_tiles; ++j) {
                const uint32_t gj = g + (uint32_t)j;
                const bool more = (j + 1 < n_tiles);
                mbar_wait(bar(B3_VF + (gj & 1u)), (gj >> 1) & 1u);
                if (!more && I.seqlen_k - j * C::BN < C::BN) {
                    zero_v_tail<C::DSLABS>(sV + (gj & 1u) * C::TILE_B, I.seqlen_k - j * C::BN, lane, 32);
                    __syncwarp();
                }
                if (more) mbar_wait(bar(B3_KF + ((gj + 1u) & 1u)), ((gj + 1u) >> 1) & 1u);
                issue_pv(0, gj, j);
                if (more) issue_qk(0, gj + 1u, j + 1, j + 2 == n_tiles);
                issue_pv(1, gj, j);
                if (more) issue_qk(1, gj + 1u, j + 1, j + 2 == n_tiles);
            }
            g += (uint32_t)n_tiles;
        }
    } else {
        // ===================================================== softmax / correction / epilogue (see the pair kernel)
        const int w = warp - 2;
        const int t = w >> 3;                                  // which Q tile
        const int half = (w >> 2) & 1;                         // which 64 key columns / which half of D
        const int quarter = warp & 3;                          // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const float sl2 = p.scale_log2;
        constexpr int HC = C::BN / 2, HD = D / 2;
        const uint32_t ts = tmem + lane_sel + C::TM_S + t * 128 + half * HC;     // my 64 score columns; my P words
        const uint32_t to = tmem + lane_sel + C::TM_O + t * 128 + half * HD;     // overwrite their first 32
        constexpr int NPOLY = SM > 0 ? SM - 1 : 0;
        volatile float* xch = reinterpret_cast<volatile float*>(smem + C::OFF_XCH) + t * 512;   // [parity][half][row]
        volatile float* lx = reinterpret_cast<volatile float*>(smem + C::OFF_LX) + t * 256;     // [half][row]
        const int pair_bar = 1 + t * 4 + quarter;              // named barrier of the 2 warps that share these rows
        uint32_t g = 0;
        Item I;
        for (int it = 0; next_item(it, I); ++it) {
            const int hq = t ? I.hq1 : I.hq0;
            const int q0 = t ? I.q01 : I.q00;
            const int prefix = I.prefix, seqlen_k = I.seqlen_k, n_tiles = I.n_tiles;
            const int qpos = p.causal ? (prefix + q0 + row) : 0x7fffffff;
            // smallest / largest query position of this warp's 32 rows: a 32-key chunk is DEAD for the whole warp when it
            // starts beyond qhi (all probabilities are exact zeros: no tcgen05.ld, no exponentials), needs the mask when
            // it reaches beyond qlo, and takes the unmasked fast path otherwise.  On a diagonal tile that is 6 dead,
            // 4 masked and 6 plain chunks of 16 instead of 16 masked ones (the diagonal tile took 5600 clk, a plain one
            // 3350: profiles/r02/trace_prefill_s2048_pair_before.log).
            const int qlo = p.causal ? (prefix + q0 + quarter * 32) : 0x3fffffff;     // qhi = qlo + 31
            float m_ref = -CUDART_INF_F, l = 0.0f;
            for (int j = 0; j < n_tiles; ++j) {
                const uint32_t gj = g + (uint32_t)j;
                const int n0 = j * C::BN + half * HC;
                const bool tr = (tid == 64) && it == 0;
                pf_stamp(p, tr, 11, j);
                mbar_wait(bar(B3_SF + t), gj & 1u);
                pf_stamp(p, tr, 0, j);
                pf_stamp(p, tr, 1, j);
                if (p.trace != nullptr && j == 0 && tid == 64 && blockIdx.x == 0 && it < 64) p.trace[15 * 64 + it] = clock64();
                tc_fence_after();
                // warp-uniform: a tile with no chunk reaching beyond qlo / seqlen_k takes the straight-line code
                const bool plain = !((n0 + HC - 1 > qlo) || (n0 + HC > seqlen_k));
                float m4[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
                if (plain) {
#pragma unroll
                    for (int c = 0; c < HC / 32; ++c) {
                        uint32_t r[32];
                        tmem_ld32(ts + c * 32, r);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; i += 8) {
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                m4[u] = fmaxf(fmaxf(m4[u], __uint_as_float(r[i + u])), __uint_as_float(r[i + 4 + u]));
                        }
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < HC / 32; ++c) {
                        const int c0 = n0 + c * 32;
                        if ((c0 > qlo + 31) || (c0 >= seqlen_k)) continue;            // dead chunk
                        uint32_t r[32];
                        tmem_ld32(ts + c * 32, r);
                        tmem_wait_ld();
                        if ((c0 + 31 > qlo) || (c0 + 32 > seqlen_k)) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                const int tok = n0 + c * 32 + i;
                                if (tok > qpos || tok >= seqlen_k) r[i] = 0xff800000u;
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 8) {
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                m4[u] = fmaxf(fmaxf(m4[u], __uint_as_float(r[i + u])), __uint_as_float(r[i + 4 + u]));
                        }
                    }
                }
                float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
                xch[(gj & 1u) * 256 + half * 128 + row] = mx;
                asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
                mx = fmaxf(mx, xch[(gj & 1u) * 256 + (half ^ 1) * 128 + row]);
                pf_stamp(p, tr, 2, j);
                float alpha = 1.0f;
                bool grow = false;
                if (j == 0) {
                    m_ref = mx;
                } else if ((mx - m_ref) * sl2 > 8.0f) {
                    alpha = fast_exp2((m_ref - mx) * sl2);
                    m_ref = mx;
                    l *= alpha;
                    grow = true;
                }
                if (j > 0 && __any_sync(0xffffffffu, grow)) {      // warp-uniform: tcgen05.ld/st are warp-collective
                    mbar_wait(bar(B3_PD + t), (gj - 1u) & 1u);     // P.V of tile j-1 retired: O stable
                    tc_fence_after();
#pragma unroll
                    for (int c = 0; c < HD / 32; ++c) {
                        uint32_t r[32];
                        tmem_ld32(to + c * 32, r);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tmem_st32(to + c * 32, r);
                    }
                }
                const float mr = (m_ref == -CUDART_INF_F) ? 0.0f : m_ref * sl2;
                float rs = 0.0f;
                if (plain) {
#pragma unroll
                    for (int c = 0; c < HC / 32; ++c) {
                        uint32_t r[32];
                        uint32_t pw[16];
                        tmem_ld32(ts + c * 32, r);
                        tmem_wait_ld();
                        rs += softmax_exp_tile<NPOLY, 32>(r, sl2, mr, pw);
                        tmem_st16(ts + c * 16, pw);
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < HC / 32; ++c) {
                        uint32_t r[32];
                        uint32_t pw[16];
                        const int c0 = n0 + c * 32;
                        if ((c0 > qlo + 31) || (c0 >= seqlen_k)) {                    // dead chunk: P = 0
#pragma unroll
                            for (int i = 0; i < 16; ++i) pw[i] = 0u;
                            tmem_st16(ts + c * 16, pw);
                            continue;
                        }
                        tmem_ld32(ts + c * 32, r);
                        tmem_wait_ld();
                        if ((c0 + 31 > qlo) || (c0 + 32 > seqlen_k)) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                const int tok = n0 + c * 32 + i;
                                if (tok > qpos || tok >= seqlen_k) r[i] = 0xff800000u;
                            }
                            rs += softmax_exp_tile<0, 32>(r, sl2, mr, pw);       // exact zeros for masked keys
                        } else {
                            rs += softmax_exp_tile<NPOLY, 32>(r, sl2, mr, pw);
                        }
                        tmem_st16(ts + c * 16, pw);
                    }
                }
                l += rs;
                pf_stamp(p, tr, 3, j);
                pf_stamp(p, tr, 4, j);
                tmem_wait_st();
                tc_fence_before();
                mbar_arrive(bar(B3_PF + t));
                pf_stamp(p, tr, 5, j);
            }
            g += (uint32_t)n_tiles;

            // epilogue: row sum across the two halves, O / l -> bf16 -> global (each thread half of D).  The leading
            // barrier also orders this item's write of lx behind the partner's read of the previous item's value.
            asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
            lx[half * 128 + row] = l;
            asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
            l += lx[(half ^ 1) * 128 + row];
            if (p.trace != nullptr && tid == 64 && blockIdx.x == 0 && it < 64) p.trace[13 * 64 + it] = clock64();
            mbar_wait(bar(B3_PD + t), (g - 1u) & 1u);
            tc_fence_after();
            if (p.trace != nullptr && tid == 64 && blockIdx.x == 0 && it < 64) p.trace[14 * 64 + it] = clock64();
            const float inv = (l > 0.0f) ? 1.0f / l : 0.0f;
            // TMEM hands every thread ITS OWN row (32 columns per tcgen05.ld): stored directly, one warp instruction
            // touches 32 different 128-byte lines — 4096 line writes per item, ~2.3k clk of LSU time with the tensor
            // pipe idle (profiles/r02/trace_prefill_s2048_persist.log: 3.6k clk from the last P.V to the end of the
            // epilogue).  Transposed through a 1 KB per-warp staging tile ([16 rows][64 B], XOR-swizzled 16-byte units:
            // conflict-free both ways) a warp instruction writes 8 rows x 64 contiguous bytes instead: 4x fewer lines.
            uint8_t* stg = smem + C::OFF_STG + (warp - 2) * 1024;
            __nv_bfloat16* obase = p.out + (int64_t)I.b * p.o_sb + (int64_t)hq * p.o_sh + half * HD;
            const int row_w0 = q0 + quarter * 32;                        // first row of this warp
#pragma unroll
            for (int c = 0; c < HD / 32; ++c) {
                uint32_t r[32];
                tmem_ld32(to + c * 32, r);
                tmem_wait_ld();
                uint4 v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    v[i].x = pack_bf16(__uint_as_float(r[i * 8 + 0]) * inv, __uint_as_float(r[i * 8 + 1]) * inv);
                    v[i].y = pack_bf16(__uint_as_float(r[i * 8 + 2]) * inv, __uint_as_float(r[i * 8 + 3]) * inv);
                    v[i].z = pack_bf16(__uint_as_float(r[i * 8 + 4]) * inv, __uint_as_float(r[i * 8 + 5]) * inv);
                    v[i].w = pack_bf16(__uint_as_float(r[i * 8 + 6]) * inv, __uint_as_float(r[i * 8 + 7]) * inv);
                }
#pragma unroll
                for (int hr = 0; hr < 2; ++hr) {
                    if ((lane >> 4) == hr) {
                        const int rr = lane & 15;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            *reinterpret_cast<uint4*>(stg + rr * 64 + ((i ^ ((rr >> 1) & 3)) << 4)) = v[i];
                    }
                    __syncwarp();
#pragma unroll
                    for (int k2 = 0; k2 < 2; ++k2) {
                        const int u = lane + 32 * k2, rr = u >> 2, seg = u & 3;
                        const uint4 w4 = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((seg ^ ((rr >> 1) & 3)) << 4));
                        const int grow = row_w0 + hr * 16 + rr;
                        if (grow < p.S)
                            *reinterpret_cast<uint4*>(obase + (int64_t)grow * p.o_ss + c * 32 + seg * 8) = w4;
                    }
                    __syncwarp();
                }
            }
            // the next item's first P.V overwrites O: it needs this thread's PF arrive, which comes after these loads
            tc_fence_before();
            if (p.trace != nullptr && tid == 64 && blockIdx.x == 0 && it < 64) p.trace[12 * 64 + it] = clock64();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"((uint32_t)C::TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)ptr;
    }
    return fn;
}

// 4-D bf16 map (d, head, row, outer) with a (64, 1, 128, 1) box, 128B swizzle.
static int make_map(CUtensorMap* map, const void* base, uint64_t d, uint64_t heads, uint64_t rows, uint64_t outer,
                    uint64_t stride_h, uint64_t stride_row, uint64_t stride_outer /* elements */) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("prefill: cuTensorMapEncodeTiled entry point not available"); return B200_E_UNSUPPORTED; }
    cuuint64_t dims[4] = {d, heads, rows, outer};
    cuuint64_t strides[3] = {stride_h * 2, stride_row * 2, stride_outer * 2};
    cuuint32_t box[4] = {64, 1, 128, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("prefill: cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return B200_E_UNSUPPORTED; }
    return 0;
}

template <class C, int SM>
static int launch_prefill_tmem(const CUtensorMap& tk, const CUtensorMap& tv, const __nv_bfloat16* q_rot,
                               const PrefillParams& p, cudaStream_t st) {
    static int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(prefill_attn_tmem_kernel<C, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { set_error("prefill: cannot opt in to %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e)); return (int)e; }
        attr_dev = dev;
    }
    PrefillParams pl = p;
    pl.nx = (p.S + C::BM - 1) / C::BM;
    pl.ny = p.Hq;
    const long long n_ctas = (long long)pl.nx * pl.ny * p.B;
    if (n_ctas > 0x7fffffffLL) { set_error("prefill: grid too large"); return B200_E_RANGE; }
    prefill_attn_tmem_kernel<C, SM><<<(unsigned)n_ctas, C::THREADS, C::SMEM, st>>>(tk, tv, q_rot, pl);
    B200_CHECK_LAUNCH();
    return 0;
}

template <class C, int SM>
static int launch_prefill_pair(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                               const PrefillParams& p, cudaStream_t st) {
    static int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(prefill_attn_pair_kernel<C, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { set_error("prefill: cannot opt in to %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e)); return (int)e; }
        attr_dev = dev;
    }
    PairParams pp;
    pp.p = p;
    pp.pair_heads = (p.G % 2 == 0) ? 1 : 0;
    const int q_tiles = (p.S + C::BM - 1) / C::BM;
    pp.p.nx = pp.pair_heads ? q_tiles : (q_tiles + 1) / 2;
    pp.p.ny = pp.pair_heads ? p.Hq / 2 : p.Hq;
    const long long n_ctas = (long long)pp.p.nx * pp.p.ny * p.B;
    if (n_ctas > 0x7fffffffLL) { set_error("prefill: grid too large"); return B200_E_RANGE; }
    prefill_attn_pair_kernel<C, SM><<<(unsigned)n_ctas, C::THREADS, C::SMEM, st>>>(tq, tk, tv, pp);
    B200_CHECK_LAUNCH();
    return 0;
}

template <class C, int SM>
static int launch_prefill_persist(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                                  const PrefillParams& p, int* counter, int n_sm, cudaStream_t st) {
    static int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(prefill_attn_persist_kernel<C, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { set_error("prefill: cannot opt in to %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e)); return (int)e; }
        attr_dev = dev;
    }
    PersistParams pp;
    pp.p = p;
    pp.pair_heads = (p.G % 2 == 0) ? 1 : 0;
    const int q_tiles = (p.S + C::BM - 1) / C::BM;
    pp.p.nx = pp.pair_heads ? q_tiles : (q_tiles + 1) / 2;
    pp.p.ny = pp.pair_heads ? p.Hq / 2 : p.Hq;
    const long long total = (long long)pp.p.nx * pp.p.ny * p.B;
    if (total > 0x3fffffffLL) { set_error("prefill: too many work items"); return B200_E_RANGE; }
    pp.total = (int)total;
    pp.counter = counter;
    const unsigned grid = (unsigned)(total < n_sm ? total : n_sm);     // one resident CTA per SM
    static const bool pdl = []() { const char* e = getenv("B200_PREFILL_PDL"); return !(e && e[0] == '0'); }();
    cudaError_t le = launch_maybe_pdl(prefill_attn_persist_kernel<C, SM>, dim3(grid), dim3(C::THREADS), (size_t)C::SMEM, st,
                                      pdl, tq, tk, tv, pp);
    if (le != cudaSuccess) { set_error("prefill launch failed: %s", cudaGetErrorString(le)); return (int)le; }
    B200_CHECK_LAUNCH();
    return 0;
}

}  // namespace b200

using namespace b200;

static long long* g_prefill_trace = nullptr;
// Debug hook (not part of the operator surface): device buffer of 16*64 int64 that the next prefill launches
// fill with clock64 stamps of CTA (0,0,0); nullptr switches tracing off.
extern "C" void b200_debug_prefill_trace(void* dev_buf) { g_prefill_trace = (long long*)dev_buf; }

extern "C" int64_t b200_prefill_workspace_bytes(int32_t B, int32_t S, int32_t Hq, int32_t D) {
    if (B <= 0 || S <= 0 || Hq <= 0 || D <= 0) return -1;
    return (int64_t)B * S * Hq * D * 2 + 256;    // rotated Q (bf16) + the persistent kernel's work-item counter
}

extern "C" int b200_paged_attn_prefill(const b200_prefill_args* a, void* stream) {
    if (!a) B200_FAIL(B200_E_NULL, "prefill: null args");
    if (!a->q || !a->k || !a->v || !a->k_cache || !a->v_cache || !a->cache_seqlens || !a->block_table || !a->out)
        B200_FAIL(B200_E_NULL, "prefill: required pointer is null");
    if ((a->rotary_cos == nullptr) != (a->rotary_sin == nullptr)) B200_FAIL(B200_E_NULL, "prefill: rotary_cos and rotary_sin must be given together");
    if (a->B <= 0 || a->S <= 0 || a->Hq <= 0 || a->Hkv <= 0) B200_FAIL(B200_E_RANGE, "prefill: B=%d S=%d Hq=%d Hkv=%d", a->B, a->S, a->Hq, a->Hkv);
    if (a->Hq % a->Hkv != 0) B200_FAIL(B200_E_SHAPE, "prefill: Hq=%d not divisible by Hkv=%d", a->Hq, a->Hkv);
    if (a->D != 64 && a->D != 128) B200_FAIL(B200_E_SHAPE, "prefill: head_dim %d unsupported (64, 128)", a->D);
    if (a->block_size <= 0 || a->block_size % 128 != 0) B200_FAIL(B200_E_SHAPE, "prefill: block_size %d must be a multiple of 128", a->block_size);
    if (a->n_blocks <= 0 || a->max_blocks_per_seq <= 0) B200_FAIL(B200_E_RANGE, "prefill: n_blocks / max_blocks_per_seq");
    if (a->B > 65535 || a->Hq > 65535) B200_FAIL(B200_E_RANGE, "prefill: grid too large");
    const bool rot = a->rotary_cos != nullptr;
    if (rot) {
        if (a->rotary_dim <= 0 || a->rotary_dim > a->D || a->rotary_dim % 16 != 0)
            B200_FAIL(B200_E_SHAPE, "prefill: rotary_dim %d must be a multiple of 16 and <= head_dim", a->rotary_dim);
        if (a->rot_rows <= 0) B200_FAIL(B200_E_RANGE, "prefill: rot_rows");
        if (!aligned16(a->rotary_cos) || !aligned16(a->rotary_sin)) B200_FAIL(B200_E_ALIGN, "prefill: rotary tables must be 16-byte aligned");
    }
    const void* ptrs[] = {a->q, a->k, a->v, a->k_cache, a->v_cache, a->out, a->workspace};
    for (const void* q : ptrs) if (q && !aligned16(q)) B200_FAIL(B200_E_ALIGN, "prefill: tensors must be 16-byte aligned");
    const int64_t strides[] = {a->q_stride_b, a->q_stride_s, a->q_stride_h, a->k_stride_b, a->k_stride_s, a->k_stride_h,
                               a->v_stride_b, a->v_stride_s, a->v_stride_h, a->o_stride_b, a->o_stride_s, a->o_stride_h,
                               a->kc_stride_blk, a->kc_stride_tok, a->kc_stride_h, a->vc_stride_blk, a->vc_stride_tok, a->vc_stride_h};
    for (int64_t s : strides) if (s % 8 != 0) B200_FAIL(B200_E_ALIGN, "prefill: strides must be multiples of 8 elements");
    const int64_t need = b200_prefill_workspace_bytes(a->B, a->S, a->Hq, a->D);
    if (!a->workspace || a->workspace_bytes < need)
        B200_FAIL(B200_E_RANGE, "prefill: workspace %lld B < required %lld B", (long long)a->workspace_bytes, (long long)need);
    cudaStream_t st = (cudaStream_t)stream;

    AppendParams ap;
    ap.q = (const __nv_bfloat16*)a->q; ap.k = (const __nv_bfloat16*)a->k; ap.v = (const __nv_bfloat16*)a->v;
    ap.k_cache = (__nv_bfloat16*)a->k_cache; ap.v_cache = (__nv_bfloat16*)a->v_cache; ap.q_rot = (__nv_bfloat16*)a->workspace;
    ap.cos = (const __nv_bfloat16*)a->rotary_cos; ap.sin = (const __nv_bfloat16*)a->rotary_sin;
    ap.seqlens = a->cache_seqlens; ap.block_table = a->block_table;
    ap.B = a->B; ap.S = a->S; ap.Hq = a->Hq; ap.Hkv = a->Hkv; ap.D = a->D;
    ap.rotary_dim = rot ? a->rotary_dim : 0; ap.interleaved = a->rotary_interleaved; ap.rot_rows = rot ? a->rot_rows : 1;
    ap.n_blocks = a->n_blocks; ap.block_size = a->block_size; ap.max_blocks_per_seq = a->max_blocks_per_seq;
    ap.q_sb = a->q_stride_b; ap.q_ss = a->q_stride_s; ap.q_sh = a->q_stride_h;
    ap.k_sb = a->k_stride_b; ap.k_ss = a->k_stride_s; ap.k_sh = a->k_stride_h;
    ap.v_sb = a->v_stride_b; ap.v_ss = a->v_stride_s; ap.v_sh = a->v_stride_h;
    ap.kc_sblk = a->kc_stride_blk; ap.kc_stok = a->kc_stride_tok; ap.kc_sh = a->kc_stride_h;
    ap.vc_sblk = a->vc_stride_blk; ap.vc_stok = a->vc_stride_tok; ap.vc_sh = a->vc_stride_h;
    ap.bt_sb = a->bt_stride_b;
    int* work_counter = reinterpret_cast<int*>((uint8_t*)a->workspace + (need - 256));
    ap.work_counter = work_counter;
    prefill_rope_append_kernel<<<dim3((unsigned)a->S, (unsigned)a->B), 128, 0, st>>>(ap);
    B200_CHECK_LAUNCH();

    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_map(&tq, a->workspace, a->D, a->Hq, a->S, a->B, a->D, (uint64_t)a->Hq * a->D, (uint64_t)a->S * a->Hq * a->D))) return rc;
    if ((rc = make_map(&tk, a->k_cache, a->D, a->Hkv, a->block_size, a->n_blocks, a->kc_stride_h, a->kc_stride_tok, a->kc_stride_blk))) return rc;
    if ((rc = make_map(&tv, a->v_cache, a->D, a->Hkv, a->block_size, a->n_blocks, a->vc_stride_h, a->vc_stride_tok, a->vc_stride_blk))) return rc;

    PrefillParams p;
    p.seqlens = a->cache_seqlens; p.block_table = a->block_table; p.out = (__nv_bfloat16*)a->out;
    p.B = a->B; p.S = a->S; p.Hq = a->Hq; p.Hkv = a->Hkv; p.G = a->Hq / a->Hkv;
    p.n_blocks = a->n_blocks; p.block_size = a->block_size; p.max_blocks_per_seq = a->max_blocks_per_seq; p.causal = a->causal;
    p.o_sb = a->o_stride_b; p.o_ss = a->o_stride_s; p.o_sh = a->o_stride_h; p.bt_sb = a->bt_stride_b;
    p.scale_log2 = a->softmax_scale * 1.4426950408889634f;
    p.nx = p.ny = 0;
    p.trace = g_prefill_trace;
    if (a->variant == 1) B200_FAIL(B200_E_RANGE, "prefill: variant 1 (operands staged in shared memory) was retired in round 2");
    int variant = a->variant;
    static int n_sm = 0;
    if (n_sm == 0) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev); }
    if (variant == 0) {
        // default: two Q tiles per work item (half the K/V traffic, ping-pong) on the persistent one-CTA-per-SM kernel
        // whenever those items still fill the GPU; the one-tile kernel for small grids
        const long long q_tiles = (a->S + 127) / 128;
        const long long pair_ctas = ((p.G % 2 == 0) ? q_tiles * (a->Hq / 2) : ((q_tiles + 1) / 2) * a->Hq) * a->B;
        if (pair_ctas >= n_sm) variant = 9;
    }
    if (variant >= 9 && variant <= 11) {  // persistent pair kernel; 9 / 10 / 11 = 25 % / 0 % / 50 % polynomial exp2
#define B200_PS_DISPATCH(V, SMV) if (variant == V) { \
        if (a->D == 128) return launch_prefill_persist<PCfg4<128>, SMV>(tq, tk, tv, p, work_counter, n_sm, st); \
        return launch_prefill_persist<PCfg4<64>, SMV>(tq, tk, tv, p, work_counter, n_sm, st); }
        B200_PS_DISPATCH(9, 2) B200_PS_DISPATCH(10, 1) B200_PS_DISPATCH(11, 3)
#undef B200_PS_DISPATCH
    }
    if (variant >= 6 && variant <= 8) {   // two Q tiles per CTA (ping-pong); 6 / 7 / 8 = 25 % / 0 % / 50 % polynomial exp2
#define B200_PP_DISPATCH(V, SMV) if (variant == V) { \
        if (a->D == 128) return launch_prefill_pair<PCfg3<128>, SMV>(tq, tk, tv, p, st); \
        return launch_prefill_pair<PCfg3<64>, SMV>(tq, tk, tv, p, st); }
        B200_PP_DISPATCH(6, 2) B200_PP_DISPATCH(7, 1) B200_PP_DISPATCH(8, 3)
#undef B200_PP_DISPATCH
    }
    const __nv_bfloat16* qr = (const __nv_bfloat16*)a->workspace;
    // variant 0 = default; 2..5 pick the softmax arithmetic explicitly (A/B runs): 2 = first scalar version,
    // 3 = packed fp32x2, 4 / 5 = packed + 25 % / 50 % of the exponentials on the FMA-pipe polynomial
    const int sm = variant == 0 ? B200_PREFILL_DEFAULT_SM : variant - 2;
    if (sm < 0 || sm > 3) B200_FAIL(B200_E_RANGE, "prefill: unknown variant %d", a->variant);
#define B200_PF_DISPATCH(SMV) case SMV: \
        if (a->D == 128) return launch_prefill_tmem<PCfg2<128>, SMV>(tk, tv, qr, p, st); \
        return launch_prefill_tmem<PCfg2<64>, SMV>(tk, tv, qr, p, st);
    switch (sm) { B200_PF_DISPATCH(0) B200_PF_DISPATCH(1) B200_PF_DISPATCH(2) B200_PF_DISPATCH(3) }
#undef B200_PF_DISPATCH
    return B200_E_RANGE;
}
