// Weight-streaming GEMM for decode-sized batches on tcgen05 tensor cores — replaces the
// torch.matmul (cuBLAS) calls of a decode layer and the elementwise launches behind them:
//   core/layer/model_graph.py:145-147  QKV projection + the separate bf16 bias add        (EPI_BIAS)
//   core/layer/model_graph.py:171      o_proj                                               (EPI_NONE)
//   core/layer/model_graph.py:183-184  gate_up projection + swiglu_gemm (kernel/swiglu.py)  (EPI_SWIGLU)
//   core/layer/model_graph.py:185      down_proj                                            (EPI_NONE)
//
// out[m, n] = sum_k x[m, k] * w[n, k]:  x [M, K] bf16 with M <= 128 rows, w [N, K] bf16 — the nn.Linear layout
// ([out, in], K contiguous).  At these M the product is HBM-bound on w (K*N*2 bytes read once, 2*M flops per
// weight), so the kernel is built around the weight stream, not the math:
//
//   * swap-AB: the WEIGHTS are the 128-row UMMA M operand (out^T[n, m] = w[n, :] . x[m, :]), the batch is the
//     UMMA N (16 ... 128 columns), accumulators live in TMEM (lane = output feature, column = batch row).
//   * one persistent CTA per SM; the (n-tile, k-block) units of the whole product are dealt out in equal
//     contiguous shares (stream-K), so every SM streams the same number of weight bytes whatever N and K are.
//   * warp 0 = TMA producer: [128 n x 64 k] weight boxes and [NB m x 64 k] activation boxes (SWIZZLE_128B,
//     UMMA-canonical K-major) through an 8-stage mbarrier ring (>= 128 KB of weights in flight per SM);
//     warp 1 = MMA issuer (one elected lane, tcgen05.mma kind::f16, 4 x K=16 per stage, tcgen05.commit frees
//     the stage); warps 2-5 = epilogue (tcgen05.ld of the 128 x NB fp32 tile, double-buffered in TMEM so the
//     next segment's MMAs overlap it).
//   * a tile whose K range was shared between CTAs is finished by ALL of them: each writes its fp32 partial
//     to the workspace, arrives on the tile's counter, and — after its own MMA work is done — reduces an
//     equal slice of the batch rows over the partials in fixed CTA order (deterministic) and applies the
//     epilogue.  No single CTA serialises a many-way reduction.
//   * epilogues reproduce the reference's rounding points: bias = bf16(bf16(acc) + bias); SwiGLU =
//     bf16(silu(bf16(gate)) * bf16(up)) with the arithmetic of swiglu_vec_kernel (weights rows interleaved
//     up_0, gate_0, up_1, gate_1, ... by the host so a pair sits in adjacent TMEM lanes).
//   * programmatic dependent launch (flag): the weight boxes of the first ring fill are issued BEFORE
//     griddepcontrol.wait — weights never depend on the previous kernel — so the ramp of this kernel overlaps
//     the tail of the one before it.
#include "common.cuh"
#include "tc05.cuh"
#include <cuda.h>

namespace b200 {

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                 uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
                 " [%0], [%1, {%3, %4}], [%2], %5;"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* map, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
                 :: "l"(map), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}

enum { EPI_NONE = 0, EPI_BIAS = 1, EPI_SWIGLU = 2 };

struct SgParams {
    const __nv_bfloat16* bias;       // [N] or null
    __nv_bfloat16* out;              // [M, N] (EPI_SWIGLU: [M, N / 2])
    float* ws_part;                  // [2 * grid] partial tiles, each [NB][128] fp32
    uint32_t* ws_cnt;                // [n_tiles]: arrival / departure counters, wrap to 0 after 2 S increments (zero between launches)
    int M, N, K;
    int64_t ldo;
    int n_tiles, kblocks;            // 128-row weight tiles, 64-element k-blocks
    long long units;                 // n_tiles * kblocks
    int epilogue, flags;
    int l2_ahead;                    // PDL form: weight boxes (16 KB) per CTA prefetched into L2 before the dependency wait
    unsigned long long* trace;       // debug only (b200_debug_stream_gemm_trace): [grid][8] %globaltimer stamps
};

// Timeline of every CTA (debug): 0 entry, 1 setup done, 2 all loads issued, 3 first stage landed, 4 last MMA committed,
// 5 last accumulator drained, 6 peers' partials complete (first shared tile), 7 exit
__device__ __forceinline__ void sg_stamp(const SgParams& p, int ev) {
    if (p.trace != nullptr) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(size_t)blockIdx.x * 8 + ev] = t;
    }
}

template <int NB_>
struct SgCfg {
    static constexpr int NB = NB_;                            // UMMA N = padded batch rows
    static constexpr int A_BYTES = 128 * 128;                 // [128 n][64 k] bf16
    static constexpr int B_BYTES = NB * 128;                  // [NB m][64 k] bf16
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (NB <= 64) ? 8 : 6;
    static constexpr int OFF_BAR = STAGES * STAGE_BYTES;
    static constexpr int NBARS = 2 * STAGES + 4;              // full[S], empty[S], acc_full[2], acc_empty[2]
    static constexpr int SMEM = OFF_BAR + NBARS * 8 + 16 + 1024;   // + tmem address word + 1 KB alignment slack
    static constexpr int THREADS = 192;                       // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
    static constexpr int TMEM_COLS = (2 * NB <= 32) ? 32 : (2 * NB <= 64) ? 64 : (2 * NB <= 128) ? 128 : 256;
};

// The work of CTA c: units [c * U / G, (c + 1) * U / G), in tile-major order; a SEGMENT is the part of one tile.
struct Seg { int tile, k0, k1; };

__device__ __forceinline__ long long sg_start(long long c, long long U, long long G) { return c * U / G; }
// the CTA whose share contains unit u
__device__ __forceinline__ int sg_owner(long long u, long long U, long long G) { return (int)(((u + 1) * G - 1) / U); }

__device__ __forceinline__ float silu_mul(float g, float u) {
    // the arithmetic of swiglu_vec_kernel (elementwise.cu): g / (1 + exp(-g)) * u
    return __fdividef(g, 1.0f + __expf(-g)) * u;
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
stream_gemm_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                   const __grid_constant__ SgParams p) {
    constexpr int NB = C::NB, STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_u32 = smem_u32(smem_raw);
    const uint32_t base = (raw_u32 + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw_u32);
    const uint32_t bars = base + C::OFF_BAR;
    auto full_bar = [&](int s) { return bars + 8u * (uint32_t)s; };
    auto empty_bar = [&](int s) { return bars + 8u * (uint32_t)(STAGES + s); };
    auto accf_bar = [&](int b) { return bars + 8u * (uint32_t)(2 * STAGES + b); };
    auto acce_bar = [&](int b) { return bars + 8u * (uint32_t)(2 * STAGES + 2 + b); };
    volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + C::NBARS * 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long U = p.units, G = gridDim.x, c = blockIdx.x;
    const long long u_begin = sg_start(c, U, G), u_end = sg_start(c + 1, U, G);
    const int KB = p.kblocks;
    const bool pdl = (p.flags & B200_GEMM_FLAG_PDL) != 0;

    if (tid == 0) {
        sg_stamp(p, 0);
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(accf_bar(b), 1); mbar_init(acce_bar(b), 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tma_prefetch_desc(&tm_w); tma_prefetch_desc(&tm_x);
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
    // the kernel after this one may start its own prologue now (its griddepcontrol.wait still waits for ALL of this grid)
    if (pdl) griddep_launch();
    if (tid == 0) sg_stamp(p, 1);

    if (warp == 0) {
        // ===================================================== TMA producer
        if (elect_one()) {
            uint64_t pol_w, pol_x;
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));   // weights: read once
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_x));    // activations: re-read by every CTA
            const int n_units = (int)(u_end - u_begin);
            int tile = (int)(u_begin / KB), kb = (int)(u_begin - (long long)tile * KB);        // the only division
            int i = 0;
            if (pdl) {
                // ring fill: the weight boxes go out before the dependency on the previous kernel resolves ...
                const int pre = n_units < STAGES ? n_units : STAGES;
                int t2 = tile, k2 = kb;
                for (int j = 0; j < pre; ++j) {
                    mbar_expect_tx(full_bar(j), C::STAGE_BYTES);
                    tma_load_2d_hint(base + j * C::STAGE_BYTES, &tm_w, full_bar(j), k2 * 64, t2 * 128, pol_w);
                    if (++k2 == KB) { k2 = 0; ++t2; }
                }
                // ... and the next boxes of this CTA's share are pulled into L2 (no shared memory needed): while the
                // kernel before this one sits in a cross-GPU reduction, HBM streams the weights it is about to need
                const int ahead = n_units - pre < (int)p.l2_ahead ? n_units - pre : (int)p.l2_ahead;
                for (int j = 0; j < ahead; ++j) {
                    tma_prefetch_l2_2d(&tm_w, k2 * 64, t2 * 128);
                    if (++k2 == KB) { k2 = 0; ++t2; }
                }
                griddep_wait();                                   // ... x (and the workspace) are only touched below this line
                for (int j = 0; j < pre; ++j) {
                    tma_load_2d_hint(base + j * C::STAGE_BYTES + C::A_BYTES, &tm_x, full_bar(j), kb * 64, 0, pol_x);
                    if (++kb == KB) { kb = 0; ++tile; }
                }
                i = pre;
            }
            int stage = i % STAGES;
            uint32_t phase = (i >= STAGES) ? 1u : 0u;             // parity of the NEXT completion of empty_bar(stage) to wait for, flipped
            for (; i < n_units; ++i) {
                if (i >= STAGES) mbar_wait(empty_bar(stage), phase ^ 1u);
                const uint32_t dst = base + stage * C::STAGE_BYTES;
                mbar_expect_tx(full_bar(stage), C::STAGE_BYTES);
                tma_load_2d_hint(dst, &tm_w, full_bar(stage), kb * 64, tile * 128, pol_w);
                tma_load_2d_hint(dst + C::A_BYTES, &tm_x, full_bar(stage), kb * 64, 0, pol_x);
                if (++kb == KB) { kb = 0; ++tile; }
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
            sg_stamp(p, 2);
        }
        __syncwarp();                                             // the idle lanes must not run ahead into the final CTA barrier
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp loops, one elected lane issues)
        constexpr uint32_t IDESC = umma_idesc_bf16(128, NB, 0, 0);    // A = weights K-major, B = activations K-major
        int stage = 0;
        uint32_t phase = 0;
        int buf = 0;
        uint32_t acc_phase[2] = {0u, 0u};
        long long u = u_begin;
        while (u < u_end) {
            const int tile = (int)(u / KB);
            const long long tile_end = (long long)(tile + 1) * KB;
            const long long seg_end = tile_end < u_end ? tile_end : u_end;
            mbar_wait(acce_bar(buf), acc_phase[buf] ^ 1u);        // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem + (uint32_t)(buf * NB);
            bool first = true;
            for (; u < seg_end; ++u) {
                mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                if (u == u_begin && lane == 0) sg_stamp(p, 3);
                if (elect_one()) {
                    const uint32_t a_base = base + stage * C::STAGE_BYTES;
                    const uint32_t b_base = a_base + C::A_BYTES;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t da = umma_desc_sw128(a_base + k * 32, 16, 1024);
                        const uint64_t db = umma_desc_sw128(b_base + k * 32, 16, 1024);
                        umma_bf16(d_tmem, da, db, IDESC, (first && k == 0) ? 0u : 1u);
                    }
                    tc_commit(empty_bar(stage));                   // stage is free once these MMAs have read it
                }
                __syncwarp();
                first = false;
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
            if (elect_one()) tc_commit(accf_bar(buf));             // segment accumulated
            __syncwarp();
            if (u >= u_end && lane == 0) sg_stamp(p, 4);
            acc_phase[buf] ^= 1u;
            buf ^= 1;
        }
    } else {
        // ===================================================== epilogue warps (TMEM lane quarter = warp % 4)
        if (pdl) griddep_wait();
        const int quarter = warp & 3;
        const int et = (warp - 2) * 32 + lane;                     // 0 .. 127
        const int row_in_tile = quarter * 32 + lane;               // TMEM lane = output feature within the tile
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        constexpr int CW = NB < 64 ? NB : 64;                      // accumulator columns held in registers at a time
        int buf = 0;
        uint32_t acc_phase[2] = {0u, 0u};
        // up to two tiles of this CTA end up as partials that need the distributed reduction
        int duty_tile[2] = {-1, -1};
        long long u = u_begin;
        while (u < u_end) {
            const int tile = (int)(u / KB);
            const long long tile_beg = (long long)tile * KB, tile_end = tile_beg + KB;
            const long long seg_end = tile_end < u_end ? tile_end : u_end;
            const bool whole = (u == tile_beg) && (seg_end == tile_end);
            const bool starts = (u == tile_beg);
            mbar_wait(accf_bar(buf), acc_phase[buf]);
            tc_fence_after();
            const uint32_t t_addr = tmem + lane_sel + (uint32_t)(buf * NB);
            const int n = tile * 128 + row_in_tile;
            float bias_v = 0.0f;
            if (whole && p.epilogue == EPI_BIAS && n < p.N) bias_v = __bfloat162float(p.bias[n]);
            // partial: fp32 tile [NB][128] into this CTA's slot (slot 1 = the segment that STARTS a tile)
            float* dst = p.ws_part + (size_t)(2 * (int)c + (starts ? 1 : 0)) * (NB * 128) + row_in_tile;
#pragma unroll
            for (int c0 = 0; c0 < NB; c0 += CW) {
                uint32_t r[CW];
#pragma unroll
                for (int q = 0; q < CW; q += 16) tmem_ld16(t_addr + c0 + q, *reinterpret_cast<uint32_t(*)[16]>(&r[q]));
                tmem_wait_ld();
                if (c0 + CW >= NB) {                               // everything is in registers: hand the accumulator back
                    tc_fence_before();
                    mbar_arrive(acce_bar(buf));
                }
                if (!whole) {
#pragma unroll
                    for (int i = 0; i < CW; ++i) dst[(c0 + i) * 128] = __uint_as_float(r[i]);
                } else if (p.epilogue == EPI_SWIGLU) {
                    // lanes 2i / 2i+1 hold up_i / gate_i of output column tile * 64 + i
#pragma unroll
                    for (int i = 0; i < CW; ++i) {
                        const float mine = bf16_round(__uint_as_float(r[i]));
                        const float other = __shfl_xor_sync(0xffffffffu, mine, 1);
                        const int m = c0 + i;
                        if (!(lane & 1) && m < p.M && n < p.N)
                            p.out[(int64_t)m * p.ldo + (n >> 1)] = __float2bfloat16_rn(silu_mul(other, mine));
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < CW; ++i) {
                        const int m = c0 + i;
                        float v = __uint_as_float(r[i]);
                        if (p.epilogue == EPI_BIAS) v = bf16_round(v) + bias_v;
                        if (m < p.M && n < p.N) p.out[(int64_t)m * p.ldo + n] = __float2bfloat16_rn(v);
                    }
                }
            }
            if (!whole) {
                duty_tile[starts ? 1 : 0] = tile;
                // publish the partial: all 128 epilogue threads have stored, then ONE release-increment of the tile's
                // counter (fire and forget: nobody waits for the atomic's round trip).  The counter wraps to 0 after 2 S
                // increments (S arrivals + S departures), so it is re-armed for the next launch without a reset.
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (et == 0) {
                    const int S = sg_owner(tile_beg + KB - 1, U, G) - sg_owner(tile_beg, U, G) + 1;
                    asm volatile("red.release.gpu.global.inc.u32 [%0], %1;" :: "l"(p.ws_cnt + tile), "r"((uint32_t)(2 * S - 1)) : "memory");
                }
            }
            acc_phase[buf] ^= 1u;
            buf ^= 1;
            u = seg_end;
        }
        if (et == 0) sg_stamp(p, 5);
        bool stamped = false;
        // ---- distributed reduction of the shared tiles: every CTA of a tile's group reduces an equal slice of
        // the batch rows over ALL partials in CTA order (deterministic) and applies the epilogue
#pragma unroll 1
        for (int d = 0; d < 2; ++d) {
            const int tile = duty_tile[d];
            if (tile < 0) continue;
            const long long tile_beg = (long long)tile * KB;
            const int c_first = sg_owner(tile_beg, U, G), c_last = sg_owner(tile_beg + KB - 1, U, G);
            const int S = c_last - c_first + 1, s_me = (int)c - c_first;
            if (et == 0) {
                uint32_t spins = 0;
                while (ld_acquire_gpu(p.ws_cnt + tile) < (uint32_t)S) {
                    if (++spins > (1u << 28)) __trap();             // protocol bug: fail loudly, never hang
                }
            }
            if (et == 0 && !stamped) { sg_stamp(p, 6); stamped = true; }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const int m_lo = (int)((long long)p.M * s_me / S), m_hi = (int)((long long)p.M * (s_me + 1) / S);
            const int n4 = (et & 31) * 4;                          // 4 consecutive features of the tile
            const int n = tile * 128 + n4;
            // Two batch rows per pass and up to 8 partials per row are requested before the first add (16 independent
            // L2 round trips in flight per thread); the adds run in fixed CTA order, zeros stand in for absent partials.
            const float* part0 = p.ws_part + (size_t)(2 * c_first + 1) * (NB * 128) + n4;     // CTA c_first: slot 2c+1
            const float* partk = p.ws_part + (size_t)(2 * (c_first + 1)) * (NB * 128) + n4;   // CTA c_first+s: slot 2c
            for (int m = m_lo + (et >> 5); m < m_hi; m += 8) {
                const bool two = (m + 4) < m_hi;
                float4 acc[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
                for (int s0 = 0; s0 < S; s0 += 8) {
                    float4 v[2][8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int s2 = s0 + j;
                        const float* src = (s2 == 0) ? part0 : partk + (size_t)(s2 - 1) * (2 * NB * 128);
                        v[0][j] = (s2 < S) ? ld_cg_f4(src + m * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
                        v[1][j] = (s2 < S && two) ? ld_cg_f4(src + (m + 4) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            acc[h].x += v[h][j].x; acc[h].y += v[h][j].y; acc[h].z += v[h][j].z; acc[h].w += v[h][j].w;
                        }
                    }
                }
                if (n >= p.N) continue;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h == 1 && !two) break;
                    const int mm = m + 4 * h;
                    const float4 a4 = acc[h];
                    if (p.epilogue == EPI_SWIGLU) {
                        const float u0 = bf16_round(a4.x), g0 = bf16_round(a4.y), u1 = bf16_round(a4.z), g1 = bf16_round(a4.w);
                        *reinterpret_cast<uint32_t*>(p.out + (int64_t)mm * p.ldo + (n >> 1)) =
                            pack_bf16(silu_mul(g0, u0), silu_mul(g1, u1));
                    } else {
                        float v4[4] = {a4.x, a4.y, a4.z, a4.w};
                        if (p.epilogue == EPI_BIAS) {
                            const uint2 bb = *reinterpret_cast<const uint2*>(p.bias + n);
                            v4[0] = bf16_round(v4[0]) + bf16lo(bb.x); v4[1] = bf16_round(v4[1]) + bf16hi(bb.x);
                            v4[2] = bf16_round(v4[2]) + bf16lo(bb.y); v4[3] = bf16_round(v4[3]) + bf16hi(bb.y);
                        }
                        uint2 o;
                        o.x = pack_bf16(v4[0], v4[1]); o.y = pack_bf16(v4[2], v4[3]);
                        *reinterpret_cast<uint2*>(p.out + (int64_t)mm * p.ldo + n) = o;
                    }
                }
            }
            // departure: my reads of the partials are done (the loads above have returned: their values were consumed)
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0)
                asm volatile("red.relaxed.gpu.global.inc.u32 [%0], %1;" :: "l"(p.ws_cnt + tile), "r"((uint32_t)(2 * S - 1)) : "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (tid == 0) sg_stamp(p, 7);
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"((uint32_t)C::TMEM_COLS) : "memory");
    }
}

static unsigned long long* g_sg_trace = nullptr;

typedef CUresult (*SgEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static SgEncodeFn sg_encode_fn() {
    static SgEncodeFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (SgEncodeFn)ptr;
    }
    return fn;
}

// 2-D bf16 map over a row-major [rows, K] matrix (row stride ld elements), box [box_rows x 64 k], 128-byte swizzle;
// rows beyond `rows` read as zeros.
static int sg_make_map(CUtensorMap* map, const void* base, int64_t rows, int64_t K, int64_t ld, int box_rows) {
    SgEncodeFn enc = sg_encode_fn();
    if (!enc) { set_error("stream_gemm: cuTensorMapEncodeTiled entry point not available"); return B200_E_UNSUPPORTED; }
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("stream_gemm: cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return B200_E_UNSUPPORTED; }
    return 0;
}

static int sg_pad_rows(int M) { return M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128; }

static int sg_sm_count() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static int sg_grid(long long units, int grid_hint) {
    long long g = grid_hint > 0 ? grid_hint : sg_sm_count();
    if (g > sg_sm_count()) g = sg_sm_count();        // all CTAs must be resident: the shared-tile reduction waits on peers
    if (g > units) g = units;
    return (int)(g < 1 ? 1 : g);
}

template <int NB>
static int sg_launch(const CUtensorMap& tw, const CUtensorMap& tx, const SgParams& p, int grid, cudaStream_t st) {
    using C = SgCfg<NB>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(stream_gemm_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { set_error("stream_gemm: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(C::THREADS);
    cfg.dynamicSmemBytes = C::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (p.flags & B200_GEMM_FLAG_PDL) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    cudaError_t e = cudaLaunchKernelEx(&cfg, stream_gemm_kernel<C>, tw, tx, p);
    if (e != cudaSuccess) { set_error("stream_gemm: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}

}  // namespace b200

using namespace b200;

// Debug only: uint64 [grid][8] device buffer that the NEXT launches stamp with %globaltimer (NULL = off).
extern "C" void b200_debug_stream_gemm_trace(void* dev_ptr) { g_sg_trace = (unsigned long long*)dev_ptr; }

extern "C" int64_t b200_stream_gemm_workspace_bytes(int M, int64_t N, int grid_hint) {
    if (M < 1 || M > 128 || N < 1) return 0;
    const int64_t n_tiles = (N + 127) / 128;
    const int64_t g = grid_hint > 0 ? grid_hint : sg_sm_count();
    // counters first (2 per tile, padded to 256 bytes), then 2 partial slots per CTA
    const int64_t cnt = ((2 * n_tiles * 4 + 255) / 256) * 256;
    return cnt + 2 * g * (int64_t)sg_pad_rows(M) * 128 * 4;
}

extern "C" int b200_stream_gemm(const void* x, const void* w, const void* bias, void* out,
                                void* workspace, int64_t workspace_bytes,
                                int M, int64_t K, int64_t N, int64_t ldx, int64_t ldw, int64_t ldo,
                                int epilogue, int flags, int grid_hint, void* stream) {
    if (!x || !w || !out) B200_FAIL(B200_E_NULL, "stream_gemm: null pointer");
    if (M < 1 || M > 128) B200_FAIL(B200_E_RANGE, "stream_gemm: M=%d (1..128 rows; larger batches are compute-bound: use a dense GEMM)", M);
    if (K < 64 || K % 64 != 0) B200_FAIL(B200_E_SHAPE, "stream_gemm: K=%lld (multiple of 64)", (long long)K);
    if (N < 8 || N % 8 != 0) B200_FAIL(B200_E_SHAPE, "stream_gemm: N=%lld (multiple of 8)", (long long)N);
    if (epilogue < EPI_NONE || epilogue > EPI_SWIGLU) B200_FAIL(B200_E_RANGE, "stream_gemm: epilogue=%d", epilogue);
    if (epilogue == EPI_BIAS && !bias) B200_FAIL(B200_E_NULL, "stream_gemm: bias epilogue without bias");
    if (!aligned16(x) || !aligned16(w) || (reinterpret_cast<uintptr_t>(out) & 7) || (bias && (reinterpret_cast<uintptr_t>(bias) & 7)))
        B200_FAIL(B200_E_ALIGN, "stream_gemm: x / w must be 16-byte aligned, out / bias 8-byte aligned");
    if (ldx % 8 || ldw % 8 || ldx < K || ldw < K) B200_FAIL(B200_E_ALIGN, "stream_gemm: ldx / ldw must be multiples of 8 elements and >= K");
    const int64_t out_cols = epilogue == EPI_SWIGLU ? N / 2 : N;
    if (ldo < out_cols || ldo % 4) B200_FAIL(B200_E_ALIGN, "stream_gemm: ldo=%lld (multiple of 4, >= %lld)", (long long)ldo, (long long)out_cols);
    SgParams p{};
    p.bias = (const __nv_bfloat16*)bias;
    p.out = (__nv_bfloat16*)out;
    p.M = M; p.N = (int)N; p.K = (int)K; p.ldo = ldo;
    p.n_tiles = (int)((N + 127) / 128);
    p.kblocks = (int)(K / 64);
    p.units = (long long)p.n_tiles * p.kblocks;
    p.epilogue = epilogue; p.flags = flags & 0xff;
    p.l2_ahead = (flags >> 8) & 0xffff;
    p.trace = g_sg_trace;
    const int grid = sg_grid(p.units, grid_hint);
    const int NB = sg_pad_rows(M);
    const int64_t cnt_bytes = ((2 * (int64_t)p.n_tiles * 4 + 255) / 256) * 256;
    const int64_t need = cnt_bytes + 2 * (int64_t)grid * NB * 128 * 4;
    if (!workspace || workspace_bytes < need || !aligned16(workspace))
        B200_FAIL(B200_E_RANGE, "stream_gemm: workspace of %lld bytes needed (got %lld)", (long long)need, (long long)workspace_bytes);
    p.ws_cnt = (uint32_t*)workspace;
    p.ws_part = (float*)((uint8_t*)workspace + cnt_bytes);
    CUtensorMap tw, tx;
    int rc = sg_make_map(&tw, w, N, K, ldw, 128);
    if (rc) return rc;
    rc = sg_make_map(&tx, x, M, K, ldx, NB);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    switch (NB) {
        case 16: return sg_launch<16>(tw, tx, p, grid, st);
        case 32: return sg_launch<32>(tw, tx, p, grid, st);
        case 64: return sg_launch<64>(tw, tx, p, grid, st);
        default: return sg_launch<128>(tw, tx, p, grid, st);
    }
}
