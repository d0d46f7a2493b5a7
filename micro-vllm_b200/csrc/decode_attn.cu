// Paged decode attention for sm_100a — replaces the flash_attn_with_kvcache
// call of the reference decode path (core/layer/model_graph.py:161-168,
// core/paged_attention.py:281-293).
//
// One launch per layer call does
//   RoPE(q, k_new at cache_seqlens[b])  ->  append k_new/v_new into the page
//   ->  split-KV flash-decoding over block_table-indexed pages
//   ->  LSE merge of the split partials by the last-arriving CTA.
//
// Work decomposition (HBM-bound, every KV byte read once per KV head):
//   grid = (num_splits, Hkv * groups, B); one CTA streams one contiguous range
//   of tokens of ONE kv head and serves all G = Hq/Hkv query heads that share it
//   (GQA packing: the G heads are the M rows of an m16n8k16 MMA tile).
//   Inside the CTA every warp owns a private cp.async ring of STAGES buffers of
//   TPS tokens (K rows + V rows, 128-bit coalesced copies, XOR-swizzled so
//   ldmatrix is bank-conflict free) and its own online-softmax state; warps
//   never synchronise with each other in the streaming loop.  Warp partials are
//   merged through shared memory, split partials through an fp32 workspace;
//   the CTA that arrives last on a per-(b, kv head) counter performs the final
//   merge, so no second kernel launch is needed and the counter is left at 0
//   (CUDA-graph replay safe).
#include "common.cuh"
#include <cstdlib>
#include <math_constants.h>

namespace b200 {

// ------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct DecodeParams {
    const __nv_bfloat16* q; const __nv_bfloat16* k_new; const __nv_bfloat16* v_new;
    __nv_bfloat16* k_cache; __nv_bfloat16* v_cache;
    const __nv_bfloat16* cos; const __nv_bfloat16* sin;
    const int32_t* seqlens; const int32_t* block_table;
    __nv_bfloat16* out; float* lse_out;
    int32_t* counters; float* ws_lse; float* ws_o;
    int B, Hq, Hkv, G, ngroups, gpg;
    int rotary_dim, interleaved, rot_rows;
    int n_blocks, block_size, max_blocks_per_seq;
    int page_shift;                      // log2(block_size) or -1 when block_size is not a power of two
    int pdl;                             // launched with programmatic stream serialization
    unsigned long long* trace;           // debug only (b200_debug_decode_trace): [CTA][8] %globaltimer stamps
    int cluster;                         // the splits of a segment form one thread-block cluster: merge through DSMEM
    int sk_ov;                           // balanced form: fixed cost of a segment, in chunks (see paged_decode_sk_kernel)
    int64_t q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, o_sb, o_sh;
    int64_t kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh, bt_sb;
    float scale, scale_log2;
};

// Timeline of every CTA of the split-KV grid (debug): 0 entry, 1 first loads issued, 2 prologue done (q rotated, new
// token staged), 3 first stage landed, 4 streaming loop done, 5 partial written, 6 exit
__device__ __forceinline__ void dec_stamp(const DecodeParams& p, int ev) {
    if (p.trace != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const size_t cta = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.trace[cta * 8 + ev] = t;
    }
}

template <int D_, int NWARPS_, int TPS_, int STAGES_, int MINB_>
struct DCfg {
    static constexpr int D = D_, NWARPS = NWARPS_, TPS = TPS_, STAGES = STAGES_, MINB = MINB_;
    static constexpr int THREADS = NWARPS * 32;
    static constexpr int CH = NWARPS * TPS;          // tokens per CTA chunk
    static constexpr int ROWB = D * 2;               // bytes per K/V row
    static constexpr int CPR = D / 8;                // 16-byte chunks per row
    static constexpr int RPI = 32 / CPR;             // rows covered by one warp-wide cp.async
    static constexpr int TILE_B = TPS * ROWB;        // bytes of the K (or V) part of a stage
    static constexpr int STAGE_B = 2 * TILE_B;
    static constexpr int KV_B = NWARPS * STAGES * STAGE_B;
    static constexpr int OPAD = D + 8;               // fp32 row pitch of the warp-merge buffer
    static constexpr int RED_B = NWARPS * 16 * OPAD * 4 + NWARPS * 16 * 2 * 4;
    // split merge through distributed shared memory: [16][D] fp32 normalised partial + [16] lse per CTA, placed BEHIND
    // the warp-merge scratch (it is written by the sink that reads that scratch)
    static constexpr int CL_OFF = (RED_B + 127) / 128 * 128;
    static constexpr int CL_B = 16 * D * 4 + 64;
    static constexpr int MAIN_B = (KV_B > CL_OFF + CL_B) ? KV_B : (CL_OFF + CL_B);
    static constexpr int Q_B = 16 * ROWB;
    static constexpr int NEW_B = 2 * ROWB;
    static constexpr int SMEM = MAIN_B + Q_B + NEW_B + 64;
    static constexpr int KS = D / 16;                // k-steps of Q.K^T
    static constexpr int NT = TPS / 8;               // score n-tiles per stage
    static constexpr int KK = TPS / 16;              // k-steps of P.V
    static constexpr int ON = D / 8;                 // output n-tiles
};

// physical byte offset of 16-byte chunk `c` of row `r` in a swizzled [rows][D] bf16 tile
template <int CPR>
__device__ __forceinline__ uint32_t swz(int r, int c) {
    return (uint32_t)(r * (CPR * 16) + ((c ^ (r & 7)) << 4));
}

// One piece of work = chunks [c0, c1) of the tokens of ONE (sequence b, kv head kvh, head group grp).
struct Piece {
    int b, kvh, grp;
    int c0, c1;            // chunk range (chunks of C::CH tokens)
    int n_chunks;          // chunks of the whole sequence (the piece with c1 == n_chunks owns the new token)
    int pos, seqlen_k;     // cache_seqlens[b], keys attended (pos + 1 with a new token)
};

// Streams one piece and merges the CTA's warps; sink(g, d, val, lse) receives the normalised partial
// output of head g of the group (val for element d, lse = natural-log sum over the piece).  Uses all
// of the CTA's shared memory; the caller must __syncthreads() before the next piece.
template <class C, class Sink>
__device__ __forceinline__ void decode_piece(const DecodeParams& p, uint8_t* smem, const Piece& pc, bool pdl, Sink&& sink) {
    constexpr int D = C::D, TPS = C::TPS, STAGES = C::STAGES, CPR = C::CPR, RPI = C::RPI;
    uint8_t* s_kv = smem;                               // [NWARPS][STAGES][K tile | V tile]
    uint8_t* s_q = smem + C::MAIN_B;                    // [16][D] bf16, swizzled
    __nv_bfloat16* s_knew = reinterpret_cast<__nv_bfloat16*>(s_q + C::Q_B);   // [D] rotated new key
    __nv_bfloat16* s_vnew = s_knew + D;                                       // [D] new value

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kvh = pc.kvh, grp = pc.grp, b = pc.b;
    const int g0 = grp * p.gpg;
    const int gcount = min(p.gpg, p.G - g0);
    const int hq0 = kvh * p.G + g0;
    const bool has_new = p.k_new != nullptr;

    const int pos = pc.pos;
    const int seqlen_k = pc.seqlen_k;
    const int n_chunks = pc.n_chunks;
    const int c0 = pc.c0;
    const int c1 = pc.c1;
    const int n_iters = max(0, c1 - c0);
    const bool owns_new = has_new && n_iters > 0 && (c1 == n_chunks);   // this piece streams the chunk holding `pos`

    const int32_t* bt = p.block_table + (int64_t)b * p.bt_sb;
    uint8_t* my_kv = s_kv + warp * (STAGES * C::STAGE_B);
    const uint32_t my_kv_u32 = smem_u32(my_kv);

    // ---- stage producer: this warp's TPS tokens of chunk c0+i -> ring slot i % STAGES
    const int ld_chunk = lane % CPR, ld_row0 = lane / CPR;
    auto issue = [&](int i) {
        if (i < n_iters) {
            const int t0 = (c0 + i) * C::CH + warp * TPS;
            if (t0 < seqlen_k) {
                const int pe = p.page_shift >= 0 ? (t0 >> p.page_shift) : (t0 / p.block_size);
                int page = (pe < p.max_blocks_per_seq) ? __ldg(bt + pe) : -1;
                const bool ok = (page >= 0) && (page < p.n_blocks);
                if (!ok) page = 0;
                const int in_page = t0 - pe * p.block_size;
                const __nv_bfloat16* kb = p.k_cache + (int64_t)page * p.kc_sblk + (int64_t)in_page * p.kc_stok + (int64_t)kvh * p.kc_sh + ld_chunk * 8;
                const __nv_bfloat16* vb = p.v_cache + (int64_t)page * p.vc_sblk + (int64_t)in_page * p.vc_stok + (int64_t)kvh * p.vc_sh + ld_chunk * 8;
                const uint32_t dst = my_kv_u32 + (i % STAGES) * C::STAGE_B;
                const int valid = ok ? (seqlen_k - t0) : 0;      // rows >= valid are zero-filled
#pragma unroll
                for (int r = 0; r < TPS / RPI; ++r) {
                    const int row = r * RPI + ld_row0;
                    const bool v = row < valid;
                    const int srow = v ? row : 0;
                    cp_async16(dst + swz<CPR>(row, ld_chunk), kb + (int64_t)srow * p.kc_stok, v ? 16u : 0u);
                }
#pragma unroll
                for (int r = 0; r < TPS / RPI; ++r) {
                    const int row = r * RPI + ld_row0;
                    const bool v = row < valid;
                    const int srow = v ? row : 0;
                    cp_async16(dst + C::TILE_B + swz<CPR>(row, ld_chunk), vb + (int64_t)srow * p.vc_stok, v ? 16u : 0u);
                }
            }
        }
        cp_async_commit();
    };

    // Start streaming before anything else: the first STAGES-1 stages only need
    // cache_seqlens and the block table.
#pragma unroll
    for (int i = 0; i < STAGES - 1; ++i) issue(i);
    dec_stamp(p, 1);

    // Programmatic dependent launch: everything above only read state that the caller
    // guarantees is older than the preceding kernel in the stream (lengths, block table, cached
    // pages); q / k_new / v_new, the workspace and every global write come after the wait.
    if (pdl && p.pdl) {
        if (p.pdl == 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }

    // ---- prologue under the in-flight copies: rotate q (and k_new), stage v_new
    const int rd = (p.cos != nullptr && has_new) ? p.rotary_dim : 0;
    const int rpos = min(pos, p.rot_rows - 1);
    const __nv_bfloat16* cosr = p.cos ? p.cos + (int64_t)rpos * (p.rotary_dim >> 1) : nullptr;
    const __nv_bfloat16* sinr = p.sin ? p.sin + (int64_t)rpos * (p.rotary_dim >> 1) : nullptr;
    if (n_iters > 0) {
        constexpr int JB = D / 16;                      // 16-element blocks per head row
        for (int t = tid; t < 16 * JB; t += C::THREADS) {
            const int row = t / JB, j = t - row * JB;
            if (row < gcount) {
                const __nv_bfloat16* src = p.q + (int64_t)b * p.q_sb + (int64_t)(hq0 + row) * p.q_sh;
                rope_block(src, j, rd, p.interleaved != 0, cosr, sinr,
                           [&](int c, const uint4& v) { *reinterpret_cast<uint4*>(s_q + swz<CPR>(row, c)) = v; });
            } else {
                const uint4 z = make_uint4(0, 0, 0, 0);
                *reinterpret_cast<uint4*>(s_q + swz<CPR>(row, 2 * j)) = z;
                *reinterpret_cast<uint4*>(s_q + swz<CPR>(row, 2 * j + 1)) = z;
            }
        }
        if (owns_new) {
            for (int t = tid; t < JB; t += C::THREADS) {
                const __nv_bfloat16* src = p.k_new + (int64_t)b * p.k_sb + (int64_t)kvh * p.k_sh;
                rope_block(src, t, rd, p.interleaved != 0, cosr, sinr,
                           [&](int c, const uint4& v) { *reinterpret_cast<uint4*>(s_knew + c * 8) = v; });
            }
            for (int t = tid; t < CPR; t += C::THREADS) {
                const __nv_bfloat16* src = p.v_new + (int64_t)b * p.v_sb + (int64_t)kvh * p.v_sh;
                *reinterpret_cast<uint4*>(s_vnew + t * 8) = ld_v4(src + t * 8);
            }
        }
    }
    __syncthreads();
    dec_stamp(p, 2);

    // ---- append: rotated k_new / raw v_new -> page block_table[b, pos / bs], slot pos % bs
    if (owns_new && grp == 0 && tid < 2 * CPR) {
        const int pe = p.page_shift >= 0 ? (pos >> p.page_shift) : (pos / p.block_size);
        const int page = (pe < p.max_blocks_per_seq) ? __ldg(bt + pe) : -1;
        if (page >= 0 && page < p.n_blocks) {
            const int slot = pos - pe * p.block_size;
            if (tid < CPR) {
                __nv_bfloat16* dst = p.k_cache + (int64_t)page * p.kc_sblk + (int64_t)slot * p.kc_stok + (int64_t)kvh * p.kc_sh;
                *reinterpret_cast<uint4*>(dst + tid * 8) = *reinterpret_cast<const uint4*>(s_knew + tid * 8);
            } else {
                const int c = tid - CPR;
                __nv_bfloat16* dst = p.v_cache + (int64_t)page * p.vc_sblk + (int64_t)slot * p.vc_stok + (int64_t)kvh * p.vc_sh;
                *reinterpret_cast<uint4*>(dst + c * 8) = *reinterpret_cast<const uint4*>(s_vnew + c * 8);
            }
        }
    }

    // ---- per-warp streaming loop
    float o[C::ON][4];
#pragma unroll
    for (int n = 0; n < C::ON; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.0f; }
    float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.0f, l1 = 0.0f;

    if (n_iters > 0) {
        uint32_t qf[C::KS][4];
        {
            const uint32_t q_u32 = smem_u32(s_q);
            const int qrow = (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
            for (int ks = 0; ks < C::KS; ++ks)
                ldsm_x4(q_u32 + swz<CPR>(qrow, ks * 2 + (lane >> 4)), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
        }
        const float sl2 = p.scale_log2;
        const int k_row = ((lane >> 4) << 3) + (lane & 7);          // token row supplied by this lane for K ldmatrix
        const int k_chk = (lane >> 3) & 1;
        const int v_row = (lane & 7) + (((lane >> 3) & 1) << 3);    // token row for V ldmatrix.trans
        const int v_chk = lane >> 4;

        for (int i = 0; i < n_iters; ++i) {
            __syncwarp();                                            // all lanes are done reading slot (i-1) % STAGES
            issue(i + STAGES - 1);
            cp_async_wait<STAGES - 1>();
            __syncwarp();
            if (i == 0) dec_stamp(p, 3);
            const int t0 = (c0 + i) * C::CH + warp * TPS;
            if (t0 >= seqlen_k) continue;                            // warp-uniform
            const uint32_t kt = my_kv_u32 + (i % STAGES) * C::STAGE_B;
            const uint32_t vt = kt + C::TILE_B;

            if (has_new && pos >= t0 && pos < t0 + TPS) {            // patch the freshly appended token into the tile
                const int row = pos - t0;
                uint8_t* kg = my_kv + (i % STAGES) * C::STAGE_B;
                if (lane < CPR) {
                    *reinterpret_cast<uint4*>(kg + swz<CPR>(row, lane)) = *reinterpret_cast<const uint4*>(s_knew + lane * 8);
                } else if (lane < 2 * CPR) {
                    const int c = lane - CPR;
                    *reinterpret_cast<uint4*>(kg + C::TILE_B + swz<CPR>(row, c)) = *reinterpret_cast<const uint4*>(s_vnew + c * 8);
                }
                __syncwarp();
            }

            // S = Q . K^T   (rows = query heads of the group, cols = tokens)
            float s[C::NT][4];
#pragma unroll
            for (int n = 0; n < C::NT; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.0f; }
#pragma unroll
            for (int ks = 0; ks < C::KS; ++ks) {
#pragma unroll
                for (int np = 0; np < C::NT / 2; ++np) {
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4(kt + swz<CPR>(np * 16 + k_row, ks * 2 + k_chk), b0, b1, b2, b3);
                    mma_bf16(s[2 * np], qf[ks], b0, b1);
                    mma_bf16(s[2 * np + 1], qf[ks], b2, b3);
                }
            }
            // mask the tail of the sequence (only the last stage of a sequence)
            if (t0 + TPS > seqlen_k) {
                const int lim = seqlen_k - t0;
#pragma unroll
                for (int n = 0; n < C::NT; ++n) {
                    const int c = n * 8 + (lane & 3) * 2;
                    if (c >= lim)     { s[n][0] = -CUDART_INF_F; s[n][2] = -CUDART_INF_F; }
                    if (c + 1 >= lim) { s[n][1] = -CUDART_INF_F; s[n][3] = -CUDART_INF_F; }
                }
            }
            // online softmax (fp32, base-2 exponent with the scale folded in)
            float mx0 = s[0][0], mx1 = s[0][2];
#pragma unroll
            for (int n = 0; n < C::NT; ++n) {
                mx0 = fmaxf(mx0, fmaxf(s[n][0], s[n][1]));
                mx1 = fmaxf(mx1, fmaxf(s[n][2], s[n][3]));
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
            const float mr0 = (mn0 == -CUDART_INF_F) ? 0.0f : mn0 * sl2;
            const float mr1 = (mn1 == -CUDART_INF_F) ? 0.0f : mn1 * sl2;
            const float a0 = fast_exp2(m0 * sl2 - mr0), a1 = fast_exp2(m1 * sl2 - mr1);
            m0 = mn0; m1 = mn1;
            float rs0 = 0.0f, rs1 = 0.0f;
#pragma unroll
            for (int n = 0; n < C::NT; ++n) {
                s[n][0] = fast_exp2(s[n][0] * sl2 - mr0);
                s[n][1] = fast_exp2(s[n][1] * sl2 - mr0);
                s[n][2] = fast_exp2(s[n][2] * sl2 - mr1);
                s[n][3] = fast_exp2(s[n][3] * sl2 - mr1);
                rs0 += s[n][0] + s[n][1];
                rs1 += s[n][2] + s[n][3];
            }
            l0 = l0 * a0 + rs0;
            l1 = l1 * a1 + rs1;
#pragma unroll
            for (int n = 0; n < C::ON; ++n) { o[n][0] *= a0; o[n][1] *= a0; o[n][2] *= a1; o[n][3] *= a1; }

            // O += P . V
#pragma unroll
            for (int kk = 0; kk < C::KK; ++kk) {
                uint32_t pa[4];
                pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
                pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
                pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
                pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
                for (int np = 0; np < C::ON / 2; ++np) {
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4_t(vt + swz<CPR>(kk * 16 + v_row, np * 2 + v_chk), b0, b1, b2, b3);
                    mma_bf16(o[2 * np], pa, b0, b1);
                    mma_bf16(o[2 * np + 1], pa, b2, b3);
                }
            }
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    }
    cp_async_wait<0>();
    dec_stamp(p, 4);
    // late trigger: the next kernel's CTAs may take over SMs from here on (their KV prefetch
    // overlaps this kernel's merge epilogue and tail)
    if (pdl && p.pdl == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();                                    // ring buffers are dead -> reuse as merge scratch

    // ---- merge the warps of this CTA
    float* r_o = reinterpret_cast<float*>(smem);                          // [NWARPS][16][OPAD]
    float* r_m = r_o + C::NWARPS * 16 * C::OPAD;                          // [NWARPS][16]
    float* r_l = r_m + C::NWARPS * 16;                                    // [NWARPS][16]
    {
        const int r0 = lane >> 2, cc = (lane & 3) * 2;
        float* wo = r_o + warp * 16 * C::OPAD;
#pragma unroll
        for (int n = 0; n < C::ON; ++n) {
            *reinterpret_cast<float2*>(wo + r0 * C::OPAD + n * 8 + cc) = make_float2(o[n][0], o[n][1]);
            *reinterpret_cast<float2*>(wo + (r0 + 8) * C::OPAD + n * 8 + cc) = make_float2(o[n][2], o[n][3]);
        }
        if ((lane & 3) == 0) {
            r_m[warp * 16 + r0] = m0; r_m[warp * 16 + r0 + 8] = m1;
            r_l[warp * 16 + r0] = l0; r_l[warp * 16 + r0 + 8] = l1;
        }
    }
    __syncthreads();

    for (int idx = tid; idx < gcount * D; idx += C::THREADS) {
        const int g = idx / D, d = idx - g * D;
        float M = -CUDART_INF_F;
#pragma unroll
        for (int w = 0; w < C::NWARPS; ++w) M = fmaxf(M, r_m[w * 16 + g]);
        const float Mr = (M == -CUDART_INF_F) ? 0.0f : M * p.scale_log2;
        float L = 0.0f, acc = 0.0f;
#pragma unroll
        for (int w = 0; w < C::NWARPS; ++w) {
            const float f = fast_exp2(r_m[w * 16 + g] * p.scale_log2 - Mr);
            L += r_l[w * 16 + g] * f;
            acc += r_o[(w * 16 + g) * C::OPAD + d] * f;
        }
        const float val = (L > 0.0f) ? acc / L : 0.0f;
        const float lse = (L > 0.0f) ? (M * p.scale + __logf(L)) : -CUDART_INF_F;
        sink(g, d, val, lse);
    }
}

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t cta_rank) {
    uint32_t remote;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(cta_rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
    return v;
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
paged_decode_kernel(const __grid_constant__ DecodeParams p) {
    constexpr int D = C::D;
    extern __shared__ __align__(128) uint8_t smem[];
    int* s_flag = reinterpret_cast<int*>(smem + C::MAIN_B + C::Q_B + C::NEW_B);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int split = blockIdx.x, S = gridDim.x;
    Piece pc;
    pc.kvh = blockIdx.y / p.ngroups; pc.grp = blockIdx.y - pc.kvh * p.ngroups;
    pc.b = blockIdx.z;
    const int b = pc.b;
    const int g0 = pc.grp * p.gpg;
    const int gcount = min(p.gpg, p.G - g0);
    const int hq0 = pc.kvh * p.G + g0;
    dec_stamp(p, 0);
    pc.pos = __ldg(p.seqlens + b);
    pc.seqlen_k = pc.pos + (p.k_new != nullptr ? 1 : 0);
    pc.n_chunks = (pc.seqlen_k + C::CH - 1) / C::CH;
    const int cps = (pc.n_chunks + S - 1) / S;
    pc.c0 = split * cps;
    pc.c1 = min(pc.n_chunks, pc.c0 + cps);
    const int64_t bh0 = (int64_t)b * p.Hq + hq0;

    float* c_o = reinterpret_cast<float*>(smem + C::CL_OFF);     // [16][D] this CTA's normalised partial (cluster form)
    float* c_lse = c_o + 16 * D;                                 // [16]
    const bool clustered = p.cluster != 0;
    decode_piece<C>(p, smem, pc, true, [&](int g, int d, float val, float lse) {
        if (S == 1) {
            p.out[(int64_t)b * p.o_sb + (int64_t)(hq0 + g) * p.o_sh + d] = __float2bfloat16_rn(val);
            if (d == 0 && p.lse_out) p.lse_out[bh0 + g] = lse;
        } else if (clustered) {
            c_o[g * D + d] = val;
            if (d == 0) c_lse[g] = lse;
        } else {
            p.ws_o[((bh0 + g) * S + split) * D + d] = val;
            if (d == 0) p.ws_lse[(bh0 + g) * S + split] = lse;
        }
    });
    dec_stamp(p, 5);
    if (S == 1) return;

    if (clustered) {
        // ---- the S CTAs of this (b, kv head, group) are ONE thread-block cluster (cluster rank = split): every CTA
        // merges an equal slice of the outputs, reading all S partials through distributed shared memory — no global
        // scratch, no atomics, no L2 round trips (the global form's last-arriver merge took 6.4 us at S = 4,
        // profiles/r02/trace_decode.log).  Same arithmetic and summation order as the global form.
        cluster_sync_all();
        const int total = gcount * D;
        const int per = (total + S - 1) / S;
        const int lo = split * per, hi = min(total, lo + per);
        const uint32_t o_addr = smem_u32(c_o), l_addr = smem_u32(c_lse);
        for (int idx = lo + tid; idx < hi; idx += C::THREADS) {
            const int g = idx / D, d = idx - g * D;
            float w[8];
            float mx = -CUDART_INF_F;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
                w[s2] = (s2 < S) ? ld_dsmem_f32(l_addr + g * 4, (uint32_t)s2) : -CUDART_INF_F;
                mx = fmaxf(mx, w[s2]);
            }
            const float mref = (mx == -CUDART_INF_F) ? 0.0f : mx;
            float sum = 0.0f;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
                w[s2] = (s2 < S) ? __expf(w[s2] - mref) : 0.0f;
                sum += w[s2];
            }
            const float inv = (sum > 0.0f) ? 1.0f / sum : 0.0f;
            float acc = 0.0f;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
                const float ws2 = w[s2] * inv;
                if (s2 < S && ws2 != 0.0f) acc += ws2 * ld_dsmem_f32(o_addr + idx * 4, (uint32_t)s2);
            }
            p.out[(int64_t)b * p.o_sb + (int64_t)(hq0 + g) * p.o_sh + d] = __float2bfloat16_rn(acc);
            if (d == 0 && p.lse_out) p.lse_out[bh0 + g] = (sum > 0.0f) ? (mref + __logf(sum)) : -CUDART_INF_F;
        }
        cluster_sync_all();                              // nobody exits while its shared memory is still being read
        dec_stamp(p, 7);
        return;
    }

    // ---- split merge by the last-arriving CTA of this (b, kv head, group)
    __syncthreads();
    if (tid == 0) {
        // acq_rel at gpu scope: releases this CTA's partials (ordered before it by the barrier,
        // fences are cumulative) and acquires the partials of every CTA that arrived earlier
        int32_t* ctr = p.counters + ((int64_t)b * gridDim.y + blockIdx.y);
        int prev;
        asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(ctr) : "memory");
        const int last = (prev == S - 1);
        if (last) *ctr = 0;                              // leave the counter reusable (graph replay)
        *s_flag = last;
    }
    __syncthreads();
    if (!*s_flag) { dec_stamp(p, 6); return; }

    float* c_w = reinterpret_cast<float*>(smem);        // [gcount][S] merge weights
    for (int idx = tid; idx < gcount * S; idx += C::THREADS) {
        const int g = idx / S, s = idx - g * S;
        c_w[idx] = __ldcg(p.ws_lse + (bh0 + g) * S + s);
    }
    __syncthreads();
    // one warp per head turns lse values into normalised weights
    for (int g = warp; g < gcount; g += C::NWARPS) {
        float mx = -CUDART_INF_F;
        for (int s = lane; s < S; s += 32) mx = fmaxf(mx, c_w[g * S + s]);
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o2));
        const float mref = (mx == -CUDART_INF_F) ? 0.0f : mx;
        float sum = 0.0f;
        for (int s = lane; s < S; s += 32) sum += __expf(c_w[g * S + s] - mref);
        sum = warp_sum(sum);
        const float inv = (sum > 0.0f) ? 1.0f / sum : 0.0f;
        for (int s = lane; s < S; s += 32) c_w[g * S + s] = __expf(c_w[g * S + s] - mref) * inv;
        if (lane == 0 && p.lse_out) p.lse_out[bh0 + g] = (sum > 0.0f) ? (mref + __logf(sum)) : -CUDART_INF_F;
    }
    __syncthreads();
    for (int idx = tid; idx < gcount * D; idx += C::THREADS) {
        const int g = idx / D, d = idx - g * D;
        const float* po = p.ws_o + (bh0 + g) * S * D + d;
        float acc = 0.0f;
        for (int s = 0; s < S; ++s) {
            const float w = c_w[g * S + s];
            if (w != 0.0f) acc += w * __ldcg(po + (int64_t)s * D);
        }
        p.out[(int64_t)b * p.o_sb + (int64_t)(hq0 + g) * p.o_sh + d] = __float2bfloat16_rn(acc);
    }
    dec_stamp(p, 7);                                    // the merging CTA of the segment
}

// ------------------------------------------------------------------ balanced persistent form
// Stream-K style scheduling for ragged / small batches: the grid is a fixed number of CTAs
// (#SMs x resident CTAs, independent of the lengths -> CUDA-graph safe) and the chunks of ALL
// (sequence, kv head, group) segments, laid end to end, are cut into equal shares, one per CTA.
// A share may cover the tail of one segment, whole segments, and the head of another; whole
// segments are written directly, the (at most two) partial pieces of a CTA go to workspace slots
// 2*cta + {0: piece starts at the CTA's range start, 1: otherwise} and the last CTA to arrive on
// the segment's counter merges them.  Every CTA derives the schedule from cache_seqlens itself.
constexpr int kSkMaxB = 512;         // prefix table of chunk counts lives in shared memory
constexpr int kSkMinChunks = 4;      // never cut shares below 4 chunks (256 tokens at CH = 64)
constexpr int kSkMaxCtas = 1024;     // workspace slots = 2 per CTA

template <class C>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
paged_decode_sk_kernel(const __grid_constant__ DecodeParams p) {
    constexpr int D = C::D;
    extern __shared__ __align__(128) uint8_t smem[];
    int* s_flag = reinterpret_cast<int*>(smem + C::MAIN_B + C::Q_B + C::NEW_B);
    int* s_pre = reinterpret_cast<int*>(smem + C::SMEM);          // [B + 1] exclusive prefix of chunk counts
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = p.B;
    const int add_new = p.k_new != nullptr ? 1 : 0;

    // Shares are cut in WEIGHT units, not chunks: a (sequence, kv head, group) segment of n chunks weighs n + OV, the
    // first OV units standing for the fixed cost of starting a piece (q load + RoPE, ring fill, warp merge, output:
    // ~3 us = 1-2 chunks of streaming at a CTA's fair share of HBM).  With plain chunk counts the CTA that drew a run of
    // 128-token sequences walked a dozen two-chunk pieces back to back and set the makespan of the launch (configs[2]:
    // 89 us against 55 us of streaming, profiles/r02/bench_n1.log).
    const int OV = p.sk_ov;
    for (int i = tid; i < B; i += C::THREADS) {
        const int len = __ldg(p.seqlens + i) + add_new;
        s_pre[i + 1] = max(1, (len + C::CH - 1) / C::CH) + OV;    // an empty sequence still owns one (masked) chunk
    }
    if (tid == 0) s_pre[0] = 0;
    __syncthreads();
    if (warp == 0) {                                               // inclusive scan, 32 entries per step
        int carry = 0;
        for (int base = 1; base <= B; base += 32) {
            const int i = base + lane;
            int v = (i <= B) ? s_pre[i] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += t;
            }
            v += carry;
            if (i <= B) s_pre[i] = v;
            carry = __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();

    const int Y = p.Hkv * p.ngroups;
    const int Wt = Y * s_pre[B];
    const int P = gridDim.x;
    const int q = max(kSkMinChunks + OV, (Wt + P - 1) / P);
    int gpos = blockIdx.x * q;
    const int gend = min(Wt, gpos + q);

    while (gpos < gend) {
        int lo = 0, hi = B - 1;                                    // largest b with Y * pre[b] <= gpos
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (Y * s_pre[mid] <= gpos) lo = mid; else hi = mid - 1;
        }
        const int b = lo;
        const int w_b = s_pre[b + 1] - s_pre[b];                   // weight of one segment of this sequence
        const int n_b = w_b - OV;                                  // its chunks
        const int rem = gpos - Y * s_pre[b];
        const int y = rem / w_b;
        const int ws = rem - y * w_b;                              // weight range [ws, we) of this segment is mine
        const int we = min(w_b, ws + (gend - gpos));
        const int cs = max(0, ws - OV), ce = max(0, we - OV);      // -> chunk range [cs, ce)
        if (ce <= cs) { gpos += we - ws; continue; }               // only the overhead zone of this segment
        const int seg0 = gpos - ws + OV;                           // global weight coordinate of the segment's chunk 0
        const int c_first = seg0 / q, c_last = (seg0 + n_b - 1) / q;
        const int npieces = c_last - c_first + 1;
        const bool whole = npieces == 1;
        const int slot = 2 * blockIdx.x + ((seg0 <= (int)blockIdx.x * q) ? 0 : 1);

        Piece pc;
        pc.b = b; pc.kvh = y / p.ngroups; pc.grp = y - pc.kvh * p.ngroups;
        pc.c0 = cs; pc.c1 = ce; pc.n_chunks = n_b;
        pc.pos = __ldg(p.seqlens + b);
        pc.seqlen_k = pc.pos + add_new;
        const int g0 = pc.grp * p.gpg;
        const int gcount = min(p.gpg, p.G - g0);
        const int hq0 = pc.kvh * p.G + g0;
        const int64_t bh0 = (int64_t)b * p.Hq + hq0;

        decode_piece<C>(p, smem, pc, false, [&](int g, int d, float val, float lse) {
            if (whole) {
                p.out[(int64_t)b * p.o_sb + (int64_t)(hq0 + g) * p.o_sh + d] = __float2bfloat16_rn(val);
                if (d == 0 && p.lse_out) p.lse_out[bh0 + g] = lse;
            } else {
                p.ws_o[((int64_t)slot * 16 + g) * D + d] = val;
                if (d == 0) p.ws_lse[slot * 16 + g] = lse;
            }
        });

        if (!whole) {
            __syncthreads();
            if (tid == 0) {
                int32_t* ctr = p.counters + ((int64_t)b * Y + y);
                int prev;
                asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(ctr) : "memory");
                const int last = (prev == npieces - 1);
                if (last) *ctr = 0;
                *s_flag = last;
            }
            __syncthreads();
            if (*s_flag) {
                // slot of piece i of this segment: CTA c_first + i; only the first piece can be a CTA's second slot
                auto slot_of = [&](int i) { return 2 * (c_first + i) + ((i == 0 && seg0 > c_first * q) ? 1 : 0); };
                float* c_w = reinterpret_cast<float*>(smem);       // [gcount][npieces] merge weights
                for (int idx = tid; idx < gcount * npieces; idx += C::THREADS) {
                    const int g = idx / npieces, i = idx - g * npieces;
                    c_w[idx] = __ldcg(p.ws_lse + slot_of(i) * 16 + g);
                }
                __syncthreads();
                for (int g = warp; g < gcount; g += C::NWARPS) {
                    float mx = -CUDART_INF_F;
                    for (int i = lane; i < npieces; i += 32) mx = fmaxf(mx, c_w[g * npieces + i]);
#pragma unroll
                    for (int o2 = 16; o2 > 0; o2 >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o2));
                    const float mref = (mx == -CUDART_INF_F) ? 0.0f : mx;
                    float sum = 0.0f;
                    for (int i = lane; i < npieces; i += 32) sum += __expf(c_w[g * npieces + i] - mref);
                    sum = warp_sum(sum);
                    const float inv = (sum > 0.0f) ? 1.0f / sum : 0.0f;
                    for (int i = lane; i < npieces; i += 32) c_w[g * npieces + i] = __expf(c_w[g * npieces + i] - mref) * inv;
                    if (lane == 0 && p.lse_out) p.lse_out[bh0 + g] = (sum > 0.0f) ? (mref + __logf(sum)) : -CUDART_INF_F;
                }
                __syncthreads();
                for (int idx = tid; idx < gcount * D; idx += C::THREADS) {
                    const int g = idx / D, d = idx - g * D;
                    float acc = 0.0f;
                    for (int i = 0; i < npieces; ++i) {
                        const float w = c_w[g * npieces + i];
                        if (w != 0.0f) acc += w * __ldcg(p.ws_o + ((int64_t)slot_of(i) * 16 + g) * D + d);
                    }
                    p.out[(int64_t)b * p.o_sb + (int64_t)(hq0 + g) * p.o_sh + d] = __float2bfloat16_rn(acc);
                }
            }
        }
        __syncthreads();                                           // shared memory is reused by the next piece
        gpos += we - ws;
    }
}

// Stand-alone merge of split partials (flash-attn's combine kernel equivalent).
__global__ void __launch_bounds__(128)
lse_combine_kernel(const float* __restrict__ o_part, const float* __restrict__ lse_part,
                   __nv_bfloat16* __restrict__ out, float* __restrict__ lse_out,
                   int Hq, int S, int D, int64_t o_sb, int64_t o_sh) {
    extern __shared__ float cw[];                       // [S]
    const int h = blockIdx.x, b = blockIdx.y;
    const int64_t bh = (int64_t)b * Hq + h;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int s = tid; s < S; s += blockDim.x) cw[s] = lse_part[bh * S + s];
    __syncthreads();
    if (tid < 32) {
        float mx = -CUDART_INF_F;
        for (int s = lane; s < S; s += 32) mx = fmaxf(mx, cw[s]);
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o2));
        const float mref = (mx == -CUDART_INF_F) ? 0.0f : mx;
        float sum = 0.0f;
        for (int s = lane; s < S; s += 32) sum += __expf(cw[s] - mref);
        sum = warp_sum(sum);
        const float inv = (sum > 0.0f) ? 1.0f / sum : 0.0f;
        for (int s = lane; s < S; s += 32) cw[s] = __expf(cw[s] - mref) * inv;
        if (lane == 0 && lse_out) lse_out[bh] = (sum > 0.0f) ? (mref + __logf(sum)) : -CUDART_INF_F;
    }
    __syncthreads();
    for (int d = tid; d < D; d += blockDim.x) {
        float acc = 0.0f;
        for (int s = 0; s < S; ++s) {
            const float w = cw[s];
            if (w != 0.0f) acc += w * o_part[(bh * S + s) * D + d];
        }
        out[(int64_t)b * o_sb + (int64_t)h * o_sh + d] = __float2bfloat16_rn(acc);
    }
}

// ------------------------------------------------------------------ host side
static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// Workspace layout: [ arrival counters (fixed capacity) | lse partials | O partials ].
// The counter region has a FIXED size and position so that it stays all-zero across launches
// of different shapes (every launch resets the counters it used); only the partials move.
constexpr int64_t kMaxCounters = 65536;
struct WsLayout { int64_t counters_off, lse_off, o_off, total; };
static WsLayout ws_layout(int B, int Hq, int D, int S) {
    WsLayout w;
    w.counters_off = 0;
    w.lse_off = kMaxCounters * 4;
    w.o_off = w.lse_off + align_up((int64_t)B * Hq * S * 4, 256);
    w.total = w.o_off + (int64_t)B * Hq * S * D * 4;
    return w;
}

// balanced form: [ arrival counters (same fixed region) | lse partials [2 * ctas][16] | O partials [2 * ctas][16][D] ]
static WsLayout ws_layout_sk(int D) {
    WsLayout w;
    w.counters_off = 0;
    w.lse_off = kMaxCounters * 4;
    w.o_off = w.lse_off + (int64_t)2 * kSkMaxCtas * 16 * 4;
    w.total = w.o_off + (int64_t)2 * kSkMaxCtas * 16 * D * 4;
    return w;
}

template <class C>
static int launch_decode_sk(const DecodeParams& p, cudaStream_t st) {
    static bool attr_done = false;
    static int attr_dev = -1, n_sm = 0;
    const int smem_bytes = C::SMEM + (int)align_up((int64_t)(p.B + 1) * 4, 16);
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done || attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(paged_decode_sk_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             C::SMEM + (kSkMaxB + 4) * 4);
        if (e != cudaSuccess) { set_error("decode: cannot opt in to %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e)); return (int)e; }
        if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sm <= 0) n_sm = 148;
        attr_done = true; attr_dev = dev;
    }
    int ctas = n_sm * C::MINB;
    if (ctas > kSkMaxCtas) ctas = kSkMaxCtas;
    // fixed cost of a piece in chunks: ~192 tokens of streaming (measured sweep, profiles/r02/bench_ragged.jsonl:
    // configs[2] 93 us at 0, 77 us at 3 chunks of 64 tokens; B200_DECODE_SK_OV overrides, 0 = plain chunk balancing)
    static const int ov_env = []() { const char* e = getenv("B200_DECODE_SK_OV"); return e ? atoi(e) : -1; }();
    DecodeParams pk = p;
    pk.sk_ov = ov_env >= 0 ? (ov_env > 64 ? 64 : ov_env) : (192 / C::CH > 1 ? 192 / C::CH : 1);
    paged_decode_sk_kernel<C><<<ctas, C::THREADS, smem_bytes, st>>>(pk);
    B200_CHECK_LAUNCH();
    return 0;
}

template <class C>
static int launch_decode(const DecodeParams& p, int S, cudaStream_t st) {
    static bool attr_done = false;      // per instantiation; the attribute is per-function, per-context
    static int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done || attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(paged_decode_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { set_error("decode: cannot opt in to %d B of shared memory: %s", C::SMEM, cudaGetErrorString(e)); return (int)e; }
        attr_done = true; attr_dev = dev;
    }
    dim3 grid((unsigned)S, (unsigned)(p.Hkv * p.ngroups), (unsigned)p.B);
    if (p.pdl || p.cluster) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid; cfg.blockDim = dim3(C::THREADS); cfg.dynamicSmemBytes = C::SMEM; cfg.stream = st;
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (p.pdl) {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        if (p.cluster) {                                 // the S splits of a segment = one cluster along x
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = (unsigned)S; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
            ++na;
        }
        cfg.attrs = attr; cfg.numAttrs = na;
        cudaError_t e = cudaLaunchKernelEx(&cfg, paged_decode_kernel<C>, p);
        if (e != cudaSuccess) { set_error("decode: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    } else {
        paged_decode_kernel<C><<<grid, C::THREADS, C::SMEM, st>>>(p);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

}  // namespace b200

using namespace b200;

static unsigned long long* g_dec_trace = nullptr;
// Debug only: uint64 [CTAs][8] device buffer that the NEXT split-KV launches stamp with %globaltimer (NULL = off).
extern "C" void b200_debug_decode_trace(void* dev_ptr) { g_dec_trace = (unsigned long long*)dev_ptr; }

extern "C" int64_t b200_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D, int32_t num_splits) {
    if (B <= 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || num_splits <= 0 || Hq % Hkv != 0) return -1;
    return ws_layout(B, Hq, D, num_splits).total;
}

extern "C" int64_t b200_decode_balanced_workspace_bytes(int32_t D) {
    if (D <= 0) return -1;
    return ws_layout_sk(D).total;
}

extern "C" int b200_paged_attn_decode(const b200_decode_args* a, void* stream) {
    if (!a) B200_FAIL(B200_E_NULL, "decode: null args");
    if (!a->q || !a->k_cache || !a->v_cache || !a->cache_seqlens || !a->block_table || !a->out)
        B200_FAIL(B200_E_NULL, "decode: required pointer is null");
    if ((a->k_new == nullptr) != (a->v_new == nullptr)) B200_FAIL(B200_E_NULL, "decode: k_new and v_new must be given together");
    if ((a->rotary_cos == nullptr) != (a->rotary_sin == nullptr)) B200_FAIL(B200_E_NULL, "decode: rotary_cos and rotary_sin must be given together");
    if (a->B <= 0 || a->Hq <= 0 || a->Hkv <= 0) B200_FAIL(B200_E_RANGE, "decode: B=%d Hq=%d Hkv=%d", a->B, a->Hq, a->Hkv);
    if (a->Hq % a->Hkv != 0) B200_FAIL(B200_E_SHAPE, "decode: Hq=%d not divisible by Hkv=%d", a->Hq, a->Hkv);
    if (a->D != 64 && a->D != 128) B200_FAIL(B200_E_SHAPE, "decode: head_dim %d unsupported (64, 128)", a->D);
    if (a->block_size <= 0 || a->block_size % 32 != 0) B200_FAIL(B200_E_SHAPE, "decode: block_size %d must be a positive multiple of 32", a->block_size);
    if (a->n_blocks <= 0 || a->max_blocks_per_seq <= 0) B200_FAIL(B200_E_RANGE, "decode: n_blocks / max_blocks_per_seq");
    if (a->num_splits < 1 || a->num_splits > 256) B200_FAIL(B200_E_RANGE, "decode: num_splits %d out of [1, 256]", a->num_splits);
    if (a->B > 65535 || a->Hkv * ((a->Hq / a->Hkv + 15) / 16) > 65535) B200_FAIL(B200_E_RANGE, "decode: grid too large");
    const bool rot = a->rotary_cos != nullptr;
    if (rot) {
        if (a->rotary_dim <= 0 || a->rotary_dim > a->D || a->rotary_dim % 16 != 0)
            B200_FAIL(B200_E_SHAPE, "decode: rotary_dim %d must be a multiple of 16 and <= head_dim", a->rotary_dim);
        if (a->rot_rows <= 0) B200_FAIL(B200_E_RANGE, "decode: rot_rows");
        if (!aligned16(a->rotary_cos) || !aligned16(a->rotary_sin)) B200_FAIL(B200_E_ALIGN, "decode: rotary tables must be 16-byte aligned");
    }
    const void* ptrs[] = {a->q, a->k_new, a->v_new, a->k_cache, a->v_cache, a->out};
    for (const void* q : ptrs) if (q && !aligned16(q)) B200_FAIL(B200_E_ALIGN, "decode: tensors must be 16-byte aligned");
    const int64_t strides[] = {a->q_stride_b, a->q_stride_h, a->k_stride_b, a->k_stride_h, a->v_stride_b, a->v_stride_h,
                               a->kc_stride_blk, a->kc_stride_tok, a->kc_stride_h, a->vc_stride_blk, a->vc_stride_tok, a->vc_stride_h};
    for (int64_t s : strides) if (s % 8 != 0) B200_FAIL(B200_E_ALIGN, "decode: strides must be multiples of 8 elements");

    const int G = a->Hq / a->Hkv;
    const int ngroups = (G + 15) / 16;
    const int S = a->num_splits;
    DecodeParams p;
    const bool balanced = (a->flags & B200_DECODE_FLAG_BALANCED) != 0;
    if (balanced) {
        if (a->B > kSkMaxB) B200_FAIL(B200_E_RANGE, "decode: balanced scheduling supports B <= %d (got %d)", kSkMaxB, a->B);
        if (a->flags & B200_DECODE_FLAG_PDL) B200_FAIL(B200_E_UNSUPPORTED, "decode: balanced scheduling has no PDL form");
        if ((int64_t)a->B * a->Hkv * ngroups > kMaxCounters)
            B200_FAIL(B200_E_RANGE, "decode: B * Hkv * groups exceeds the %lld split-merge counters", (long long)kMaxCounters);
        const WsLayout w = ws_layout_sk(a->D);
        if (!a->workspace || a->workspace_bytes < w.total)
            B200_FAIL(B200_E_RANGE, "decode: workspace %lld B < required %lld B", (long long)a->workspace_bytes, (long long)w.total);
        uint8_t* ws = (uint8_t*)a->workspace;
        p.counters = (int32_t*)(ws + w.counters_off);
        p.ws_lse = (float*)(ws + w.lse_off);
        p.ws_o = (float*)(ws + w.o_off);
    } else if (S > 1 && S <= 8 && !(a->flags & B200_DECODE_FLAG_NO_CLUSTER)) {
        p.counters = nullptr; p.ws_lse = nullptr; p.ws_o = nullptr;      // merged through distributed shared memory
    } else if (S > 1) {
        if ((int64_t)a->B * a->Hkv * ngroups > kMaxCounters)
            B200_FAIL(B200_E_RANGE, "decode: B * Hkv * groups = %lld exceeds the %lld split-merge counters",
                      (long long)a->B * a->Hkv * ngroups, (long long)kMaxCounters);
        const WsLayout w = ws_layout(a->B, a->Hq, a->D, S);
        if (!a->workspace || a->workspace_bytes < w.total)
            B200_FAIL(B200_E_RANGE, "decode: workspace %lld B < required %lld B", (long long)a->workspace_bytes, (long long)w.total);
        uint8_t* ws = (uint8_t*)a->workspace;
        p.counters = (int32_t*)(ws + w.counters_off);
        p.ws_lse = (float*)(ws + w.lse_off);
        p.ws_o = (float*)(ws + w.o_off);
    } else {
        p.counters = nullptr; p.ws_lse = nullptr; p.ws_o = nullptr;
    }
    p.q = (const __nv_bfloat16*)a->q; p.k_new = (const __nv_bfloat16*)a->k_new; p.v_new = (const __nv_bfloat16*)a->v_new;
    p.k_cache = (__nv_bfloat16*)a->k_cache; p.v_cache = (__nv_bfloat16*)a->v_cache;
    p.cos = (const __nv_bfloat16*)a->rotary_cos; p.sin = (const __nv_bfloat16*)a->rotary_sin;
    p.seqlens = a->cache_seqlens; p.block_table = a->block_table;
    p.out = (__nv_bfloat16*)a->out; p.lse_out = a->softmax_lse;
    p.B = a->B; p.Hq = a->Hq; p.Hkv = a->Hkv; p.G = G; p.ngroups = ngroups; p.gpg = (G + ngroups - 1) / ngroups;
    p.rotary_dim = rot ? a->rotary_dim : 0; p.interleaved = a->rotary_interleaved; p.rot_rows = rot ? a->rot_rows : 1;
    p.n_blocks = a->n_blocks; p.block_size = a->block_size; p.max_blocks_per_seq = a->max_blocks_per_seq;
    p.page_shift = -1;
    for (int sft = 5; sft < 31; ++sft) if ((1 << sft) == a->block_size) p.page_shift = sft;
    p.trace = balanced ? nullptr : g_dec_trace;
    p.cluster = (!balanced && S > 1 && S <= 8 && !(a->flags & B200_DECODE_FLAG_NO_CLUSTER)) ? 1 : 0;
    p.pdl = (a->flags & B200_DECODE_FLAG_PDL) ? ((a->flags & B200_DECODE_FLAG_PDL_LATE) ? 2 : 1) : 0;
    p.q_sb = a->q_stride_b; p.q_sh = a->q_stride_h; p.k_sb = a->k_stride_b; p.k_sh = a->k_stride_h;
    p.v_sb = a->v_stride_b; p.v_sh = a->v_stride_h; p.o_sb = a->o_stride_b; p.o_sh = a->o_stride_h;
    p.kc_sblk = a->kc_stride_blk; p.kc_stok = a->kc_stride_tok; p.kc_sh = a->kc_stride_h;
    p.vc_sblk = a->vc_stride_blk; p.vc_stok = a->vc_stride_tok; p.vc_sh = a->vc_stride_h;
    p.bt_sb = a->bt_stride_b;
    p.scale = a->softmax_scale;
    p.scale_log2 = a->softmax_scale * 1.4426950408889634f;
    cudaStream_t st = (cudaStream_t)stream;

    if (balanced) {
        if (a->D == 128) {
            switch (a->variant) {
                case 0:
                case 6: return launch_decode_sk<DCfg<128, 4, 16, 2, 3>>(p, st);
                case 1: return launch_decode_sk<DCfg<128, 4, 16, 3, 2>>(p, st);
                case 5: return launch_decode_sk<DCfg<128, 2, 16, 4, 3>>(p, st);
                case 8: return launch_decode_sk<DCfg<128, 2, 32, 3, 2>>(p, st);
                default: B200_FAIL(B200_E_RANGE, "decode: variant %d has no balanced form", a->variant);
            }
        }
        switch (a->variant) {
            case 0:
            case 1: return launch_decode_sk<DCfg<64, 4, 32, 3, 2>>(p, st);
            case 2: return launch_decode_sk<DCfg<64, 4, 16, 4, 3>>(p, st);
            default: B200_FAIL(B200_E_RANGE, "decode: variant %d has no balanced form", a->variant);
        }
    }
    if (a->D == 128) {
        switch (a->variant) {
            case 1: return launch_decode<DCfg<128, 4, 16, 3, 2>>(p, S, st);   //  96 KB, 2 CTA/SM
            case 2: return launch_decode<DCfg<128, 4, 16, 4, 1>>(p, S, st);   // 128 KB
            case 3: return launch_decode<DCfg<128, 4, 32, 3, 1>>(p, S, st);   // 192 KB
            case 4: return launch_decode<DCfg<128, 8, 16, 3, 1>>(p, S, st);   // 192 KB
            case 5: return launch_decode<DCfg<128, 2, 16, 4, 3>>(p, S, st);   //  64 KB, 3 CTA/SM
            case 0:
            case 6: return launch_decode<DCfg<128, 4, 16, 2, 3>>(p, S, st);   //  64 KB, 3 CTA/SM (default: best on the B200 sweep)
            case 7: return launch_decode<DCfg<128, 8, 16, 2, 1>>(p, S, st);   // 128 KB
            case 8: return launch_decode<DCfg<128, 2, 32, 3, 2>>(p, S, st);   //  96 KB, 2 CTA/SM
            case 9: return launch_decode<DCfg<128, 2, 32, 2, 3>>(p, S, st);   //  64 KB, 3 CTA/SM
            case 10: return launch_decode<DCfg<128, 4, 32, 2, 1>>(p, S, st);  // 128 KB
            case 11: return launch_decode<DCfg<128, 2, 16, 3, 4>>(p, S, st);  //  48 KB, 4 CTA/SM
            default: B200_FAIL(B200_E_RANGE, "decode: unknown variant %d", a->variant);
        }
    } else {
        switch (a->variant) {
            case 0:
            case 1: return launch_decode<DCfg<64, 4, 32, 3, 2>>(p, S, st);    //  96 KB
            case 2: return launch_decode<DCfg<64, 4, 16, 4, 3>>(p, S, st);    //  64 KB
            case 3: return launch_decode<DCfg<64, 4, 32, 4, 1>>(p, S, st);    // 128 KB
            default: B200_FAIL(B200_E_RANGE, "decode: unknown variant %d", a->variant);
        }
    }
}

extern "C" int b200_lse_combine(const float* o_part, const float* lse_part, void* out, float* lse_out,
                                int32_t B, int32_t Hq, int32_t S, int32_t D,
                                int64_t o_stride_b, int64_t o_stride_h, void* stream) {
    if (!o_part || !lse_part || !out) B200_FAIL(B200_E_NULL, "lse_combine: null pointer");
    if (B <= 0 || Hq <= 0 || S <= 0 || D <= 0 || S > 4096) B200_FAIL(B200_E_RANGE, "lse_combine: bad sizes");
    lse_combine_kernel<<<dim3((unsigned)Hq, (unsigned)B), 128, S * sizeof(float), (cudaStream_t)stream>>>(
        o_part, lse_part, (__nv_bfloat16*)out, lse_out, Hq, S, D, o_stride_b, o_stride_h);
    B200_CHECK_LAUNCH();
    return 0;
}
