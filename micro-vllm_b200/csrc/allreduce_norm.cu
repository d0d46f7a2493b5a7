// One-shot all-reduce over NVLink peer memory fused with the residual add and the RMSNorm
// that follow it on the decode path:
//   all_reduce(attn_out) -> rmsnorm_residual_gemm        (reference: core/layer/model_graph.py:210-216)
//   all_reduce(mlp_out).add_(residual) -> rmsnorm_(h)     (reference: core/layer/model_graph.py:219, :143, :222)
//   all_reduce itself                                     (reference: core/parallel_config.py:65-69)
//
// Every rank launches the same kernel.  One CTA owns whole rows: it (1) signals "my partial sums
// are written" into every peer's flag array and waits for every peer's signal (release/acquire at
// system scope, a monotonically increasing epoch instead of a reset, so a captured CUDA graph can be
// replayed forever), (2) loads its rows from all `world` buffers with 128-bit loads — world x 16 B
// in flight per thread, the remote ones straight over NVLink 5 / NVSwitch — and sums them in fp32
// in rank order, so every rank computes bit-identical results, (3) adds the residual, keeps the row in
// registers for the sum of squares and writes res_out and y.  The decode activations are tiny
// ([64, 8192] bf16 = 1 MiB), so the one-shot form (every rank reads everything, 1 barrier) beats a
// reduce-scatter/all-gather pair (2 barriers); the host side falls back to NCCL above a few MiB.
//
// There is no trailing barrier: the caller alternates between two input buffers, and the barrier of
// call k+1 is the proof that every peer finished reading the buffer of call k (b200_abi.h).
#include "common.cuh"

namespace b200 {

#define B200_LAUNCH_AR(kernel, grid, threads, st, pdl, ...) do { \
    cudaError_t le__ = launch_maybe_pdl(kernel, dim3(grid), dim3(threads), 0, st, pdl, __VA_ARGS__); \
    if (le__ != cudaSuccess) { set_error("allreduce launch failed: %s", cudaGetErrorString(le__)); return (int)le__; } } while (0)

struct ArParams {
    const __nv_bfloat16* x[B200_AR_MAX_WORLD];
    uint32_t* flags[B200_AR_MAX_WORLD];
    uint32_t* epoch;
    const __nv_bfloat16* r;
    const void* w;
    __nv_bfloat16* y;
    __nv_bfloat16* res_out;
    int64_t rows, n, sx, sr, sy, sres;
    float eps;
    int rank, ss_rounded;
    long long timeout_ns;
    __nv_bfloat16* yp[B200_AR_MAX_WORLD];   // row-sharded form: every rank's y buffer (entry `rank` is local)
    int rows_per_rank;                      // row-sharded form: rank r owns rows [r * rows_per_rank, ...)
    uint8_t* xstage[B200_AR_MAX_WORLD];     // LL form: every rank's staging area for pushed partial rows
    uint8_t* ystage[B200_AR_MAX_WORLD];     // LL form: every rank's staging area for pushed normalised rows
    unsigned long long* trace;              // optional [grid][4] %globaltimer stamps: start, barrier passed, row reduced, end
    const __nv_bfloat16* x_mc;              // NVLS form: multicast address of the x slot (multimem.ld_reduce)
    __nv_bfloat16* y_mc;                    // NVLS form: multicast address of the y buffer (multimem.st)
    uint32_t* flags_mc;                     // NVLS form: multicast address of the flag region (multimem.red)
    int pdl;                                // launched with programmatic stream serialization: wait for the previous kernel
    uint8_t* ystage_mc;                     // NVLS form with the packet tail: multicast address of the y staging
};

__device__ __forceinline__ void stamp(const ArParams& p, int slot) {
    if (p.trace != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(size_t)blockIdx.x * 4 + slot] = t;
    }
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// Flags that announce data written by an EARLIER kernel of the sender's stream need no fence: a kernel boundary
// already made those writes visible in the sender's L2, which is where peers read them.  System-scope
// release / acquire on every flag costs 3-6 us when 64 CTAs fence at once (profiles/r01/allreduce_tp2.json),
// so those flags are relaxed system-scope (volatile) stores and polls; only flags that publish data written
// by THIS kernel (the row-sharded form's "done" flag) are release stores.
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// peer data: never served from a stale L1 line (the same addresses carry new data every step)
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}

template <typename WT> struct ArW;
template <> struct ArW<__nv_bfloat16> {
    static __device__ __forceinline__ void load8(const void* w, int64_t e, float (&o)[8]) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>((const __nv_bfloat16*)w + e));
        const uint32_t* p = &v.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) { o[2 * i] = bf16lo(p[i]); o[2 * i + 1] = bf16hi(p[i]); }
    }
};
template <> struct ArW<float> {
    static __device__ __forceinline__ void load8(const void* w, int64_t e, float (&o)[8]) {
        const float4 a = __ldg(reinterpret_cast<const float4*>((const float*)w + e));
        const float4 b = __ldg(reinterpret_cast<const float4*>((const float*)w + e + 4));
        o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
    }
};

__device__ __forceinline__ long long now_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// bounded wait for flag >= e (signed distance: the epoch may wrap).  The bound is wall time (%globaltimer,
// ns), not SM cycles.  A peer that never arrives sets the error word (epoch[B200_AR_MAX_CTAS]) instead of
// hanging the GPU; the host side polls that word (PeerComm.poll_error) and raises.
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t e, const ArParams& p) {
    if ((int32_t)(ld_relaxed_sys(flag) - e) >= 0) return;
    const long long t0 = now_ns();
    while ((int32_t)(ld_relaxed_sys(flag) - e) < 0) {
        if (now_ns() - t0 > p.timeout_ns) {
            atomicExch(p.epoch + B200_AR_MAX_CTAS, 1u);
            break;
        }
    }
}

// TS = row-sharded ("two-shot") form: grid = rows; the CTA of row i on the rank that OWNS row i waits for
// every peer's "x row ready" flag, reduces + normalises the row and pushes y to every rank, then raises
// the row's "done" flag on every rank; the CTAs of row i on the other ranks only raise their ready flag
// and wait for the done flag.  NVLink traffic per rank drops from (world-1) * rows * n to
// (world-1) * rows/world * n in each direction, for one extra flag hop.
template <int WS, int VPT, typename WT, bool TS>
__global__ void __launch_bounds__(1024)
allreduce_residual_rmsnorm_kernel(const __grid_constant__ ArParams p) {
    __shared__ float red[32];
    const int bid = blockIdx.x;
    const int T = blockDim.x;
    const int nvec = (int)(p.n >> 3);
    // the kernel behind this one (a PDL-launched weight-streaming GEMM) may start fetching weights while this one
    // sits in the cross-GPU exchange; its own griddepcontrol.wait still waits for this whole grid
    pdl_release_next();
    if (p.pdl) pdl_wait_prev();
    stamp(p, 0);

    if (TS) {
        const uint32_t e = p.epoch[bid] + 1u;
        const int owner = bid / p.rows_per_rank;
        uint32_t* ready = p.flags[owner] + bid * B200_AR_MAX_WORLD;                       // on the owner: [row][src rank]
        if (owner != p.rank) {
            if (threadIdx.x != 0) return;
            st_relaxed_sys(ready + p.rank, e);                                            // my partial sums of this row are written (previous kernel)
            wait_flag(p.flags[p.rank] + B200_AR_MAX_CTAS * B200_AR_MAX_WORLD + bid, e, p); // owner pushed y (and is done with my x)
            p.epoch[bid] = e;
            stamp(p, 3);
            return;
        }
        if (threadIdx.x < WS && threadIdx.x != p.rank) wait_flag(ready + threadIdx.x, e, p);
        __syncthreads();
        if (threadIdx.x == 0) p.epoch[bid] = e;
    } else if (WS > 1) {
        const uint32_t e = p.epoch[bid] + 1u;
        if (threadIdx.x < WS) {
            const int peer = threadIdx.x;
            // x was written by the GEMM kernel that precedes this one in the stream (see st_relaxed_sys); the
            // peer's volatile data loads follow its poll in program order behind the CTA barrier
            st_relaxed_sys(p.flags[peer] + bid * B200_AR_MAX_WORLD + p.rank, e);
            wait_flag(p.flags[p.rank] + bid * B200_AR_MAX_WORLD + peer, e, p);
        }
        __syncthreads();
        if (threadIdx.x == 0) p.epoch[bid] = e;
    }

    stamp(p, 1);
    const bool has_r = p.r != nullptr;
    const bool norm = p.w != nullptr;
    for (int64_t row = bid; row < p.rows; row += gridDim.x) {
        float ss = 0.0f;
        float f[VPT][8];
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                // world x 16 B in flight per thread (VPT = 1 covers n <= 8192: the whole row is in flight at once)
                uint4 xv[WS];
                uint4 rv = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int k = 0; k < WS; ++k) {
                    const __nv_bfloat16* src = p.x[k] + row * p.sx + (int64_t)v * 8;
                    xv[k] = (WS > 1) ? ld_peer_v4(src) : ld_v4(src);
                }
                if (has_r) rv = ld_v4(p.r + row * p.sr + (int64_t)v * 8);
                uint4 ro;
                uint32_t* pro = &ro.x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float a = bf16lo((&xv[0].x)[k]), b = bf16hi((&xv[0].x)[k]);
#pragma unroll
                    for (int q = 1; q < WS; ++q) {
                        a += bf16lo((&xv[q].x)[k]);
                        b += bf16hi((&xv[q].x)[k]);
                    }
                    if (WS > 1) {                     // what all_reduce hands back is a bf16 tensor
                        const uint32_t s = pack_bf16(a, b);
                        a = bf16lo(s); b = bf16hi(s);
                    }
                    if (has_r) {
                        a += bf16lo((&rv.x)[k]);
                        b += bf16hi((&rv.x)[k]);
                    }
                    const uint32_t pk = pack_bf16(a, b);
                    pro[k] = pk;
                    const float ar = bf16lo(pk), br = bf16hi(pk);
                    if (p.ss_rounded) { ss += ar * ar; ss += br * br; }
                    else              { ss += a * a;   ss += b * b; }
                    f[i][2 * k] = ar; f[i][2 * k + 1] = br;
                }
                st_na_v4(p.res_out + row * p.sres + (int64_t)v * 8, ro);
            }
        }
        if (!norm) continue;
        // block reduction (all threads of the CTA take this path together: `norm` is uniform)
        {
            const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
            const int nw = (T + 31) >> 5;
            ss = warp_sum(ss);
            if (nw > 1) {
                __syncthreads();                      // red[] may still be read by the previous row
                if (lane == 0) red[wid] = ss;
                __syncthreads();
                float t = (lane < nw) ? red[lane] : 0.0f;
                ss = warp_sum(t);
            }
        }
        const float rrms = rsqrtf(ss / (float)p.n + p.eps);
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                float wf[8];
                ArW<WT>::load8(p.w, (int64_t)v * 8, wf);
                uint4 o;
                uint32_t* po = &o.x;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    po[k] = pack_bf16(f[i][2 * k] * rrms * wf[2 * k], f[i][2 * k + 1] * rrms * wf[2 * k + 1]);
                if (TS) {
#pragma unroll
                    for (int k = 0; k < WS; ++k)                                           // local + (world-1) remote stores
                        *reinterpret_cast<uint4*>(p.yp[k] + row * p.sy + (int64_t)v * 8) = o;
                } else {
                    st_na_v4(p.y + row * p.sy + (int64_t)v * 8, o);
                }
            }
        }
    }
    stamp(p, 2);
    if (TS) {
        __syncthreads();                                     // every thread's pushes are ordered before the release below
        if (threadIdx.x < WS && threadIdx.x != p.rank)
            st_release_sys(p.flags[threadIdx.x] + B200_AR_MAX_CTAS * B200_AR_MAX_WORLD + bid, p.epoch[bid]);
    }
    stamp(p, 3);
}

// ------------------------------------------------------------------ LL push form (flag-in-data, no fences)
// Measured on 2 x B200 (profiles/r01/allreduce_tp2.json): a system-scope release/acquire flag hop costs
// 3-6 us when 64 CTAs fence at once and a remote LOAD round trip over NVLink ~3.4 us, so the pull forms
// above spend ~9 us per call before moving a byte.  This form never loads remote memory and never
// fences: data travels as 16-byte packets {d0, epoch, d1, epoch} written with remote STORES (8-byte
// halves are single-copy atomic: a packet whose two flags equal the call's epoch carries valid data —
// NCCL's LL protocol), and receivers poll packets in their OWN memory.
//   CTA of row i on a rank that does not own row i:  pushes its partial row to the owner's staging area,
//       then polls its own y staging for the owner's normalised row and writes it out as plain bf16.
//   CTA of row i on the owner: polls the world-1 staged rows, sums in rank order with its own row, adds its
//       residual row, normalises, writes res_out / y locally and pushes y as packets to every rank.
// Two one-way hops per call, (world-1)/world of an activation (x 2 for the flags) per direction per rank.
// The epoch is ONE counter per rank, advanced by the last CTA of every call, so stale packets (older calls,
// other row partitions) can never match.  The residual stream stays row-sharded as in the sharded form.
// Packets go out as WEAK 16-byte stores: they are fire-and-forget (volatile / relaxed.sys stores of one thread
// were measured to retire one NVLink round trip apiece: 5.5 us for 4 packets), the hardware writes each aligned
// 8-byte half in one piece, and the flag travels inside the packet, so no ordering between packets is needed.
__device__ __forceinline__ void st_ll(void* p, uint32_t d0, uint32_t d1, uint32_t e) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(d0), "r"(e), "r"(d1), "r"(e) : "memory");
}
__device__ __forceinline__ bool ll_ready(const uint4& v, uint32_t e) { return v.y == e && v.w == e; }

// One 16-byte vector (8 bf16) from each of N sources in order; source `self` (if 0 <= self < N) is a plain local
// row, the others are LL-staged rows (2 packets).  Every load is issued first, late packets are re-polled.
template <int N>
__device__ __forceinline__ void ll_gather(const uint8_t* (&src)[N], int self, uint32_t e, const ArParams& p, uint4 (&out)[N]) {
    uint4 pk[N][2];
#pragma unroll
    for (int s = 0; s < N; ++s) {
        if (s == self) { pk[s][0] = ld_v4(src[s]); pk[s][1] = pk[s][0]; }
        else           { pk[s][0] = ld_peer_v4(src[s]); pk[s][1] = ld_peer_v4(src[s] + 16); }
    }
#pragma unroll
    for (int s = 0; s < N; ++s) {
        if (s == self) { out[s] = pk[s][0]; continue; }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (!ll_ready(pk[s][h], e)) {
                const long long t0 = now_ns();
                do {
                    pk[s][h] = ld_peer_v4(src[s] + 16 * h);
                    if (now_ns() - t0 > p.timeout_ns) { atomicExch(p.epoch + B200_AR_MAX_CTAS, 1u); break; }
                } while (!ll_ready(pk[s][h], e));
            }
        }
        out[s] = make_uint4(pk[s][0].x, pk[s][0].z, pk[s][1].x, pk[s][1].z);
    }
}

template <int WS, int VPT, typename WT>
__global__ void __launch_bounds__(512)
allreduce_ll_kernel(const __grid_constant__ ArParams p) {
    __shared__ float red[32];
    const int row = blockIdx.x;                              // grid = rows
    const int T = blockDim.x;
    const int nvec = (int)(p.n >> 3);
    // the kernel behind this one (a PDL-launched weight-streaming GEMM) may start fetching weights while this one
    // sits in the cross-GPU exchange; its own griddepcontrol.wait still waits for this whole grid
    pdl_release_next();
    if (p.pdl) pdl_wait_prev();
    stamp(p, 0);
    const uint32_t e = p.epoch[B200_AR_MAX_CTAS + 1] + 1u;   // one epoch per call, the same on every rank
    const int R = p.rows_per_rank;
    const int owner = row / R, j = row - owner * R;
    const size_t row_stage = (size_t)nvec * 32;              // bytes of one staged row (2 packets per 16-byte vector)

    if (owner != p.rank) {
        // ---- push my partial row to the owner, then receive the normalised row
        uint8_t* dst = p.xstage[owner] + ((size_t)p.rank * R + j) * row_stage;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                const uint4 x = ld_v4(p.x[p.rank] + (int64_t)row * p.sx + (int64_t)v * 8);
                st_ll(dst + (size_t)v * 32, x.x, x.y, e);
                st_ll(dst + (size_t)v * 32 + 16, x.z, x.w, e);
            }
        }
        stamp(p, 1);
        const uint8_t* ysrc = p.ystage[p.rank] + (size_t)row * row_stage;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                const uint8_t* srcs[1] = {ysrc + (size_t)v * 32};
                uint4 o[1];
                ll_gather<1>(srcs, -1, e, p, o);
                st_na_v4(p.y + (int64_t)row * p.sy + (int64_t)v * 8, o[0]);
            }
        }
        stamp(p, 2);
    } else {
        // ---- owner: gather the staged partial rows (local memory), reduce in rank order, normalise, push y
        const bool has_r = p.r != nullptr;
        float ss = 0.0f;
        float f[VPT][8];
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                const uint8_t* srcs[WS];
#pragma unroll
                for (int s = 0; s < WS; ++s)
                    srcs[s] = (s == p.rank)
                        ? reinterpret_cast<const uint8_t*>(p.x[p.rank] + (int64_t)row * p.sx + (int64_t)v * 8)
                        : p.xstage[p.rank] + ((size_t)s * R + j) * row_stage + (size_t)v * 32;
                uint4 rv = make_uint4(0u, 0u, 0u, 0u);
                if (has_r) rv = ld_v4(p.r + (int64_t)row * p.sr + (int64_t)v * 8);
                uint4 got[WS];
                ll_gather<WS>(srcs, p.rank, e, p, got);
                uint4 ro;
                uint32_t* pro = &ro.x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float a = bf16lo((&got[0].x)[k]), b = bf16hi((&got[0].x)[k]);
#pragma unroll
                    for (int q = 1; q < WS; ++q) {           // fixed rank order 0 .. world-1 (the one-shot form's order)
                        a += bf16lo((&got[q].x)[k]);
                        b += bf16hi((&got[q].x)[k]);
                    }
                    const uint32_t s2 = pack_bf16(a, b);     // what all_reduce hands back is a bf16 tensor
                    a = bf16lo(s2); b = bf16hi(s2);
                    if (has_r) { a += bf16lo((&rv.x)[k]); b += bf16hi((&rv.x)[k]); }
                    const uint32_t pk = pack_bf16(a, b);
                    pro[k] = pk;
                    const float ar = bf16lo(pk), br = bf16hi(pk);
                    if (p.ss_rounded) { ss += ar * ar; ss += br * br; }
                    else              { ss += a * a;   ss += b * b; }
                    f[i][2 * k] = ar; f[i][2 * k + 1] = br;
                }
                st_na_v4(p.res_out + (int64_t)row * p.sres + (int64_t)v * 8, ro);
            }
        }
        stamp(p, 1);
        {
            const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
            const int nw = (T + 31) >> 5;
            ss = warp_sum(ss);
            if (nw > 1) {
                if (lane == 0) red[wid] = ss;
                __syncthreads();
                float t = (lane < nw) ? red[lane] : 0.0f;
                ss = warp_sum(t);
            }
        }
        const float rrms = rsqrtf(ss / (float)p.n + p.eps);
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                float wf[8];
                ArW<WT>::load8(p.w, (int64_t)v * 8, wf);
                uint4 o;
                uint32_t* po = &o.x;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    po[k] = pack_bf16(f[i][2 * k] * rrms * wf[2 * k], f[i][2 * k + 1] * rrms * wf[2 * k + 1]);
#pragma unroll
                for (int k = 0; k < WS; ++k) {
                    if (k == p.rank) continue;
                    uint8_t* dst = p.ystage[k] + (size_t)row * row_stage + (size_t)v * 32;
                    st_ll(dst, o.x, o.y, e);
                    st_ll(dst + 16, o.z, o.w, e);
                }
                st_na_v4(p.y + (int64_t)row * p.sy + (int64_t)v * 8, o);
            }
        }
        stamp(p, 2);
    }
    // the last CTA of the call advances the epoch (every CTA has read it before it can arrive here)
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicAdd(p.epoch + B200_AR_MAX_CTAS + 2, 1u);
        if (done == gridDim.x - 1) {
            p.epoch[B200_AR_MAX_CTAS + 2] = 0u;
            __threadfence();
            p.epoch[B200_AR_MAX_CTAS + 1] = e;
        }
    }
    stamp(p, 3);
}


// ------------------------------------------------------------------ NVLS form (in-switch reduction)
// All ranks map ONE multicast object over their symmetric slabs (cuMulticast*, set up by the host side).  A
// multimem.ld_reduce on a multicast address makes the NVSwitch read the same offset from every GPU and hand back the
// SUM (fp32 accumulation in the switch, one rounding to bf16); a multimem.st / multimem.red writes / adds the operand
// on every GPU at once.  One call is therefore
//     1 x multimem.red  (x of this rank is written)         -> every rank's `ready` counter
//     rank r, for its rows [r*R, (r+1)*R):  multimem.ld_reduce the row  (one switch round trip instead of world-1
//         peer reads), add ITS residual row, normalise, write res_out locally, multimem.st y to EVERY rank
//     1 x multimem.red.release per CTA (my rows are stored) -> every rank's `done` counter; wait for all CTAs of all ranks
// Counters are monotonic; the targets live in local memory and are advanced by the last CTA of every call, so
// a captured CUDA graph replays forever and calls of different row counts can be mixed (every rank makes the
// same sequence of calls).  Row ownership and the row-sharded residual contract are those of ROW_SHARDED.  All
// ranks get y from ONE computation, so they are bit-identical.  After a call returns on any rank every owner
// has finished reading x, so the x slot may be rewritten at once (no slot alternation needed).
#define B200_AR_NVLS_READY_WORD (B200_AR_MAX_CTAS * (B200_AR_MAX_WORLD + 1))         // u32 index in the flag region
#define B200_AR_NVLS_DONE_WORD  (B200_AR_NVLS_READY_WORD + 32)                         // 128 bytes apart
#define B200_AR_EPOCH_NVLS_READY (B200_AR_MAX_CTAS + 4)                                // local targets / arrival counter
#define B200_AR_EPOCH_NVLS_DONE  (B200_AR_MAX_CTAS + 5)
#define B200_AR_EPOCH_NVLS_ARRIVE (B200_AR_MAX_CTAS + 6)
#define B200_AR_EPOCH_NVLS_CALL (B200_AR_MAX_CTAS + 10)                                // packet-tail form: call counter = packet flag

__device__ __forceinline__ void multimem_red_add_relaxed(uint32_t* mc, uint32_t v) {
    asm volatile("multimem.red.relaxed.sys.global.add.u32 [%0], %1;" :: "l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_release(uint32_t* mc, uint32_t v) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" :: "l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
// WEAK multicast store: fire-and-forget like the LL form's packets (a relaxed.sys store makes the issuing thread wait
// for the NVLink round trip; the data is published either by the release that follows the CTA barrier or by the flag
// words inside the packet itself)
__device__ __forceinline__ void multimem_st_v4(void* mc, const uint4& v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// ONE = the one-hop form: EVERY rank asks the switch for the sum of EVERY row (the switch reduces the same slot once
// per requester — 1 MB per GPU at decode sizes, far below NVLink bandwidth), adds its REPLICATED residual, normalises
// and stores y locally.  No second hop and no trailing barrier: like the one-shot peer-load form, consecutive calls
// must alternate x slots, and res_out / residual are full tensors on every rank.
// PKT = the two-hop form with a PACKET tail: the owner stores y as 16-byte packets {d0, e, d1, e} (e = call counter)
// into every rank's y staging with plain peer stores, and every rank's CTAs poll their LOCAL staging and unpack it into the plain y.
// That replaces "multimem.red.release per CTA (a system-scope fence behind 16 KB of multicast stores) + wait for all
// 8 x grid arrivals" — 6.7 us of the 16.4 us kernel at 8 ranks (profiles/r02/allreduce_tp8.json) — by one one-way hop.
template <int VPT, typename WT, bool ONE, bool PKT>
__global__ void __launch_bounds__(1024)
allreduce_nvls_kernel(const __grid_constant__ ArParams p, const int world) {
    __shared__ float red[32];
    const int T = blockDim.x;
    const int nvec = (int)(p.n >> 3);
    // the kernel behind this one (a PDL-launched weight-streaming GEMM) may start fetching weights while this one
    // sits in the cross-GPU exchange; its own griddepcontrol.wait still waits for this whole grid
    pdl_release_next();
    if (p.pdl) pdl_wait_prev();
    stamp(p, 0);
    const uint32_t ready_t = p.epoch[B200_AR_EPOCH_NVLS_READY] + (uint32_t)world;
    const uint32_t done_t = p.epoch[B200_AR_EPOCH_NVLS_DONE] + (uint32_t)world * gridDim.x;
    // packet flag of this call; the top bit keeps it apart from the LL form's call numbers (same staging region)
    const uint32_t pkt_n = p.epoch[B200_AR_EPOCH_NVLS_CALL] + 1u;
    const uint32_t pkt_e = 0x80000000u | pkt_n;
    const size_t row_stage = (size_t)nvec * 32;             // bytes of one staged row (2 packets per 16-byte vector)
    uint32_t* my_flags = p.flags[p.rank];
    if (threadIdx.x == 0) {
        // x was written by the kernel that precedes this one in the stream: no fence needed (see st_relaxed_sys)
        if (blockIdx.x == 0) multimem_red_add_relaxed(p.flags_mc + B200_AR_NVLS_READY_WORD, 1u);
        wait_flag(my_flags + B200_AR_NVLS_READY_WORD, ready_t, p);
    }
    __syncthreads();
    stamp(p, 1);
    // a CTA walks rows j, j + grid, ... of this rank's share: the grid is capped so that every CTA of the call is
    // resident at once (the done barrier below waits for ALL of them)
    for (int64_t j = blockIdx.x; j < (ONE ? p.rows : (int64_t)p.rows_per_rank); j += gridDim.x) {
        const int64_t row = ONE ? j : (int64_t)p.rank * p.rows_per_rank + j;
        if (row >= p.rows) break;
        const bool has_r = p.r != nullptr;
        float ss = 0.0f;
        float f[VPT][8];
        uint4 sv[VPT], rv[VPT];
#pragma unroll
        for (int i = 0; i < VPT; ++i) {                       // every switch round trip of the row in flight at once
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                sv[i] = multimem_ld_reduce_bf16x8(p.x_mc + row * p.sx + (int64_t)v * 8);
                rv[i] = has_r ? ld_v4(p.r + row * p.sr + (int64_t)v * 8) : make_uint4(0u, 0u, 0u, 0u);
            }
        }
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                uint4 ro;
                uint32_t* pro = &ro.x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    // the switch hands back bf16(sum over ranks): what all_reduce returns
                    float a = bf16lo((&sv[i].x)[k]), b = bf16hi((&sv[i].x)[k]);
                    if (has_r) { a += bf16lo((&rv[i].x)[k]); b += bf16hi((&rv[i].x)[k]); }
                    const uint32_t pk = pack_bf16(a, b);
                    pro[k] = pk;
                    const float ar = bf16lo(pk), br = bf16hi(pk);
                    if (p.ss_rounded) { ss += ar * ar; ss += br * br; }
                    else              { ss += a * a;   ss += b * b; }
                    f[i][2 * k] = ar; f[i][2 * k + 1] = br;
                }
                st_na_v4(p.res_out + row * p.sres + (int64_t)v * 8, ro);
            }
        }
        {
            const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
            const int nw = (T + 31) >> 5;
            ss = warp_sum(ss);
            if (nw > 1) {
                __syncthreads();                              // red[] may still be read for the previous row
                if (lane == 0) red[wid] = ss;
                __syncthreads();
                float t = (lane < nw) ? red[lane] : 0.0f;
                ss = warp_sum(t);
            }
        }
        const float rrms = rsqrtf(ss / (float)p.n + p.eps);
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int v = threadIdx.x + i * T;
            if (v < nvec) {
                float wf[8];
                ArW<WT>::load8(p.w, (int64_t)v * 8, wf);
                uint4 o;
                uint32_t* po = &o.x;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    po[k] = pack_bf16(f[i][2 * k] * rrms * wf[2 * k], f[i][2 * k + 1] * rrms * wf[2 * k + 1]);
                if constexpr (ONE) {
                    st_na_v4(p.y + row * p.sy + (int64_t)v * 8, o);
                } else if constexpr (PKT) {
                    // unicast packets to every rank (own copy included): measured on B200s, multicast stores need ~6 us
                    // to land on every GPU, plain peer stores of the LL form ~1 us (profiles/r02/allreduce_tp2.json)
                    const size_t off = (size_t)row * row_stage + (size_t)v * 32;
#pragma unroll
                    for (int k = 0; k < B200_AR_MAX_WORLD; ++k) {
                        if (k < world) {
                            uint8_t* dst = p.ystage[(k + p.rank) % world] + off;
                            st_ll(dst, o.x, o.y, pkt_e);
                            st_ll(dst + 16, o.z, o.w, pkt_e);
                        }
                    }
                } else {
                    multimem_st_v4(p.y_mc + row * p.sy + (int64_t)v * 8, o);  // lands in every rank's y buffer
                }
            }
        }
    }
    stamp(p, 2);
    if constexpr (PKT) {
        // every rank unpacks ALL rows from its local staging (this rank's CTAs share them); a packet is complete when
        // both flag words carry this call's number
        const uint8_t* mine = p.ystage[p.rank];
        for (int64_t row = blockIdx.x; row < p.rows; row += gridDim.x) {
#pragma unroll
            for (int i = 0; i < VPT; ++i) {
                const int v = threadIdx.x + i * T;
                if (v < nvec) {
                    const uint8_t* src = mine + (size_t)row * row_stage + (size_t)v * 32;
                    uint4 a = ld_peer_v4(src), b = ld_peer_v4(src + 16);
                    if (!ll_ready(a, pkt_e) || !ll_ready(b, pkt_e)) {
                        const long long t0 = now_ns();
                        do {
                            a = ld_peer_v4(src); b = ld_peer_v4(src + 16);
                            if (now_ns() - t0 > p.timeout_ns) { atomicExch(p.epoch + B200_AR_MAX_CTAS, 1u); break; }
                        } while (!ll_ready(a, pkt_e) || !ll_ready(b, pkt_e));
                    }
                    st_na_v4(p.y + row * p.sy + (int64_t)v * 8, make_uint4(a.x, a.z, b.x, b.z));
                }
            }
        }
    }
    __syncthreads();                                         // every thread's multimem.st is ordered before the release below
    if (threadIdx.x == 0) {
        if constexpr (!ONE && !PKT) {
            multimem_red_add_release(p.flags_mc + B200_AR_NVLS_DONE_WORD, 1u);
            // all CTAs of all ranks have stored their rows (and finished reading x)
            if ((int32_t)(ld_acquire_sys(my_flags + B200_AR_NVLS_DONE_WORD) - done_t) < 0) {
                const long long t0 = now_ns();
                while ((int32_t)(ld_acquire_sys(my_flags + B200_AR_NVLS_DONE_WORD) - done_t) < 0) {
                    if (now_ns() - t0 > p.timeout_ns) { atomicExch(p.epoch + B200_AR_MAX_CTAS, 1u); break; }
                }
            }
        }
        // the last CTA of the call advances the targets (every CTA read them before it could arrive here)
        const unsigned arrived = atomicAdd(p.epoch + B200_AR_EPOCH_NVLS_ARRIVE, 1u);
        if (arrived == gridDim.x - 1) {
            p.epoch[B200_AR_EPOCH_NVLS_ARRIVE] = 0u;
            p.epoch[B200_AR_EPOCH_NVLS_READY] = ready_t;
            if constexpr (!ONE && !PKT) p.epoch[B200_AR_EPOCH_NVLS_DONE] = done_t;
            if constexpr (PKT) p.epoch[B200_AR_EPOCH_NVLS_CALL] = pkt_n & 0x7fffffffu;
        }
    }
    stamp(p, 3);
}

template <typename WT, bool ONE, bool PKT>
static int launch_nvls(const ArParams& p, int world, cudaStream_t st) {
    const int64_t nvec = p.n / 8;
    const int64_t my_rows = ONE ? p.rows : (int64_t)p.rows_per_rank;
    const unsigned grid = (unsigned)(my_rows < 128 ? my_rows : 128);   // all resident at once (148 SMs)
    if (nvec <= 1024) {
        int threads = (int)((nvec + 31) & ~31LL);
        B200_LAUNCH_AR((allreduce_nvls_kernel<1, WT, ONE, PKT>), grid, threads, st, p.pdl != 0, p, world);
    } else {
        int threads = (int)((((nvec + 1) / 2) + 31) & ~31LL);
        B200_LAUNCH_AR((allreduce_nvls_kernel<2, WT, ONE, PKT>), grid, threads, st, p.pdl != 0, p, world);
    }
    B200_CHECK_LAUNCH();
    return 0;
}


// ------------------------------------------------------------------ column all-gather (vocab-parallel lm_head)
// The reference keeps lm_head replicated under tensor parallelism (models/adapter.py:25): every rank streams the whole
// [vocab, hidden] matrix — 2.1 GB at the 70B shape, 0.33 ms of a 2.6 ms TP=8 decode step.  Here every rank multiplies
// by ITS vocab / world rows and this kernel assembles the [rows, vocab] logits on every rank: each rank writes its
// [rows, cols_local] block at column offset rank * cols_local of every rank's gather buffer (ONE multimem.st per 16
// bytes when the multicast mapping exists, else world peer stores), then arrives on every rank's counter with release
// semantics and waits until all CTAs of all ranks have arrived: when the kernel ends on a rank, its buffer is complete.
// Counters are monotonic (CUDA-graph replay safe).  WAR: a rank may overwrite a peer's buffer only after that peer
// consumed the previous contents; the decode step guarantees it (the 2 * layers reductions between two gathers are
// barriers that every rank enters after its consumer kernels, in stream order).
#define B200_AG_DONE_WORD (B200_AR_NVLS_READY_WORD + 48)
#define B200_AR_EPOCH_AG_DONE (B200_AR_MAX_CTAS + 8)
#define B200_AR_EPOCH_AG_ARRIVE (B200_AR_MAX_CTAS + 9)

struct AgParams {
    const __nv_bfloat16* src;
    __nv_bfloat16* dst[B200_AR_MAX_WORLD];
    __nv_bfloat16* dst_mc;
    uint32_t* flags[B200_AR_MAX_WORLD];
    uint32_t* flags_mc;
    uint32_t* epoch;
    int64_t rows, cols_local, ss, sd;
    int rank, world, pdl;
    long long timeout_ns;
};

__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(512)
allgather_cols_kernel(const __grid_constant__ AgParams p) {
    pdl_release_next();
    if (p.pdl) pdl_wait_prev();
    const uint32_t done_t = p.epoch[B200_AR_EPOCH_AG_DONE] + (uint32_t)p.world * gridDim.x;
    const int64_t vec_per_row = p.cols_local >> 3;
    const int64_t total = p.rows * vec_per_row;
    const int64_t col0 = (int64_t)p.rank * p.cols_local;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / vec_per_row, v = i - row * vec_per_row;
        const uint4 x = ld_v4(p.src + row * p.ss + v * 8);
        const int64_t off = row * p.sd + col0 + v * 8;
        if (p.dst_mc != nullptr) {
            multimem_st_v4(p.dst_mc + off, x);
        } else {
            for (int r = 0; r < p.world; ++r) st_na_v4(p.dst[(r + p.rank) % p.world] + off, x);
        }
    }
    __syncthreads();                                         // every thread's stores are ordered before the release below
    if (threadIdx.x == 0) {
        if (p.flags_mc != nullptr) {
            multimem_red_add_release(p.flags_mc + B200_AG_DONE_WORD, 1u);
        } else {
            for (int r = 0; r < p.world; ++r) red_release_sys_add(p.flags[r] + B200_AG_DONE_WORD, 1u);
        }
        const uint32_t* mine = p.flags[p.rank] + B200_AG_DONE_WORD;
        if ((int32_t)(ld_acquire_sys(mine) - done_t) < 0) {
            const long long t0 = now_ns();
            while ((int32_t)(ld_acquire_sys(mine) - done_t) < 0) {
                if (now_ns() - t0 > p.timeout_ns) { atomicExch(p.epoch + B200_AR_MAX_CTAS, 1u); break; }
            }
        }
        const unsigned arrived = atomicAdd(p.epoch + B200_AR_EPOCH_AG_ARRIVE, 1u);
        if (arrived == gridDim.x - 1) {                      // every CTA read the target before it could arrive here
            p.epoch[B200_AR_EPOCH_AG_ARRIVE] = 0u;
            p.epoch[B200_AR_EPOCH_AG_DONE] = done_t;
        }
    }
}

template <int WS, typename WT>
static int launch_ll(const ArParams& p, cudaStream_t st) {
    const int64_t nvec = p.n / 8;
    const unsigned grid = (unsigned)p.rows;
    if (nvec <= 512) {
        const int threads = (int)((nvec + 31) & ~31LL);
        B200_LAUNCH_AR((allreduce_ll_kernel<WS, 1, WT>), grid, threads, st, p.pdl != 0, p);
    } else if (nvec <= 1024) {
        const int threads = (int)(((nvec + 1) / 2 + 31) & ~31LL);
        B200_LAUNCH_AR((allreduce_ll_kernel<WS, 2, WT>), grid, threads, st, p.pdl != 0, p);
    } else {
        const int threads = (int)(((nvec + 3) / 4 + 31) & ~31LL);
        B200_LAUNCH_AR((allreduce_ll_kernel<WS, 4, WT>), grid, threads, st, p.pdl != 0, p);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

template <typename WT>
static int dispatch_ll(int world, const ArParams& p, cudaStream_t st) {
    switch (world) {
        case 2: return launch_ll<2, WT>(p, st);
        case 4: return launch_ll<4, WT>(p, st);
        case 8: return launch_ll<8, WT>(p, st);
    }
    B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: LL form needs world 2, 4 or 8 (got %d)", world);
}

template <int WS, typename WT, bool TS>
static int launch_ar(const ArParams& p, cudaStream_t st) {
    const int64_t nvec = p.n / 8;
    const unsigned grid = (unsigned)(p.rows < B200_AR_MAX_CTAS ? p.rows : B200_AR_MAX_CTAS);
    if (nvec <= 1024) {
        int threads = (int)((nvec + 31) & ~31LL);
        if (threads < 32 * ((WS + 31) / 32)) threads = 32;
        B200_LAUNCH_AR((allreduce_residual_rmsnorm_kernel<WS, 1, WT, TS>), grid, threads, st, p.pdl != 0, p);
    } else {
        int threads = (int)((nvec + 1) / 2);
        threads = (threads + 31) & ~31;
        B200_LAUNCH_AR((allreduce_residual_rmsnorm_kernel<WS, 2, WT, TS>), grid, threads, st, p.pdl != 0, p);
    }
    B200_CHECK_LAUNCH();
    return 0;
}

template <typename WT>
static int dispatch_ws(int world, bool sharded, const ArParams& p, cudaStream_t st) {
    if (sharded) {
        switch (world) {
            case 2: return launch_ar<2, WT, true>(p, st);
            case 4: return launch_ar<4, WT, true>(p, st);
            case 8: return launch_ar<8, WT, true>(p, st);
        }
        B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: row-sharded form needs world 2, 4 or 8 (got %d)", world);
    }
    switch (world) {
        case 1: return launch_ar<1, WT, false>(p, st);
        case 2: return launch_ar<2, WT, false>(p, st);
        case 4: return launch_ar<4, WT, false>(p, st);
        case 8: return launch_ar<8, WT, false>(p, st);
    }
    B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: world=%d (1, 2, 4 or 8)", world);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_allreduce_residual_rmsnorm(const b200_ar_args* a, void* stream) {
    if (!a) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: null args");
    if (a->world < 1 || a->world > B200_AR_MAX_WORLD || a->rank < 0 || a->rank >= a->world)
        B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: rank=%d world=%d", a->rank, a->world);
    if (!a->res_out) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: res_out is NULL");
    if ((a->weight == nullptr) != (a->y == nullptr))
        B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: weight and y must both be given or both be NULL");
    if (a->rows < 0 || a->n <= 0) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: rows=%lld n=%lld",
                                            (long long)a->rows, (long long)a->n);
    if (a->n % 8 != 0 || a->n > 16384)
        B200_FAIL(B200_E_SHAPE, "allreduce_residual_rmsnorm: n=%lld (multiple of 8, <= 16384)", (long long)a->n);
    ArParams p{};
    for (int r = 0; r < a->world; ++r) {
        if (!a->x_peers[r]) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: x_peers[%d] is NULL", r);
        if (!aligned16(a->x_peers[r])) B200_FAIL(B200_E_ALIGN, "allreduce_residual_rmsnorm: x_peers[%d] not 16-byte aligned", r);
        p.x[r] = (const __nv_bfloat16*)a->x_peers[r];
        if (a->world > 1) {
            if (!a->flag_peers[r]) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: flag_peers[%d] is NULL", r);
            p.flags[r] = (uint32_t*)a->flag_peers[r];
        }
    }
    if (a->world > 1 && !a->epoch) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: epoch is NULL");
    if (!aligned16(a->res_out) || (a->residual && !aligned16(a->residual)) || (a->y && !aligned16(a->y)) ||
        (a->weight && !aligned16(a->weight)))
        B200_FAIL(B200_E_ALIGN, "allreduce_residual_rmsnorm: pointers must be 16-byte aligned");
    if (a->stride_x % 8 || a->stride_res % 8 || (a->residual && a->stride_r % 8) || (a->y && a->stride_y % 8))
        B200_FAIL(B200_E_ALIGN, "allreduce_residual_rmsnorm: row strides must be multiples of 8 elements");
    const bool sharded = (a->mode & B200_AR_MODE_ROW_SHARDED) != 0;
    if (sharded) {
        if (a->world < 2) B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: row-sharded form needs world > 1");
        if (a->rows > B200_AR_MAX_CTAS)
            B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: row-sharded form handles at most %d rows (got %lld)",
                      B200_AR_MAX_CTAS, (long long)a->rows);
        if (!a->weight) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: row-sharded form needs the norm (weight, y_peers)");
        for (int r = 0; r < a->world; ++r) {
            if (!a->y_peers[r] || !aligned16(a->y_peers[r]))
                B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: y_peers[%d] missing or unaligned", r);
            p.yp[r] = (__nv_bfloat16*)a->y_peers[r];
        }
        if (a->y_peers[a->rank] != a->y) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: y_peers[rank] must be y");
        p.rows_per_rank = (int)((a->rows + a->world - 1) / a->world);
        if (p.rows_per_rank < 1) p.rows_per_rank = 1;
    }
    const bool ll = (a->mode & B200_AR_MODE_LL) != 0;
    if (ll) {
        if (sharded) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: choose one of ROW_SHARDED / LL");
        if (a->world < 2) B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: LL form needs world > 1");
        if (a->rows > 65535) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: LL form handles at most 65535 rows");
        if (!a->weight) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: LL form needs the norm (weight, y)");
        for (int r = 0; r < a->world; ++r) {
            if (!a->xstage_peers[r] || !a->ystage_peers[r] || !aligned16(a->xstage_peers[r]) || !aligned16(a->ystage_peers[r]))
                B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: staging pointers of rank %d missing or unaligned", r);
            p.xstage[r] = (uint8_t*)a->xstage_peers[r];
            p.ystage[r] = (uint8_t*)a->ystage_peers[r];
        }
        p.rows_per_rank = (int)((a->rows + a->world - 1) / a->world);
        if (p.rows_per_rank < 1) p.rows_per_rank = 1;
    }
    const bool nvls1 = (a->mode & B200_AR_MODE_NVLS_ONEHOP) != 0;
    const bool nvls = (a->mode & B200_AR_MODE_NVLS) != 0 || nvls1;
    if (nvls) {
        if (sharded || ll) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: choose one of ROW_SHARDED / LL / NVLS");
        if (a->world < 2) B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: NVLS form needs world > 1");
        if (!a->weight) B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: NVLS form needs the norm (weight, y)");
        if (nvls1 && (a->mode & B200_AR_MODE_NVLS)) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: choose one of NVLS / NVLS_ONEHOP");
        if (!a->x_mc || !a->flags_mc || !aligned16(a->x_mc) ||
            (!nvls1 && !(a->mode & B200_AR_FLAG_PACKET_TAIL) && (!a->y_mc || !aligned16(a->y_mc))))
            B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: NVLS form needs the multicast addresses of x, y and the flags");
        if (a->rows > 65535LL * a->world) B200_FAIL(B200_E_RANGE, "allreduce_residual_rmsnorm: NVLS form: too many rows");
        p.x_mc = (const __nv_bfloat16*)a->x_mc;
        p.y_mc = (__nv_bfloat16*)a->y_mc;
        p.flags_mc = (uint32_t*)a->flags_mc;
        p.rows_per_rank = (int)((a->rows + a->world - 1) / a->world);
        if (p.rows_per_rank < 1) p.rows_per_rank = 1;
    }
    if (a->rows == 0) return 0;
    p.epoch = (uint32_t*)a->epoch;
    p.r = (const __nv_bfloat16*)a->residual;
    p.w = a->weight;
    p.y = (__nv_bfloat16*)a->y;
    p.res_out = (__nv_bfloat16*)a->res_out;
    p.rows = a->rows; p.n = a->n;
    p.sx = a->stride_x; p.sr = a->stride_r; p.sy = a->stride_y; p.sres = a->stride_res;
    p.eps = a->eps; p.rank = a->rank; p.ss_rounded = a->ss_rounded ? 1 : 0;
    const int ms = a->timeout_ms > 0 ? a->timeout_ms : 2000;
    p.timeout_ns = (long long)ms * 1000000LL;
    p.trace = (unsigned long long*)a->trace;
    p.pdl = (a->mode & B200_AR_FLAG_PDL) ? 1 : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (nvls1) {
        if (a->w_dtype == B200_DT_F32) return launch_nvls<float, true, false>(p, a->world, st);
        if (a->w_dtype == B200_DT_BF16) return launch_nvls<__nv_bfloat16, true, false>(p, a->world, st);
        B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: weight dtype %d", a->w_dtype);
    }
    if (nvls && (a->mode & B200_AR_FLAG_PACKET_TAIL)) {
        for (int r = 0; r < a->world; ++r) {
            if (!a->ystage_peers[r] || !aligned16(a->ystage_peers[r]))
                B200_FAIL(B200_E_NULL, "allreduce_residual_rmsnorm: the packet tail needs ystage_peers[%d]", r);
            p.ystage[r] = (uint8_t*)a->ystage_peers[r];
        }
        if (a->w_dtype == B200_DT_F32) return launch_nvls<float, false, true>(p, a->world, st);
        if (a->w_dtype == B200_DT_BF16) return launch_nvls<__nv_bfloat16, false, true>(p, a->world, st);
        B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: weight dtype %d", a->w_dtype);
    }
    if (nvls) {
        if (a->w_dtype == B200_DT_F32) return launch_nvls<float, false, false>(p, a->world, st);
        if (a->w_dtype == B200_DT_BF16) return launch_nvls<__nv_bfloat16, false, false>(p, a->world, st);
        B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: weight dtype %d", a->w_dtype);
    }
    if (ll) {
        if (a->w_dtype == B200_DT_F32) return dispatch_ll<float>(a->world, p, st);
        if (a->w_dtype == B200_DT_BF16) return dispatch_ll<__nv_bfloat16>(a->world, p, st);
        B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: weight dtype %d", a->w_dtype);
    }
    if (a->weight && a->w_dtype == B200_DT_F32) return dispatch_ws<float>(a->world, sharded, p, st);
    if (!a->weight || a->w_dtype == B200_DT_BF16) return dispatch_ws<__nv_bfloat16>(a->world, sharded, p, st);
    B200_FAIL(B200_E_UNSUPPORTED, "allreduce_residual_rmsnorm: weight dtype %d", a->w_dtype);
}

extern "C" int b200_allgather_cols(const b200_gather_args* a, void* stream) {
    if (!a) B200_FAIL(B200_E_NULL, "allgather_cols: null args");
    if (a->world != 2 && a->world != 4 && a->world != 8) B200_FAIL(B200_E_UNSUPPORTED, "allgather_cols: world=%d (2, 4, 8)", a->world);
    if (a->rank < 0 || a->rank >= a->world) B200_FAIL(B200_E_RANGE, "allgather_cols: rank=%d", a->rank);
    if (!a->src || !a->epoch) B200_FAIL(B200_E_NULL, "allgather_cols: src / epoch is NULL");
    if (a->rows < 0 || a->cols_local <= 0 || a->cols_local % 8) B200_FAIL(B200_E_SHAPE, "allgather_cols: rows=%lld cols_local=%lld (multiple of 8)", (long long)a->rows, (long long)a->cols_local);
    if (a->stride_src % 8 || a->stride_dst % 8 || a->stride_src < a->cols_local || a->stride_dst < a->cols_local * a->world)
        B200_FAIL(B200_E_ALIGN, "allgather_cols: strides must be multiples of 8 elements and cover the row");
    if (!aligned16(a->src)) B200_FAIL(B200_E_ALIGN, "allgather_cols: src not 16-byte aligned");
    AgParams p{};
    for (int r = 0; r < a->world; ++r) {
        if (!a->dst_peers[r] || !a->flag_peers[r]) B200_FAIL(B200_E_NULL, "allgather_cols: dst_peers[%d] / flag_peers[%d] is NULL", r, r);
        if (!aligned16(a->dst_peers[r])) B200_FAIL(B200_E_ALIGN, "allgather_cols: dst_peers[%d] not 16-byte aligned", r);
        p.dst[r] = (__nv_bfloat16*)a->dst_peers[r];
        p.flags[r] = (uint32_t*)a->flag_peers[r];
    }
    if ((a->dst_mc == nullptr) != (a->flags_mc == nullptr)) B200_FAIL(B200_E_NULL, "allgather_cols: dst_mc and flags_mc go together");
    if (a->dst_mc && !aligned16(a->dst_mc)) B200_FAIL(B200_E_ALIGN, "allgather_cols: dst_mc not 16-byte aligned");
    if (a->rows == 0) return 0;
    p.src = (const __nv_bfloat16*)a->src;
    p.dst_mc = (__nv_bfloat16*)a->dst_mc;
    p.flags_mc = (uint32_t*)a->flags_mc;
    p.epoch = (uint32_t*)a->epoch;
    p.rows = a->rows; p.cols_local = a->cols_local; p.ss = a->stride_src; p.sd = a->stride_dst;
    p.rank = a->rank; p.world = a->world; p.pdl = (a->flags & 1) ? 1 : 0;
    p.timeout_ns = (long long)(a->timeout_ms > 0 ? a->timeout_ms : 2000) * 1000000LL;
    const int64_t vecs = a->rows * (a->cols_local / 8);
    int64_t grid = (vecs + 512 * 4 - 1) / (512 * 4);        // ~4 vectors per thread
    if (grid < 1) grid = 1;
    if (grid > 64) grid = 64;                                // all CTAs resident at once: they wait on one another
    cudaError_t e = launch_maybe_pdl(allgather_cols_kernel, dim3((unsigned)grid), dim3(512), 0, (cudaStream_t)stream, p.pdl != 0, p);
    if (e != cudaSuccess) { set_error("allgather_cols: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    B200_CHECK_LAUNCH();
    return 0;
}

// ------------------------------------------------------------------ peer-memory plumbing (init time)
extern "C" int b200_comm_alloc(int64_t bytes, void** dev_ptr) {
    if (!dev_ptr) B200_FAIL(B200_E_NULL, "comm_alloc: null out pointer");
    if (bytes <= 0) B200_FAIL(B200_E_RANGE, "comm_alloc: bytes=%lld", (long long)bytes);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, (size_t)bytes);
    if (e != cudaSuccess) { set_error("comm_alloc: cudaMalloc(%lld): %s", (long long)bytes, cudaGetErrorString(e)); return (int)e; }
    e = cudaMemset(p, 0, (size_t)bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p); set_error("comm_alloc: memset: %s", cudaGetErrorString(e)); return (int)e; }
    *dev_ptr = p;
    return 0;
}

extern "C" int b200_comm_free(void* dev_ptr) {
    if (!dev_ptr) return 0;
    cudaError_t e = cudaFree(dev_ptr);
    if (e != cudaSuccess) { set_error("comm_free: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}

extern "C" int b200_ipc_export(const void* dev_ptr, void* handle_out) {
    if (!dev_ptr || !handle_out) B200_FAIL(B200_E_NULL, "ipc_export: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "handle size");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr));
    if (e != cudaSuccess) { set_error("ipc_export: %s", cudaGetErrorString(e)); return (int)e; }
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}

extern "C" int b200_ipc_import(const void* handle, void** peer_ptr) {
    if (!handle || !peer_ptr) B200_FAIL(B200_E_NULL, "ipc_import: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { set_error("ipc_import: %s", cudaGetErrorString(e)); return (int)e; }
    *peer_ptr = p;
    return 0;
}

extern "C" int b200_ipc_close(void* peer_ptr) {
    if (!peer_ptr) return 0;
    cudaError_t e = cudaIpcCloseMemHandle(peer_ptr);
    if (e != cudaSuccess) { set_error("ipc_close: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}
