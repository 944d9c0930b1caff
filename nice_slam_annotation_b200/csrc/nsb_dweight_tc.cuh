// tcgen05 version of the weight-gradient kernel (included by nsb_decoder.cu after k_dweight's definitions).
//
// d(W) = G^T X summed over points, from the same 16-point records k_dweight reads (RecCfg in nsb_layout.h).  Here the
// sum over points is the K dimension of tensor-core products whose operands are the records themselves:
//  * a record's 32-channel atoms ([16 points][32 channels], 32-byte chunks XOR-swizzled with point & 3) are exactly
//    tcgen05's MN-major tf32 operand layout (SWIZZLE_128B_BASE32B; scripts/probe_mn_decode.cu), so one bulk
//    asynchronous copy (cp.async.bulk -> mbarrier) lands a record in shared memory as a ready operand image;
//  * adjacent atoms stack along M / N, and two wide products per record cover every 32 x 32 block at once:
//        D1[128 x 128] += [h1 h2 h3 h4]^T [da1 da2 da4 da3]       (W1, W2, W4 and the h part of W3 on the diagonal)
//        D2[128 x 224] += [c e0 e1 e2]^T [da3 da0 dh0..dh4]       (W3 / W0 embedding parts, the five fc_c)
//    accumulated in tensor memory over all records of the CTA (352 columns) -- the off-diagonal blocks are wasted
//    work on a pipe that is otherwise idle, traded for 12 instead of ~70 MMA instructions per record;
//  * 3xTF32: the tensor core truncates an fp32 operand to tf32 (measured), so the record is its own "hi" image and
//    the split warps only write lo = x - trunc(x) (exact) into a second buffer: hi*hi + lo*hi + hi*lo;
//  * the small products -- the output layer, d(B) of the embedding, all bias sums -- are fp32 FMAs of the split
//    warps, kept in registers for the whole kernel;
//  * warp roles: warp 0 = producer (bulk copies, 3-deep ring), warp 1 = MMA issuer, warps 2..17 = split / small
//    products / final flush (one atomic per element per CTA, like k_dweight).
// Decoders with a 32-channel feature input (middle, colour); the fine decoder (64) keeps k_dweight.
#pragma once
#include "nsb_tc.cuh"

namespace nsb {

constexpr int DWT_SPLIT_WARPS = 16;
constexpr int DWT_SPLIT = 32 * DWT_SPLIT_WARPS;
constexpr int DWT_THREADS = 64 + DWT_SPLIT;
constexpr int DWT_RING = 3;
constexpr int DWT_D1 = 0, DWT_D2 = 128;            // TMEM columns

struct DwTcOffs {          // flat-gradient offsets of the canonical parameters (floats)
    int w[5], b[5], wc[5], bc[5], wo, bo, emb;
};

template <int KIND>
struct DwTcSmem {          // float offsets in dynamic shared memory (base aligned to 1024 B)
    using RC = RecCfg<KIND>;
    static_assert(DecCfg<KIND>::C_IN == HID && DecCfg<KIND>::EMBED, "k_dweight_tc: 32-channel feature input");
    static constexpr int X_LO = 8 * RC::ATOM;                  // c, e0..e2, h1..h4
    static constexpr int G_LO = 10 * RC::ATOM;                 // d(a) slots 0..4, d(h1)..d(h5)
    static constexpr int S_RAW = 0;
    static constexpr int S_LO = S_RAW + DWT_RING * RC::REC;
    static constexpr int S_LIVE = S_LO + 2 * (X_LO + G_LO);
    static constexpr int S_BAR = S_LIVE + DW_LIVE_MAX;         // mbarriers (8 B each)
    static constexpr int S_END = S_BAR + 32;
    static_assert(RC::REC % 128 == 0, "records keep the 512-byte atom alignment");
    static_assert(RC::R_H == RC::R_C + 4 * RC::ATOM && RC::R_DH == RC::R_DA + 5 * RC::ATOM, "operand stacks are adjacent");
};

__device__ __forceinline__ uint64_t dwt_desc(uint32_t saddr) {        // MN-major atom stack, see the header
    return tc::smem_desc(saddr, 2048u, 512u) | (1ull << 61);
}
__host__ __device__ constexpr uint32_t dwt_idesc(int N) { return tc::idesc_tf32(128, N) | (1u << 15) | (1u << 16); }

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(tc::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     tc::smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(tc::smem_u32(bar))
                 : "memory");
}

// one small-product output of a split thread: sum over the 16 points of U[r] * V[r] (U == 1 for a bias sum)
struct DwtSlot {
    int u_base, u_stride, u_col, u_sw;     // u_base < 0: U == 1
    int v_base, v_col;                     // V always lives in an atom: v_base + r * 32 + (v_col ^ swizzle(r))
    int out;                               // flat gradient index, < 0: no output
};

template <int KIND>
__device__ __forceinline__ DwtSlot dwt_slot(int idx, const DwTcOffs& O) {
    using C = DecCfg<KIND>;
    using RC = RecCfg<KIND>;
    DwtSlot s{-1, 0, 0, 0, 0, 0, -1};
    if (idx < 160) {                                   // b_i = sum d(a_i)
        const int i = idx >> 5, n = idx & 31;
        s.v_base = RC::da(i); s.v_col = n; s.out = O.b[i] + n;
    } else if (idx < 320) {                            // bc_i = sum d(h_{i+1})
        const int i = (idx - 160) >> 5, n = idx & 31;
        s.v_base = RC::R_DH + i * RC::ATOM; s.v_col = n; s.out = O.bc[i] + n;
    } else if (idx < 320 + C::NOUT) {                  // bo
        const int o = idx - 320;
        s.v_base = RC::R_DO; s.v_col = o; s.out = O.bo + o;
    } else if (idx < 352) {
        // unused (keeps the ranges warp aligned)
    } else if (idx < 352 + 32 * C::NOUT) {             // Wo[o][m] = sum d(out)[o] h5[m]
        const int o = (idx - 352) >> 5, m = idx & 31;
        s.u_base = RC::R_DO; s.u_stride = HID; s.u_col = o; s.u_sw = 1;
        s.v_base = RC::R_H + 4 * RC::ATOM; s.v_col = m; s.out = O.wo + o * HID + m;
    } else if (idx >= 480 && idx < 480 + 3 * EMB_PAD) {   // d(B)[a][f] = sum p[a] d(e)[f] cos
        const int a = (idx - 480) / EMB_PAD, f = (idx - 480) % EMB_PAD;
        if (f < EMB) {
            s.u_base = RC::R_P; s.u_stride = 8; s.u_col = a; s.u_sw = 0;
            s.v_base = RC::R_DE + (f >> 5) * RC::ATOM; s.v_col = f & 31; s.out = O.emb + a * EMB + f;
        }
    }
    return s;
}
constexpr int DWT_SLOTS = 480 + 3 * EMB_PAD;           // 768 outputs -> two rounds of 512 threads (second: 256)

template <int KIND>
__device__ void dweight_tc_body(const DwParams& P, const DwTcOffs& O, float* sm) {
    using C = DecCfg<KIND>;
    using RC = RecCfg<KIND>;
    using S = DwTcSmem<KIND>;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __shared__ uint32_t s_tmem;
    __shared__ int s_nlive;
    __shared__ int s_warp_cnt[DWT_THREADS / 32];
    int* s_live = reinterpret_cast<int*>(sm + S::S_LIVE);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + S::S_BAR);
    uint64_t* full = bars;                   // [3] record landed
    uint64_t* done = bars + 3;               // [3] the MMAs that read the record (and its lo image) have completed
    uint64_t* lo_ready = bars + 6;           // [2] lo image written
    if (warp == 1) tc::tmem_alloc<512>(&s_tmem);
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i) { tc::mbar_init(full + i, 1); tc::mbar_init(done + i, 1); }
        for (int i = 0; i < 2; ++i) tc::mbar_init(lo_ready + i, DWT_SPLIT);
        tc::mbar_init_fence();
    }
    pdl_wait();                              // flags and records come from the kernels before
    // balanced shares of the live records, as in k_dweight
    const long long T_ = P.ntiles;
    const int per_thread = (int)((T_ + DWT_THREADS - 1) / DWT_THREADS);
    const long long my0 = (long long)threadIdx.x * per_thread;
    int mine = 0;
    for (int j = 0; j < per_thread; ++j) {
        const long long i = my0 + j;
        if (i < T_ && P.flags[i] != 0) ++mine;
    }
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(FULL, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp_cnt[warp] = incl;
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    int before = incl - mine, total = 0;
    for (int w = 0; w < DWT_THREADS / 32; ++w) {
        const int c = s_warp_cnt[w];
        if (w < warp) before += c;
        total += c;
    }
    const long long lo_r = (long long)blockIdx.x * total / gridDim.x, hi_r = (long long)(blockIdx.x + 1) * total / gridDim.x;
    {
        int rank = before;
        for (int j = 0; j < per_thread; ++j) {
            const long long i = my0 + j;
            if (i < T_ && P.flags[i] != 0) {
                if (rank >= lo_r && rank < hi_r && rank - lo_r < DW_LIVE_MAX) s_live[rank - lo_r] = (int)i;
                ++rank;
            }
        }
    }
    if (threadIdx.x == 0) s_nlive = (int)((hi_r - lo_r) < DW_LIVE_MAX ? (hi_r - lo_r) : DW_LIVE_MAX);
    __syncthreads();
    const int nlive = s_nlive;
    const uint32_t tbase = s_tmem;
    float* raw = sm + S::S_RAW;
    float* lo = sm + S::S_LO;
    constexpr uint32_t REC_BYTES = RC::REC * 4;

    if (warp == 0) {
        // ---- producer ------------------------------------------------------------------------------------------
        if (lane == 0) {
            for (int i = 0; i < nlive; ++i) {
                const int s = i % DWT_RING;
                if (i >= DWT_RING) tc::mbar_wait(done + s, (uint32_t)((i / DWT_RING - 1) & 1));
                mbar_expect_tx(full + s, REC_BYTES);
                const char* src = reinterpret_cast<const char*>(P.scratch + (size_t)s_live[i] * RC::REC);
                char* dst = reinterpret_cast<char*>(raw + s * RC::REC);
                constexpr uint32_t PART = REC_BYTES / 3;
                static_assert(PART % 16 == 0 && PART * 3 == REC_BYTES, "bulk copy granularity");
#pragma unroll
                for (int c = 0; c < 3; ++c) bulk_g2s(dst + c * PART, src + c * PART, PART, full + s);
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer ----------------------------------------------------------------------------------------
        if (tc::elect_one()) {
            for (int i = 0; i < nlive; ++i) {
                const int s = i % DWT_RING, b = i & 1;
                tc::mbar_wait(full + s, (uint32_t)((i / DWT_RING) & 1));
                tc::mbar_wait(lo_ready + b, (uint32_t)((i >> 1) & 1));
                tc::fence_after_sync();
                const uint32_t r0 = tc::smem_u32(raw + s * RC::REC), l0 = tc::smem_u32(lo + b * (S::X_LO + S::G_LO));
                const uint32_t x2h = r0 + 4u * RC::R_C, x1h = r0 + 4u * RC::R_H, g1h = r0 + 4u * RC::R_DA,
                               g2h = r0 + 4u * (RC::R_DA + 3 * RC::ATOM);
                const uint32_t x2l = l0, x1l = l0 + 4u * (4 * RC::ATOM), g1l = l0 + 4u * S::X_LO,
                               g2l = l0 + 4u * (S::X_LO + 3 * RC::ATOM);
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    const uint32_t k = 1024u * ks;
                    const bool acc = i > 0 || ks > 0;
                    tc::mma_ss(tbase + DWT_D1, dwt_desc(x1l + k), dwt_desc(g1h + k), dwt_idesc(128), acc);
                    tc::mma_ss(tbase + DWT_D1, dwt_desc(x1h + k), dwt_desc(g1l + k), dwt_idesc(128), true);
                    tc::mma_ss(tbase + DWT_D1, dwt_desc(x1h + k), dwt_desc(g1h + k), dwt_idesc(128), true);
                    tc::mma_ss(tbase + DWT_D2, dwt_desc(x2l + k), dwt_desc(g2h + k), dwt_idesc(224), acc);
                    tc::mma_ss(tbase + DWT_D2, dwt_desc(x2h + k), dwt_desc(g2l + k), dwt_idesc(224), true);
                    tc::mma_ss(tbase + DWT_D2, dwt_desc(x2h + k), dwt_desc(g2h + k), dwt_idesc(224), true);
                }
                tc::mma_commit(done + s);
            }
        }
        __syncwarp();
    } else {
        // ---- split warps: lo images, small products ---------------------------------------------------------------
        const int t = threadIdx.x - 64;
        const DwtSlot s0 = dwt_slot<KIND>(t, O);
        const DwtSlot s1 = t + DWT_SPLIT < DWT_SLOTS ? dwt_slot<KIND>(t + DWT_SPLIT, O) : DwtSlot{-1, 0, 0, 0, 0, 0, -1};
        float acc0 = 0.f, acc1 = 0.f;
        auto small = [&](const DwtSlot& sl, const float* R, float& acc) {
            if (sl.out < 0) return;
            float a = 0.f;
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int sw = (r & 3) << 3;
                const float v = R[sl.v_base + r * HID + (sl.v_col ^ sw)];
                const float u = sl.u_base < 0 ? 1.f : R[sl.u_base + r * sl.u_stride + (sl.u_col ^ (sl.u_sw ? sw : 0))];
                a = fmaf(u, v, a);
            }
            acc += a;
        };
        for (int i = 0; i < nlive; ++i) {
            const int s = i % DWT_RING, b = i & 1;
            tc::mbar_wait(full + s, (uint32_t)((i / DWT_RING) & 1));
            const float* R = raw + s * RC::REC;
            small(s0, R, acc0);
            small(s1, R, acc1);
            if (i >= 2) tc::mbar_wait(done + (i - 2) % DWT_RING, (uint32_t)(((i - 2) / DWT_RING) & 1));   // lo slot b is free
            float* L = lo + b * (S::X_LO + S::G_LO);
            constexpr int NX4 = S::X_LO / 4, NG4 = S::G_LO / 4;
#pragma unroll 2
            for (int q = t; q < NX4 + NG4; q += DWT_SPLIT) {
                const float4 x = q < NX4 ? reinterpret_cast<const float4*>(R + RC::R_C)[q]
                                         : reinterpret_cast<const float4*>(R + RC::R_DA)[q - NX4];
                float4 l;
                l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
                l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
                l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u);
                l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u);
                reinterpret_cast<float4*>(L)[q] = l;
            }
            tc::fence_async_smem();
            mbar_arrive(lo_ready + b);
        }
        if (s0.out >= 0 && acc0 != 0.f) atomicAdd(P.dflat + s0.out, acc0);
        if (s1.out >= 0 && acc1 != 0.f) atomicAdd(P.dflat + s1.out, acc1);
        // ---- flush of the accumulators (after the last record's MMAs) ---------------------------------------------
        if (nlive > 0 && warp < 10) {
            const int il = nlive - 1;
            tc::mbar_wait(done + il % DWT_RING, (uint32_t)((il / DWT_RING) & 1));
            tc::fence_after_sync();
            const int q = warp & 3;                                   // TMEM lane quadrant of this warp
            const uint32_t lane_base = tc::tmem_addr(tbase, 32 * q, 0);
            auto add_block = [&](int col0, int out0, int ld_n, bool row_ok) {
                // 32 columns (n) of this thread's row: flat[out0 + n * ld_n]
#pragma unroll
                for (int c = 0; c < 32; c += 16) {
                    float v[16];
                    tc::tmem_ld16(lane_base + col0 + c, v);
                    tc::tmem_ld_wait();
                    if (row_ok) {
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            if (v[j] != 0.f) atomicAdd(P.dflat + out0 + (c + j) * ld_n, v[j]);
                    }
                }
            };
            if (warp < 6) {
                // D1: rows of quadrant q are h_{q+1}; its layer's gradient block (columns: da1, da2, da4, da3)
                const int layer = q + 1;
                const int col0 = DWT_D1 + (layer == 3 ? 96 : (layer == 4 ? 64 : 32 * q));
                const int in = C::in_dim(layer);
                add_block(col0, O.w[layer] + (layer == 3 ? EMB : 0) + lane, in, true);
            } else {
                if (q == 0) {                                          // rows c: the five fc_c
                    for (int i = 0; i < 5; ++i) add_block(DWT_D2 + 64 + 32 * i, O.wc[i] + lane, C::C_IN, true);
                } else {                                               // rows e (features 32 (q - 1) + lane)
                    const int f = 32 * (q - 1) + lane;
                    add_block(DWT_D2, O.w[3] + f, C::in_dim(3), f < EMB);
                    add_block(DWT_D2 + 32, O.w[0] + f, C::in_dim(0), f < EMB);
                }
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<512>(tbase);
}

struct DwTcParams {
    DwParams p;
    DwTcOffs off;
    int kind;
};

__global__ void __launch_bounds__(DWT_THREADS, 1) k_dweight_tc(const DwTcParams P) {
    extern __shared__ __align__(16) float dwt_sm_raw[];
    // the operand atoms need their 512-byte alignment: align the base by hand (the launch asks for 1 KB of slack)
    float* dwt_sm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(dwt_sm_raw) + 1023) & ~(uintptr_t)1023);
    pdl_launch_dependents();
    if (P.kind == KIND_COLOR) dweight_tc_body<KIND_COLOR>(P.p, P.off, dwt_sm);
    else dweight_tc_body<KIND_MIDDLE>(P.p, P.off, dwt_sm);
}

}  // namespace nsb
