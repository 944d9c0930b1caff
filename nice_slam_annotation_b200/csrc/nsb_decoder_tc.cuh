// tcgen05 version of the decoder forward (included by nsb_decoder.cu after the shared helpers).
//
// Same mathematics as k_decoders_forward -- trilinear gather, fp32 Fourier embedding, the
// five-layer MLP with per-layer feature injection, 3xTF32 products with fp32 accumulation --
// mapped on the 5th-generation tensor cores:
//  * a CTA (512 threads) is bound to one decoder, keeps its weights in shared memory as tf32
//    hi / lo operand images (UMMA K-major, no swizzle) and loops over 128-point tiles;
//  * point i of the tile == TMEM lane i is served by FOUR threads (warps q, q+4, q+8, q+12 all
//    address lane quadrant q): each computes a quarter of the point's embedding e = sin(p B)
//    and, after every layer, a quarter of the activation row h_i, and writes them (split into
//    hi / lo) to tensor memory, where they are the A operands of the next MMAs (TS form);
//    sixteen warps per SM instead of four is what hides the sinf / gather / TMEM latencies;
//  * the grid features are gathered cooperatively (8 lanes x 16 B per 128-byte voxel line, so
//    every request is a fully used line) into a shared-memory operand tile (SS form);
//  * a seventeenth warp issues the MMAs: the embedding GEMM for layers 0 and 3 (N = 64) and the
//    feature GEMM for all five layers (N = 160) up front, then the serial chain of 32 x 32
//    layers.  The workers arrive at a named barrier when a group's operands are in place
//    (bar.arrive: they do not wait) and the issuing warp, waiting there, issues the group and commits
//    it to an mbarrier on which the workers wait before they read the accumulator;
//  * accumulators live in tensor memory and are read back with tcgen05.ld for the bias / ReLU /
//    feature-add epilogue of each layer.
//
// TMEM columns (all 512): e hi [0,96) | e lo [96,192) | acc E (layers 0, 3) [192,256) |
// acc C (5 x 32) [256,416) | acc H [416,448) (the 16 output columns alias it) | h hi [448,480) |
// h lo [480,512).
#pragma once
#include "nsb_tc.cuh"

namespace nsb {

constexpr int TC_THREADS = 512;                 // worker threads (the backward's chain threads use the same constant)
constexpr int TC_PARTS = TC_THREADS / 128;      // threads per point
constexpr int TCF_THREADS = TC_THREADS + 32;    // forward CTA: workers + one MMA-issuing warp
// named barriers of the forward (0 is __syncthreads)
constexpr int FB_WORK = 1;                      // the 512 workers among themselves
constexpr int FB_LAYER = 2;                     // workers arrive (operands of the next MMA group are in place), the issuer waits
constexpr int FB_INPUT = 3;                     // workers arrive (next tile's inputs in place, accumulators read), the issuer waits
__device__ __forceinline__ void nbar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void nbar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
constexpr int TM_EHI = 0, TM_ELO = 96, TM_ACCE = 192, TM_ACCC = 256, TM_ACCH = 416, TM_ACCO = 416;
constexpr int TM_HHI = 448, TM_HLO = 480;
constexpr int TRI_WORDS = 12;

template <int KIND>
struct TcSmem {      // float offsets in dynamic shared memory
    using T = TcCfg<KIND>;
    static constexpr int S_CONST = 0;                          // T_CONST_END floats
    static constexpr int S_BV = 640;                           // embedding matrix as [96] x float4
    static constexpr int S_WHI = 1024;                         // 128-byte aligned
    static constexpr int S_WLO = S_WHI + T::W_FWD;
    static constexpr int S_CHI = S_WLO + T::W_FWD;             // feature tile hi: 128 x 32, K-major image
    static constexpr int S_CLO = S_CHI + 128 * CDIM;
    static constexpr int S_TRI = S_CLO + 128 * CDIM;           // per point and level: voxel base / steps / coordinates
    static constexpr int S_PT = S_TRI + (KIND == KIND_FINE ? 2 : 1) * 128 * TRI_WORDS;   // per point: p (fp32)
    static constexpr int S_END = S_PT + 128 * 4;
    static_assert(T::T_CONST_END <= S_BV, "constants overflow their slot");
};

// 3xTF32 product D (+)= A W^T with A in tensor memory: per K step of 8, lo*hi + hi*lo + hi*hi.
// b_hi / b_lo are descriptors of an R-row K-major operand at K = 0; the next K step is 2R
// 16-byte units further (nsb_tc.cuh).
__device__ __forceinline__ void mma3_ts(uint32_t d, uint32_t a_hi, uint32_t a_lo, uint64_t b_hi, uint64_t b_lo, int R,
                                        int ksteps, uint32_t idesc, bool accumulate) {
    for (int ks = 0; ks < ksteps; ++ks) {
        const uint64_t off = (uint64_t)(ks * 2 * R);
        tc::mma_ts(d, a_lo + 8 * ks, b_hi + off, idesc, accumulate || ks > 0);
        tc::mma_ts(d, a_hi + 8 * ks, b_lo + off, idesc, true);
        tc::mma_ts(d, a_hi + 8 * ks, b_hi + off, idesc, true);
    }
}
// same with A in shared memory (128-row K-major image)
__device__ __forceinline__ void mma3_ss(uint32_t d, uint64_t a_hi, uint64_t a_lo, uint64_t b_hi, uint64_t b_lo, int R,
                                        int ksteps, uint32_t idesc, bool accumulate) {
    for (int ks = 0; ks < ksteps; ++ks) {
        const uint64_t off = (uint64_t)(ks * 2 * R), aoff = (uint64_t)(ks * 2 * 128);
        tc::mma_ss(d, a_lo + aoff, b_hi + off, idesc, accumulate || ks > 0);
        tc::mma_ss(d, a_hi + aoff, b_lo + off, idesc, true);
        tc::mma_ss(d, a_hi + aoff, b_hi + off, idesc, true);
    }
}

// split 8 values and store them as the hi / lo A-operand columns [col, col + 8) of this lane
__device__ __forceinline__ void st_split8(uint32_t lane_base, int col_hi, int col_lo, const float (&v)[8]) {
    float h[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t hi, lo;
        split_tf32(v[j], hi, lo);
        h[j] = __uint_as_float(hi);
        l[j] = __uint_as_float(lo);
    }
    tc::tmem_st8(lane_base + col_hi, h);
    tc::tmem_st8(lane_base + col_lo, l);
}

// Cooperative trilinear gather of one grid level for the tile's 128 points into the feature
// operand tile.  One thread per point publishes the voxel addressing (same arithmetic as
// tri_setup / gather8, so the features are bit-identical to the mma.sync path); after a CTA
// barrier 8 lanes fetch the 8 x 16 B of every corner line: 4 points per warp per round, 16 warps.
__device__ __forceinline__ void tc_publish_tri(const GridDev& G, const float (&u)[3], float* s_tri, int row) {
    Tri t = tri_setup(u, G);
    float* e = s_tri + row * TRI_WORDS;
    reinterpret_cast<int4*>(e)[0] = make_int4(t.base, t.sx, t.sy, t.sz);
    reinterpret_cast<float4*>(e)[1] = make_float4(t.ix, t.iy, t.iz, t.x0);
    reinterpret_cast<float2*>(e)[4] = make_float2(t.y0, t.z0);
}
// xrec (optional): weight-gradient records of the tile's eight 16-point sub-tiles (stride rec_stride floats),
// already offset to the c section and the level's first column: the fp32 features are deposited there
// for k_dweight, so that the backward does not have to recompute them (RecCfg in nsb_layout.h)
__device__ __forceinline__ void tc_gather_level(const GridDev& G, const float* s_tri, float* chi, float* clo,
                                                int warp, int lane, float* xrec = nullptr, int rec_stride = 0,
                                                int c_in = 0, int rows_ok = 128) {
    const int sub = lane & 7;
#pragma unroll
    for (int r = 0; r < 128 / (4 * (TC_THREADS / 32)); ++r) {
        const int q = (r * (TC_THREADS / 32) + warp) * 4 + (lane >> 3);
        const float* e = s_tri + q * TRI_WORDS;
        const int4 a = reinterpret_cast<const int4*>(e)[0];
        const float4 b = reinterpret_cast<const float4*>(e)[1];
        const float2 c2 = reinterpret_cast<const float2*>(e)[4];
        Tri t;
        t.base = a.x; t.sx = a.y; t.sy = a.z; t.sz = a.w;
        t.ix = b.x; t.iy = b.y; t.iz = b.z; t.x0 = b.w; t.y0 = c2.x; t.z0 = c2.y;
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
            v[k] = __ldg(reinterpret_cast<const float4*>(G.data + (size_t)tri_offset(t, k) * CDIM + 4 * sub));
        float f[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float w = tri_weight(t, k);
            f[0] = fmaf(v[k].x, w, f[0]); f[1] = fmaf(v[k].y, w, f[1]);
            f[2] = fmaf(v[k].z, w, f[2]); f[3] = fmaf(v[k].w, w, f[3]);
        }
        uint32_t h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_tf32(f[j], h[j], l[j]);
        const int idx = tc::kmajor_index(q, 4 * sub, 128);
        *reinterpret_cast<uint4*>(chi + idx) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(clo + idx) = make_uint4(l[0], l[1], l[2], l[3]);
        if (xrec && q < rows_ok) {
            const int rr = q & 15;
            *reinterpret_cast<float4*>(xrec + (size_t)(q >> 4) * rec_stride + rr * c_in + rec_sw(rr, 4 * sub)) =
                make_float4(f[0], f[1], f[2], f[3]);
        }
    }
}

// the point of row `row` of a tile: fp32 position (embedding input) and normalised grid coordinate
template <int KIND>
__device__ __forceinline__ void tc_load_point(const PointsDev& q, const DecTask& T, long long tile, int row,
                                              float (&p32)[3], float (&u)[3], bool& ok, long long& i) {
    i = tile * 128 + row;
    ok = i < q.P;
    const long long ii = ok ? i : q.P - 1;
    if (q.pu) {                 // evaluated once per point by k_sample_z (NsbPoints::point_cache)
        const float4 a4 = __ldg(reinterpret_cast<const float4*>(q.pu + 8 * ii)), b4 = __ldg(reinterpret_cast<const float4*>(q.pu + 8 * ii) + 1);
        p32[0] = a4.x; p32[1] = a4.y; p32[2] = a4.z;
        u[0] = b4.x; u[1] = b4.y; u[2] = b4.z;
    } else if (q.ro == nullptr && q.p32 != nullptr) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            p32[a] = q.p32[3 * ii + a];
            u[a] = __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(p32[a], (float)T.lo[a]), (float)T.ext[a]), 2.f), 1.0f);
        }
    } else {
        double p[3];
        if (q.ro) ray_point(q, ii, p);
        else { p[0] = q.p64[3 * ii]; p[1] = q.p64[3 * ii + 1]; p[2] = q.p64[3 * ii + 2]; }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            p32[a] = (float)p[a];
            u[a] = (float)__dsub_rn(__dmul_rn(__ddiv_rn(__dsub_rn(p[a], T.lo[a]), T.ext[a]), 2.0), 1.0);
        }
    }
}

// one 256-bit store (sm_100: STG.256) of eight consecutive floats at a 32-byte aligned address: a whole sector
// per request, where two float4 stores cost the L2 two partial-sector writes
__device__ __forceinline__ void st_global_v8(float* d, const float (&v)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "l"(d), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}

// 8 embedding features [c0, c0 + 8) of this thread's point -> tensor memory (hi / lo)
// (Bv: [feature] -> (B[0][f], B[1][f], B[2][f], 0); same operation order as phase())
// erow (optional): the e section (3 atoms, rec_e) of this point's record, r its row index there
__device__ __forceinline__ void tc_embed8(const float (&p32)[3], const float4* Bv, uint32_t lane_base, int c0,
                                          float* erow = nullptr, int r = 0) {
    float e[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 b = Bv[c0 + j];
        e[j] = sin_embed(fmaf(p32[2], b.z, fmaf(p32[1], b.y, p32[0] * b.x)));
    }
    st_split8(lane_base, TM_EHI + c0, TM_ELO + c0, e);
    if (erow) {
        st_global_v8(erow + rec_e(r, c0), e);
    }
}

// Software pipeline: the inputs of tile n+1 (point, voxel addressing, feature gather, embedding)
// are produced in the gaps of tile n's layer chain -- between issuing a layer's MMAs and waiting
// for them -- so the ALU / memory work of one tile hides the tensor-core round trips of the
// other.  The feature tile and the e columns are free for tile n+1 as soon as tile n's up-front
// GEMMs have completed; h has its own columns for that reason.
template <int KIND>
__device__ void tc_forward_cta(const StageParams& P, const DecTask& T, int task, float* sm) {
    using C = DecCfg<KIND>;
    using TT = TcCfg<KIND>;
    using S = TcSmem<KIND>;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __shared__ long long s_tile[2];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quad = warp & 3, part = warp >> 2;               // TMEM lane quadrant, share of the row's work
    const int row = 32 * quad + lane;                          // point of the tile == TMEM lane
    const float* blob = T.packed + C::TOTAL;                   // tcgen05 section of the packed blob
    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) {
        tc::mbar_init(&bar, 1);
        tc::mbar_init_fence();
        s_tile[0] = (long long)atomicAdd(P.ctr + task, 1u);
        s_tile[1] = (long long)atomicAdd(P.ctr + task, 1u);
    }
    const bool worker = tid < TC_THREADS;       // warp 16 only issues MMAs
    for (int i = tid; worker && i < TT::T_CONST_END; i += TC_THREADS) sm[S::S_CONST + i] = __ldg(blob + i);
    if (tid < EMB_PAD)
        reinterpret_cast<float4*>(sm + S::S_BV)[tid] =
            make_float4(__ldg(blob + TT::T_B + tid), __ldg(blob + TT::T_B + EMB_PAD + tid),
                        __ldg(blob + TT::T_B + 2 * EMB_PAD + tid), 0.f);
#pragma unroll 4
    for (int i = tid; worker && i < TT::W_FWD / 4; i += TC_THREADS) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(blob + TT::T_CONST_END) + i);
        uint32_t h[4], l[4];
        split_tf32(w.x, h[0], l[0]); split_tf32(w.y, h[1], l[1]);
        split_tf32(w.z, h[2], l[2]); split_tf32(w.w, h[3], l[3]);
        reinterpret_cast<uint4*>(sm + S::S_WHI)[i] = make_uint4(h[0], h[1], h[2], h[3]);
        reinterpret_cast<uint4*>(sm + S::S_WLO)[i] = make_uint4(l[0], l[1], l[2], l[3]);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    pdl_wait();                 // weights, TMEM and counters are set up; the points come from the kernel before
    const uint32_t tbase = tmem_slot;
    const uint32_t lane_base = tc::tmem_addr(tbase, 32 * quad, 0);
    const float4* Bv = reinterpret_cast<const float4*>(sm + S::S_BV);
    float4* s_pt = reinterpret_cast<float4*>(sm + S::S_PT);
    const float* bias = sm + S::S_CONST + TT::T_BIAS;
    const float* biasc = sm + S::S_CONST + TT::T_BC;
    float* s_tri = sm + S::S_TRI;
    // operand descriptors at K = 0
    const uint32_t whi = tc::smem_u32(sm + S::S_WHI), wlo = tc::smem_u32(sm + S::S_WLO);
    auto wdesc = [&](uint32_t base, int off_floats, int R) {
        return tc::kmajor_desc(base + 4u * (uint32_t)(off_floats - TT::T_CONST_END), R, 0);
    };
    const uint64_t d_chi = tc::kmajor_desc(tc::smem_u32(sm + S::S_CHI), 128, 0);
    const uint64_t d_clo = tc::kmajor_desc(tc::smem_u32(sm + S::S_CLO), 128, 0);
    constexpr uint32_t ID64 = tc::idesc_tf32(128, 64), ID160 = tc::idesc_tf32(128, 160), ID32 = tc::idesc_tf32(128, 32),
                       ID16 = tc::idesc_tf32(128, 16);
    constexpr int EPP = EMB_PAD / TC_PARTS;                    // embedding features per thread (24)
    constexpr int HPP = HID / TC_PARTS;                        // hidden columns per thread (8)
    static_assert(EPP == 24 && HPP == 8, "per-thread shares are whole 8-column TMEM accesses");
    uint32_t par = 0;                                          // mbarrier phase parity
    const long long ntiles = (P.pts.P + 127) / 128;
    // X sections of the weight-gradient records (e, c, h1..h5), written here when the backward of these
    // points wants d(W) of this decoder: its backward then skips the recomputation (DecTask::dw_scratch)
    using RC = RecCfg<KIND>;
    float* const xrec = T.dw_scratch;
    const long long rows_total = (P.pts.P + 15) / 16 * 16;     // whole 16-point records (padded rows hold finite values)
    auto rec_row = [&](long long t_, int section, int width) -> float* {      // this thread's row of a section
        const long long i_ = t_ * 128 + row;
        if (!xrec || i_ >= rows_total) return nullptr;
        return xrec + (size_t)(i_ >> 4) * RC::REC + section + (row & 15) * width;
    };
    auto rec_tile = [&](long long t_, int section) -> float* {                // first record of a 128-point tile
        return xrec ? xrec + (size_t)(t_ * 8) * RC::REC + section : nullptr;
    };
    auto rows_ok = [&](long long t_) -> int {
        const long long left = rows_total - t_ * 128;
        return left >= 128 ? 128 : (left > 0 ? (int)left : 0);
    };
#ifdef NSB_TC_PROFILE
    long long tq[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
#define TQ(k) { const long long tn_ = clock64(); tq[k] += tn_ - tlast; tlast = tn_; }
#else
#define TQ(k)
#endif
    // the up-front GEMMs of a tile: features x all five fc_c (N = 160) and e x [W0 ; W3e] (N = 64)
    auto issue_input_gemms = [&]() {
        mma3_ss(tbase + TM_ACCC, d_chi, d_clo, wdesc(whi, TT::T_WC, 160), wdesc(wlo, TT::T_WC, 160), 160, CDIM / 8,
                ID160, false);
        mma3_ts(tbase + TM_ACCE, tbase + TM_EHI, tbase + TM_ELO, wdesc(whi, TT::T_WE, 64), wdesc(wlo, TT::T_WE, 64), 64,
                EMB_PAD / 8, ID64, false);
        tc::mma_commit(&bar);
    };

    long long tile = s_tile[0];
    int slot = 0;                      // s_tile[slot] is the current tile, s_tile[slot ^ 1] the next
    if (!worker) {
        // ---- the issuing warp: waits for the workers' arrival at a named barrier, issues the MMA group they prepared
        // and commits it to `bar`.  It follows the same tile sequence as the workers (s_tile), one group behind them;
        // the workers never wait at a CTA barrier for the issue and none of them pays its instruction time.
        auto issue = [&](auto&& f) {
            if (tc::elect_one()) {
                tc::fence_after_sync();
                f();
            }
            __syncwarp();
        };
        if (tile < ntiles) {
            nbar_sync(FB_INPUT, TCF_THREADS);
            issue(issue_input_gemms);
        }
        while (tile < ntiles) {
            if constexpr (KIND == KIND_FINE) {
                nbar_sync(FB_LAYER, TCF_THREADS);
                issue([&]() {
                    const uint64_t koff = (uint64_t)(4 * 2 * 160);     // K columns 32..63: 4 K steps further
                    mma3_ss(tbase + TM_ACCC, d_chi, d_clo, wdesc(whi, TT::T_WC, 160) + koff, wdesc(wlo, TT::T_WC, 160) + koff,
                            160, CDIM / 8, ID160, true);
                    tc::mma_commit(&bar);
                });
            }
#pragma unroll 1
            for (int L = 0; L < 5; ++L) {
                nbar_sync(FB_LAYER, TCF_THREADS);
                issue([&]() {
                    if (L < 4) {
                        const int off = L == 0 ? TT::T_W1 : (L == 1 ? TT::T_W2 : (L == 2 ? TT::T_W3H : TT::T_W4));
                        // layer 3 accumulates on top of its embedding part (the skip connection [e, h])
                        mma3_ts(tbase + (L == 2 ? TM_ACCE + 32 : TM_ACCH), tbase + TM_HHI, tbase + TM_HLO, wdesc(whi, off, 32),
                                wdesc(wlo, off, 32), 32, HID / 8, ID32, L == 2);
                    } else {
                        mma3_ts(tbase + TM_ACCO, tbase + TM_HHI, tbase + TM_HLO, wdesc(whi, TT::T_WO, TT::NO),
                                wdesc(wlo, TT::T_WO, TT::NO), TT::NO, HID / 8, ID16, false);
                    }
                    tc::mma_commit(&bar);
                });
            }
            nbar_sync(FB_INPUT, TCF_THREADS);          // the workers published s_tile[slot] (tile after next) before it
            const long long nxt = s_tile[slot ^ 1];
            if (nxt < ntiles) issue(issue_input_gemms);
            tile = nxt;
            slot ^= 1;
        }
        tc::fence_before_sync();
        __syncthreads();
        return;
    }
    // ---- prologue: inputs and up-front GEMMs of the first tile ------------------------------------------
    // one thread per point (part 0) evaluates the fp64 point and publishes it for the other three
    auto publish_point = [&](long long t) {
        if (part == 0) {
            float p32[3], u[3];
            bool ok_;
            long long i_;
            tc_load_point<KIND>(P.pts, T, t, row, p32, u, ok_, i_);
            s_pt[row] = make_float4(p32[0], p32[1], p32[2], 0.f);

            tc_publish_tri(T.own, u, s_tri, row);
            if constexpr (KIND == KIND_FINE) tc_publish_tri(T.aux, u, s_tri + 128 * TRI_WORDS, row);
        }
    };
    if (tile < ntiles) {
        publish_point(tile);
        nbar_sync(FB_WORK, TC_THREADS);
        tc_gather_level(T.own, s_tri, sm + S::S_CHI, sm + S::S_CLO, warp, lane, rec_tile(tile, RC::R_C), RC::REC, C::C_IN,
                        rows_ok(tile));
        {
            const float4 pp = s_pt[row];
            const float p32[3] = {pp.x, pp.y, pp.z};
            float* er = rec_row(tile, RC::R_E, 0);
#pragma unroll
            for (int j = 0; j < 3; ++j) tc_embed8(p32, Bv, lane_base, EPP * part + 8 * j, er, row & 15);
        }
        tc::tmem_st_wait();
        tc::fence_async_smem();
        tc::fence_before_sync();
        nbar_arrive(FB_INPUT, TCF_THREADS);
    }
    TQ(0)      // set-up + prologue
    while (tile < ntiles) {
        tc::mbar_wait(&bar, par);          // up-front GEMMs of this tile
        par ^= 1;
        tc::fence_after_sync();
        TQ(1)  // wait for the up-front GEMMs
        if constexpr (KIND == KIND_FINE) {
            // second half of the feature GEMM: the middle-grid feature (channels 32..63), no gradient
            tc_gather_level(T.aux, s_tri + 128 * TRI_WORDS, sm + S::S_CHI, sm + S::S_CLO, warp, lane,
                            xrec ? rec_tile(tile, RC::R_C) + 32 : nullptr, RC::REC, C::C_IN, rows_ok(tile));
            tc::fence_async_smem();
            tc::fence_before_sync();
            nbar_arrive(FB_LAYER, TCF_THREADS);
            tc::mbar_wait(&bar, par);
            par ^= 1;
            tc::fence_after_sync();
        }
        TQ(2)  // fine: second feature pass
        // ---- next tile: point and voxel addressing (published before the chain's first barrier) ----
        const long long nxt = s_tile[slot ^ 1];
        const bool has_next = nxt < ntiles;
        unsigned after_next = 0u;                   // the tile after next, fetched now, published in the L = 0 slot
        if (tid == 0) after_next = atomicAdd(P.ctr + task, 1u);
        if (has_next) publish_point(nxt);
        TQ(3)  // next tile's point + publish
        // ---- the layer chain: every thread finishes 8 columns of its row ------------------------------
        const int hc = HPP * part;
        unsigned relu_lo = 0, relu_hi = 0;
#pragma unroll 1
        for (int L = 0; L < 5; ++L) {
            const int acc_col = L == 0 ? TM_ACCE : (L == 3 ? TM_ACCE + 32 : TM_ACCH);
            float a[8], cc[8];
            tc::tmem_ld8(lane_base + acc_col + hc, a);
            tc::tmem_ld8(lane_base + TM_ACCC + 32 * L + hc, cc);
            tc::tmem_ld_wait();
            unsigned sign = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float pre = a[j] + bias[32 * L + hc + j];
                sign |= (pre > 0.f ? 1u : 0u) << j;
                a[j] = (fmaxf(pre, 0.f) + biasc[32 * L + hc + j]) + cc[j];
            }
            if (L < 4) relu_lo |= sign << (8 * L);      // signs of features 8 part .. 8 part + 7, one byte per layer
            else relu_hi = sign;
            st_split8(lane_base, TM_HHI + hc, TM_HLO + hc, a);
            if (float* hr = rec_row(tile, RC::R_H + L * 16 * HID, HID)) {        // h_{L+1}: input of the next layer's d(W)
                st_global_v8(hr + rec_sw(row & 15, hc), a);
            }
            tc::tmem_st_wait();
            tc::fence_before_sync();
            TQ(4)  // layer epilogue (ld, alu, st)
            nbar_arrive(FB_LAYER, TCF_THREADS);         // my columns of h are in place: the issuing warp takes it from here
            TQ(5)  // arrive
            if (L == 0) {
                // the tile after next: its slot holds the current tile index, which everyone (the issuing warp too)
                // has in a register by now; readers are ordered behind this store by the barriers that follow
                if (tid == 0) s_tile[slot] = (long long)after_next;
                nbar_sync(FB_WORK, TC_THREADS);         // next tile's points / voxel addressing are published
            }
            TQ(6)  // MMA issue (warp 0)
            // ---- a slice of the next tile's inputs while the tensor core works ------------------------
            if (has_next) {
                if (L == 0) {
                    if (!(P.flags & 0x20000000u))      // NSB_DEBUG_FWD=2: timing experiment, wrong results
                        tc_gather_level(T.own, s_tri, sm + S::S_CHI, sm + S::S_CLO, warp, lane, rec_tile(nxt, RC::R_C),
                                        RC::REC, C::C_IN, rows_ok(nxt));
                } else if (L <= 3 && !(P.flags & 0x10000000u)) {      // NSB_DEBUG_FWD=1: no embedding
                    const float4 pp = s_pt[row];
                    const float p32n[3] = {pp.x, pp.y, pp.z};
                    tc_embed8(p32n, Bv, lane_base, EPP * part + 8 * (L - 1), rec_row(nxt, RC::R_E, 0), row & 15);
                }
            }
            TQ(7)  // next tile's slice
            tc::mbar_wait(&bar, par);
            par ^= 1;
            tc::fence_after_sync();
            TQ(8)  // wait for the layer's MMAs
        }
        if (T.relu) {      // 8 bytes per (point, part): layers 0..4 (+ padding), one coalesced 4 KB store per tile
            const long long i = tile * 128 + row;
            if (i < P.pts.P) reinterpret_cast<uint2*>(T.relu)[i * 4 + part] = make_uint2(relu_lo, relu_hi);
        }
        if (part == 0) {
            float o[8];
            tc::tmem_ld8(lane_base + TM_ACCO, o);
            tc::tmem_ld_wait();
            const float* bo = sm + S::S_CONST + TT::T_BO;
            const long long i = tile * 128 + row;
            if (i < P.pts.P) {
                if constexpr (C::NOUT == 1) T.out[i] = o[0] + bo[0];
                else *reinterpret_cast<float4*>(T.out + 4 * i) = make_float4(o[0] + bo[0], o[1] + bo[1], o[2] + bo[2], o[3] + bo[3]);
            }
        }
        // this tile's tensor-memory reads and the next tile's operand writes are complete before
        // the next tile's up-front GEMMs overwrite the accumulators
        tc::tmem_st_wait();
        tc::fence_async_smem();
        tc::fence_before_sync();
        TQ(9)  // output
        nbar_arrive(FB_INPUT, TCF_THREADS);
        TQ(10) // end of tile
        tile = nxt;
        slot ^= 1;
    }
#ifdef NSB_TC_PROFILE
    if ((tid == 0 || tid == 200) && (int)blockIdx.x == T.cta_begin && P.prof)
        for (int k = 0; k < 12; ++k) P.prof[(task * 2 + (tid != 0)) * 12 + k] = tq[k];
#endif
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tbase);
}

__global__ void __launch_bounds__(TCF_THREADS, 1)
k_decoders_forward_tc(const __grid_constant__ StageParams P) {
    extern __shared__ __align__(16) float sm[];
    pdl_launch_dependents();     // the compositing kernel may be scheduled as CTAs of this kernel retire
    int k = 0;
    while (k + 1 < P.ntask && (int)blockIdx.x >= P.task[k + 1].cta_begin) ++k;
    const DecTask& T = P.task[k];
    switch (T.kind) {
        case KIND_MIDDLE: tc_forward_cta<KIND_MIDDLE>(P, T, k, sm); break;
        case KIND_FINE: tc_forward_cta<KIND_FINE>(P, T, k, sm); break;
        case KIND_COLOR: tc_forward_cta<KIND_COLOR>(P, T, k, sm); break;
        default: break;
    }
    release_counters(P.ctr);
}

}  // namespace nsb
