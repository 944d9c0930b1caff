// tcgen05 version of the decoder backward (included by nsb_decoder.cu after nsb_decoder_tc.cuh).
//
// Adjoint of MLP.forward (src/conv_onet/models/decoder.py:190-219) + the trilinear scatter (adjoint of
// F.grid_sample, decoder.py:179-188) for decoders whose forward (k_decoders_forward_tc) saved the ReLU signs and,
// when their weights train, deposited the layer inputs in the weight-gradient records: nothing is recomputed.
// Same mathematics as k_decoders_backward (3xTF32 products, fp32 accumulation), mapped like the forward:
//  * a CTA (512 threads) is bound to one decoder and loops over 128-point tiles; point i of the tile == TMEM lane i
//    is served by four threads (part = warp / 4), each owning 8 of the 32 hidden columns -- exactly the (point,
//    part) cells the forward wrote the ReLU sign bytes for;
//  * with da_i = d(a_i) (da_5 := d(out)), layer i of the adjoint chain is ONE MMA group with A = da_i (tensor
//    memory, hi / lo) and N = 64:  [ dh_{i-1}' | dc_i ] = da_i [ W_i^T | (W_i Wc_{i-1})^T ]  -- the feature gradient
//    dc = sum_i dh_i' Wc_{i-1} is folded into the chain through the precomputed products (TcCfg::T_BG*), so a layer
//    costs 12 MMA instructions instead of 24; every group writes its own 64 accumulator columns, the five dc parts
//    are summed in the epilogue;
//  * d(e) = da_0 W_0[:, :93] + da_3 W_3[:, :93] (N = 96) only when somebody needs it: d(point) (tracking, BA) or
//    d(B) of a decoder that trains;
//  * scatter: FOUR EXTRA WARPS do nothing else.  The chain threads leave d(c) of a tile and its voxel addressing in a
//    double-buffered shared-memory image and go on to the next tile; the scatter warps add it to the grid gradient
//    with red.global.add.v4.f32, 8 lanes x 16 B per 128-byte voxel line (whole lines and sectors, 4 608 000 sectors
//    per 48 000-point step = the algorithmic count).  The scatter is bound by the SM's write path to L2 (~131 KB per
//    tile), the chain by tensor-core round trips: they overlap instead of alternating (named barriers 2..5);
//  * weight gradients: the G sections of the records (d(a_i), d(h_i), d(out), d(e) cos, p) are written from the
//    epilogue registers in the layout k_dweight reads; 16-point sub-tiles whose incoming gradient is zero are
//    flagged dead and skipped, whole tiles of zeros are skipped altogether.
//
// TMEM columns: da hi [0,32) | da lo [32,64) | group 5,4,3,2,1 accumulators [64 + 64 g', +64) (dh' in the first 32,
// dc part in the last 32) | d(e) [384,480).
#pragma once
#include "nsb_decoder_tc.cuh"

namespace nsb {

constexpr int TB_DAHI = 0, TB_DALO = 32, TB_GRP = 64, TB_DE = 384;
constexpr int TCB_SCATTER_WARPS = 4;
constexpr int TCB_THREADS = TC_THREADS + 32 * TCB_SCATTER_WARPS;     // 512 chain threads + the scatter warps
constexpr int BAR_CHAIN = 1, BAR_FULL = 2, BAR_EMPTY = 4;            // named barriers (+ buffer index for FULL / EMPTY)
__device__ __forceinline__ int nbar_or(int id, int n, int pred) {
    int r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 p, %3, 0;\n\t"
        "bar.red.or.pred q, %1, %2, p;\n\t"
        "selp.b32 %0, 1, 0, q;\n\t"
        "}\n"
        : "=r"(r)
        : "r"(id), "r"(n), "r"(pred)
        : "memory");
    return r;
}
__device__ __forceinline__ constexpr int tb_group(int i) { return TB_GRP + 64 * (5 - i); }     // i = 5..1

template <int KIND>
struct TbSmem {      // float offsets in dynamic shared memory
    using T = TcCfg<KIND>;
    static constexpr int S_BV = 0;                             // embedding matrix as [96] x float4
    static constexpr int S_WHI = 384;                          // 128-byte aligned
    static constexpr int S_WLO = S_WHI + T::W_BWD;
    static constexpr int S_TRI = S_WLO + T::W_BWD;             // x2: per point: clipped coordinate / floor / border flags
    static constexpr int S_PT = S_TRI + 2 * 128 * TRI_WORDS;   // x2: per point: p (fp32), valid
    static constexpr int S_DP = S_PT + 2 * 128 * 4;            // per part and point: partial d(point) [4][128][4]
    static constexpr int S_DC = S_DP + 4 * 128 * 4;            // x2: d(c) of the tile [128][32], 16-byte chunks XOR-swizzled by row
    static constexpr int S_END = S_DC + 2 * 128 * CDIM;
};

// one axis of a point: fp32 position (embedding input) and normalised grid coordinate, as tc_load_point evaluates them
template <int KIND>
__device__ __forceinline__ void tb_point_axis(const PointsDev& q, const DecTask& T, long long ii, int a, float& p32, float& u) {
    if (q.ro == nullptr && q.p32 != nullptr) {
        p32 = q.p32[3 * ii + a];
        u = __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(p32, (float)T.lo[a]), (float)T.ext[a]), 2.f), 1.0f);
    } else {
        double p;
        if (q.ro) {
            const long long r = ii / q.S;
            p = __dadd_rn((double)q.ro[r * q.so + a], __dmul_rn((double)q.rd[r * q.sd + a], q.z[ii]));
        } else {
            p = q.p64[3 * ii + a];
        }
        p32 = (float)p;
        u = (float)__dsub_rn(__dmul_rn(__ddiv_rn(__dsub_rn(p, T.lo[a]), T.ext[a]), 2.0), 1.0);
    }
}
// per point in shared memory: [ix iy iz | x0 y0 z0 | mult flags x y z] (one axis written by each of three parts)
__device__ __forceinline__ Tri tb_read_tri(const float* e, const GridDev& G, int& mbits) {
    Tri t;
    t.ix = e[0]; t.iy = e[1]; t.iz = e[2];
    t.x0 = e[3]; t.y0 = e[4]; t.z0 = e[5];
    mbits = (e[6] != 0.f ? 1 : 0) | (e[7] != 0.f ? 2 : 0) | (e[8] != 0.f ? 4 : 0);
    const int x0 = (int)t.x0, y0 = (int)t.y0, z0 = (int)t.z0;      // as tri_setup
    t.base = (z0 * G.Y + y0) * G.X + x0;
    t.sx = (x0 + 1 <= G.X - 1) ? 1 : 0;
    t.sy = (y0 + 1 <= G.Y - 1) ? G.X : 0;
    t.sz = (z0 + 1 <= G.Z - 1) ? G.X * G.Y : 0;
    t.mx = t.my = t.mz = 0.f;
    return t;
}

__device__ __forceinline__ void red_add8(float* addr, const float (&v)[8], float w) {
    red_add_v4(addr, w * v[0], w * v[1], w * v[2], w * v[3]);
    red_add_v4(addr + 4, w * v[4], w * v[5], w * v[6], w * v[7]);
}

template <int KIND>
__device__ void tc_backward_cta(const StageParams& P, const DecTask& T, int task, float* sm) {
    using C = DecCfg<KIND>;
    using TT = TcCfg<KIND>;
    using S = TbSmem<KIND>;
    using RC = RecCfg<KIND>;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __shared__ long long s_tile[2];
    __shared__ long long s_cmd[2];                             // tile whose image is in buffer b, or -1: no more tiles
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quad = warp & 3, part = warp >> 2;
    const int row = 32 * quad + lane;
    const float* blob = T.packed + C::TOTAL;
    const bool want_grid = (P.flags & NSB_GRAD_GRID) && T.own.grad != nullptr;
    const bool want_point = (P.flags & NSB_GRAD_POINT) && T.d_point != nullptr;
    float* const rec_base = T.dw_scratch;                      // weight-gradient records (X sections deposited by the forward)
    const bool need_de = want_point || rec_base != nullptr;
    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) {
        tc::mbar_init(&bar, 1);
        tc::mbar_init_fence();
        s_tile[0] = (long long)atomicAdd(P.ctr + task, 1u);
        s_tile[1] = (long long)atomicAdd(P.ctr + task, 1u);
    }
    if (tid < EMB_PAD)
        reinterpret_cast<float4*>(sm + S::S_BV)[tid] =
            make_float4(__ldg(blob + TT::T_B + tid), __ldg(blob + TT::T_B + EMB_PAD + tid),
                        __ldg(blob + TT::T_B + 2 * EMB_PAD + tid), 0.f);
    // the d(e) blocks are the last 2 x 3072 floats: skipped when nobody needs d(e)
    const int w_floats = need_de ? TT::W_BWD : TT::T_BE0 - TT::T_FWD_END;
#pragma unroll 4
    for (int i = tid; i < w_floats / 4; i += TCB_THREADS) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(blob + TT::T_FWD_END) + i);
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
    pdl_wait();                 // weights, TMEM and counters are set up; everything else comes from the kernels before
    const uint32_t tbase = tmem_slot;
    const uint32_t lane_base = tc::tmem_addr(tbase, 32 * quad, 0);
    const float4* Bv = reinterpret_cast<const float4*>(sm + S::S_BV);
    float4* const s_pt0 = reinterpret_cast<float4*>(sm + S::S_PT);
    float* const s_tri0 = sm + S::S_TRI;
    float4* s_dp = reinterpret_cast<float4*>(sm + S::S_DP);
    float4* const s_dc0 = reinterpret_cast<float4*>(sm + S::S_DC);
    // ---- the scatter warps -------------------------------------------------------------------------------------
    if (warp >= TC_THREADS / 32) {
        if (want_grid) {
            const int sub = lane & 7, sw = warp - TC_THREADS / 32;
            for (int b = 0;; b ^= 1) {
                nbar_sync(BAR_FULL + b, TCB_THREADS);              // image b is complete
                if (s_cmd[b] < 0) break;
                const float4* s_pt = s_pt0 + b * 128;
                const float* s_tri = s_tri0 + b * 128 * TRI_WORDS;
                const float4* s_dc = s_dc0 + b * 128 * 8;
#pragma unroll 1
                for (int r = 0; r < 128 / (4 * TCB_SCATTER_WARPS); ++r) {
                    const int q = (r * TCB_SCATTER_WARPS + sw) * 4 + (lane >> 3);
                    if (s_pt[q].w == 0.f) continue;                     // beyond the batch
                    int mb;
                    const Tri t = tb_read_tri(s_tri + q * TRI_WORDS, T.own, mb);
                    const float4 v = s_dc[q * 8 + (sub ^ (q & 7))];
                    // a point whose d(c) is exactly zero (behind the surface the transmittance underflows) adds nothing:
                    // no reductions, and its lines do not travel in the exchange.  Vote of the point's 8 lanes.
                    const bool nz = v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f;
                    if (!__any_sync(0xffu << (lane & 24), nz)) continue;
                    float* g0 = T.own.grad + 4 * sub;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float w = tri_weight(t, k);
                        float* line = g0 + (size_t)tri_offset(t, k) * CDIM;
                        if (!(P.flags & 0x80000000u))      // NSB_DEBUG_NO_SCATTER (timing experiments only)
                            red_add_v4(line, w * v.x, w * v.y, w * v.z, w * v.w);
                    }
                    if (P.pts.marks) {      // ray-sharded mapping: these 8 lines travel in the exchange.  Lane sub marks corner
                                            // sub with a plain one-byte store (idempotent: no atomics, nothing waits for memory)
                        const size_t d = (size_t)(reinterpret_cast<const char*>(T.own.grad + (size_t)tri_offset(t, sub) * CDIM) -
                                                  reinterpret_cast<const char*>(P.pts.marks_base));
                        reinterpret_cast<unsigned char*>(P.pts.marks)[128 + (d >> 7)] = 1;
                    }
                }
                nbar_arrive(BAR_EMPTY + b, TCB_THREADS);           // image b may be overwritten
            }
        }
        tc::fence_before_sync();
        __syncthreads();
        return;
    }
    // ---- the chain threads -------------------------------------------------------------------------------------
    int nlive = 0;                                                 // live tiles so far: image buffer = nlive & 1
    const uint32_t whi = tc::smem_u32(sm + S::S_WHI), wlo = tc::smem_u32(sm + S::S_WLO);
    auto wdesc = [&](uint32_t base, int off_floats, int R) {
        return tc::kmajor_desc(base + 4u * (uint32_t)(off_floats - TT::T_FWD_END), R, 0);
    };
    constexpr uint32_t ID64 = tc::idesc_tf32(128, 64), ID96 = tc::idesc_tf32(128, 96);
    constexpr int EPP = EMB_PAD / TC_PARTS, HPP = HID / TC_PARTS;
    const int hc = HPP * part;
    uint32_t par = 0;
    const long long ntiles = (P.pts.P + 127) / 128;
    long long tile = s_tile[0];
    int slot = 0;
#ifdef NSB_TC_PROFILE
    long long tq[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
#define TQB(k) { const long long tn_ = clock64(); tq[k] += tn_ - tlast; tlast = tn_; }
#else
#define TQB(k)
#endif
    TQB(0)     // set-up
    // incoming gradient d(out) of my point (every part reads it: 16 B, L1 serves three of four), fetched one tile ahead
    float4 dv_next = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tile * 128 + row < P.pts.P) dv_next = __ldg(reinterpret_cast<const float4*>(P.d_raw) + tile * 128 + row);
    while (tile < ntiles) {
        const long long nxt = s_tile[slot ^ 1];
        const long long i = tile * 128 + row;
        const bool ok = i < P.pts.P;
        const float4 dv = dv_next;
        dv_next = make_float4(0.f, 0.f, 0.f, 0.f);
        if (nxt * 128 + row < P.pts.P) dv_next = __ldg(reinterpret_cast<const float4*>(P.d_raw) + nxt * 128 + row);
        float dout[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if constexpr (C::NOUT == 1) dout[0] = dv.w;
        else { dout[0] = dv.x; dout[1] = dv.y; dout[2] = dv.z; }      // the colour decoder's 4th output is overwritten
        const bool live_pt = dout[0] != 0.f || dout[1] != 0.f || dout[2] != 0.f;
        const unsigned live_bits = __ballot_sync(FULL, live_pt);      // lanes 0..15: sub-tile 2 quad, 16..31: 2 quad + 1
        const bool sub_live = ((lane < 16 ? live_bits & 0xffffu : live_bits >> 16) != 0u);
        const int tile_live = nbar_or(BAR_CHAIN, TC_THREADS, live_bits != 0u);      // also: every chain thread is done with the previous tile
        TQB(1) // tile top: d(out), liveness vote
        if (tid == 0) s_tile[slot] = (long long)atomicAdd(P.ctr + task, 1u);     // the tile after next
        float* rec = nullptr;                                          // my 16-point record (live sub-tiles only)
        if (rec_base && part == 0 && (lane & 15) == 0 && i < (P.pts.P + 15) / 16 * 16)
            T.dw_flags[(i >> 4) - P.tile_begin] = sub_live ? 1 : 0;
        if (rec_base && sub_live) rec = rec_base + (size_t)((i >> 4) - P.tile_begin) * RC::REC;
        if (!tile_live) {
            if (want_point && part == 0 && ok) {
                T.d_point[3 * i] = 0.f; T.d_point[3 * i + 1] = 0.f; T.d_point[3 * i + 2] = 0.f;
            }
            nbar_sync(BAR_CHAIN, TC_THREADS);      // thread 0's refill of s_tile[slot] is read at the top of the next iteration
            tile = nxt;
            slot ^= 1;
            continue;
        }
        // ---- prologue: d(out) as the first A operand, the point and its voxel addressing -----------------------
        const int r16 = row & 15;
        const int buf = nlive & 1;
        float4* s_pt = s_pt0 + buf * 128;
        float* s_tri = s_tri0 + buf * 128 * TRI_WORDS;
        float4* s_dc = s_dc0 + buf * 128 * 8;
        if (want_grid && nlive >= 2) nbar_sync(BAR_EMPTY + buf, TCB_THREADS);      // the scatter warps are done with this image
        ++nlive;
        if (part == 0) st_split8(lane_base, TB_DAHI, TB_DALO, dout);
        if (part < 3) {          // parts 0..2: one axis of the point each (the fp64 evaluation is the slow part)
            const long long ii = ok ? i : P.pts.P - 1;
            float p32a, ua, ic, f0, mult;
            if (P.pts.pu) { p32a = P.pts.pu[8 * ii + part]; ua = P.pts.pu[8 * ii + 4 + part]; }      // left by the forward
            else tb_point_axis<KIND>(P.pts, T, ii, part, p32a, ua);
            tri_axis(ua, part == 0 ? T.own.X : (part == 1 ? T.own.Y : T.own.Z), ic, f0, mult);
            float* e = s_tri + row * TRI_WORDS;
            e[part] = ic; e[3 + part] = f0; e[6 + part] = mult;
            reinterpret_cast<float*>(s_pt + row)[part] = p32a;
        } else {
            reinterpret_cast<float*>(s_pt + row)[3] = ok ? 1.f : 0.f;
        }
        if (rec) {
            float o8[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o8[j] = part == 0 ? dout[j] : 0.f;
            st_global_v8(rec + RC::R_DO + r16 * HID + rec_sw(r16, hc), o8);
        }
        // ReLU sign bytes of my (point, part): layers 0..3 in .x, layer 4 in .y (written by the forward)
        uint2 rb = make_uint2(0u, 0u);
        if (ok) rb = __ldg(reinterpret_cast<const uint2*>(T.relu) + i * 4 + part);
        tc::tmem_st_wait();
        tc::fence_before_sync();
        TQB(2) // prologue: first operand, point axis
        nbar_sync(BAR_CHAIN, TC_THREADS);
        TQB(3) // prologue barrier
        if (rec && part == 1) {      // record section p [16][8]
            const float4 pp = s_pt[row];
            const float pr[8] = {pp.x, pp.y, pp.z, 0.f, 0.f, 0.f, 0.f, 0.f};
            st_global_v8(rec + RC::R_P + r16 * 8, pr);
        }
        if (warp == 0 && tc::elect_one()) {
            tc::fence_after_sync();
            mma3_ts(tbase + tb_group(5), tbase + TB_DAHI, tbase + TB_DALO, wdesc(whi, TT::T_BG5, 64), wdesc(wlo, TT::T_BG5, 64),
                    64, 1, ID64, false);
            tc::mma_commit(&bar);
        }
        // ---- the chain: round i applies the ReLU of layer i to d(h_{i+1}) = first half of group i + 1 ---------------
#pragma unroll 1
        for (int i = 4; i >= 0; --i) {
            TQB(4) // issue (warp 0) / nothing
            tc::mbar_wait(&bar, par);
            par ^= 1;
            tc::fence_after_sync();
            TQB(5) // wait for the layer's MMA group
            float a[8];
            tc::tmem_ld8(lane_base + tb_group(i + 1) + hc, a);
            tc::tmem_ld_wait();
            if (rec) st_global_v8(rec + RC::R_DH + i * 16 * HID + r16 * HID + rec_sw(r16, hc), a);      // d(h_{i+1})
            const unsigned bits = i < 4 ? (rb.x >> (8 * i)) : rb.y;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (!((bits >> j) & 1u)) a[j] = 0.f;
            if (rec) st_global_v8(rec + RC::da(i) + r16 * HID + rec_sw(r16, hc), a);      // d(a_i)
            const bool issue = i >= 1 || need_de;                      // uniform over the CTA
            if (issue) st_split8(lane_base, TB_DAHI + hc, TB_DALO + hc, a);
            tc::tmem_st_wait();
            tc::fence_before_sync();
            TQB(6) // layer epilogue: ld, mask, records, st
            nbar_sync(BAR_CHAIN, TC_THREADS);
            TQB(7) // layer barrier
            if (issue && warp == 0 && tc::elect_one()) {
                tc::fence_after_sync();
                if (i >= 1) {
                    const int off = i == 1 ? TT::T_BG1 : (i == 2 ? TT::T_BG2 : (i == 3 ? TT::T_BG3 : TT::T_BG4));
                    mma3_ts(tbase + tb_group(i), tbase + TB_DAHI, tbase + TB_DALO, wdesc(whi, off, 64), wdesc(wlo, off, 64),
                            64, HID / 8, ID64, false);
                }
                if (need_de && (i == 3 || i == 0)) {                   // e enters layers 0 and 3 (the skip [e, h])
                    const int off = i == 3 ? TT::T_BE3 : TT::T_BE0;
                    mma3_ts(tbase + TB_DE, tbase + TB_DAHI, tbase + TB_DALO, wdesc(whi, off, EMB_PAD), wdesc(wlo, off, EMB_PAD),
                            EMB_PAD, HID / 8, ID96, i == 0);
                }
                tc::mma_commit(&bar);
            }
        }
        if (need_de) {              // the d(e) part of layer 0 is the last group in flight
            tc::mbar_wait(&bar, par);
            par ^= 1;
            tc::fence_after_sync();
        }
        TQB(8) // last issue + wait for d(e)
        // ---- d(c): the five parts, my 8 channels -------------------------------------------------------------
        float dc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int g5 = 5; g5 >= 1; --g5) {
            float t8[8];
            tc::tmem_ld8(lane_base + tb_group(g5) + 32 + hc, t8);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) dc[j] += t8[j];
        }
        float dp[3] = {0.f, 0.f, 0.f};
        if (want_grid) {     // my 8 channels as two 16-byte chunks of the tile's d(c) image (chunk index XOR row: no bank conflicts)
            s_dc[row * 8 + ((2 * part) ^ (row & 7))] = make_float4(dc[0], dc[1], dc[2], dc[3]);
            s_dc[row * 8 + ((2 * part + 1) ^ (row & 7))] = make_float4(dc[4], dc[5], dc[6], dc[7]);
        }
        if (ok && want_point) {
            int mbits;
            const Tri t = tb_read_tri(s_tri + row * TRI_WORDS, T.own, mbits);
            float gx = 0.f, gy = 0.f, gz = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
                const float wx = dx ? (t.ix - t.x0) : ((t.x0 + 1.f) - t.ix);
                const float wy = dy ? (t.iy - t.y0) : ((t.y0 + 1.f) - t.iy);
                const float wz = dz ? (t.iz - t.z0) : ((t.z0 + 1.f) - t.iz);
                const size_t off = (size_t)tri_offset(t, k) * CDIM + hc;
                {                      // ATen grid_sampler_3d_backward: d(ix) = sum_corners +-(other two weights) <g, v>
                    const float4* gp = reinterpret_cast<const float4*>(T.own.data + off);
                    const float4 v0 = __ldg(gp), v1 = __ldg(gp + 1);
                    const float s8 = v0.x * dc[0] + v0.y * dc[1] + v0.z * dc[2] + v0.w * dc[3] + v1.x * dc[4] + v1.y * dc[5] +
                                     v1.z * dc[6] + v1.w * dc[7];
                    gx += (dx ? s8 : -s8) * wy * wz;
                    gy += (dy ? s8 : -s8) * wx * wz;
                    gz += (dz ? s8 : -s8) * wx * wy;
                }
            }
            // border rule: no coordinate gradient on a clamped axis
            dp[0] = gx * ((mbits & 1) ? (float)(T.own.X - 1) / 2 : 0.f) * (float)(2.0 / T.ext[0]);
            dp[1] = gy * ((mbits & 2) ? (float)(T.own.Y - 1) / 2 : 0.f) * (float)(2.0 / T.ext[1]);
            dp[2] = gz * ((mbits & 4) ? (float)(T.own.Z - 1) / 2 : 0.f) * (float)(2.0 / T.ext[2]);
        }
        TQB(9) // d(c) sum, shared-memory image, coordinate gradient
        // ---- hand the image (d(c), voxel addressing, validity) of this tile to the scatter warps -----------------------
        if (want_grid) {
            if (tid == 0) s_cmd[buf] = tile;
            __threadfence_block();
            nbar_arrive(BAR_FULL + buf, TCB_THREADS);
        }
        TQB(10) // hand-over to the scatter warps
        // ---- d(e) cos(p B): -> d(point) through B, and the record section d(B) is computed from -----------------
        if (need_de) {
            const float4 pp = s_pt[row];
#pragma unroll 1
            for (int j = 0; j < EPP / 8; ++j) {
                const int c0 = EPP * part + 8 * j;
                float d8[8];
                tc::tmem_ld8(lane_base + TB_DE + c0, d8);
                tc::tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 b = Bv[c0 + q];
                    const float v = d8[q] * cos_embed(fmaf(pp.z, b.z, fmaf(pp.y, b.y, pp.x * b.x)));
                    d8[q] = v;
                    dp[0] = fmaf(v, b.x, dp[0]);
                    dp[1] = fmaf(v, b.y, dp[1]);
                    dp[2] = fmaf(v, b.z, dp[2]);
                }
                if (rec) st_global_v8(rec + RC::R_DE + rec_e(r16, c0), d8);
            }
        }
        if (want_point) {              // the four parts of a point add up through shared memory
            s_dp[part * 128 + row] = make_float4(dp[0], dp[1], dp[2], 0.f);
            nbar_sync(BAR_CHAIN, TC_THREADS);
            if (part == 0 && ok) {
                const float4 a0 = s_dp[row], a1 = s_dp[128 + row], a2 = s_dp[256 + row], a3 = s_dp[384 + row];
                T.d_point[3 * i] = (a0.x + a1.x) + (a2.x + a3.x);
                T.d_point[3 * i + 1] = (a0.y + a1.y) + (a2.y + a3.y);
                T.d_point[3 * i + 2] = (a0.z + a1.z) + (a2.z + a3.z);
            }
        }
        tc::fence_before_sync();       // my tensor-memory reads are complete before the next tile's MMAs (after its barrier)
        TQB(11) // d(e) epilogue, d(point)
        tile = nxt;
        slot ^= 1;
    }
    if (want_grid) {               // no more tiles: release the scatter warps (they wait on the next image)
        const int buf = nlive & 1;
        if (nlive >= 2) nbar_sync(BAR_EMPTY + buf, TCB_THREADS);
        if (tid == 0) s_cmd[buf] = -1;
        __threadfence_block();
        nbar_arrive(BAR_FULL + buf, TCB_THREADS);
    }
#ifdef NSB_TC_PROFILE
    if ((tid == 0 || tid == 200) && (int)blockIdx.x == T.cta_begin && P.prof)
        for (int k = 0; k < 12; ++k) P.prof[(task * 2 + (tid != 0)) * 12 + k] = tq[k];
#endif
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tbase);
}

__global__ void __launch_bounds__(TCB_THREADS, 1)
k_decoders_backward_tc(const __grid_constant__ StageParams P) {
    extern __shared__ __align__(16) float sm[];
    pdl_launch_dependents();     // k_dweight / k_ray_grad_reduce may be scheduled as CTAs of this kernel retire
    int k = 0;
    while (k + 1 < P.ntask && (int)blockIdx.x >= P.task[k + 1].cta_begin) ++k;
    const DecTask& T = P.task[k];
    if (blockIdx.x == 0 && threadIdx.x == 0 && P.pts.marks) P.pts.marks[0] = 1u;      // this launch marks what it scatters
    switch (T.kind) {
        case KIND_MIDDLE: tc_backward_cta<KIND_MIDDLE>(P, T, k, sm); break;
        case KIND_FINE: tc_backward_cta<KIND_FINE>(P, T, k, sm); break;
        case KIND_COLOR: tc_backward_cta<KIND_COLOR>(P, T, k, sm); break;
        default: break;
    }
    release_counters(P.ctr);
}

}  // namespace nsb
