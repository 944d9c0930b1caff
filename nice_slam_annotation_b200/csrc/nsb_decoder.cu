// K2 + K3 (+ their part of K5): trilinear grid gather, Gaussian-Fourier embedding and the
// NICE decoder MLPs, forward and backward, as ONE persistent kernel per stage.
//
// Replaces, per point: MLP.sample_grid_feature / F.grid_sample (decoder.py:179-188,273-283),
// GaussianFourierFeatureTransform (decoder.py:26-30), MLP.forward / MLP_no_xyz.forward
// (decoder.py:190-219,285-297) and the stage mixing of NICE.forward (decoder.py:336-370),
// plus their autograd backward.
//
// Design (B200, sm_100a):
//  * a CTA is bound to ONE decoder of the stage (the stage's decoders run side by side in one
//    launch on disjoint CTA ranges); it stages that decoder's packed weights in shared memory
//    once and then loops over 16-point tiles, one tile per warp per iteration;
//  * a warp owns 16 consecutive points.  Lane (g,t) (g = lane/4, t = lane%4) owns points g
//    and g+8 and, of each gathered 32-channel voxel line, channels 8t..8t+7: every corner
//    fetch is two 128-bit loads per lane, four lanes cover the 128-byte line, and the result
//    lands DIRECTLY in mma A-fragment registers (slot map "gather", nsb_layout.h);
//  * all layers run on the tensor cores as mma.sync.m16n8k8 tf32 with the 3xTF32 error
//    compensation (fp32-class accuracy: parity with the reference's fp32 SGEMM needs it,
//    SURVEY.md 7.3-1); the accumulator of one layer is re-used as the A operand of the next
//    by permuting the k index of the packed weights ("chain" slot map), so activations never
//    leave registers;
//  * the Fourier embedding is evaluated in plain fp32 FMA + sinf/cosf (tf32 would destroy
//    the phase, |p.B| reaches hundreds of radians);
//  * backward recomputes the forward per tile (keeping only the ReLU sign bits), runs the
//    transposed chain, turns d(e) into d(p) on the fly, scatters d(c) to the grid gradient
//    with 128-bit vector reductions (red.global.add.v4.f32) and re-gathers the corners for
//    the coordinate gradient.  Tiles whose incoming gradient is exactly zero are skipped.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "nsb_common.cuh"

namespace nsb {


struct GridDev {
    const float* data;
    float* grad;
    int X, Y, Z;
};

struct DecTask {
    int kind;
    int cta_begin, cta_count;
    const float* packed;
    GridDev own;      // grid whose feature feeds the decoder (gradient flows)
    GridDev aux;      // FINE: the middle grid, sampled without gradient (decoder.py:199-201)
    double lo[3], ext[3];
    float* out;       // forward: [P] (1 output) or [P,4]
    float* d_point;   // backward: [P,3] plane or NULL
    float* dw_scratch;         // backward: weight-gradient records of this launch's tiles, or NULL
    unsigned char* dw_flags;   // one byte per tile: record written
    unsigned* relu;            // [P][8] ReLU sign words of this decoder (forward writes, backward reads) or NULL
    int x_saved;               // backward: the forward deposited the X sections (e, c, h1..h5) of the records
};

struct StageParams {
    PointsDev pts;
    DecTask task[3];
    int ntask;
    const float* d_raw;   // backward: [P,4]
    unsigned flags;
    long long tile_begin, tile_end;   // tiles of this launch (backward is chunked for d(W))
    unsigned* ctr;                    // [4]: next tile of task 0..2 (relative to tile_begin), CTAs done
    long long* prof;                  // NSB_TC_PROFILE builds: per-phase cycle sums of one CTA per task
};

// Tiles are handed out dynamically: a warp takes the next tile of its CTA's decoder with one
// atomic (fetched one tile ahead, so the round trip is hidden).  Work per tile is uneven --
// the backward skips tiles whose incoming gradient is zero -- and a static round-robin left
// the slowest CTA running for 2x the average (ncu, profiles/r01b).
__device__ __forceinline__ long long next_tile(unsigned* ctr, int lane) {
    unsigned v = 0;
    if (lane == 0) v = atomicAdd(ctr, 1u);
    return (long long)__shfl_sync(FULL, v, 0);
}
// last CTA out re-arms the counters for the next launch that uses this slot
__device__ __forceinline__ void release_counters(unsigned* ctr) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(ctr + 3, 1u) == gridDim.x - 1) {
            ctr[0] = 0; ctr[1] = 0; ctr[2] = 0;
            __threadfence();
            ctr[3] = 0;
        }
    }
}

// ---- per-thread view of its two points -------------------------------------------------------
struct Rows {
    float p[2][3];   // world point, fp32 (embedding input: decoder.py:204 p.float())
    float u[2][3];   // normalised to [-1,1] (common.py:280-295), fp32 as grid_sample sees it
    bool valid[2];
    long long idx[2];
};

template <int KIND>
__device__ __forceinline__ void load_rows(const PointsDev& q, const DecTask& T, long long tile,
                                          int lane, Rows& R) {
    long long i = tile * 16 + (lane & 15);
    bool ok = i < q.P;
    long long ii = ok ? i : q.P - 1;
    float p32[3], u[3];
    if (q.ro == nullptr && q.p32 != nullptr) {
        // fp32 points (Mesher / regulation callers): normalised in fp32
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            p32[a] = q.p32[3 * ii + a];
            u[a] = __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(p32[a], (float)T.lo[a]), (float)T.ext[a]), 2.f), 1.0f);
        }
    } else {
        double p[3];
        if (q.ro) {
            ray_point(q, ii, p);
        } else {
            p[0] = q.p64[3 * ii]; p[1] = q.p64[3 * ii + 1]; p[2] = q.p64[3 * ii + 2];
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            p32[a] = (float)p[a];
            u[a] = (float)__dsub_rn(__dmul_rn(__ddiv_rn(__dsub_rn(p[a], T.lo[a]), T.ext[a]), 2.0), 1.0);
        }
    }
    const int g = lane >> 2;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int src = g + 8 * r;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            R.p[r][a] = __shfl_sync(FULL, p32[a], src);
            R.u[r][a] = __shfl_sync(FULL, u[a], src);
        }
        R.valid[r] = __shfl_sync(FULL, (int)ok, src) != 0;
        R.idx[r] = tile * 16 + src;
    }
}

// ---- trilinear, ATen grid_sampler_3d semantics (bilinear, border, align_corners=True) ----------
struct Tri {
    float ix, iy, iz;      // clipped un-normalised coordinate
    float x0, y0, z0;      // floor
    int base, sx, sy, sz;  // voxel index of the (x0,y0,z0) corner and +1 steps (0 at the border)
    float mx, my, mz;      // d(ix)/d(u) with the border rule (0 when clipped)
};

__device__ __forceinline__ void tri_axis(float u, int S, float& ic, float& f0, float& mult) {
    float v = ((u + 1.f) / 2) * (float)(S - 1);
    mult = (v <= 0.f || v >= (float)(S - 1)) ? 0.f : (float)(S - 1) / 2;
    ic = fminf((float)(S - 1), fmaxf(v, 0.f));
    f0 = floorf(ic);
}

__device__ __forceinline__ Tri tri_setup(const float (&u)[3], const GridDev& G) {
    Tri t;
    tri_axis(u[0], G.X, t.ix, t.x0, t.mx);
    tri_axis(u[1], G.Y, t.iy, t.y0, t.my);
    tri_axis(u[2], G.Z, t.iz, t.z0, t.mz);
    int x0 = (int)t.x0, y0 = (int)t.y0, z0 = (int)t.z0;
    t.base = (z0 * G.Y + y0) * G.X + x0;
    t.sx = (x0 + 1 <= G.X - 1) ? 1 : 0;
    t.sy = (y0 + 1 <= G.Y - 1) ? G.X : 0;
    t.sz = (z0 + 1 <= G.Z - 1) ? G.X * G.Y : 0;
    return t;
}

__device__ __forceinline__ float tri_weight(const Tri& T, int k) {
    float wx = (k & 1) ? (T.ix - T.x0) : ((T.x0 + 1.f) - T.ix);
    float wy = (k & 2) ? (T.iy - T.y0) : ((T.y0 + 1.f) - T.iy);
    float wz = (k & 4) ? (T.iz - T.z0) : ((T.z0 + 1.f) - T.iz);
    return (wx * wy) * wz;
}

__device__ __forceinline__ int tri_offset(const Tri& T, int k) {
    return T.base + ((k & 1) ? T.sx : 0) + ((k & 2) ? T.sy : 0) + ((k & 4) ? T.sz : 0);
}

// 8 channels (8t .. 8t+7) of the trilinear feature of one point
__device__ __forceinline__ void gather8(const float* __restrict__ grid, const Tri& T, int t,
                                        float (&f)[8]) {
    float4 v[8][2];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4* p = reinterpret_cast<const float4*>(grid + (size_t)tri_offset(T, k) * CDIM + 8 * t);
        v[k][0] = __ldg(p);
        v[k][1] = __ldg(p + 1);
    }
#pragma unroll
    for (int c = 0; c < 8; ++c) f[c] = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float w = tri_weight(T, k);
        f[0] = fmaf(v[k][0].x, w, f[0]); f[1] = fmaf(v[k][0].y, w, f[1]);
        f[2] = fmaf(v[k][0].z, w, f[2]); f[3] = fmaf(v[k][0].w, w, f[3]);
        f[4] = fmaf(v[k][1].x, w, f[4]); f[5] = fmaf(v[k][1].y, w, f[5]);
        f[6] = fmaf(v[k][1].z, w, f[6]); f[7] = fmaf(v[k][1].w, w, f[7]);
    }
}

// features of the thread's two points -> 4 A-fragment slabs ("gather" slot map).  The row
// loop is rolled (one copy of the 16-load body in the instruction stream).
__device__ __forceinline__ void gather_level(const GridDev& G, const Rows& R, int t,
                                             float (*c)[4]) {
#pragma unroll 1
    for (int r = 0; r < 2; ++r) {
        float f[8];
        const float u[3] = {r ? R.u[1][0] : R.u[0][0], r ? R.u[1][1] : R.u[0][1], r ? R.u[1][2] : R.u[0][2]};
        Tri tr = tri_setup(u, G);
        gather8(G.data, tr, t, f);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (r == 0) { c[j][0] = f[2 * j]; c[j][2] = f[2 * j + 1]; }
            else { c[j][1] = f[2 * j]; c[j][3] = f[2 * j + 1]; }
        }
    }
}

// ---- Fourier embedding -------------------------------------------------------------------------
// e = sin(p @ B) (decoder.py:26-30).  The 16 x 96 tile of a warp lives in a per-warp shared
// buffer, each lane writing exactly the A-fragment elements it later reads (slot map
// "natural": rows g / g+8, features 8j+t / 8j+t+4), XOR-swizzled so that both the stores and
// the fragment loads are bank-conflict free.  Keeping e out of registers lets the 12-slab
// GEMMs of layers 0 and 3 be ROLLED loops: the kernel body stays small enough for the
// instruction cache (the fully unrolled first version stalled 3-5 cycles per issue on
// instruction fetch) and 48 registers per thread are freed for a third warp per scheduler.
constexpr int E_BUF = 16 * EMB_PAD;      // floats per warp

__device__ __forceinline__ int esw(int row, int col) { return row * EMB_PAD + (col ^ ((row & 7) << 2)); }

__device__ __forceinline__ float phase(const float (&p)[3], const float* Bs, int f) {
    return fmaf(p[2], Bs[2 * EMB_PAD + f], fmaf(p[1], Bs[EMB_PAD + f], p[0] * Bs[f]));
}

__device__ __forceinline__ void embed_to_smem(const Rows& R, const float* Bs, int lane, float* es) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll 1
    for (int j = 0; j < E_SLABS; ++j) {
        const int f0 = 8 * j + t, f1 = f0 + 4;
        es[esw(g, f0)] = sin_cw(phase(R.p[0], Bs, f0));
        es[esw(g + 8, f0)] = sin_cw(phase(R.p[1], Bs, f0));
        es[esw(g, f1)] = sin_cw(phase(R.p[0], Bs, f1));
        es[esw(g + 8, f1)] = sin_cw(phase(R.p[1], Bs, f1));
    }
}

// acc[4][4] += e(16 x 96) * B, A fragments streamed from the warp's e buffer (rolled)
__device__ __forceinline__ void egemm(float (&acc)[4][4], const float* es,
                                      const float* __restrict__ Bs, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const float4* B4 = reinterpret_cast<const float4*>(Bs);
#pragma unroll 2
    for (int j = 0; j < E_SLABS; ++j) {
        const int f0 = 8 * j + t, f1 = f0 + 4;
        uint32_t ah[4], al[4];
        split_tf32(es[esw(g, f0)], ah[0], al[0]);
        split_tf32(es[esw(g + 8, f0)], ah[1], al[1]);
        split_tf32(es[esw(g, f1)], ah[2], al[2]);
        split_tf32(es[esw(g + 8, f1)], ah[3], al[3]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float4 b = B4[(j * 2 + h) * 32 + lane];
            uint32_t bh[4], bl[4];
            split_tf32(b.x, bh[0], bl[0]);
            split_tf32(b.y, bh[1], bl[1]);
            split_tf32(b.z, bh[2], bl[2]);
            split_tf32(b.w, bh[3], bl[3]);
            mma_tf32(acc[2 * h], al, bh[0], bh[1]);
            mma_tf32(acc[2 * h], ah, bl[0], bl[1]);
            mma_tf32(acc[2 * h], ah, bh[0], bh[1]);
            mma_tf32(acc[2 * h + 1], al, bh[2], bh[3]);
            mma_tf32(acc[2 * h + 1], ah, bl[2], bl[3]);
            mma_tf32(acc[2 * h + 1], ah, bh[2], bh[3]);
        }
    }
}

// ---- MLP pieces ------------------------------------------------------------------------------
__device__ __forceinline__ void bias_init(float (&acc)[4][4], const float* b, int t) {
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        float2 v = *reinterpret_cast<const float2*>(b + 8 * n + 2 * t);
        acc[n][0] = v.x; acc[n][1] = v.y; acc[n][2] = v.x; acc[n][3] = v.y;
    }
}
__device__ __forceinline__ void bias_add(float (&acc)[4][4], const float* b, int t) {
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        float2 v = *reinterpret_cast<const float2*>(b + 8 * n + 2 * t);
        acc[n][0] += v.x; acc[n][1] += v.y; acc[n][2] += v.x; acc[n][3] += v.y;
    }
}
// ReLU in place; returns the 16 sign bits (bit n*4+q)
__device__ __forceinline__ uint32_t relu_(float (&acc)[4][4]) {
    uint32_t bits = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            bool pos = acc[n][q] > 0.f;
            bits |= (pos ? 1u : 0u) << (n * 4 + q);
            acc[n][q] = pos ? acc[n][q] : 0.f;
        }
    return bits;
}
__device__ __forceinline__ void mask_(float (&acc)[4][4], uint32_t m) {
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (!((m >> (n * 4 + q)) & 1u)) acc[n][q] = 0.f;
}
__device__ __forceinline__ void zero_(float (&acc)[4][4]) {
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[n][q] = 0.f;
}

// ReLU sign bits of the five layers, 16 bits each, in three registers
struct Masks {
    uint32_t m01, m23, m4;
    __device__ __forceinline__ void set(int i, uint32_t bits) {
        if (i == 0) m01 = bits;
        else if (i == 1) m01 |= bits << 16;
        else if (i == 2) m23 = bits;
        else if (i == 3) m23 |= bits << 16;
        else m4 = bits;
    }
    __device__ __forceinline__ uint32_t get(int i) const {
        const uint32_t w = i < 2 ? m01 : (i < 4 ? m23 : m4);
        return (i & 1) ? (w >> 16) : (w & 0xffffu);
    }
};

// blob offsets of layer i as functions (runtime i: the layer loops are rolled)
template <int KIND>
struct Off {
    using C = DecCfg<KIND>;
    static constexpr int XS = C::EMBED ? E_SLABS : 4;      // slabs of the input-side part of layers 0 / 3
    __device__ __forceinline__ static int X(int i) { return i == 0 ? C::OFF_L0 : C::OFF_L3; }
    __device__ __forceinline__ static int H(int i) {
        return i == 1 ? C::OFF_L1 : (i == 2 ? C::OFF_L2 : (i == 3 ? C::OFF_L3 + XS * C::SZ_SLAB : C::OFF_L4));
    }
    __device__ __forceinline__ static int Cc(int i) {
        return i == 0 ? C::OFF_C0 : (i == 1 ? C::OFF_C1 : (i == 2 ? C::OFF_C2 : (i == 3 ? C::OFF_C3 : C::OFF_C4)));
    }
    __device__ __forceinline__ static int TL(int i) {
        return i == 1 ? C::OFF_TL1 : (i == 2 ? C::OFF_TL2 : (i == 3 ? C::OFF_TL3 : C::OFF_TL4));
    }
    __device__ __forceinline__ static int TX(int i) { return i == 0 ? C::OFF_TL0 : C::OFF_TL3X; }
    __device__ __forceinline__ static int TC(int i) {
        return i == 0 ? C::OFF_TC0 : (i == 1 ? C::OFF_TC1 : (i == 2 ? C::OFF_TC2 : (i == 3 ? C::OFF_TC3 : C::OFF_TC4)));
    }
};

// C fragments [4 n-tiles] of a 16 x 32 tile -> [16][32] record section (rec_sw columns)
__device__ __forceinline__ void store_c32(float* dst, const float (&acc)[4][4], int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const int col = rec_sw(g, 8 * n + 2 * t);          // (g + 8) & 3 == g & 3
        *reinterpret_cast<float2*>(dst + g * HID + col) = make_float2(acc[n][0], acc[n][1]);
        *reinterpret_cast<float2*>(dst + (g + 8) * HID + col) = make_float2(acc[n][2], acc[n][3]);
    }
}

// MLP.forward (EMBED) / MLP_no_xyz.forward (COARSE) for one 16-point tile.  es: the warp's
// embedding buffer (unused for COARSE), c: feature slabs.  out: C fragment of the padded
// 8-column output tile (columns 2t, 2t+1 for rows g, g+8).
template <int KIND>
__device__ __forceinline__ void mlp_forward(const float* sm, const float* es, int lane,
                                            const float (&c)[DecCfg<KIND>::CS][4], float (&out)[4],
                                            Masks& mk, float* rec = nullptr) {
    using C = DecCfg<KIND>;
    using O = Off<KIND>;
    const int t = lane & 3;
    float acc[4][4], h[4][4];
#pragma unroll 1
    for (int i = 0; i < 5; ++i) {
        bias_init(acc, sm + C::OFF_BIAS + 32 * i, t);
        if (i == 0 || i == 3) {                     // input-side part; layer 3 is the skip [x, h]
            if constexpr (C::EMBED) egemm(acc, es, sm + O::X(i), lane);
            else gemm3<4, 4>(acc, c, sm + O::X(i), lane);
        }
        if (i > 0) gemm3<4, 4>(acc, h, sm + O::H(i), lane);
        mk.set(i, relu_(acc));
        if constexpr (C::EMBED) {
            bias_add(acc, sm + C::OFF_BC + 32 * i, t);
            gemm3<C::CS, 4>(acc, c, sm + O::Cc(i), lane);
        }
        chain<4>(h, acc);
        if (rec) store_c32(rec + RecCfg<KIND>::R_H + i * 16 * HID, acc, lane);     // h_{i+1}
    }
    float2 bo = *reinterpret_cast<const float2*>(sm + C::OFF_BO + 2 * t);
    out[0] = bo.x; out[1] = bo.y; out[2] = bo.x; out[3] = bo.y;
    gemm3_n1<4>(out, h, sm + C::OFF_OUT, lane);
}

__device__ __forceinline__ void stage_weights(float* sm, const float* __restrict__ packed, int n) {
    const float4* src = reinterpret_cast<const float4*>(packed);
    float4* dst = reinterpret_cast<float4*>(sm);
    for (int i = threadIdx.x; i < n / 4; i += blockDim.x) dst[i] = __ldg(src + i);
    __syncthreads();
}

// floats [begin, end) of the blob to the same offsets of shared memory (both multiples of 4); no barrier
__device__ __forceinline__ void stage_range(float* sm, const float* __restrict__ packed, int begin, int end) {
    const float4* src = reinterpret_cast<const float4*>(packed);
    float4* dst = reinterpret_cast<float4*>(sm);
    for (int i = begin / 4 + threadIdx.x; i < end / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// inputs of a tile: rows, embedding (to shared), features (to registers)
template <int KIND>
__device__ __forceinline__ void tile_inputs(const PointsDev& q, const DecTask& T, const float* sm,
                                            float* es, long long tile, int lane, Rows& R,
                                            float (&c)[DecCfg<KIND>::CS][4]) {
    using C = DecCfg<KIND>;
    const int t = lane & 3;
    load_rows<KIND>(q, T, tile, lane, R);
    gather_level(T.own, R, t, c);
    if constexpr (KIND == KIND_FINE) gather_level(T.aux, R, t, c + 4);
    if constexpr (C::EMBED) embed_to_smem(R, sm + C::OFF_B, lane, es);
}

template <int KIND>
__device__ void decoder_forward_cta(const StageParams& P, const DecTask& T, int task, float* sm) {
    using C = DecCfg<KIND>;
    stage_weights(sm, T.packed, C::FWD_END);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 3;
    float* es = sm + C::FWD_END + warp * E_BUF;
    const long long ntiles = (P.pts.P + 15) / 16;
    unsigned* ctr = P.ctr + task;
    for (long long tile = next_tile(ctr, lane), nxt; tile < ntiles; tile = nxt) {
        nxt = next_tile(ctr, lane);
        Rows R;
        float c[C::CS][4];
        tile_inputs<KIND>(P.pts, T, sm, es, tile, lane, R, c);
        float out[4];
        Masks mk;
        mlp_forward<KIND>(sm, es, lane, c, out, mk);
        if constexpr (C::NOUT == 1) {
            if (t == 0) {
                if (R.valid[0]) T.out[R.idx[0]] = out[0];
                if (R.valid[1]) T.out[R.idx[1]] = out[2];
            }
        } else {
            if (t < 2) {
                if (R.valid[0]) *reinterpret_cast<float2*>(T.out + 4 * R.idx[0] + 2 * t) = make_float2(out[0], out[1]);
                if (R.valid[1]) *reinterpret_cast<float2*>(T.out + 4 * R.idx[1] + 2 * t) = make_float2(out[2], out[3]);
            }
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1)
k_decoders_forward(const __grid_constant__ StageParams P) {
    extern __shared__ __align__(16) float sm[];
    int k = 0;
    while (k + 1 < P.ntask && (int)blockIdx.x >= P.task[k + 1].cta_begin) ++k;
    const DecTask& T = P.task[k];
    switch (T.kind) {
        case KIND_COARSE: decoder_forward_cta<KIND_COARSE>(P, T, k, sm); break;
        case KIND_MIDDLE: decoder_forward_cta<KIND_MIDDLE>(P, T, k, sm); break;
        case KIND_FINE: decoder_forward_cta<KIND_FINE>(P, T, k, sm); break;
        default: decoder_forward_cta<KIND_COLOR>(P, T, k, sm); break;
    }
    release_counters(P.ctr);
}

// ---- backward -------------------------------------------------------------------------------
// d(a) (A fragments, chain map) -> d(e) = d(a) W^T two n-tiles at a time (rolled over the 6
// n-tile pairs).  e enters layers 3 and 0: the layer-3 call (last == false) parks its part in
// the warp's embedding buffer (e itself is dead by then), the layer-0 call adds its own and
// consumes the sum: e = sin(p.B) so d(p_a) += d(e_f) cos(p.B_f) B[a][f] -- one cosine per
// feature instead of two.
__device__ __forceinline__ void embed_grad(const float (&a)[4][4], const float* __restrict__ Ws,
                                           const Rows& R, const float* Bs, int lane, float (&dp)[2][3],
                                           float* es, bool last, float* des = nullptr) {
    const int t = lane & 3, g = lane >> 2;
    uint32_t ah[4][4], al[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) split_tf32(a[j][i], ah[j][i], al[j][i]);
    const float4* B4 = reinterpret_cast<const float4*>(Ws);
#pragma unroll 1
    for (int h = 0; h < E_SLABS / 2; ++h) {
        float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float4 b = B4[(j * (E_SLABS / 2) + h) * 32 + lane];
            uint32_t bh[4], bl[4];
            split_tf32(b.x, bh[0], bl[0]);
            split_tf32(b.y, bh[1], bl[1]);
            split_tf32(b.z, bh[2], bl[2]);
            split_tf32(b.w, bh[3], bl[3]);
            mma_tf32(d0, al[j], bh[0], bh[1]);
            mma_tf32(d0, ah[j], bl[0], bl[1]);
            mma_tf32(d0, ah[j], bh[0], bh[1]);
            mma_tf32(d1, al[j], bh[2], bh[3]);
            mma_tf32(d1, ah[j], bl[2], bl[3]);
            mma_tf32(d1, ah[j], bh[2], bh[3]);
        }
#pragma unroll
        for (int nn = 0; nn < 2; ++nn)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int r = q >> 1, f = 8 * (2 * h + nn) + 2 * t + (q & 1);
                float* slot = es + esw(g + 8 * r, f);
                const float d = nn ? d1[q] : d0[q];
                if (!last) { *slot = d; continue; }
                const float v = (d + *slot) * cos_cw(phase(R.p[r], Bs, f));
                if (des) des[rec_e(g + 8 * r, f)] = v;   // d(B)[a][f] = sum_points p_a v
                dp[r][0] = fmaf(v, Bs[f], dp[r][0]);
                dp[r][1] = fmaf(v, Bs[EMB_PAD + f], dp[r][1]);
                dp[r][2] = fmaf(v, Bs[2 * EMB_PAD + f], dp[r][2]);
            }
    }
}

// scatter-add d(c) (8 channels of one point) to the 8 corners; optionally the coordinate
// gradient (ATen grid_sampler_3d_backward: d(ix) = sum_corners +-(other two weights) <g, v>)
__device__ __forceinline__ void grid_backward(bool scatter, bool coord, const GridDev& G,
                                              const float (&u)[3], const double* ext, int t,
                                              const float (&dc)[8], float (&dp)[3]) {
    Tri T = tri_setup(u, G);
    float gx = 0.f, gy = 0.f, gz = 0.f;
#pragma unroll 1
    for (int k = 0; k < 8; ++k) {
        const int dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
        const float wx = dx ? (T.ix - T.x0) : ((T.x0 + 1.f) - T.ix);
        const float wy = dy ? (T.iy - T.y0) : ((T.y0 + 1.f) - T.iy);
        const float wz = dz ? (T.iz - T.z0) : ((T.z0 + 1.f) - T.iz);
        const size_t off = (size_t)tri_offset(T, k) * CDIM + 8 * t;
        if (scatter) {
            const float w = (wx * wy) * wz;
            red_add_v4(G.grad + off, w * dc[0], w * dc[1], w * dc[2], w * dc[3]);
            red_add_v4(G.grad + off + 4, w * dc[4], w * dc[5], w * dc[6], w * dc[7]);
        }
        if (coord) {
            const float4* p = reinterpret_cast<const float4*>(G.data + off);
            float4 a = __ldg(p), b = __ldg(p + 1);
            float s = a.x * dc[0] + a.y * dc[1] + a.z * dc[2] + a.w * dc[3] + b.x * dc[4] +
                      b.y * dc[5] + b.z * dc[6] + b.w * dc[7];
            gx += (dx ? s : -s) * wy * wz;
            gy += (dy ? s : -s) * wx * wz;
            gz += (dz ? s : -s) * wx * wy;
        }
    }
    if (coord) {
        dp[0] += gx * T.mx * (float)(2.0 / ext[0]);
        dp[1] += gy * T.my * (float)(2.0 / ext[1]);
        dp[2] += gz * T.mz * (float)(2.0 / ext[2]);
    }
}

template <int KIND>
__device__ void decoder_backward_cta(const StageParams& P, const DecTask& T, int task, float* sm) {
    using C = DecCfg<KIND>;
    using O = Off<KIND>;
    // a decoder that never recomputes its forward here (ReLU signs saved, layer inputs deposited) needs the
    // constants and the transposed (backward) operand blocks only: 62 of the 131 KB
    if (T.relu && (!T.dw_scratch || T.x_saved)) {
        stage_range(sm, T.packed, 0, C::OFF_FWD);
        stage_range(sm, T.packed, C::OFF_BWD, C::TOTAL);
        __syncthreads();
    } else {
        stage_weights(sm, T.packed, C::TOTAL);
    }
    pdl_wait();                 // everything below reads what the kernel before this one wrote
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 3;
    float* es = sm + C::TOTAL + warp * E_BUF;
    const long long ntiles = (P.pts.P + 15) / 16;
    const bool want_grid = (P.flags & NSB_GRAD_GRID) && T.own.grad != nullptr;
    const bool want_point = (P.flags & NSB_GRAD_POINT) && T.d_point != nullptr;
    using RC = RecCfg<KIND>;
    (void)ntiles;
    unsigned* ctr = P.ctr + task;
    // Tile queue, fetched TWO tiles ahead: the counter's answer (an L2 atomic round trip) is consumed one whole
    // tile after it was asked for, instead of right away.  Warps over-fetch up to two tickets past the end.
    long long tile = P.tile_begin + next_tile(ctr, lane);
    unsigned ticket_next = 0, ticket_after = 0;
    if (lane == 0) ticket_next = atomicAdd(ctr, 1u);
    for (long long nxt; tile < P.tile_end; tile = nxt) {
        if (lane == 0) ticket_after = atomicAdd(ctr, 1u);
        nxt = P.tile_begin + (long long)__shfl_sync(FULL, ticket_next, 0);
        ticket_next = ticket_after;
        // the next tile's first reads (incoming gradient, ReLU signs) and this tile's LAST reads (sample depths of
        // its rays: the point evaluation is scheduled at its first use, the scatter at the end of the tile) go to L1
        // now, so that neither waits for L2 when its turn comes
        if (lane < 8) {
            const void* a = nullptr;
            const long long last = P.pts.P - 1;                       // every address is that of a valid point
            auto pt = [&](long long i) { return i < last ? i : last; };
            if (lane < 2) a = P.d_raw + 4 * pt(nxt * 16 + 8 * lane);                       // 16 B per point
            else if (lane < 6) a = T.relu ? T.relu + 8 * pt(nxt * 16 + 4 * (lane - 2)) : nullptr;   // 32 B per point
            else if (P.pts.z) a = P.pts.z + pt(lane == 6 ? tile * 16 : nxt * 16);          // 8 B per point
            if (a) prefetch_l1(a);
        }
        float* rec = T.dw_scratch ? T.dw_scratch + (size_t)(tile - P.tile_begin) * RC::REC : nullptr;
        // incoming gradient first: most tiles behind the surface carry exact zeros
        float dA[1][4];
        {
            const long long i0 = tile * 16 + (lane >> 2), i1 = i0 + 8;
            float v0 = 0.f, v1 = 0.f;
            if (C::NOUT == 1) {
                if (t == 0) {
                    if (i0 < P.pts.P) v0 = P.d_raw[4 * i0 + 3];
                    if (i1 < P.pts.P) v1 = P.d_raw[4 * i1 + 3];
                }
            } else {
                if (t < 3) {
                    if (i0 < P.pts.P) v0 = P.d_raw[4 * i0 + t];
                    if (i1 < P.pts.P) v1 = P.d_raw[4 * i1 + t];
                }
            }
            dA[0][0] = v0; dA[0][1] = v1; dA[0][2] = 0.f; dA[0][3] = 0.f;
            if (!__any_sync(FULL, v0 != 0.f || v1 != 0.f)) {
                if (rec && lane == 0) T.dw_flags[tile - P.tile_begin] = 0;
                if (want_point && t == 0) {
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        if (i0 < P.pts.P) T.d_point[3 * i0 + a] = 0.f;
                        if (i1 < P.pts.P) T.d_point[3 * i1 + a] = 0.f;
                    }
                }
                continue;
            }
        }
        Rows R;
        Masks mk;
        if (T.relu && (!rec || T.x_saved)) {
            // the forward saved the ReLU signs: no recomputation (gather, embedding, six layers).
            // Saved as bytes [point][part n = feature / 8][layer]; lane (g, t) needs bits 2t, 2t + 1 of
            // byte (n, layer) of points g and g + 8 for its C-fragment columns 8n + 2t (+1)
            load_rows<KIND>(P.pts, T, tile, lane, R);
            const int g = lane >> 2;
            const long long i0 = tile * 16 + g, i1 = i0 + 8;
            const uint4* src = reinterpret_cast<const uint4*>(T.relu);
            const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
            const uint4 a0 = i0 < P.pts.P ? __ldg(src + 2 * i0) : z4, b0 = i0 < P.pts.P ? __ldg(src + 2 * i0 + 1) : z4;
            const uint4 a1 = i1 < P.pts.P ? __ldg(src + 2 * i1) : z4, b1 = i1 < P.pts.P ? __ldg(src + 2 * i1 + 1) : z4;
            const unsigned lo0[4] = {a0.x, a0.z, b0.x, b0.z}, hi0[4] = {a0.y, a0.w, b0.y, b0.w};
            const unsigned lo1[4] = {a1.x, a1.z, b1.x, b1.z}, hi1[4] = {a1.y, a1.w, b1.y, b1.w};
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                unsigned bits = 0;
#pragma unroll
                for (int n = 0; n < 4; ++n) {
                    const unsigned y0 = i < 4 ? lo0[n] >> (8 * i) : hi0[n], y1 = i < 4 ? lo1[n] >> (8 * i) : hi1[n];
                    bits |= (((y0 >> (2 * t)) & 3u) | (((y1 >> (2 * t)) & 3u) << 2)) << (4 * n);
                }
                mk.set(i, bits);
            }
        } else {
            float c[C::CS][4];
            float out[4];
            tile_inputs<KIND>(P.pts, T, sm, es, tile, lane, R, c);
            mlp_forward<KIND>(sm, es, lane, c, out, mk, rec);
            if (rec) {      // layer inputs that live outside the MLP loop: e, c, p; and d(out)
                const int g = lane >> 2;
                if (lane == 0) T.dw_flags[tile - P.tile_begin] = 1;
                if constexpr (C::EMBED) {
#pragma unroll 1
                    for (int j = 0; j < E_SLABS; ++j) {
                        const int f0 = 8 * j + t, f1 = f0 + 4;
                        rec[RC::R_E + rec_e(g, f0)] = es[esw(g, f0)];
                        rec[RC::R_E + rec_e(g + 8, f0)] = es[esw(g + 8, f0)];
                        rec[RC::R_E + rec_e(g, f1)] = es[esw(g, f1)];
                        rec[RC::R_E + rec_e(g + 8, f1)] = es[esw(g + 8, f1)];
                    }
                }
#pragma unroll
                for (int L = 0; L < C::CS / 4; ++L) {       // channels 8t..8t+7 of grid level L
                    float* c0 = rec + RC::R_C + g * C::C_IN + rec_sw(g, 32 * L + 8 * t);
                    float* c1 = rec + RC::R_C + (g + 8) * C::C_IN + rec_sw(g, 32 * L + 8 * t);
                    const int j0 = 4 * L;
                    *reinterpret_cast<float4*>(c0) = make_float4(c[j0][0], c[j0][2], c[j0 + 1][0], c[j0 + 1][2]);
                    *reinterpret_cast<float4*>(c0 + 4) = make_float4(c[j0 + 2][0], c[j0 + 2][2], c[j0 + 3][0], c[j0 + 3][2]);
                    *reinterpret_cast<float4*>(c1) = make_float4(c[j0][1], c[j0][3], c[j0 + 1][1], c[j0 + 1][3]);
                    *reinterpret_cast<float4*>(c1 + 4) = make_float4(c[j0 + 2][1], c[j0 + 2][3], c[j0 + 3][1], c[j0 + 3][3]);
                }
            }
        }
        if (rec) {          // sections both paths write: the live flag, p and d(out)
            const int g = lane >> 2;
            if (lane == 0) T.dw_flags[tile - P.tile_begin] = 1;
                // p: [16][8] (cols 0..2), d(out): [16][32] (cols < NOUT, rest zero)
            {
                float* p0 = rec + RC::R_P + g * 8, * p1 = rec + RC::R_P + (g + 8) * 8;
                if (t == 0) {
                    *reinterpret_cast<float4*>(p0) = make_float4(R.p[0][0], R.p[0][1], R.p[0][2], 0.f);
                    *reinterpret_cast<float4*>(p1) = make_float4(R.p[1][0], R.p[1][1], R.p[1][2], 0.f);
                } else if (t == 1) {
                    *reinterpret_cast<float4*>(p0 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
                    *reinterpret_cast<float4*>(p1 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                // gather the (up to 4) output gradients of the row from lanes t = 0..3
                float d0[4], d1[4];
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    d0[o] = __shfl_sync(FULL, dA[0][0], (lane & ~3) | o);
                    d1[o] = __shfl_sync(FULL, dA[0][1], (lane & ~3) | o);
                }
                // logical columns 8t..8t+7 of the row; the row's data (cols 0..3) is lane t == 0's
                float* o0 = rec + RC::R_DO + g * HID + rec_sw(g, 8 * t);
                float* o1 = rec + RC::R_DO + (g + 8) * HID + rec_sw(g, 8 * t);
                const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                *reinterpret_cast<float4*>(o0) = t == 0 ? make_float4(d0[0], d0[1], d0[2], d0[3]) : z4;
                *reinterpret_cast<float4*>(o1) = t == 0 ? make_float4(d1[0], d1[1], d1[2], d1[3]) : z4;
                *reinterpret_cast<float4*>(o0 + 4) = z4;
                *reinterpret_cast<float4*>(o1 + 4) = z4;
            }
        }
        float dp[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
        float dh[4][4], a[4][4], dc[4][4];
        zero_(dh);
        zero_(dc);
        gemm3<1, 4>(dh, dA, sm + C::OFF_TOUT, lane);                  // d(h5)
#pragma unroll 1
        for (int i = 4; i >= 0; --i) {
            chain<4>(a, dh);                                           // d(h_{i+1})
            if constexpr (C::EMBED) {
                gemm3<4, 4>(dc, a, sm + O::TC(i), lane);
                if (rec) store_c32(rec + RC::R_DH + i * 16 * HID, dh, lane);
            }
            mask_(dh, mk.get(i));                                      // through the ReLU
            if (rec) store_c32(rec + RC::da(i), dh, lane);
            chain<4>(a, dh);
            if (i == 3 || i == 0) {                                    // input-side part of the layer
                if constexpr (C::EMBED) {
                    if (want_point || rec)
                        embed_grad(a, sm + O::TX(i), R, sm + C::OFF_B, lane, dp, es, i == 0,
                                   rec ? rec + RC::R_DE : nullptr);
                } else {
                    gemm3<4, 4>(dc, a, sm + O::TX(i), lane);
                }
            }
            if (i > 0) {
                zero_(dh);
                gemm3<4, 4>(dh, a, sm + O::TL(i), lane);               // d(h_i)
            }
        }
        // d(c): gather column map -> lane holds channels 8t..8t+7 of rows g, g+8
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            float d8[8], g3[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 8; ++k) d8[k] = r ? dc[k >> 1][2 + (k & 1)] : dc[k >> 1][k & 1];
            const float u[3] = {r ? R.u[1][0] : R.u[0][0], r ? R.u[1][1] : R.u[0][1], r ? R.u[1][2] : R.u[0][2]};
            if ((r ? R.valid[1] : R.valid[0]) && (want_grid || want_point))
                grid_backward(want_grid, want_point, T.own, u, T.ext, t, d8, g3);
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                if (r == 0) dp[0][ax] += g3[ax]; else dp[1][ax] += g3[ax];
            }
        }
        if (want_point) {
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int ax = 0; ax < 3; ++ax) {
                    float v = dp[r][ax];
                    v += __shfl_xor_sync(FULL, v, 1);
                    v += __shfl_xor_sync(FULL, v, 2);
                    if (t == 0 && R.valid[r]) T.d_point[3 * R.idx[r] + ax] = v;
                }
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1)
k_decoders_backward(const __grid_constant__ StageParams P) {
    extern __shared__ __align__(16) float sm[];
    pdl_launch_dependents();     // k_dweight / k_ray_grad_reduce may be scheduled as CTAs of this kernel retire
    int k = 0;
    while (k + 1 < P.ntask && (int)blockIdx.x >= P.task[k + 1].cta_begin) ++k;
    const DecTask& T = P.task[k];
#ifdef NSB_TC_PROFILE
    const long long t_start = clock64();
#endif
    switch (T.kind) {
        case KIND_COARSE: decoder_backward_cta<KIND_COARSE>(P, T, k, sm); break;
        case KIND_MIDDLE: decoder_backward_cta<KIND_MIDDLE>(P, T, k, sm); break;
        case KIND_FINE: decoder_backward_cta<KIND_FINE>(P, T, k, sm); break;
        default: decoder_backward_cta<KIND_COLOR>(P, T, k, sm); break;
    }
#ifdef NSB_TC_PROFILE
    if (P.prof) {      // per decoder: slowest and summed warp run time (cycles), warps
        const long long dt = clock64() - t_start;
        if ((threadIdx.x & 31) == 0) {
            atomicMax((unsigned long long*)(P.prof + 36 + 3 * k), (unsigned long long)dt);
            atomicAdd((unsigned long long*)(P.prof + 36 + 3 * k + 1), (unsigned long long)dt);
            atomicAdd((unsigned long long*)(P.prof + 36 + 3 * k + 2), 1ull);
        }
    }
#endif
    release_counters(P.ctr);
}

// ---- weight gradients ------------------------------------------------------------------------
// d(W) = G^T X summed over points, for every linear layer of one decoder, from the records
// the backward kernel wrote (nsb_layout.h RecCfg).  A CTA owns a contiguous range of tiles,
// compacts the live ones (flag set) and streams their records through a two-stage shared-
// memory ring with 16-byte async copies (a record is one contiguous, pre-swizzled block, so
// the copy is linear and the transposed fragment reads are bank-conflict free).  Each warp
// runs its "units" -- a unit is one 32 x 8 block of some d(W) (two m-tiles, one n-tile),
// 3xTF32, accumulators in registers for the whole kernel -- then adds them to the flat
// gradient with one atomic per element per CTA.  Bias gradients are column sums of G.
constexpr int DW_UNIT_INTS = 12;
constexpr int DW_UPW = 8;          // max units per warp
constexpr int DW_MAX_UNITS = DW_UPW * WARPS;
constexpr int DW_MAX_SEC = 26;
constexpr int DW_MAX_BIAS = THREADS;
constexpr int DW_LIVE_MAX = 256;   // tiles per CTA range

struct DwUnit {       // all offsets in floats of the record
    int a_off, a_w, a_col0;      // G section, row stride, first row (f') of the 32-row block
    int b_off, b_w, b_col0;      // X section, row stride, first column of the n-tile
    int out_off, ld_row, ld_col; // flat gradient index = out_off + row*ld_row + col*ld_col
    int rows_valid, cols_valid;
    int sw;                      // bit 0: G section columns are rec_sw-swizzled, bit 1: X section
};
struct DwSec { int g_off, width; };
struct DwBias { int s_off, s_w, col, out_idx; };

struct DwTables {
    DwUnit unit[DW_MAX_UNITS];
    DwSec sec[DW_MAX_SEC];
    DwBias bias[DW_MAX_BIAS];
    int nunit, nsec, nbias, rec;
};

struct DwParams {
    const float* scratch;
    const unsigned char* flags;
    long long ntiles;
    const DwTables* tab;     // device copy
    float* dflat;
};

constexpr int DW_HEAD_FLOATS = DW_MAX_UNITS * DW_UNIT_INTS + DW_LIVE_MAX;   // unit table + live list

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gmem_src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}

__global__ void __launch_bounds__(THREADS, 1) k_dweight(const DwParams P) {
    extern __shared__ __align__(16) float sm[];
    pdl_launch_dependents();
    const DwTables& TB = *P.tab;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    int* s_unit = reinterpret_cast<int*>(sm);
    int* s_live = s_unit + DW_MAX_UNITS * DW_UNIT_INTS;
    float* ring = sm + DW_HEAD_FLOATS;
    __shared__ int s_nlive;
    const int nunit = TB.nunit, nbias = TB.nbias, rec = TB.rec;
    const int upw = (nunit + WARPS - 1) / WARPS;             // units per warp (<= DW_UPW)
    for (int i = threadIdx.x; i < nunit * DW_UNIT_INTS; i += THREADS)
        s_unit[i] = reinterpret_cast<const int*>(TB.unit)[i];
    DwBias mybias{0, 0, 0, 0};
    if ((int)threadIdx.x < nbias) mybias = TB.bias[threadIdx.x];
    pdl_wait();                  // the tables are constants; flags and records come from the backward kernel
    // Balanced shares of the LIVE tiles: every CTA scans all flags (a few KB), ranks the live
    // tiles globally and keeps those ranked [b L / G, (b + 1) L / G) -- live tiles cluster along
    // rays, so equal index ranges left some CTAs with three times the mean (r01 profile).
    // Order is by tile index: the summation order is deterministic.
    __shared__ int s_warp_cnt[WARPS];
    const long long T_ = P.ntiles;
    const int per_thread = (int)((T_ + THREADS - 1) / THREADS);
    const long long my0 = (long long)threadIdx.x * per_thread;
    int mine = 0;
    for (int j = 0; j < per_thread; ++j) {
        const long long i = my0 + j;
        if (i < T_ && P.flags[i] != 0) ++mine;
    }
    int incl = mine;                                           // inclusive scan over the CTA
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(FULL, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp_cnt[warp] = incl;
    __syncthreads();
    int before = incl - mine, total = 0;
    for (int w = 0; w < WARPS; ++w) {
        const int c = s_warp_cnt[w];
        if (w < warp) before += c;
        total += c;
    }
    const long long lo = (long long)blockIdx.x * total / gridDim.x, hi = (long long)(blockIdx.x + 1) * total / gridDim.x;
    {
        int rank = before;
        for (int j = 0; j < per_thread; ++j) {
            const long long i = my0 + j;
            if (i < T_ && P.flags[i] != 0) {
                if (rank >= lo && rank < hi && rank - lo < DW_LIVE_MAX) s_live[rank - lo] = (int)i;
                ++rank;
            }
        }
    }
    if (threadIdx.x == 0) s_nlive = (int)((hi - lo) < DW_LIVE_MAX ? (hi - lo) : DW_LIVE_MAX);
    __syncthreads();
    const int nlive = s_nlive;
    float acc[DW_UPW][2][4];
#pragma unroll
    for (int k = 0; k < DW_UPW; ++k)
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[k][m][q] = 0.f;
    float bias_acc = 0.f;
    const int rec4 = rec >> 2;
    auto issue = [&](int i) {
        const float* src = P.scratch + (size_t)s_live[i] * rec;
        float* dst = ring + (i & 1) * rec;
        for (int c = threadIdx.x; c < rec4; c += THREADS) cp_async16(dst + 4 * c, src + 4 * c);
    };
    if (nlive > 0) issue(0);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    for (int i = 0; i < nlive; ++i) {
        if (i + 1 < nlive) issue(i + 1);
        asm volatile("cp.async.commit_group;\n" ::: "memory");
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        __syncthreads();                                     // record i has landed for every thread
        const float* R = ring + (i & 1) * rec;
        if ((int)threadIdx.x < nbias) {
            float s = 0.f;
#pragma unroll
            for (int pt = 0; pt < 16; ++pt) s += R[mybias.s_off + pt * mybias.s_w + rec_sw(pt, mybias.col)];
            bias_acc += s;
        }
        // a warp owns CONSECUTIVE units: they mostly share their G section (the A operand), which
        // is then loaded and split once per tile instead of once per unit
        int prev_a = -1, prev_c = -1;
        uint32_t ah[2][2][4], al[2][2][4];
#pragma unroll
        for (int k = 0; k < DW_UPW; ++k) {
            const int u = warp * upw + k;
            if (k < upw && u < nunit) {
                const int* U = s_unit + u * DW_UNIT_INTS;
                const int a_off = U[0], a_w = U[1], a_col0 = U[2], b_off = U[3], b_w = U[4], b_col0 = U[5];
                const int swa = (U[11] & 1) ? (t << 3) : 0, swb = (U[11] & 2) ? (t << 3) : 0;
                if (a_off != prev_a || a_col0 != prev_c) {      // uniform over the warp
                    prev_a = a_off; prev_c = a_col0;
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const float* Ar = R + a_off + (8 * ks + t) * a_w;   // rows 8ks+t, +4: (row & 3) == t
#pragma unroll
                        for (int m = 0; m < 2; ++m) {
                            const int c0 = (a_col0 + g + 16 * m) ^ swa, c1 = (a_col0 + g + 16 * m + 8) ^ swa;
                            split_tf32(Ar[c0], ah[ks][m][0], al[ks][m][0]);
                            split_tf32(Ar[c1], ah[ks][m][1], al[ks][m][1]);
                            split_tf32(Ar[4 * a_w + c0], ah[ks][m][2], al[ks][m][2]);
                            split_tf32(Ar[4 * a_w + c1], ah[ks][m][3], al[ks][m][3]);
                        }
                    }
                }
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    const float* Bp = R + b_off + (8 * ks + t) * b_w + ((b_col0 + g) ^ swb);
                    uint32_t bh[2], bl[2];
                    split_tf32(Bp[0], bh[0], bl[0]);
                    split_tf32(Bp[4 * b_w], bh[1], bl[1]);
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        mma_tf32(acc[k][m], al[ks][m], bh[0], bh[1]);
                        mma_tf32(acc[k][m], ah[ks][m], bl[0], bl[1]);
                        mma_tf32(acc[k][m], ah[ks][m], bh[0], bh[1]);
                    }
                }
            }
        }
        __syncthreads();                                     // stage (i & 1) may be refilled
    }
    // flush: C fragment (row 16m + g (+8), col 2t (+1)) of unit u
#pragma unroll
    for (int k = 0; k < DW_UPW; ++k) {
        const int u = warp * upw + k;
        if (k < upw && u < nunit) {
            const int* U = s_unit + u * DW_UNIT_INTS;
            // table entries in registers: the atomics below would otherwise force a reload per element
            const int out_off = U[6], ld_row = U[7], ld_col = U[8], rows_valid = U[9], cols_valid = U[10];
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int row = 16 * m + g + 8 * (q >> 1), col = 2 * t + (q & 1);
                    if (row < rows_valid && col < cols_valid && acc[k][m][q] != 0.f)
                        atomicAdd(P.dflat + out_off + row * ld_row + col * ld_col, acc[k][m][q]);
                }
        }
    }
    if ((int)threadIdx.x < nbias && bias_acc != 0.f) atomicAdd(P.dflat + mybias.out_idx, bias_acc);
}

// host: flat offsets of the canonical parameters of one decoder kind
template <int KIND>
static void canonical_offsets(int (&off)[32]) {
    using C = DecCfg<KIND>;
    int acc_off = 0;
    for (int i = 0; i < C::N_PARAM; ++i) {
        off[i] = acc_off;
        int numel;
        if (i < 5) numel = HID * C::in_dim(i);
        else if (i < 10) numel = HID;
        else if (C::EMBED) numel = i < 15 ? HID * C::C_IN : (i < 20 ? HID : (i == 20 ? C::NOUT * HID : (i == 21 ? C::NOUT : 3 * EMB)));
        else numel = i == 10 ? HID : 1;
        acc_off += numel;
    }
}

// host: tables of one decoder kind
template <int KIND>
static void build_dw_tables(DwTables& T) {
    using C = DecCfg<KIND>;
    using RC = RecCfg<KIND>;
    memset(&T, 0, sizeof(T));
    int off[32];
    canonical_offsets<KIND>(off);
    // sections of the record (the shared-memory copy is a verbatim image of it)
    auto add_sec = [&](int g_off, int width) {
        DwSec& s = T.sec[T.nsec++];
        s.g_off = g_off; s.width = width;
        return g_off;
    };
    int sE[3] = {-1, -1, -1}, sH[5], sC, sP = -1, sDA[5], sDH[5], sDO, sDE[3] = {-1, -1, -1};
    if (C::EMBED) for (int j = 0; j < 3; ++j) sE[j] = add_sec(RC::R_E + j * RC::ATOM, HID);
    for (int i = 0; i < 5; ++i) sH[i] = add_sec(RC::R_H + i * 16 * HID, HID);
    sC = add_sec(RC::R_C, C::C_IN);
    if (C::EMBED) sP = add_sec(RC::R_P, 8);
    for (int i = 0; i < 5; ++i) sDA[i] = add_sec(RC::da(i), HID);
    if (C::EMBED) for (int i = 0; i < 5; ++i) sDH[i] = add_sec(RC::R_DH + i * 16 * HID, HID);
    sDO = add_sec(RC::R_DO, HID);
    if (C::EMBED) for (int j = 0; j < 3; ++j) sDE[j] = add_sec(RC::R_DE + j * RC::ATOM, HID);
    T.rec = RC::REC;
    auto add_unit = [&](int a_off, int a_width, int a_col0, int b_off, int b_width, int b_col0, int out_off,
                        int ld_row, int ld_col, int rows_valid, int cols_valid) {
        DwUnit& u = T.unit[T.nunit++];
        u.a_off = a_off; u.a_w = a_width; u.a_col0 = a_col0;
        u.b_off = b_off; u.b_w = b_width; u.b_col0 = b_col0;
        u.out_off = out_off; u.ld_row = ld_row; u.ld_col = ld_col;
        u.rows_valid = rows_valid; u.cols_valid = cols_valid;
        u.sw = (a_width >= 32 ? 1 : 0) | (b_width >= 32 ? 2 : 0);
    };
    // X parts of layer i: (section, width, number of n-tiles, valid features, first input column)
    for (int i = 0; i < 5; ++i) {
        const int in = C::in_dim(i);
        auto part = [&](int sec, int width, int feats, int col_base) {
            for (int n = 0; n * 8 < feats; ++n) {
                const int valid = feats - 8 * n < 8 ? feats - 8 * n : 8;
                add_unit(sDA[i], HID, 0, sec, width, 8 * n, off[C::P_W + i] + col_base + 8 * n, in, 1, HID, valid);
            }
        };
        if (C::EMBED) {
            auto e_parts = [&]() {
                for (int j = 0; j < 3; ++j) part(sE[j], HID, EMB - 32 * j < 32 ? EMB - 32 * j : 32, 32 * j);
            };
            if (i == 0) e_parts();
            else if (i == 3) { e_parts(); part(sH[2], HID, HID, EMB); }
            else part(sH[i - 1], HID, HID, 0);
            for (int n = 0; n * 8 < C::C_IN; ++n)
                add_unit(sDH[i], HID, 0, sC, C::C_IN, 8 * n, off[C::P_WC + i] + 8 * n, C::C_IN, 1, HID, 8);
        } else {
            if (i == 0) part(sC, CDIM, CDIM, 0);
            else if (i == 3) { part(sC, CDIM, CDIM, 0); part(sH[2], HID, HID, CDIM); }
            else part(sH[i - 1], HID, HID, 0);
        }
    }
    for (int n = 0; n < 4; ++n)       // output layer: rows = outputs
        add_unit(sDO, HID, 0, sH[4], HID, 8 * n, off[C::P_WO] + 8 * n, HID, 1, C::NOUT, 8);
    if (C::EMBED)
        for (int m0 = 0; m0 < EMB_PAD; m0 += 32)   // d(B)[a][f]: rows = features f, cols = axes a
            add_unit(sDE[m0 / 32], HID, 0, sP, 8, 0, off[C::P_EMB] + m0, 1, EMB, (EMB - m0 < 32 ? EMB - m0 : 32), 3);
    auto add_bias = [&](int sec, int width, int col, int out_idx) {
        DwBias& b = T.bias[T.nbias++];
        b.s_off = sec; b.s_w = width; b.col = col; b.out_idx = out_idx;
    };
    for (int i = 0; i < 5; ++i)
        for (int f = 0; f < HID; ++f) {
            add_bias(sDA[i], HID, f, off[C::P_B + i] + f);
            if (C::EMBED) add_bias(sDH[i], HID, f, off[C::P_BC + i] + f);
        }
    for (int o = 0; o < C::NOUT; ++o) add_bias(sDO, HID, o, off[C::P_BO] + o);
}

}  // namespace nsb
#include "nsb_decoder_tc.cuh"
#include "nsb_decoder_bwd_tc.cuh"
#include "nsb_dweight_tc.cuh"
namespace nsb {

template <int KIND>
static DwTcOffs dw_tc_offsets() {
    using C = DecCfg<KIND>;
    int off[32];
    canonical_offsets<KIND>(off);
    DwTcOffs o{};
    for (int i = 0; i < 5; ++i) {
        o.w[i] = off[C::P_W + i]; o.b[i] = off[C::P_B + i]; o.wc[i] = off[C::P_WC + i]; o.bc[i] = off[C::P_BC + i];
    }
    o.wo = off[C::P_WO]; o.bo = off[C::P_BO]; o.emb = off[C::P_EMB];
    o.total = off[C::P_EMB] + 3 * EMB;
    return o;
}

// ---- host side -------------------------------------------------------------------------------
int dweight_grid_cap_take();
static int g_num_sms = 0;
static int g_use_tc = 1;         // NSB_TC=0 selects the mma.sync kernels (A/B comparisons)
static int g_use_tc_bwd = 1;     // NSB_TC_BWD=0 keeps the mma.sync backward under a tcgen05 forward
static int g_use_dw_tc = 1;      // NSB_DW_TC=0 keeps the mma.sync weight-gradient kernel (middle / colour decoders)
static float g_xsaved_cost = 0.62f;  // backward cost of a decoder whose records' X sections the forward deposited: no
                                     // recomputation, but the G sections are still written (NSB_XSAVED_COST)
static float g_xfwd_cost = 1.4f;     // forward cost of depositing them (NSB_XFWD_COST)
static float g_tcb_dw_cost = 22.f;   // tcgen05 backward: cost of a decoder that writes records / d(e), others 10 (NSB_TCB_DW_COST)
static float g_saved_cost = 0.3f; // backward cost of a decoder that skips the recomputation, relative (NSB_SAVED_COST; swept on B200, r01)
static bool g_attr_done = false;
constexpr int CTR_SLOTS = 256;                   // launches in flight before a counter slot is reused
__device__ unsigned g_tile_ctr[CTR_SLOTS * 4];
static unsigned* g_ctr_dev = nullptr;
struct DwTables;
static DwTables* g_dw_tables_dev = nullptr;
static unsigned g_ctr_next = 0;
static unsigned* next_counters() { return g_ctr_dev + 4 * (g_ctr_next++ % CTR_SLOTS); }
__device__ DwTables g_dw_tables[4];
static DwTables g_dw_host[4];

static int blob_floats(int kind, bool bwd) {
    switch (kind) {
        case KIND_COARSE: return bwd ? DecCfg<KIND_COARSE>::TOTAL : DecCfg<KIND_COARSE>::FWD_END;
        case KIND_MIDDLE: return bwd ? DecCfg<KIND_MIDDLE>::TOTAL : DecCfg<KIND_MIDDLE>::FWD_END;
        case KIND_FINE: return bwd ? DecCfg<KIND_FINE>::TOTAL : DecCfg<KIND_FINE>::FWD_END;
        default: return bwd ? DecCfg<KIND_COLOR>::TOTAL : DecCfg<KIND_COLOR>::FWD_END;
    }
}

static int init_device() {
    if (g_attr_done) return 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error((int)e, "cudaGetDevice: %s", cudaGetErrorString(e));
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return set_error((int)e, "device attribute: %s", cudaGetErrorString(e));
    g_num_sms = sms;
    const int max_smem = (DecCfg<KIND_FINE>::TOTAL + WARPS * E_BUF) * 4;
    e = cudaFuncSetAttribute(k_decoders_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_decoders_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_decoders_forward_tc, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 TcSmem<KIND_FINE>::S_END * 4);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_decoders_backward_tc, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 TbSmem<KIND_FINE>::S_END * 4);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_dweight_tc, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 DwTcSmem<KIND_COLOR>::S_END * 4 + 1024);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    if (const char* v = getenv("NSB_TC")) g_use_tc = atoi(v);
    if (const char* v = getenv("NSB_DW_TC")) g_use_dw_tc = atoi(v);
    if (const char* v = getenv("NSB_TC_BWD")) g_use_tc_bwd = atoi(v);
    if (const char* v = getenv("NSB_SAVED_COST")) g_saved_cost = (float)atof(v);
    if (const char* v = getenv("NSB_XSAVED_COST")) g_xsaved_cost = (float)atof(v);
    if (const char* v = getenv("NSB_XFWD_COST")) g_xfwd_cost = (float)atof(v);
    if (const char* v = getenv("NSB_TCB_DW_COST")) g_tcb_dw_cost = (float)atof(v);
    build_dw_tables<KIND_COARSE>(g_dw_host[0]);
    build_dw_tables<KIND_MIDDLE>(g_dw_host[1]);
    build_dw_tables<KIND_FINE>(g_dw_host[2]);
    build_dw_tables<KIND_COLOR>(g_dw_host[3]);
    e = cudaMemcpyToSymbol(g_dw_tables, g_dw_host, sizeof(g_dw_host));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_dweight, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (DW_HEAD_FLOATS + 2 * REC_MAX) * 4);
    if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&g_ctr_dev, g_tile_ctr);
    if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&g_dw_tables_dev, g_dw_tables);
    if (e != cudaSuccess) return set_error((int)e, "d(W) tables: %s", cudaGetErrorString(e));
    g_attr_done = true;
    return 0;
}

// decoders of a stage (decoder.py:336-370) and their relative cost per point
static int stage_tasks(const NsbScene& sc, int stage, bool bwd, StageParams& sp, float* out_occ0,
                       float* out_occ1, float* out_rgb, float* d_point, long long P, unsigned dw_mask = 0,
                       unsigned x_mask = 0, bool tc_bwd = false) {
    int kinds[3], n = 0;
    float cost[3];
    switch (stage) {
        case NSB_STAGE_COARSE: kinds[n] = KIND_COARSE; cost[n++] = 12.f; break;
        case NSB_STAGE_MIDDLE: kinds[n] = KIND_MIDDLE; cost[n++] = 30.f; break;
        case NSB_STAGE_FINE:
            kinds[n] = KIND_MIDDLE; cost[n++] = 30.f;
            kinds[n] = KIND_FINE; cost[n++] = 41.f;
            break;
        case NSB_STAGE_COLOR:
            kinds[n] = KIND_MIDDLE; cost[n++] = 30.f;
            kinds[n] = KIND_FINE; cost[n++] = 41.f;
            kinds[n] = KIND_COLOR; cost[n++] = 31.f;
            break;
        default:
            if (stage >= NSB_STAGE_SINGLE && stage < NSB_STAGE_SINGLE + NSB_NUM_DEC) {
                kinds[n] = stage - NSB_STAGE_SINGLE; cost[n++] = 1.f;     // one decoder on its own
                break;
            }
            return set_error(NSB_EINVAL, "bad stage %d", stage);
    }
    // measured cycles per 128-point tile of the tcgen05 forward (in-kernel timers, r01): the fine
    // decoder pays a second feature pass
    if (!bwd && g_use_tc)
        for (int i = 0; i < n; ++i)
            cost[i] = kinds[i] == KIND_MIDDLE ? 15.f : (kinds[i] == KIND_FINE ? 21.5f : (kinds[i] == KIND_COLOR ? 16.f : cost[i]));
    if (!bwd && g_use_tc)                      // forward that also deposits the X sections of the d(W) records
        for (int i = 0; i < n; ++i)
            if (x_mask & (1u << kinds[i])) cost[i] *= g_xfwd_cost;
    // backward, measured warp run times per decoder (r01): fine 1.16x middle; writing the
    // weight-gradient records costs a decoder another 1.35x
    // a decoder without a weight gradient whose forward saved its ReLU signs skips the recomputation
    const bool saved = sp.pts.relu != nullptr && g_use_tc;
    if (bwd && tc_bwd)      // tcgen05 backward: one chain per decoder; records + d(e) make a trained decoder dearer
        for (int i = 0; i < n; ++i) cost[i] = (dw_mask & (1u << kinds[i])) ? g_tcb_dw_cost : 10.f;
    else if (bwd)
        for (int i = 0; i < n; ++i) {
            if (kinds[i] == KIND_FINE) cost[i] = 35.f;
            if (dw_mask & (1u << kinds[i])) cost[i] *= (saved && (x_mask & (1u << kinds[i]))) ? g_xsaved_cost * g_saved_cost / 0.3f : 1.35f;
            else if (saved && kinds[i] != KIND_COARSE) cost[i] *= g_saved_cost;
        }
    const long long tiles = (P + 15) / 16;
    const long long want_ll = (tiles + WARPS - 1) / WARPS;   // CTAs that give every warp one tile
    const int want = want_ll > (1 << 20) ? (1 << 20) : (int)want_ll;
    float total = 0.f;
    for (int i = 0; i < n; ++i) total += cost[i];
    // one persistent CTA per SM, split between the decoders in proportion to their cost
    int share_of[3], used = 0;
    float frac[3];
    for (int i = 0; i < n; ++i) {
        float s = g_num_sms * cost[i] / total;
        share_of[i] = (int)s;
        frac[i] = s - share_of[i];
        if (share_of[i] < 1) { share_of[i] = 1; frac[i] = 0.f; }
        used += share_of[i];
    }
    while (used < g_num_sms) {
        int best = 0;
        for (int i = 1; i < n; ++i) if (frac[i] > frac[best]) best = i;
        share_of[best]++; frac[best] = -1.f; used++;
    }
    // tcgen05 forward: a CTA runs whole 128-point tiles, and at 1000 rays a decoder has only 375 of them, 6-11 per
    // CTA -- the kernel lasts max_i ceil(tiles / share_i) * cost_i, which the proportional split misses by up
    // to a tile (measured: 35/55/58 CTAs 115 us, 38/54/56 CTAs 109 us).  Exact search over the splits, cached.
    if ((!bwd || tc_bwd) && g_use_tc && n >= 2 && !getenv(bwd ? "NSB_BWD_SHARES" : "NSB_FWD_SHARES")) {
        struct Key { long long tiles; int n, sms; float c[3]; int share[3]; };
        static thread_local Key cache[8];
        static thread_local int cache_next = 0;
        const long long t128 = (P + 127) / 128;
        const Key* hit = nullptr;
        for (const Key& k : cache)
            if (k.tiles == t128 && k.n == n && k.sms == g_num_sms && k.c[0] == cost[0] && k.c[1] == cost[1] &&
                (n < 3 || k.c[2] == cost[2])) hit = &k;
        if (!hit && t128 >= n && t128 <= 64ll * g_num_sms) {
            auto span = [&](int i, int sh) { return (float)((t128 + sh - 1) / sh) * cost[i]; };
            float best = INFINITY, best_sum = INFINITY;
            int bs[3] = {share_of[0], share_of[1], n == 3 ? share_of[2] : 0};
            for (int a = 1; a <= g_num_sms - (n - 1); ++a) {
                const float sa = span(0, a);
                if (sa > best) continue;
                const int b_hi = n == 3 ? g_num_sms - a - 1 : g_num_sms - a;
                for (int b = n == 3 ? 1 : b_hi; b <= b_hi; ++b) {
                    const float sb = span(1, b), sc3 = n == 3 ? span(2, g_num_sms - a - b) : 0.f;
                    const float m = fmaxf(sa, fmaxf(sb, sc3)), sum = sa + sb + sc3;
                    if (m < best || (m == best && sum < best_sum)) {
                        best = m; best_sum = sum; bs[0] = a; bs[1] = b; bs[2] = n == 3 ? g_num_sms - a - b : 0;
                    }
                }
            }
            Key& k = cache[cache_next++ % 8];
            k.tiles = t128; k.n = n; k.sms = g_num_sms;
            for (int i = 0; i < 3; ++i) { k.c[i] = i < n ? cost[i] : 0.f; k.share[i] = bs[i]; }
            hit = &k;
        }
        if (hit) for (int i = 0; i < n; ++i) share_of[i] = hit->share[i];
    }
    // tuning hook: NSB_BWD_SHARES="a,b,c" / NSB_FWD_SHARES forces the CTA split of a 3-decoder stage
    if (n == 3)
        if (const char* e = getenv(bwd ? "NSB_BWD_SHARES" : "NSB_FWD_SHARES")) {
            int a = 0, b = 0, c = 0;
            if (sscanf(e, "%d,%d,%d", &a, &b, &c) == 3 && a > 0 && b > 0 && c > 0 && a + b + c <= g_num_sms) {
                share_of[0] = a; share_of[1] = b; share_of[2] = c;
            }
        }
    int begin = 0;
    for (int i = 0; i < n; ++i) {
        DecTask& T = sp.task[i];
        const int k = kinds[i];
        T.kind = k;
        if (!sc.packed[k]) return set_error(NSB_EINVAL, "stage %d needs packed decoder %d", stage, k);
        T.packed = sc.packed[k];
        const NsbGrid& G = sc.grid[k];
        if (!G.data) return set_error(NSB_EINVAL, "stage %d needs grid %d", stage, k);
        T.own = GridDev{G.data, G.grad, G.X, G.Y, G.Z};
        T.aux = GridDev{nullptr, nullptr, 1, 1, 1};
        if (k == KIND_FINE) {
            const NsbGrid& M = sc.grid[KIND_MIDDLE];
            if (!M.data) return set_error(NSB_EINVAL, "fine decoder needs the middle grid");
            T.aux = GridDev{M.data, nullptr, M.X, M.Y, M.Z};
        }
        for (int a = 0; a < 3; ++a) {
            const double lo = k == KIND_COARSE ? sc.coarse_bound[a][0] : sc.bound[a][0];
            const double hi = k == KIND_COARSE ? sc.coarse_bound[a][1] : sc.bound[a][1];
            T.lo[a] = lo;
            T.ext[a] = hi - lo;
        }
        T.out = (k == KIND_COLOR) ? out_rgb : (k == KIND_FINE ? out_occ1 : out_occ0);
        if (!bwd && !T.out) return set_error(NSB_EINVAL, "missing output plane for decoder %d", k);
        T.d_point = d_point ? d_point + (size_t)i * P * 3 : nullptr;
        T.dw_scratch = nullptr;
        T.dw_flags = nullptr;
        T.relu = (saved && k != KIND_COARSE) ? sp.pts.relu + (size_t)i * P * 8 : nullptr;
        T.x_saved = (bwd && saved && (x_mask & (1u << k))) ? 1 : 0;
        int share = share_of[i];
        if (share > want) share = want;
        T.cta_begin = begin;
        T.cta_count = share;
        begin += share;
    }
    sp.ntask = n;
    return begin;
}

// dynamic shared memory of a launch in floats: the largest blob + one e buffer per warp
static int max_blob(const StageParams& sp, bool bwd) {
    int m = 0;
    for (int i = 0; i < sp.ntask; ++i) {
        int b = blob_floats(sp.task[i].kind, bwd);
        if (b > m) m = b;
    }
    return m + WARPS * E_BUF;
}

}  // namespace nsb

using namespace nsb;

// bytes of one tile's record + flag, sized for the largest decoder
static const int64_t kDwTileBytes = (int64_t)REC_MAX * 4 + 1;
// Records of a whole batch are kept when it has at most kDwMaxTiles tiles (4 Mi points, 12.5 GB per trained decoder):
// the forward can then deposit the layer inputs and the tcgen05 backward runs.  Larger batches go through in chunks of
// kDwChunkTiles (512 Ki points, 1.56 GB; measured faster than larger chunks: 49.9 vs 58.4 ms per step at 256 k rays).
static const int64_t kDwMaxTiles = 262144;
static const int64_t kDwChunkTiles = 32768;

// Weight-gradient scratch layout: [records of trained decoder 0 | 1 | ... | flags ...], `chunk` tiles per
// decoder (the backward runs in chunks when the scratch is smaller than the batch).  Returns chunk (0:
// too small) and the number of trained decoders of the stage.
static long long dw_layout(const int* kinds, int ntask, unsigned mask, int64_t scratch_bytes, long long ntiles, int& ndw) {
    ndw = 0;
    for (int i = 0; i < ntask; ++i)
        if (mask & (1u << kinds[i])) ++ndw;
    if (!ndw) return ntiles;
    long long chunk = ((scratch_bytes - 256) / ndw) / kDwTileBytes;
    return chunk > ntiles ? ntiles : chunk;
}
// The forward deposits the X sections of the records (and the backward then skips the recomputation of the
// trained decoders) when the points carry the backward's scratch and it holds the whole batch at once.
static bool x_deposit_ok(const NsbPoints& pts, const StageParams& sp, long long ntiles) {
    if (!g_use_tc || !pts.x_scratch || !pts.x_mask || !sp.pts.relu) return false;
    int kinds[3], ndw;
    for (int i = 0; i < sp.ntask; ++i) {
        kinds[i] = sp.task[i].kind;
        if (kinds[i] == KIND_COARSE) return false;
    }
    return dw_layout(kinds, sp.ntask, pts.x_mask, pts.x_scratch_bytes, ntiles, ndw) == ntiles && ndw > 0;
}

extern "C" int nsb_decoders_forward(const NsbPoints* pts, const NsbScene* scene, int stage,
                                    float* out_occ0, float* out_occ1, float* out_rgb,
                                    void* stream) {
    if (!pts || !scene) return set_error(NSB_EINVAL, "null argument");
    if (int rc = init_device()) return rc;
    StageParams sp{};
    sp.pts = to_dev(*pts);
    if (sp.pts.P <= 0) return 0;
    if (!sp.pts.ro && !sp.pts.p64 && !sp.pts.p32) return set_error(NSB_EINVAL, "no points given");
    int ctas = stage_tasks(*scene, stage, false, sp, out_occ0, out_occ1, out_rgb, nullptr, sp.pts.P);
    if (ctas < 0) return ctas;
    if (x_deposit_ok(*pts, sp, (sp.pts.P + 15) / 16)) {
        ctas = stage_tasks(*scene, stage, false, sp, out_occ0, out_occ1, out_rgb, nullptr, sp.pts.P, 0u, pts->x_mask);
        const long long nt = (sp.pts.P + 15) / 16;
        char* base = (char*)(((uintptr_t)pts->x_scratch + 255) & ~(uintptr_t)255);
        int k = 0;
        for (int i = 0; i < sp.ntask; ++i)
            if (pts->x_mask & (1u << sp.task[i].kind)) sp.task[i].dw_scratch = (float*)(base + (size_t)(k++) * nt * REC_MAX * 4);
    }
    sp.d_raw = nullptr;
    sp.flags = 0;
    {
        static const int dbg = getenv("NSB_DEBUG_FWD") ? atoi(getenv("NSB_DEBUG_FWD")) : 0;   // timing experiments only
        sp.flags |= (unsigned)(dbg & 3) << 28;
    }
    sp.tile_begin = 0;
    sp.tile_end = (sp.pts.P + 15) / 16;
    sp.ctr = next_counters();
#ifdef NSB_TC_PROFILE
    static long long* prof_dev = nullptr;
    if (!prof_dev) cudaMalloc(&prof_dev, 72 * sizeof(long long));
    sp.prof = prof_dev;
#endif
    bool tc = g_use_tc != 0;
    int tc_smem = 0;
    for (int i = 0; i < sp.ntask; ++i) {
        const int k = sp.task[i].kind;
        if (k == KIND_COARSE) tc = false;
        const int need = k == KIND_FINE ? TcSmem<KIND_FINE>::S_END : TcSmem<KIND_MIDDLE>::S_END;
        if (need > tc_smem) tc_smem = need;
    }
    timing_mark(0, 0, (cudaStream_t)stream);
    const bool pdl = pdl_take();
    if (tc) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(TCF_THREADS);
        cfg.dynamicSmemBytes = (size_t)tc_smem * 4; cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, k_decoders_forward_tc, sp);
    } else k_decoders_forward<<<ctas, THREADS, max_blob(sp, false) * 4, (cudaStream_t)stream>>>(sp);
    timing_mark(0, 1, (cudaStream_t)stream);
#ifdef NSB_TC_PROFILE
    if (getenv("NSB_TC_PROFILE_PRINT")) {
        long long h[72];
        cudaMemcpy(h, prof_dev, sizeof(h), cudaMemcpyDeviceToHost);
        const char* names[12] = {"setup+prologue", "wait upfront", "fine pass2", "next point", "epilogue", "barrier", "mma issue",
                                 "next slice", "wait layer", "output", "end barrier", ""};
        for (int t = 0; t < sp.ntask; ++t)
            for (int w = 0; w < 2; ++w) {
                fprintf(stderr, "[tcprof] task %d kind %d ctas %d %s:", t, sp.task[t].kind, sp.task[t].cta_count, w ? "thread200" : "thread0");
                for (int k = 0; k < 11; ++k) fprintf(stderr, " %s=%lld", names[k], h[(t * 2 + w) * 12 + k]);
                fprintf(stderr, "\n");
            }
    }
#endif
    return check_launch(tc ? "k_decoders_forward_tc" : "k_decoders_forward");
}


extern "C" int64_t nsb_dweight_scratch_bytes(int kind, int64_t P) {
    (void)kind;
    int64_t tiles = (P + 15) / 16;
    if (tiles > kDwMaxTiles) tiles = kDwChunkTiles;
    if (tiles < 1) tiles = 1;
    return tiles * kDwTileBytes + 256;
}

extern "C" int nsb_decoders_backward(const NsbPoints* pts, const NsbScene* scene, int stage,
                                     const float* d_raw, unsigned flags, unsigned dweight_mask,
                                     float* const* d_weight_host, float* d_point, void* scratch,
                                     int64_t scratch_bytes, void* stream) {
    if (!pts || !scene || !d_raw) return set_error(NSB_EINVAL, "null argument");
    if (int rc = init_device()) return rc;
    const bool defer_dw = (flags & NSB_FLAG_DEFER_DWEIGHT) != 0, only_dw = (flags & NSB_FLAG_ONLY_DWEIGHT) != 0;
    flags &= ~(NSB_FLAG_DEFER_DWEIGHT | NSB_FLAG_ONLY_DWEIGHT);
    const int dw_cap = dweight_grid_cap_take();
    StageParams sp{};
    sp.pts = to_dev(*pts);
    if (sp.pts.P <= 0) return 0;
    if ((flags & NSB_GRAD_POINT) && !d_point) return set_error(NSB_EINVAL, "d_point is NULL");
    const unsigned dwm = (flags & NSB_GRAD_WEIGHT) ? dweight_mask : 0u;
    int ctas = stage_tasks(*scene, stage, true, sp, nullptr, nullptr, nullptr,
                           (flags & NSB_GRAD_POINT) ? d_point : nullptr, sp.pts.P, dwm);
    if (ctas < 0) return ctas;
    // did the forward of these points deposit the X sections of the records into THIS scratch?
    bool x_dep = false;
    if (dwm && pts->x_scratch == scratch && pts->x_scratch_bytes == scratch_bytes && pts->x_mask == dwm &&
        x_deposit_ok(*pts, sp, (sp.pts.P + 15) / 16)) {
        x_dep = true;
        ctas = stage_tasks(*scene, stage, true, sp, nullptr, nullptr, nullptr,
                           (flags & NSB_GRAD_POINT) ? d_point : nullptr, sp.pts.P, dwm, dwm);
        if (ctas < 0) return ctas;
    }
    // tcgen05 backward: every decoder of the launch has its ReLU signs from the forward, and the ones whose weights
    // train found their layer inputs deposited (nothing to recompute); otherwise the mma.sync kernel
    bool tc_bwd = g_use_tc && g_use_tc_bwd && sp.pts.relu != nullptr && (dwm == 0 || x_dep);
    for (int i = 0; i < sp.ntask; ++i)
        if (sp.task[i].kind == KIND_COARSE || !sp.task[i].relu) tc_bwd = false;
    if (tc_bwd) {
        ctas = stage_tasks(*scene, stage, true, sp, nullptr, nullptr, nullptr,
                           (flags & NSB_GRAD_POINT) ? d_point : nullptr, sp.pts.P, dwm, x_dep ? dwm : 0u, true);
        if (ctas < 0) return ctas;
    }
    sp.d_raw = d_raw;
    sp.flags = flags;
    if (tc_bwd && getenv("NSB_DEBUG_NO_SCATTER")) sp.flags |= 0x80000000u;      // timing experiments: wrong gradients
    if (sp.pts.marks && !tc_bwd)      // this path does not mark the lines it scatters into: the exchange has to scan
        cudaMemsetAsync(sp.pts.marks, 0, sizeof(unsigned), (cudaStream_t)stream);
    const long long ntiles = (sp.pts.P + 15) / 16;
    // decoders whose parameters want a gradient
    int ndw = 0;
    if (flags & NSB_GRAD_WEIGHT)
        for (int i = 0; i < sp.ntask; ++i)
            if (dweight_mask & (1u << sp.task[i].kind)) {
                if (!d_weight_host || !d_weight_host[sp.task[i].kind])
                    return set_error(NSB_EINVAL, "d_weight[%d] is NULL", sp.task[i].kind);
                ++ndw;
            }
    long long chunk = ntiles;
    if (ndw) {
        if (!scratch) return set_error(NSB_EINVAL, "weight gradients need a scratch buffer");
        const int64_t share = (scratch_bytes - 256) / ndw;
        chunk = share / kDwTileBytes;
        if (chunk < 1) return set_error(NSB_EINVAL, "scratch too small: %lld bytes for %d decoders",
                                        (long long)scratch_bytes, ndw);
        if (chunk >= ntiles) chunk = ntiles;
        else if (chunk > kDwChunkTiles) chunk = kDwChunkTiles;
        // carve: [records of decoder 0 | records of decoder 1 | ... | flags ...]
        char* base = (char*)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
        int k = 0;
        for (int i = 0; i < sp.ntask; ++i)
            if (dweight_mask & (1u << sp.task[i].kind)) {
                sp.task[i].dw_scratch = (float*)(base + (size_t)k * chunk * REC_MAX * 4);
                sp.task[i].dw_flags = (unsigned char*)(base + (size_t)ndw * chunk * REC_MAX * 4 + (size_t)k * chunk);
                ++k;
            }
    }
    const int smem = max_blob(sp, true) * 4;
#ifdef NSB_TC_PROFILE
    static long long* bprof_dev = nullptr;
    if (!bprof_dev) cudaMalloc(&bprof_dev, 72 * sizeof(long long));
    cudaMemsetAsync(bprof_dev, 0, 72 * sizeof(long long), (cudaStream_t)stream);
    sp.prof = bprof_dev;
#endif
    const bool pdl = pdl_take();
    if ((defer_dw || only_dw) && ndw && chunk != ntiles)
        return set_error(NSB_EINVAL, "two-phase backward needs a scratch that holds the whole batch");
    for (long long t0 = 0; t0 < ntiles; t0 += chunk) {
        sp.tile_begin = t0;
        sp.tile_end = t0 + chunk < ntiles ? t0 + chunk : ntiles;
        sp.ctr = next_counters();
        if (!only_dw) timing_mark(1, 0, (cudaStream_t)stream);
        if (!only_dw) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(tc_bwd ? TCB_THREADS : THREADS);
            cfg.dynamicSmemBytes = tc_bwd ? (size_t)TbSmem<KIND_FINE>::S_END * 4 : (size_t)smem;
            cfg.stream = (cudaStream_t)stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = (pdl && t0 == 0) ? 1 : 0;
            if (tc_bwd) cudaLaunchKernelEx(&cfg, k_decoders_backward_tc, sp);
            else cudaLaunchKernelEx(&cfg, k_decoders_backward, sp);
        }
        if (!only_dw) {
            timing_mark(1, 1, (cudaStream_t)stream);
            if (int rc = check_launch(tc_bwd ? "k_decoders_backward_tc" : "k_decoders_backward")) return rc;
        }
#ifdef NSB_TC_PROFILE
        if (getenv("NSB_TC_PROFILE_PRINT")) {
            long long h[72];
            cudaMemcpy(h, bprof_dev, sizeof(h), cudaMemcpyDeviceToHost);
            const char* nm[12] = {"setup", "top", "prologue", "prol barrier", "issue", "wait mma", "epilogue", "barrier",
                                  "wait de", "dc+coord", "scatter", "de+dp"};
            for (int t = 0; t < sp.ntask && tc_bwd; ++t)
                for (int w = 0; w < 2; ++w) {
                    fprintf(stderr, "[tcbprof] task %d kind %d ctas %d %s:", t, sp.task[t].kind, sp.task[t].cta_count, w ? "thread200" : "thread0");
                    for (int k = 0; k < 12; ++k) fprintf(stderr, " %s=%lld", nm[k], h[(t * 2 + w) * 12 + k]);
                    fprintf(stderr, "\n");
                }
            for (int t = 0; t < sp.ntask && !tc_bwd; ++t)
                fprintf(stderr, "[bwdprof] task %d kind %d ctas %d: slowest warp %lld cycles, mean %lld, warps %lld\n", t,
                        sp.task[t].kind, sp.task[t].cta_count, h[36 + 3 * t], h[36 + 3 * t + 1] / (h[36 + 3 * t + 2] ? h[36 + 3 * t + 2] : 1),
                        h[36 + 3 * t + 2]);
        }
#endif
        for (int i = 0; i < sp.ntask && ndw && !defer_dw; ++i) {
            const DecTask& T = sp.task[i];
            if (!T.dw_scratch) continue;
            DwParams dp;
            dp.scratch = T.dw_scratch;
            dp.flags = T.dw_flags;
            dp.ntiles = sp.tile_end - sp.tile_begin;
            dp.tab = g_dw_tables_dev + T.kind;
            dp.dflat = d_weight_host[T.kind];
            // every CTA takes at most DW_LIVE_MAX live tiles (its share is <= ntiles / grid + 1)
            long long grid = dp.ntiles < g_num_sms ? dp.ntiles : g_num_sms;
            const long long need = (dp.ntiles + DW_LIVE_MAX - 2) / (DW_LIVE_MAX - 1);
            if (grid < need) grid = need;
            if (dw_cap > 0 && grid > dw_cap && dw_cap >= need) grid = dw_cap;
            timing_mark(2, 0, (cudaStream_t)stream);
            const bool dw_tc = g_use_tc && g_use_dw_tc && (T.kind == KIND_COLOR || T.kind == KIND_MIDDLE);
            if (dw_tc) {
                // one wave whatever the batch: a CTA walks its strided share of the records, accumulators stay in TMEM
                grid = dp.ntiles < g_num_sms ? dp.ntiles : g_num_sms;
                if (dw_cap > 0 && grid > dw_cap) grid = dw_cap;
                DwTcParams tp;
                tp.p = dp;
                tp.kind = T.kind;
                static const int dbg = getenv("NSB_DWT_DEBUG") ? atoi(getenv("NSB_DWT_DEBUG")) : 0;
                tp.debug = dbg;
                tp.off = T.kind == KIND_COLOR ? dw_tc_offsets<KIND_COLOR>() : dw_tc_offsets<KIND_MIDDLE>();
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3((unsigned)((grid + DWT_CLUSTER - 1) / DWT_CLUSTER * DWT_CLUSTER));   // idle CTAs add zeros
                cfg.blockDim = dim3(DWT_THREADS);
                cfg.dynamicSmemBytes = (size_t)DwTcSmem<KIND_COLOR>::S_END * 4 + 1024;
                cfg.stream = (cudaStream_t)stream;
                cudaLaunchAttribute at[2];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = DWT_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[1].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = DWT_CLUSTER > 1 ? at : at + 1;
                cfg.numAttrs = (DWT_CLUSTER > 1 ? 1 : 0) + (pdl_enabled() ? 1 : 0);
                cudaLaunchKernelEx(&cfg, k_dweight_tc, tp);
            } else {
                launch_pdl(k_dweight, dim3((unsigned)grid), dim3(THREADS), (size_t)(DW_HEAD_FLOATS + 2 * g_dw_host[T.kind].rec) * 4, (cudaStream_t)stream, dp);
            }
            timing_mark(2, 1, (cudaStream_t)stream);
            if (int rc = check_launch(dw_tc ? "k_dweight_tc" : "k_dweight")) return rc;
        }
    }
    return 0;
}
