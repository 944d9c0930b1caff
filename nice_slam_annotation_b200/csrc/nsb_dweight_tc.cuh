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
//  * the bias gradients (column sums of G) fall out of the split warps' loads; the output layer and d(B) of the
//    embedding are fp32 FMAs of four more warps; all kept in registers for the whole kernel;
//  * warp roles: warp 0 = producer (bulk copies into a ring of records), warp 1 = MMA issuer, warps 2..19 = split
//    (one atom each), warps 20..23 = small products;
//  * flush: every CTA assembles its partial gradient in flat parameter order in shared memory and adds it to global
//    memory with coalesced reductions, each CTA starting at a different element (148 CTAs reducing into the same
//    line at the same moment serialise in L2: scalar bias sums straight from registers cost 16 us that way).
//    Tried and dropped (DWT_CLUSTER = 4): adding the four images of a cluster up through distributed shared memory
//    first -- 37 instead of 148 reductions per element, but a cluster of four 217 KB CTAs needs four free SMs of one
//    GPC at once and cannot trickle in behind the backward kernel: the kernel went from 36 to 56 us.
// Decoders with a 32-channel feature input (middle, colour); the fine decoder (64) keeps k_dweight.
#pragma once
#include "nsb_tc.cuh"

namespace nsb {

constexpr int DWT_SPLIT_WARPS = 18;                // one per atom that needs a lo image (8 of X, 10 of G)
constexpr int DWT_SMALL_WARPS = 4;                 // d(B) (3 warps: one feature per lane), output layer (1 warp)
constexpr int DWT_WORKERS = 32 * (DWT_SPLIT_WARPS + DWT_SMALL_WARPS);
constexpr int DWT_THREADS = 64 + DWT_WORKERS;
constexpr int DWT_RING = 4;                        // records in flight or in use per CTA
constexpr int DWT_LO = 1;                          // lo image buffers
constexpr int DWT_D1 = 0, DWT_D2 = 128;            // TMEM columns
constexpr int DWT_CLUSTER = 1;                     // 4: cluster-wide pre-reduction of the flush (measured slower, see below)

struct DwTcOffs {          // flat-gradient offsets of the canonical parameters (floats)
    int w[5], b[5], wc[5], bc[5], wo, bo, emb, total;
};

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ float ld_cluster_f32(const float* local_smem, uint32_t cta) {
    uint32_t a;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(a) : "r"(tc::smem_u32(local_smem)), "r"(cta));
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(a) : "memory");
    return v;
}

template <int KIND>
struct DwTcSmem {          // float offsets in dynamic shared memory (base aligned to 1024 B)
    using RC = RecCfg<KIND>;
    static_assert(DecCfg<KIND>::C_IN == HID && DecCfg<KIND>::EMBED, "k_dweight_tc: 32-channel feature input");
    static constexpr int X_LO = 8 * RC::ATOM;                  // c, e0..e2, h1..h4
    static constexpr int G_LO = 10 * RC::ATOM;                 // d(a) slots 0..4, d(h1)..d(h5)
    static constexpr int S_RAW = 0;
    static constexpr int S_LO = S_RAW + DWT_RING * RC::REC;
    static constexpr int S_LIVE = S_LO + DWT_LO * (X_LO + G_LO);
    static constexpr int S_BAR = S_LIVE + 16;                  // mbarriers (8 B each)
    static constexpr int S_END = S_BAR + 2 * (2 * DWT_RING + DWT_LO) + 2;
    static_assert(RC::REC % 128 == 0, "records keep the 512-byte atom alignment");
    static_assert(RC::R_H == RC::R_C + 4 * RC::ATOM && RC::R_DH == RC::R_DA + 5 * RC::ATOM, "operand stacks are adjacent");
};

__device__ __forceinline__ uint64_t dwt_desc(uint32_t saddr) {        // MN-major atom stack, see the header
    return tc::smem_desc(saddr, 2048u, 512u) | (1ull << 61);
}
__host__ __device__ constexpr uint32_t dwt_idesc(int N) { return tc::idesc_tf32(128, N) | (1u << 15) | (1u << 16); }

using tc::bulk_g2s;
using tc::mbar_arrive;
using tc::mbar_expect_tx;

template <int KIND>
__device__ void dweight_tc_body(const DwParams& P, const DwTcOffs& O, float* sm, int debug) {
    using C = DecCfg<KIND>;
    using RC = RecCfg<KIND>;
    using S = DwTcSmem<KIND>;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __shared__ uint32_t s_tmem;
    __shared__ int s_nlive;
    __shared__ int s_warp_cnt[DWT_THREADS / 32];
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + S::S_BAR);
    uint64_t* full = bars;                   // [RING] record landed
    uint64_t* done = bars + DWT_RING;        // [RING] the MMAs that read the record (and its lo image) have completed
    uint64_t* lo_ready = bars + 2 * DWT_RING;   // [LO] lo image written, small products have read the record
    if (warp == 1) tc::tmem_alloc<512>(&s_tmem);
    if (threadIdx.x == 0) {
        for (int i = 0; i < DWT_RING; ++i) { tc::mbar_init(full + i, 1); tc::mbar_init(done + i, 1); }
        for (int i = 0; i < DWT_LO; ++i) tc::mbar_init(lo_ready + i, DWT_WORKERS);
        tc::mbar_init_fence();
    }
    pdl_wait();                              // flags and records come from the kernels before
    // CTA b takes the records T - 1 - (b + k G), k = 0, 1, ..., that are live: strided, so that all CTAs work on the same
    // age band at any time (and get statistically equal shares), and newest first -- the backward wrote its sections
    // in tile order, the last ones are the likeliest to be still in L2.  Only the count is needed up front; the
    // producer warp finds the live ones as it goes (32 flags per ballot).
    const long long T_ = P.ntiles;
    const int G_ = (int)gridDim.x, b_ = (int)blockIdx.x;
    const int cand = T_ > b_ ? (int)((T_ - 1 - b_) / G_) + 1 : 0;
    {
        int mine = 0;
        for (int k = threadIdx.x; k < cand; k += DWT_THREADS)
            if (P.flags[T_ - 1 - (b_ + (long long)k * G_)] != 0) ++mine;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(FULL, mine, d);
        if (lane == 0) s_warp_cnt[warp] = mine;
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (threadIdx.x == 0) {
        int total = 0;
        for (int w = 0; w < DWT_THREADS / 32; ++w) total += s_warp_cnt[w];
        s_nlive = total;
    }
    __syncthreads();
    const int nlive = s_nlive;
    const uint32_t tbase = s_tmem;
    float* raw = sm + S::S_RAW;
    float* lo = sm + S::S_LO;
    constexpr uint32_t REC_BYTES = RC::REC * 4;

    // the flat partial gradient of this CTA is assembled over the record ring once the last MMAs have completed
    float* img = raw;
    auto finish_loops = [&]() {
        if (nlive > 0) {
            const int il = nlive - 1;
            tc::mbar_wait(done + il % DWT_RING, (uint32_t)((il / DWT_RING) & 1));
            tc::fence_after_sync();
        }
        __syncthreads();
        for (int i = threadIdx.x; i < O.total; i += DWT_THREADS) img[i] = 0.f;
        __syncthreads();
    };
    if (warp == 0) {
        // ---- producer (whole warp: 32 candidate flags per ballot, lane 0 issues the copies) --------------------------
        int k0 = 0;                                  // first candidate of the current 32-candidate window
        unsigned live_bits = 0u;                     // live candidates of the window not yet taken
        bool have_window = false;
        for (int i = 0; i < nlive; ++i) {
            while (live_bits == 0u) {
                if (have_window) k0 += 32;
                have_window = true;
                const int k = k0 + lane;
                const bool lv = k < cand && P.flags[T_ - 1 - (b_ + (long long)k * G_)] != 0;
                live_bits = __ballot_sync(FULL, lv);
            }
            const int bit = __ffs(live_bits) - 1;
            live_bits &= live_bits - 1u;
            const long long rec_index = T_ - 1 - (b_ + (long long)(k0 + bit) * G_);
            if (lane == 0) {
                const int s = i % DWT_RING;
                if (i >= DWT_RING) tc::mbar_wait(done + s, (uint32_t)((i / DWT_RING - 1) & 1));
                mbar_expect_tx(full + s, REC_BYTES);
                const char* src = reinterpret_cast<const char*>(P.scratch + (size_t)rec_index * RC::REC);
                char* dst = reinterpret_cast<char*>(raw + s * RC::REC);
                constexpr uint32_t PART = REC_BYTES / 3;
                static_assert(PART % 16 == 0 && PART * 3 == REC_BYTES, "bulk copy granularity");
#pragma unroll
                for (int c = 0; c < 3; ++c) bulk_g2s(dst + c * PART, src + c * PART, PART, full + s);
            }
            __syncwarp();
        }
        __syncwarp();
        finish_loops();
    } else if (warp == 1) {
        // ---- MMA issuer ----------------------------------------------------------------------------------------
        if (tc::elect_one()) {
            for (int i = 0; i < nlive; ++i) {
                const int s = i % DWT_RING, b = i % DWT_LO;
                tc::mbar_wait(full + s, (uint32_t)((i / DWT_RING) & 1));
                tc::mbar_wait(lo_ready + b, (uint32_t)((i / DWT_LO) & 1));
                tc::fence_after_sync();
                const uint32_t r0 = tc::smem_u32(raw + s * RC::REC), l0 = tc::smem_u32(lo + b * (S::X_LO + S::G_LO));
                const uint32_t x2h = r0 + 4u * RC::R_C, x1h = r0 + 4u * RC::R_H, g1h = r0 + 4u * RC::R_DA,
                               g2h = r0 + 4u * (RC::R_DA + 3 * RC::ATOM);
                const uint32_t x2l = l0, x1l = l0 + 4u * (4 * RC::ATOM), g1l = l0 + 4u * S::X_LO,
                               g2l = l0 + 4u * (S::X_LO + 3 * RC::ATOM);
#pragma unroll
                for (int ks = 0; ks < ((debug & 4) ? 0 : 2); ++ks) {
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
        finish_loops();
    } else if (warp < 2 + DWT_SPLIT_WARPS) {
        // ---- split warps: warp w owns atom w of the lo image (X atoms 0..7 = c, e0..e2, h1..h4; G atoms 8..17 = the
        // d(a) slots and d(h1)..d(h5)).  Lane (pc, rm) = (lane & 7, lane >> 3) handles the 16-byte piece at physical
        // position pc of the four rows r = rm + 4 j: under the record's swizzle these hold the SAME four channels
        // (logical piece pc ^ (rm << 1)), so the bias gradients -- column sums of the G atoms -- fall out of the same
        // loads as four running sums per lane.
        const int aw = warp - 2, pc = lane & 7, rm = lane >> 3;
        const bool is_g = aw >= 8;
        const int src_off = is_g ? RC::R_DA + (aw - 8) * RC::ATOM : RC::R_C + aw * RC::ATOM;
        const int piece = rm * HID + 4 * pc;                       // float offset of my piece in row rm
        float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int i = 0; i < nlive; ++i) {
            const int s = i % DWT_RING, b = i % DWT_LO;
            tc::mbar_wait(full + s, (uint32_t)((i / DWT_RING) & 1));
            const float* R = raw + s * RC::REC + src_off + piece;
            float4 x[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = *reinterpret_cast<const float4*>(R + 4 * j * HID);
            if (i >= DWT_LO) tc::mbar_wait(done + (i - DWT_LO) % DWT_RING, (uint32_t)(((i - DWT_LO) / DWT_RING) & 1));   // lo slot b is free
            float* L = lo + b * (S::X_LO + S::G_LO) + aw * RC::ATOM + piece;
            if (!(debug & 2)) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float4 l;
                    l.x = x[j].x - __uint_as_float(__float_as_uint(x[j].x) & 0xFFFFE000u);
                    l.y = x[j].y - __uint_as_float(__float_as_uint(x[j].y) & 0xFFFFE000u);
                    l.z = x[j].z - __uint_as_float(__float_as_uint(x[j].z) & 0xFFFFE000u);
                    l.w = x[j].w - __uint_as_float(__float_as_uint(x[j].w) & 0xFFFFE000u);
                    *reinterpret_cast<float4*>(L + 4 * j * HID) = l;
                    bsum.x += x[j].x; bsum.y += x[j].y; bsum.z += x[j].z; bsum.w += x[j].w;
                }
            }
            tc::fence_async_smem();
            mbar_arrive(lo_ready + b);
        }
        // ---- my shares of the flat partial gradient (image in shared memory, see below) ------------------------------
        finish_loops();
        if (is_g) {
            // bias gradients: the four row classes of a logical piece sit in four lanes, each adds its own share
            const int g = aw - 8;                                   // G atom: d(a) slot g (< 5) or d(h_{g-4})
            int out;
            if (g < 5) { const int layer = g == 4 ? 0 : (g == 3 ? 3 : (g == 2 ? 4 : g + 1)); out = O.b[layer]; }
            else out = O.bc[g - 5];
            const int ch = 4 * (pc ^ (rm << 1));
            atomicAdd(img + out + ch, bsum.x);
            atomicAdd(img + out + ch + 1, bsum.y);
            atomicAdd(img + out + ch + 2, bsum.z);
            atomicAdd(img + out + ch + 3, bsum.w);
        }
        if (nlive > 0 && warp < 10) {
            const int q = warp & 3;                                   // TMEM lane quadrant of this warp
            const uint32_t lane_base = tc::tmem_addr(tbase, 32 * q, 0);
            auto put_block = [&](int col0, int out0, int ld_n, bool row_ok) {
                // 32 columns (n) of this thread's row: img[out0 + n * ld_n]
#pragma unroll
                for (int c = 0; c < 32; c += 16) {
                    float v[16];
                    tc::tmem_ld16(lane_base + col0 + c, v);
                    tc::tmem_ld_wait();
                    if (row_ok) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) img[out0 + (c + j) * ld_n] = v[j];
                    }
                }
            };
            if (warp < 6) {
                // D1: rows of quadrant q are h_{q+1}; its layer's gradient block (columns: da1, da2, da4, da3)
                const int layer = q + 1;
                const int col0 = DWT_D1 + (layer == 3 ? 96 : (layer == 4 ? 64 : 32 * q));
                put_block(col0, O.w[layer] + (layer == 3 ? EMB : 0) + lane, C::in_dim(layer), true);
            } else {
                if (q == 0) {                                          // rows c: the five fc_c
                    for (int i = 0; i < 5; ++i) put_block(DWT_D2 + 64 + 32 * i, O.wc[i] + lane, C::C_IN, true);
                } else {                                               // rows e (features 32 (q - 1) + lane)
                    const int f = 32 * (q - 1) + lane;
                    put_block(DWT_D2, O.w[3] + f, C::in_dim(3), f < EMB);
                    put_block(DWT_D2 + 32, O.w[0] + f, C::in_dim(0), f < EMB);
                }
            }
        }
    } else {
        // ---- small products in fp32, one output row per lane, accumulated in registers over all records ------------
        const int sw_ = warp - 2 - DWT_SPLIT_WARPS;                 // 0..2: d(B) features 32 sw_ + lane; 3: output layer
        float acc[4] = {0.f, 0.f, 0.f, 0.f}, accb[4] = {0.f, 0.f, 0.f, 0.f};
        for (int i = 0; i < nlive; ++i) {
            const int s = i % DWT_RING, b = i % DWT_LO;
            tc::mbar_wait(full + s, (uint32_t)((i / DWT_RING) & 1));
            const float* R = raw + s * RC::REC;
            if (!(debug & 1)) {
                if (sw_ < 3) {        // d(B)[a][f] = sum_r p[r][a] * (d(e) cos)[r][f]
                    const float* V = R + RC::R_DE + sw_ * RC::ATOM;
#pragma unroll
                    for (int r = 0; r < 16; ++r) {
                        const float v = V[r * HID + (lane ^ ((r & 3) << 3))];
                        const float4 p = *reinterpret_cast<const float4*>(R + RC::R_P + r * 8);
                        acc[0] = fmaf(p.x, v, acc[0]); acc[1] = fmaf(p.y, v, acc[1]); acc[2] = fmaf(p.z, v, acc[2]);
                    }
                } else {              // Wo[o][m] = sum_r d(out)[r][o] * h5[r][m]; bo[o] = sum_r d(out)[r][o]
                    const float* V = R + RC::R_H + 4 * RC::ATOM;
#pragma unroll
                    for (int r = 0; r < 16; ++r) {
                        const int sw = (r & 3) << 3;
                        const float v = V[r * HID + (lane ^ sw)];
                        const float4 d = *reinterpret_cast<const float4*>(R + RC::R_DO + r * HID + sw);   // outputs 0..3
                        acc[0] = fmaf(d.x, v, acc[0]); acc[1] = fmaf(d.y, v, acc[1]);
                        acc[2] = fmaf(d.z, v, acc[2]); acc[3] = fmaf(d.w, v, acc[3]);
                        accb[0] += d.x; accb[1] += d.y; accb[2] += d.z; accb[3] += d.w;
                    }
                }
            }
            mbar_arrive(lo_ready + b);                              // done reading the record
        }
        finish_loops();
        if (sw_ < 3) {
            const int f = 32 * sw_ + lane;
            if (f < EMB)
#pragma unroll
                for (int a = 0; a < 3; ++a) img[O.emb + a * EMB + f] = acc[a];
        } else {
#pragma unroll
            for (int o = 0; o < C::NOUT; ++o) {
                img[O.wo + o * HID + lane] = acc[o];
                if (lane == 0) img[O.bo + o] = accb[o];
            }
        }
    }
    // ---- flush: cluster-wide sum of the partial gradients, one coalesced reduction per element -------------------------
    tc::fence_before_sync();
    __syncthreads();
    if constexpr (DWT_CLUSTER > 1) cluster_sync_all();
    {
        const int per = (O.total + DWT_CLUSTER - 1) / DWT_CLUSTER;
        const int beg = DWT_CLUSTER > 1 ? (int)cluster_ctarank() * per : 0, end = beg + per < O.total ? beg + per : O.total;
        // CTAs start at different elements: 148 CTAs reducing into the same line at the same moment serialise in L2
        const int span = end - beg, rot = (int)(((long long)blockIdx.x * span) / gridDim.x) & ~31;
        for (int k = (int)threadIdx.x; k < span; k += DWT_THREADS) {
            int idx = k + rot;
            idx = beg + (idx >= span ? idx - span : idx);
            float v = 0.f;
            if constexpr (DWT_CLUSTER > 1) {
#pragma unroll
                for (uint32_t c = 0; c < DWT_CLUSTER; ++c) v += ld_cluster_f32(img + idx, c);
            } else {
                v = img[idx];
            }
            if (v != 0.f && !(debug & 8)) atomicAdd(P.dflat + idx, v);
        }
    }
    if constexpr (DWT_CLUSTER > 1) cluster_sync_all();     // nobody leaves while a neighbour still reads its image
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<512>(tbase);
}

struct DwTcParams {
    DwParams p;
    DwTcOffs off;
    int kind;
    int debug;      // NSB_DWT_DEBUG (timing experiments, wrong results): 1 no small products, 2 no lo pass, 4 no MMAs, 8 no flush
};

__global__ void __launch_bounds__(DWT_THREADS, 1) k_dweight_tc(const DwTcParams P) {
    extern __shared__ __align__(16) float dwt_sm_raw[];
    // the operand atoms need their 512-byte alignment: align the base by hand (the launch asks for 1 KB of slack)
    float* dwt_sm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(dwt_sm_raw) + 1023) & ~(uintptr_t)1023);
    pdl_launch_dependents();
    if (P.kind == KIND_COLOR) dweight_tc_body<KIND_COLOR>(P.p, P.off, dwt_sm, P.debug);
    else dweight_tc_body<KIND_MIDDLE>(P.p, P.off, dwt_sm, P.debug);
}

}  // namespace nsb
