// k_gemm3x: chained / batched 3xTF32 GEMM on tcgen05 (see nsb_gemm.cuh for the contract).
//
// CTA (GemmCfg): epilogue warps (TMEM lane quadrant = warp & 3; with eight of them two share a quadrant and
// take alternate 16-column groups), one MMA-issuing warp (one thread), in PACKED mode a warp that fetches the
// packed B images with bulk asynchronous copies (one thread: it runs ahead of the chain dependency the
// producers wait for, so the weights of the next item are in place when its rows become readable), and
// eight producer warps in four groups.  Shared memory: a ring of 4 stages x { A hi, A lo (128 x 16), B hi, B lo
// (256 x 16) } = 4 x 48 KB of operand images (MN mode: 8 stages of K = 8, two per producer group).  Tensor
// memory: two 256-column accumulators, so the epilogue of one item overlaps the MMAs of the next.
// What the structure answers to (profiles/r02zz_gemm3x_ablation.txt): the tensor pipe, L2 and DRAM were never
// the limit of this kernel -- the instruction streams of the producer warps and the load/store unit were.
#include "nsb_gemm.cuh"

#include <cstdlib>

#include "nsb_common.cuh"
#include "nsb_tc.cuh"

namespace nsb {
namespace {

constexpr int G_STAGES = 4;
constexpr int G_KC = 16;                                          // k per stage (two MMA K steps)
constexpr int G_A_FLOATS = 128 * G_KC;                            // one image (hi or lo) of A
constexpr int G_B_FLOATS = 256 * G_KC;
constexpr int G_STAGE_FLOATS = 2 * G_A_FLOATS + 2 * G_B_FLOATS;   // 48 KB
constexpr int G_PROD_WARPS = 8;
constexpr int G_SMEM = G_STAGES * G_STAGE_FLOATS * 4 + 1024;
// MN mode (weight gradients) cuts the same 192 KB into eight stages of K = 8, two per producer group, so that a group
// has the copies of its next chunk in flight while it finishes the current one
constexpr int MN_STAGES = 8, MN_KC = 8;
constexpr int MN_A_FLOATS = 128 * MN_KC, MN_B_FLOATS = 256 * MN_KC;
constexpr int MN_STAGE_FLOATS = 2 * MN_A_FLOATS + 2 * MN_B_FLOATS;   // 24 KB
constexpr int G_MAX_STAGES = MN_STAGES;
static_assert(MN_STAGES * MN_STAGE_FLOATS == G_STAGES * G_STAGE_FLOATS, "both rings fill the same shared memory");

// How the operands of a batch reach shared memory: one kernel instantiation each (all layers of a batch agree), so
// that the register budget of one path is not set by the others.
//   LEGACY  K-major images, both operands split by the producers through registers (any addressing)
//   PACKED  K-major images, B's pre-split images fetched by bulk copies (a dedicated warp), A by the producers
//   MN      both operands row-contiguous: MN-major images
enum : int { G_MODE_LEGACY = 0, G_MODE_PACKED = 1, G_MODE_MN = 2 };
template <int MODE>
struct GemmCfg {
    static constexpr int EPI_WARPS = MODE == G_MODE_MN ? 4 : 8;      // weight gradients: one epilogue per 128 chunks
    static constexpr int MMA_WARP = EPI_WARPS;
    static constexpr int BULK_WARP = MODE == G_MODE_PACKED ? EPI_WARPS + 1 : -1;
    static constexpr int PROD_WARP0 = EPI_WARPS + (MODE == G_MODE_PACKED ? 2 : 1);
    static constexpr int THREADS = (PROD_WARP0 + G_PROD_WARPS) * 32;
    static constexpr int EPI_SPLIT = EPI_WARPS / 4;                  // warps sharing a TMEM lane quadrant
};
inline bool layer_mn_capable(const GemmLayer& L) { return L.a_rc && L.b_rc && (L.N & 31) == 0; }

struct Item {
    int layer;
    int dep;              // the item whose output this one reads is `dep` items back (chain), else 2
    long long m0, k0, k1;
};

// The sequence of work items of one CTA; every role (producer, MMA, epilogue) walks the same one.
struct ItemIter {
    int b, G;
    int n_i, i0, l, h;        // chain: tiles of this CTA, current pair, layer, half
    long long g, total;       // independent: global item index

    __device__ void init(const GemmBatch& gb) {
        b = blockIdx.x; G = gridDim.x;
        n_i = gb.m_tiles > b ? (gb.m_tiles - b + G - 1) / G : 0;
        i0 = 0; l = 0; h = 0;
        g = b; total = 0;
        if (!gb.chain)
            for (int i = 0; i < gb.n_layers; ++i) total += (long long)gb.L[i].m_tiles * gb.L[i].n_split;
    }
    __device__ bool next(const GemmBatch& gb, Item& it) {
        if (gb.chain) {
            for (;;) {
                if (i0 >= n_i) return false;
                const int np = n_i - i0 < 2 ? n_i - i0 : 2;
                if (l < gb.n_layers) {
                    if (h < np) {
                        it.layer = l; it.dep = np;
                        it.m0 = (long long)(b + (i0 + h) * G) * 128;
                        it.k0 = 0; it.k1 = gb.L[l].K;
                        ++h;
                        return true;
                    }
                    h = 0; ++l;
                } else {
                    l = 0; i0 += 2;
                }
            }
        }
        while (g < total) {
            long long r = g;
            int li = 0;
            while (r >= (long long)gb.L[li].m_tiles * gb.L[li].n_split) { r -= (long long)gb.L[li].m_tiles * gb.L[li].n_split; ++li; }
            const GemmLayer& L = gb.L[li];
            // the row tiles of one K split are neighbours in the item order: CTAs b and b + 1 read the same rows of B at
            // the same time, so the second read is served by L2
            const long long mt = r % L.m_tiles, sp = r / L.m_tiles;
            g += G;
            it.layer = li; it.dep = 2;
            it.m0 = mt * 128;
            it.k0 = sp * L.k_per_split;
            it.k1 = it.k0 + L.k_per_split < L.K ? it.k0 + L.k_per_split : L.K;
            if (it.k1 > it.k0) return true;
        }
        return false;
    }
};

using tc::mbar_arrive;
// plain (coherent) global loads: in a chain the A operand was written by this CTA moments ago
__device__ __forceinline__ float ld_f32(const float* p) {
    float v;
    asm volatile("ld.global.f32 %0, [%1];\n" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_v4(const float* p) {
    float4 v;
    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// four consecutive k of row `row` (zeros outside [.., rows_valid) x [.., k_end))
__device__ __forceinline__ float4 load_quad(const float* base, long long ld, int rc, bool vec_ok, long long row,
                                            long long rows_valid, long long k, long long k_end) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row >= rows_valid || k >= k_end) return v;
    if (rc == 0) {
        const float* p = base + row * ld + k;
        if (vec_ok && k + 3 < k_end) return ld_v4(p);
        v.x = ld_f32(p);
        if (k + 1 < k_end) v.y = ld_f32(p + 1);
        if (k + 2 < k_end) v.z = ld_f32(p + 2);
        if (k + 3 < k_end) v.w = ld_f32(p + 3);
    } else {
        const float* p = base + k * ld + row;
        v.x = ld_f32(p);
        if (k + 1 < k_end) v.y = ld_f32(p + ld);
        if (k + 2 < k_end) v.z = ld_f32(p + 2 * ld);
        if (k + 3 < k_end) v.w = ld_f32(p + 3 * ld);
    }
    return v;
}

__device__ __forceinline__ void split_store(float* hi_img, float* lo_img, int idx, const float4& v) {
    uint32_t h0, h1, h2, h3, l0, l1, l2, l3;
    split_tf32(v.x, h0, l0); split_tf32(v.y, h1, l1); split_tf32(v.z, h2, l2); split_tf32(v.w, h3, l3);
    *reinterpret_cast<uint4*>(hi_img + idx) = make_uint4(h0, h1, h2, h3);
    *reinterpret_cast<uint4*>(lo_img + idx) = make_uint4(l0, l1, l2, l3);
}

__device__ __forceinline__ bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// row-contiguous operand: four consecutive rows r .. r + 3 at reduction index k (zeros outside)
__device__ __forceinline__ float4 load_rquad(const float* base, long long ld, bool vec_ok, long long r, long long rows_valid,
                                             long long k, long long k_end) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r >= rows_valid || k >= k_end) return v;
    const float* p = base + k * ld + r;
    if (vec_ok && r + 3 < rows_valid) return ld_v4(p);
    v.x = ld_f32(p);
    if (r + 1 < rows_valid) v.y = ld_f32(p + 1);
    if (r + 2 < rows_valid) v.z = ld_f32(p + 2);
    if (r + 3 < rows_valid) v.w = ld_f32(p + 3);
    return v;
}

// 16-byte asynchronous copy global -> shared (LDGSTS), the last 16 - src_bytes bytes zero-filled
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const float* gmem_src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_dst), "l"(gmem_src), "r"(src_bytes) : "memory");
}
// lo image of four values already in the hi image (raw fp32 there: the tensor core truncates hi itself)
__device__ __forceinline__ float4 lo_pass(const float* hi_img, float* lo_img, int idx) {
    const float4 v = *reinterpret_cast<const float4*>(hi_img + idx);
    uint32_t h, l0, l1, l2, l3;
    split_tf32(v.x, h, l0); split_tf32(v.y, h, l1); split_tf32(v.z, h, l2); split_tf32(v.w, h, l3);
    *reinterpret_cast<uint4*>(lo_img + idx) = make_uint4(l0, l1, l2, l3);
    return v;
}

// B operand images of every layer, chunk by chunk: [chunk][hi N x 16 | lo N x 16], K-major core matrices,
// zero outside N_valid x K -- what the GEMM's bulk copies land in a stage as is
__global__ void __launch_bounds__(256) k_gemm_pack_b(const __grid_constant__ GemmBatch gb) {
    const GemmLayer& L = gb.L[blockIdx.y];
    const int N = L.N, per = N * G_KC;
    const long long chunks = (L.K + G_KC - 1) / G_KC;
    for (long long c = blockIdx.x; c < chunks; c += gridDim.x) {
        float* dst = L.b_packed + c * (2LL * per);
        for (int e = threadIdx.x; e < per; e += blockDim.x) {
            int n, kk;
            if (L.b_rc) { kk = e / N; n = e - kk * N; } else { n = e >> 4; kk = e & 15; }
            const long long k = c * G_KC + kk;
            float v = 0.f;
            if (n < L.N_valid && k < L.K) v = L.b_rc ? L.B[k * L.ldb + n] : L.B[(long long)n * L.ldb + k];
            uint32_t hi, lo;
            split_tf32(v, hi, lo);
            const int idx = tc::kmajor_index(n, kk, N);
            dst[idx] = __uint_as_float(hi);
            dst[per + idx] = __uint_as_float(lo);
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(GemmCfg<MODE>::THREADS, 1) k_gemm3x(const __grid_constant__ GemmBatch gb) {
    using Cfg = GemmCfg<MODE>;
    constexpr int G_EPI_WARPS = Cfg::EPI_WARPS, G_MMA_WARP = Cfg::MMA_WARP, G_BULK_WARP = Cfg::BULK_WARP,
                  G_PROD_WARP0 = Cfg::PROD_WARP0, G_THREADS = Cfg::THREADS, EPI_SPLIT = Cfg::EPI_SPLIT;
    constexpr bool PACKED = MODE == G_MODE_PACKED, MN = MODE == G_MODE_MN;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* stages = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar_full[G_MAX_STAGES], bar_empty[G_MAX_STAGES], bar_acc_full[2], bar_acc_empty[2];
    __shared__ __align__(8) uint64_t bar_epi_done[2];     // chain: the epilogue of an item has stored its rows
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(16) float s_bias[GEMM_MAX_LAYERS][256];
    __shared__ __align__(16) float s_stage[G_EPI_WARPS][32 * 16];      // epilogue transposition tiles

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        // a stage is full when its two producer warps have arrived (+ the bulk copy's expect_tx arrival)
        for (int s = 0; s < G_MAX_STAGES; ++s) { tc::mbar_init(&bar_full[s], 64 + ((PACKED && !(gb.debug & 8)) ? 1 : 0)); tc::mbar_init(&bar_empty[s], 1); }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&bar_acc_full[a], 1);
            tc::mbar_init(&bar_acc_empty[a], G_EPI_WARPS * 32);
            tc::mbar_init(&bar_epi_done[a], G_EPI_WARPS * 32);
        }
        tc::mbar_init_fence();
    }
    for (int i = tid; i < gb.n_layers * 256; i += G_THREADS) {
        const GemmLayer& L = gb.L[i >> 8];
        const int n = i & 255;
        s_bias[i >> 8][n] = (L.bias && n < L.N_valid) ? __ldg(L.bias + n) : 0.f;
    }
    if (warp == G_MMA_WARP) tc::tmem_alloc<512>(&tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tbase = tmem_slot;

    ItemIter iter;
    iter.init(gb);
    Item it;

    if (warp < G_EPI_WARPS) {
        // ------------------------------------------------ epilogue: TMEM -> registers -> shared (transpose) -> global
        // tcgen05.ld hands a thread 16 columns of ITS row; stored like that, a warp's request touches 32 different
        // 128-byte lines (16 bytes of each) and the load/store unit takes 32 passes per request -- measured as a third
        // of the forward chain.  So a 16-column group goes through a 32 x 16 staging tile (16-byte chunks XOR-swizzled,
        // conflict-free both ways) and leaves as 4 requests of 8 rows x 64 contiguous bytes; the ReLU-adjoint mask is
        // read and the split-K sums are reduced (red.global.add.v4) in the same coalesced shape.
        const int quad = warp & 3, half = warp >> 2;       // half < EPI_SPLIT
        float* T = s_stage[warp];
        const int t_st = lane * 16, t_sx = (lane >> 1) & 3;                 // writer: row = lane, chunk q -> q ^ t_sx
        const int q_ld = lane & 3;                                          // reader: row 8 j + (lane >> 2), chunk q_ld
        for (int i = 0; iter.next(gb, it); ++i) {
            const GemmLayer& L = gb.L[it.layer];
            const float* bias = s_bias[it.layer];
            const int acc = i & 1;
            const long long Mv = gb.chain ? gb.M : L.M;
            const long long m_base = it.m0 + 32 * quad + (lane >> 2);       // + 8 j: the rows this thread stores
            const bool cvec = (L.ldc & 3) == 0 && aligned16(L.C);
            const bool mvec = L.mask && (L.ldmask & 3) == 0 && aligned16(L.mask);
            const bool bias_epi = L.epi == GEMM_EPI_BIAS_RELU || L.epi == GEMM_EPI_BIAS;
            // ReLU-adjoint mask of a 16-column group (the saved activations): fetched one group ahead of the
            // accumulator, the first one before the MMAs of the item have even finished
            const bool mpre = L.epi == GEMM_EPI_MASK && mvec;
            float4 mk[4], mkn[4];
            auto load_mask = [&](int n0, float4 (&mm)[4]) {
                const int n = n0 + 4 * q_ld;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const long long m = m_base + 8 * j;
                    mm[j] = (mpre && n0 < L.N && n + 3 < L.N_valid && m < Mv) ? ld_v4(L.mask + m * L.ldmask + n)
                                                                              : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            };
            load_mask(16 * half, mk);
            tc::mbar_wait(&bar_acc_full[acc], (uint32_t)((i >> 1) & 1));
            tc::fence_after_sync();
            const uint32_t tacc = tc::tmem_addr(tbase, 32 * quad, acc * 256);
            if (16 * half >= L.N) {                         // no column group for this warp (N = 16)
                tc::fence_before_sync();
                mbar_arrive(&bar_acc_empty[acc]);
            }
            for (int n0 = 16 * half; n0 < L.N; n0 += 16 * EPI_SPLIT) {
                float v[16];
                load_mask(n0 + 16 * EPI_SPLIT, mkn);
                tc::tmem_ld16(tacc + n0, v);
                tc::tmem_ld_wait();
                if (n0 + 16 * EPI_SPLIT >= L.N) {           // my share of the accumulator is drained: hand it back
                    tc::fence_before_sync();
                    mbar_arrive(&bar_acc_empty[acc]);
                }
                if (bias_epi) {                             // per column: while the thread still holds a row
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 b4 = *reinterpret_cast<const float4*>(bias + n0 + 4 * q);     // zeros beyond N_valid
                        v[4 * q] += b4.x; v[4 * q + 1] += b4.y; v[4 * q + 2] += b4.z; v[4 * q + 3] += b4.w;
                    }
                    if (L.epi == GEMM_EPI_BIAS_RELU) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
                    }
                }
                __syncwarp();                               // the previous group has been read out of the tile
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    *reinterpret_cast<float4*>(T + t_st + 4 * (q ^ t_sx)) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                __syncwarp();
                const int n = n0 + 4 * q_ld;
                const bool full4 = n + 3 < L.N_valid;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = 8 * j + (lane >> 2);
                    const float4 t4 = *reinterpret_cast<const float4*>(T + r * 16 + 4 * (q_ld ^ ((r >> 1) & 3)));
                    const float4 mk_j = mk[j];
                    mk[j] = mkn[j];
                    const long long m = m_base + 8 * j;
                    if (m >= Mv || n >= L.N_valid || (gb.debug & 1)) continue;     // (debug 1: timing experiment, no global traffic)
                    float o[4] = {t4.x, t4.y, t4.z, t4.w};
                    float* cp = L.C + m * L.ldc + n;
                    if (L.epi == GEMM_EPI_MASK) {
                        if (mvec && full4) {
                            o[0] = mk_j.x > 0.f ? o[0] : 0.f; o[1] = mk_j.y > 0.f ? o[1] : 0.f;
                            o[2] = mk_j.z > 0.f ? o[2] : 0.f; o[3] = mk_j.w > 0.f ? o[3] : 0.f;
                        } else {
                            const float* mp = L.mask + m * L.ldmask + n;
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (n + e < L.N_valid) o[e] = ld_f32(mp + e) > 0.f ? o[e] : 0.f;
                        }
                    }
                    if (L.epi == GEMM_EPI_ATOMIC) {
                        if (cvec && full4) {
                            red_add_v4(cp, o[0], o[1], o[2], o[3]);
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (n + e < L.N_valid) atomicAdd(cp + e, o[e]);
                        }
                    } else if (cvec && full4) {
                        *reinterpret_cast<float4*>(cp) = make_float4(o[0], o[1], o[2], o[3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (n + e < L.N_valid) cp[e] = o[e];
                    }
                }
            }
            mbar_arrive(&bar_epi_done[acc]);       // release: this item's C rows are visible to the producers
        }
    } else if (warp == G_MMA_WARP) {
        // ------------------------------------------------ MMA issue (one thread)
        if (lane == 0) {
            uint32_t itc = 0;
            for (int i = 0; iter.next(gb, it); ++i) {
                const GemmLayer& L = gb.L[it.layer];
                const int acc = i & 1;
                tc::mbar_wait(&bar_acc_empty[acc], (uint32_t)(((i >> 1) & 1) ^ 1));
                tc::fence_after_sync();
                if constexpr (MN) {
                    // ring of eight K = 8 stages: one K step (three MMAs) per stage; rows beyond the reduction are zero
                    const uint32_t idesc = tc::idesc_tf32(128, L.N) | tc::IDESC_A_MN | tc::IDESC_B_MN;
                    const uint32_t dacc = tbase + (uint32_t)(acc * 256);
                    const int ns = (int)((it.k1 - it.k0 + MN_KC - 1) / MN_KC);
                    const uint32_t s0a = tc::smem_u32(stages);
                    const uint64_t dA = tc::mnmajor_desc(s0a, 1024u), dB = tc::mnmajor_desc(s0a + 2 * MN_A_FLOATS * 4, 1024u);
                    const uint64_t a_lo_off = (MN_A_FLOATS * 4) >> 4, b_lo_off = (uint64_t)(L.N * MN_KC * 4) >> 4;
                    for (int s = 0; s < ns; ++s, ++itc) {
                        const int st = itc % MN_STAGES;
                        tc::mbar_wait(&bar_full[st], (itc / MN_STAGES) & 1);
                        tc::fence_after_sync();
                        const uint64_t so = (uint64_t)st * ((MN_STAGE_FLOATS * 4) >> 4);
                        const uint64_t ah = dA + so, bh = dB + so;
                        tc::mma_ss(dacc, ah + a_lo_off, bh, idesc, s > 0);
                        tc::mma_ss(dacc, ah, bh + b_lo_off, idesc, true);
                        tc::mma_ss(dacc, ah, bh, idesc, true);
                        tc::mma_commit(&bar_empty[st]);
                    }
                    tc::mma_commit(&bar_acc_full[acc]);
                    continue;
                }
                constexpr bool mn = false;
                const uint32_t idesc = tc::idesc_tf32(128, L.N) | (mn ? (tc::IDESC_A_MN | tc::IDESC_B_MN) : 0u);
                const uint32_t dacc = tbase + (uint32_t)(acc * 256);
                const int ns = (int)((it.k1 - it.k0 + G_KC - 1) / G_KC);
                // operand descriptors of stage 0 at k = 0; the address field counts 16-byte units, so a stage, the lo
                // image and the second K step are plain additions (one thread issues everything: keep its loop short)
                const uint32_t s0a = tc::smem_u32(stages);
                const uint64_t dA = mn ? tc::mnmajor_desc(s0a) : tc::kmajor_desc(s0a, 128, 0);
                const uint64_t dB = mn ? tc::mnmajor_desc(s0a + 2 * G_A_FLOATS * 4) : tc::kmajor_desc(s0a + 2 * G_A_FLOATS * 4, L.N, 0);
                const uint64_t a_lo_off = (G_A_FLOATS * 4) >> 4, b_lo_off = (uint64_t)(L.N * G_KC * 4) >> 4;
                const uint64_t a_k8 = mn ? (1024 >> 4) : (2 * 16 * 128) >> 4;              // 8 k further
                const uint64_t b_k8 = mn ? (1024 >> 4) : (uint64_t)(2 * (L.N >> 3) * 128) >> 4;
                const int k_last = (int)(it.k1 - it.k0 - (long long)(ns - 1) * G_KC);     // valid k of the last chunk
                bool accum = false;
                for (int s = 0; s < ns; ++s, ++itc) {
                    const int st = itc % G_STAGES;
                    tc::mbar_wait(&bar_full[st], (itc / G_STAGES) & 1);
                    tc::fence_after_sync();
                    const uint64_t so = (uint64_t)st * ((G_STAGE_FLOATS * 4) >> 4);
                    const uint64_t ah = dA + so, bh = dB + so, al = ah + a_lo_off, bl = bh + b_lo_off;
                    tc::mma_ss(dacc, al, bh, idesc, accum);
                    tc::mma_ss(dacc, ah, bl, idesc, true);
                    tc::mma_ss(dacc, ah, bh, idesc, true);
                    accum = true;
                    if (s + 1 < ns || k_last > 8) {
                        tc::mma_ss(dacc, al + a_k8, bh + b_k8, idesc, true);
                        tc::mma_ss(dacc, ah + a_k8, bl + b_k8, idesc, true);
                        tc::mma_ss(dacc, ah + a_k8, bh + b_k8, idesc, true);
                    }
                    tc::mma_commit(&bar_empty[st]);        // stage reusable once these MMAs have read it
                }
                tc::mma_commit(&bar_acc_full[acc]);
            }
        }
        __syncwarp();
    } else if (PACKED && warp == G_BULK_WARP) {
        // ------------------------------------------------ packed B images: one bulk copy per stage (one thread)
        if (lane == 0 && !(gb.debug & 8)) {
            uint32_t itc = 0;
            for (int i = 0; iter.next(gb, it); ++i) {
                const GemmLayer& L = gb.L[it.layer];
                const uint32_t bytes = (uint32_t)L.N * G_KC * 2 * 4;
                const int ns = (int)((it.k1 - it.k0 + G_KC - 1) / G_KC);
                const float* src = L.b_packed + (it.k0 / G_KC) * (2LL * L.N * G_KC);
                float* dst = stages + 2 * G_A_FLOATS;
                for (int s = 0; s < ns; ++s, ++itc) {
                    const int st = itc % G_STAGES;
                    tc::mbar_wait(&bar_empty[st], ((itc / G_STAGES) & 1) ^ 1);
                    tc::mbar_expect_tx(&bar_full[st], bytes);
                    tc::bulk_g2s(dst + (size_t)st * G_STAGE_FLOATS, src + (long long)s * (2LL * L.N * G_KC), bytes, &bar_full[st]);
                }
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------ producers: global fp32 -> tf32 hi/lo images
        // Group g (two warps) owns shared-memory stage g: it stages every 4th K chunk, so four chunks
        // are in flight and a group has four MMA-stage times to cover its L2 round trips.
        const int pw = warp - G_PROD_WARP0;         // 0..7
        const int grp = pw >> 1, w2 = pw & 1;
        uint32_t itc_base = 0;                      // K chunks of all previous items
        for (int i = 0; iter.next(gb, it); ++i) {
            const GemmLayer& L = gb.L[it.layer];
            const int ns = (int)((it.k1 - it.k0 + G_KC - 1) / G_KC);
            const int s0 = (grp - (int)(itc_base & 3) + 4) & 3;       // this group's first chunk of the item
            if (gb.chain && it.layer > 0 && s0 < ns) {
                // the rows this item reads were written by the epilogue of item i - dep (same CTA)
                const int j = i - it.dep;
                tc::mbar_wait(&bar_epi_done[j & 1], (uint32_t)((j >> 1) & 1));
            }
            const long long Mv = gb.chain ? gb.M : L.M;
            const bool avec = (L.lda & 3) == 0 && aligned16(L.A), bvec = (L.ldb & 3) == 0 && aligned16(L.B);
            const int N = L.N, n8 = N >> 3, n_rb = (N + 31) >> 5;
            float* a_hi = stages + (size_t)grp * G_STAGE_FLOATS;
            float* a_lo = a_hi + G_A_FLOATS;
            float* b_hi = a_lo + G_A_FLOATS;
            float* b_lo = b_hi + N * G_KC;
            if constexpr (MN) {
                // ---- both operands row-contiguous: MN-major images (atoms [8 k][32 rows]), one 16-byte piece along the rows
                // per lane and k row.  Group g takes the chunks c = g (mod 4) of the global chunk sequence, i.e. ring
                // stages g and g + 4 in turn; warp w2 of the group takes the k rows w2, w2 + 2, w2 + 4, w2 + 6 of a chunk.
                const int nsm = (int)((it.k1 - it.k0 + MN_KC - 1) / MN_KC);
                const int s0m = (grp - (int)(itc_base & 3) + 4) & 3;
                float csum[4] = {0.f, 0.f, 0.f, 0.f};
                const long long am = it.m0 + 4 * lane;
                const int a_idx = (lane >> 3) * 256 + ((lane & 7) >> 1) * 8 + (lane & 1) * 4;      // + k * 32, chunk ^ (k & 3)
                const int nbq = (N + 127) >> 7;                         // float4 per lane and k row of B
                // bytes of this lane's 16-byte pieces that are inside the operands (-1: no such piece)
                const long long a_left = Mv - am;
                const int a_bytes = a_left >= 4 ? 16 : (a_left > 0 ? 4 * (int)a_left : 0);
                auto piece_bytes = [&](int n) { const int left = L.N_valid - n; return n >= N ? -1 : (left >= 4 ? 16 : (left > 0 ? 4 * left : 0)); };
                const int b_bytes0 = piece_bytes(4 * lane), b_bytes1 = nbq > 1 ? piece_bytes(128 + 4 * lane) : -1;
                auto stage_of = [&](int s) { return (int)((itc_base + (uint32_t)s) % MN_STAGES); };
                auto wait_empty = [&](int s) {
                    const uint32_t c = itc_base + (uint32_t)s;
                    tc::mbar_wait(&bar_empty[c % MN_STAGES], ((c / MN_STAGES) & 1) ^ 1);
                };
                if (avec && bvec) {
                    // 16-byte asynchronous copies (LDGSTS) straight into the hi images -- raw fp32, the tensor core truncates
                    // -- then the lo images x - trunc(x) of the same pieces from shared memory.  Software pipeline: the
                    // copies of the group's next chunk are issued before it waits for the current one.
                    auto issue = [&](int s) {
                        wait_empty(s);
                        float* st_base = stages + (size_t)stage_of(s) * MN_STAGE_FLOATS;
                        const uint32_t sa = tc::smem_u32(st_base), sb = sa + 2 * MN_A_FLOATS * 4;
                        const long long kb = it.k0 + (long long)s * MN_KC;
                        const float* pa = L.A + (kb + w2) * L.lda + am;
                        const float* pb = L.B + (kb + w2) * L.ldb + 4 * lane;
                        const long long k_left = it.k1 - kb - w2;                  // rows w2, w2 + 2, .. below it are valid
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int kk = w2 + 2 * r;
                            const uint32_t o = 4u * (uint32_t)((a_idx ^ ((kk & 3) << 3)) + kk * 32);
                            const bool kv = 2 * r < k_left;
                            cp_async16(sa + o, (kv && a_bytes) ? pa : L.A, kv ? a_bytes : 0);
                            if (b_bytes0 >= 0) cp_async16(sb + o, (kv && b_bytes0) ? pb : L.B, kv ? b_bytes0 : 0);
                            if (b_bytes1 >= 0) cp_async16(sb + o + 4u * 1024u, (kv && b_bytes1) ? pb + 128 : L.B, kv ? b_bytes1 : 0);
                            pa += 2 * L.lda;
                            pb += 2 * L.ldb;
                        }
                        asm volatile("cp.async.commit_group;\n" ::: "memory");
                    };
                    if (s0m < nsm) issue(s0m);
                    for (int s = s0m; s < nsm; s += 4) {
                        if (s + 4 < nsm) {
                            issue(s + 4);
                            asm volatile("cp.async.wait_group 1;\n" ::: "memory");
                        } else {
                            asm volatile("cp.async.wait_group 0;\n" ::: "memory");
                        }
                        float* a_hi8 = stages + (size_t)stage_of(s) * MN_STAGE_FLOATS;
                        float* a_lo8 = a_hi8 + MN_A_FLOATS;
                        float* b_hi8 = a_lo8 + MN_A_FLOATS;
                        float* b_lo8 = b_hi8 + N * MN_KC;
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int kk = w2 + 2 * r;
                            const int idx = (a_idx ^ ((kk & 3) << 3)) + kk * 32;
                            const float4 v = lo_pass(a_hi8, a_lo8, idx);
                            csum[0] += v.x; csum[1] += v.y; csum[2] += v.z; csum[3] += v.w;
                            if (b_bytes0 >= 0) lo_pass(b_hi8, b_lo8, idx);
                            if (b_bytes1 >= 0) lo_pass(b_hi8, b_lo8, idx + 1024);
                        }
                        tc::fence_async_smem();
                        mbar_arrive(&bar_full[stage_of(s)]);
                    }
                } else {
                    // operands that are not 16-byte addressable: through registers, one chunk at a time
                    for (int s = s0m; s < nsm; s += 4) {
                        const long long kb = it.k0 + (long long)s * MN_KC;
                        float4 va[4], vb[4][2];
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int kk = w2 + 2 * r;
                            va[r] = load_rquad(L.A, L.lda, avec, am, Mv, kb + kk, it.k1);
#pragma unroll
                            for (int q = 0; q < 2; ++q)
                                if (q < nbq) vb[r][q] = load_rquad(L.B, L.ldb, bvec, 128 * q + 4 * lane, 128 * q + 4 * lane < N ? L.N_valid : 0, kb + kk, it.k1);
                        }
                        wait_empty(s);
                        float* a_hi8 = stages + (size_t)stage_of(s) * MN_STAGE_FLOATS;
                        float* a_lo8 = a_hi8 + MN_A_FLOATS;
                        float* b_hi8 = a_lo8 + MN_A_FLOATS;
                        float* b_lo8 = b_hi8 + N * MN_KC;
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int kk = w2 + 2 * r;
                            const int idx = (a_idx ^ ((kk & 3) << 3)) + kk * 32;
                            split_store(a_hi8, a_lo8, idx, va[r]);
                            csum[0] += va[r].x; csum[1] += va[r].y; csum[2] += va[r].z; csum[3] += va[r].w;
#pragma unroll
                            for (int q = 0; q < 2; ++q)
                                if (q < nbq && 128 * q + 4 * lane < N) split_store(b_hi8, b_lo8, idx + q * 1024, vb[r][q]);
                        }
                        tc::fence_async_smem();
                        mbar_arrive(&bar_full[stage_of(s)]);
                    }
                }
                if (L.colsum && s0m < nsm) {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        if (am + c < Mv) atomicAdd(L.colsum + am + c, csum[c]);
                }
                itc_base += (uint32_t)nsm;
            } else {
            float csum[2] = {0.f, 0.f};
            for (int s = s0; s < ns; s += G_STAGES) {
                const uint32_t itc = itc_base + (uint32_t)s;
                const long long kb = it.k0 + (long long)s * G_KC;
                float4 va[8], vb[4];
                // A (8 quads per thread) and the first quarter of B in flight, then the rest of B quarter by quarter
                if (gb.debug & 2) {                 // timing experiment: no A production
                } else if (L.a_rc == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        va[j] = load_quad(L.A, L.lda, 0, avec, it.m0 + (w2 + 2 * j) * 8 + (lane & 7), Mv, kb + 4 * (lane >> 3), it.k1);
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int itx = w2 + 2 * j;
                        va[j] = load_quad(L.A, L.lda, 1, false, it.m0 + (itx & 3) * 32 + lane, Mv, kb + 4 * (itx >> 2), it.k1);
                        csum[j & 1] += (va[j].x + va[j].y) + (va[j].z + va[j].w);
                    }
                }
                auto store_a = [&]() {
                    if (L.a_rc == 0) {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            split_store(a_hi, a_lo, ((lane >> 3) * 16 + (w2 + 2 * j)) * 32 + (lane & 7) * 4, va[j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int itx = w2 + 2 * j, r = (itx & 3) * 32 + lane, kq = itx >> 2;
                            split_store(a_hi, a_lo, (kq * 16 + (r >> 3)) * 32 + (r & 7) * 4, va[j]);
                        }
                    }
                };
                if constexpr (PACKED) {             // B arrives by bulk copy (warp G_BULK_WARP)
                    tc::mbar_wait(&bar_empty[grp], ((itc / G_STAGES) & 1) ^ 1);
                    if (!(gb.debug & 2)) store_a();
                } else {
#pragma unroll
                    for (int part = 0; part < 4; ++part) {
                        if (L.b_rc == 0) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int rg = w2 + 2 * (j + 4 * part);
                                if (rg < n8) vb[j] = load_quad(L.B, L.ldb, 0, bvec, rg * 8 + (lane & 7), L.N_valid, kb + 4 * (lane >> 3), it.k1);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int itx = w2 + 2 * (j + 4 * part);
                                if (itx < 4 * n_rb) vb[j] = load_quad(L.B, L.ldb, 1, false, (itx % n_rb) * 32 + lane, L.N_valid, kb + 4 * (itx / n_rb), it.k1);
                            }
                        }
                        if (part == 0) {
                            tc::mbar_wait(&bar_empty[grp], ((itc / G_STAGES) & 1) ^ 1);
                            store_a();
                        }
                        if (L.b_rc == 0) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int rg = w2 + 2 * (j + 4 * part);
                                if (rg < n8) split_store(b_hi, b_lo, ((lane >> 3) * n8 + rg) * 32 + (lane & 7) * 4, vb[j]);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int itx = w2 + 2 * (j + 4 * part);
                                if (itx < 4 * n_rb) {
                                    const int r = (itx % n_rb) * 32 + lane, kq = itx / n_rb;
                                    if (r < N) split_store(b_hi, b_lo, (kq * n8 + (r >> 3)) * 32 + (r & 7) * 4, vb[j]);
                                }
                            }
                        }
                    }
                }
                tc::fence_async_smem();
                mbar_arrive(&bar_full[grp]);
            }
            itc_base += (uint32_t)ns;
            if (L.colsum && L.a_rc) {
#pragma unroll
                for (int q = 0; q < 2; ++q) {       // rows of itx = w2 + 2 j: row block (w2 + 2 (j & 1)) & 3
                    const long long r = it.m0 + ((w2 + 2 * q) & 3) * 32 + lane;
                    if (r < Mv) atomicAdd(L.colsum + r, csum[q]);
                }
            }
            }       // K-major modes
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == G_MMA_WARP) tc::tmem_dealloc<512>(tbase);
}

}  // namespace

static int g_gemm_sms = 0;

int gemm3x_launch(GemmBatch& gb, cudaStream_t stream) {
    if (gb.n_layers < 1 || gb.n_layers > GEMM_MAX_LAYERS) return set_error(NSB_EINVAL, "gemm: %d layers", gb.n_layers);
    long long items = 0;
    for (int i = 0; i < gb.n_layers; ++i) {
        GemmLayer& L = gb.L[i];
        if (!L.A || !L.B || !L.C) return set_error(NSB_EINVAL, "gemm layer %d: null operand", i);
        if (L.N < 16 || L.N > 256 || (L.N & 15) || L.N_valid < 1 || L.N_valid > L.N)
            return set_error(NSB_EINVAL, "gemm layer %d: bad width %d (%d valid)", i, L.N, L.N_valid);
        if (L.K < 1) return set_error(NSB_EINVAL, "gemm layer %d: empty reduction", i);
        if ((L.epi == GEMM_EPI_BIAS || L.epi == GEMM_EPI_BIAS_RELU) && !L.bias) return set_error(NSB_EINVAL, "gemm layer %d: no bias", i);
        if (L.epi == GEMM_EPI_MASK && !L.mask) return set_error(NSB_EINVAL, "gemm layer %d: no mask", i);
        if (L.colsum && !L.a_rc) return set_error(NSB_EINVAL, "gemm layer %d: colsum needs a row-contiguous A", i);
        if (gb.chain) {
            L.n_split = 1; L.k_per_split = (L.K + 15) / 16 * 16;
        } else {
            if (L.M < 1) return set_error(NSB_EINVAL, "gemm layer %d: no rows", i);
            L.m_tiles = (int)((L.M + 127) / 128);
            if (L.n_split < 1) L.n_split = 1;
            // the tensor-memory accumulator truncates on every update (measured: ~3e-8 relative bias per
            // MMA, towards zero), so one accumulation run is kept to <= 1024 reduction steps
            if (L.epi == GEMM_EPI_ATOMIC && (L.K + L.n_split - 1) / L.n_split > 1024) L.n_split = (int)((L.K + 1023) / 1024);
            L.k_per_split = ((L.K + L.n_split - 1) / L.n_split + 15) / 16 * 16;
            if (L.n_split > 1 && L.epi != GEMM_EPI_ATOMIC) return set_error(NSB_EINVAL, "gemm layer %d: split-K needs the atomic epilogue", i);
            items += (long long)L.m_tiles * L.n_split;
        }
    }
    if (gb.chain) {
        if (gb.M < 1) return 0;
        gb.m_tiles = (int)((gb.M + 127) / 128);
        items = gb.m_tiles;
    }
    if (!g_gemm_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_gemm_sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaFuncSetAttribute(k_gemm3x<G_MODE_LEGACY>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm3x<G_MODE_PACKED>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm3x<G_MODE_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM);
        if (e != cudaSuccess) { g_gemm_sms = 0; return set_error((int)e, "cudaFuncSetAttribute(k_gemm3x): %s", cudaGetErrorString(e)); }
    }
    int n_packed = 0;
    long long max_chunks = 1;
    for (int i = 0; i < gb.n_layers; ++i) {
        const GemmLayer& L = gb.L[i];
        if (!L.b_packed) continue;
        if (reinterpret_cast<uintptr_t>(L.b_packed) & 15u) return set_error(NSB_EINVAL, "gemm layer %d: b_packed is not 16-byte aligned", i);
        ++n_packed;
        if ((L.K + G_KC - 1) / G_KC > max_chunks) max_chunks = (L.K + G_KC - 1) / G_KC;
    }
    if (n_packed && n_packed != gb.n_layers) return set_error(NSB_EINVAL, "gemm: b_packed on %d of %d layers", n_packed, gb.n_layers);
    gb.packed = n_packed ? 1 : 0;
    static const int dbg = [] { const char* e = getenv("NSB_GEMM_DEBUG"); return e ? atoi(e) : 0; }();
    gb.debug = dbg;
    if (gb.packed) {
        k_gemm_pack_b<<<dim3((unsigned)(max_chunks < 64 ? max_chunks : 64), (unsigned)gb.n_layers), 256, 0, stream>>>(gb);
        if (int rc = check_launch("k_gemm_pack_b")) return rc;
    }
    const int grid = (int)(items < g_gemm_sms ? items : g_gemm_sms);
    bool all_mn = !gb.packed;
    for (int i = 0; i < gb.n_layers; ++i) all_mn = all_mn && layer_mn_capable(gb.L[i]);
    if (gb.packed) k_gemm3x<G_MODE_PACKED><<<grid, GemmCfg<G_MODE_PACKED>::THREADS, G_SMEM, stream>>>(gb);
    else if (all_mn) k_gemm3x<G_MODE_MN><<<grid, GemmCfg<G_MODE_MN>::THREADS, G_SMEM, stream>>>(gb);
    else k_gemm3x<G_MODE_LEGACY><<<grid, GemmCfg<G_MODE_LEGACY>::THREADS, G_SMEM, stream>>>(gb);
    return check_launch("k_gemm3x");
}

}  // namespace nsb

using namespace nsb;

// test hook: one GEMM through the chained kernel (tests/test_gemm3x.py)
extern "C" int64_t nsb_gemm3x_pack_floats(int N_valid, int64_t K) { return gemm3x_pack_floats((N_valid + 15) / 16 * 16, K); }

extern "C" int nsb_gemm3x(const float* A, int64_t lda, int a_rc, const float* B, int64_t ldb, int b_rc, float* C,
                          int64_t ldc, int64_t M, int N_valid, int64_t K, int epi, const float* bias,
                          const float* mask, int64_t ldmask, float* colsum, int n_split, float* b_scratch,
                          void* stream) {
    GemmBatch gb{};
    gb.n_layers = 1;
    gb.chain = 0;
    GemmLayer& L = gb.L[0];
    L.A = A; L.lda = lda; L.a_rc = a_rc; L.B = B; L.ldb = ldb; L.b_rc = b_rc; L.C = C; L.ldc = ldc;
    L.M = M; L.N_valid = N_valid; L.N = (N_valid + 15) / 16 * 16; L.K = K; L.epi = epi; L.bias = bias;
    L.mask = mask; L.ldmask = ldmask; L.colsum = colsum; L.n_split = n_split; L.b_packed = b_scratch;
    return gemm3x_launch(gb, (cudaStream_t)stream);
}
