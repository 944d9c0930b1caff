// Gradient exchange of the sharded mapping step over NVLink / NVSwitch peer memory.
//
// Mapping rays shard across the GPUs of one box; every rank scatters its partial gradients into its
// copy of ONE flat bucket that lives in symmetric memory (every rank's copy is mapped into every
// process, plus -- when the fabric offers NVLS -- one multicast mapping of all copies).  The kernels
// here replace "NCCL all-reduce, then the same optimiser step on every rank":
//
//   k_exchange<false>   two-shot all-reduce.  Rank r owns the r-th slice of the range: it reads that
//                       slice of every copy (multimem.ld_reduce: the switch adds the copies; without
//                       NVLS: one 16-byte peer load per copy) and stores the sum into every copy
//                       (multimem.st, or one peer store per copy).
//   k_exchange<true>    the same reduction fused with the mapper's masked Adam (src/Mapper.py:462-482,
//                       633-642): the owner keeps the Adam moments of ITS slice only, updates the
//                       parameters of the slice and stores the new parameters into every rank's copy
//                       of the parameter buffer.  The gradient sum never travels back and the
//                       optimiser state is sharded world-ways.
//
// Ranks meet at two block-wise barriers through words of a symmetric signal pad (CTA b of every
// rank exchanges flags with CTA b of every other rank): one before the first peer read (every rank
// has finished its backward scatter), one after the last peer store.  A flag is a compare-and-swap
// 0 -> 1 by the sender and 1 -> 0 by the receiver, so the pad is all zero again when the kernel
// ends and the launch can be replayed inside a CUDA graph.  A wait that lasts longer than
// ~20 s sets *error and gives up, so a lost peer cannot hang the GPU.
#include <stdlib.h>
#include <cstdio>
#include <cstdlib>
#include "nsb_common.cuh"

namespace nsb {

constexpr int X_THREADS = 512;
constexpr int X_MAX_BLOCKS = 148;       // upper bound of the CTAs of one exchange (sizes the signal pad)
constexpr int X_UNROLL = 8;             // 16-byte requests in flight per thread

struct PeersDev {
    int rank, world;
    float* bucket[NSB_MAX_PEERS];
    float* param[NSB_MAX_PEERS];
    float* bucket_mc;
    float* param_mc;
    unsigned* signal[NSB_MAX_PEERS];
    unsigned* touched[NSB_MAX_PEERS];
    float* scalar[NSB_MAX_PEERS];
    int* error;
    unsigned* marks;          // NsbPeers::line_marks (local): word 0 valid flag, byte 128 + l = line l was touched
    long long marks_units;    // leading units the marks cover
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// CTA-wise barrier over all ranks: thread t < world talks to rank t.  `slot` selects this CTA's
// words of the pad for this phase: word slot + s of rank d's pad is written by rank s only.
// Epoch flags: the n-th barrier on a slot posts the number n into the peer's word (a store, no NVLink round
// trip) and spins on its own word until the peer's n has arrived; the count of what this rank has sent lives
// in the second half of its own pad.  Every rank runs the same sequence of barriers on a slot.
constexpr int SIGNAL_WORDS = 2 * X_MAX_BLOCKS * NSB_MAX_PEERS;
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_barrier(const PeersDev& g, int slot) {
    __syncthreads();
    const int t = threadIdx.x;
    if (t < g.world && t != g.rank) {
        const unsigned long long t0 = globaltimer_ns();
        bool ok = true;
        unsigned* sent = g.signal[g.rank] + SIGNAL_WORDS + slot + t;
        const unsigned e = *sent + 1u;
        *sent = e;
        __threadfence_system();
        st_release_sys(g.signal[t] + slot + g.rank, e);
        while ((int)(ld_acquire_sys(g.signal[g.rank] + slot + t) - e) < 0)
            if (globaltimer_ns() - t0 > 20000000000ull) { ok = false; break; }
        __threadfence_system();
        if (!ok && g.error) *g.error = 1;
    }
    __syncthreads();
}

__device__ __forceinline__ float4 mc_ld_reduce(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// peer copies are read past the (incoherent) L1
__device__ __forceinline__ float4 peer_ld(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_st(float* p, const float4& v) {
    asm volatile("st.volatile.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ float4 reduce_at(const PeersDev& g, long long off) {
    if (g.bucket_mc) return mc_ld_reduce(g.bucket_mc + off);
    float4 s = peer_ld(g.bucket[0] + off);
    for (int r = 1; r < g.world; ++r) {                         // fixed rank order: every slice sums the same way
        const float4 v = peer_ld(g.bucket[r] + off);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    return s;
}

struct AdamDev {
    float* m;                  // moments of the OWNED slice: element i of the range lives at m[i - lo4*4]
    float* v;
    const unsigned char* mask; // voxel mask of the whole range (one byte per 32 floats) or null
    int* state2;
    const float* lr;
    float b1, b2, eps;
    long long p_off;           // float offset of the range inside the parameter buffers
};

__device__ __forceinline__ float adam1(float& p, float g, float& m, float& v, float b1, float b2, float eps,
                                       float step_size, float bc2_sqrt) {
    m += (g - m) * (1.f - b1);
    v = v * b2 + (1.f - b2) * g * g;
    p -= step_size * (m / (sqrtf(v) / bc2_sqrt + eps));
    return p;
}

// range = floats [off, off + 4 * n4) of the bucket; rank r owns vectors [r * per, min(n4, (r + 1) * per))
template <bool ADAM>
__global__ void __launch_bounds__(X_THREADS) k_exchange(const __grid_constant__ PeersDev g, long long off,
                                                         long long n4, long long per, const __grid_constant__ AdamDev a) {
    const int nb = gridDim.x;
    peer_barrier(g, (0 * X_MAX_BLOCKS + blockIdx.x) * NSB_MAX_PEERS);
    const long long lo = g.rank * per, hi = min(n4, lo + per);
    float step_size = 0.f, bc2_sqrt = 1.f;
    if (ADAM) {                                                 // same scalars as k_adam (nsb_render.cu)
        const int t = a.state2[0] + 1;
        const double bc1 = 1.0 - pow((double)a.b1, (double)t), bc2 = 1.0 - pow((double)a.b2, (double)t);
        step_size = (float)((double)a.lr[0] / bc1);
        bc2_sqrt = (float)sqrt(bc2);
    }
    const long long stride = (long long)nb * X_THREADS;
    for (long long i0 = lo + (long long)blockIdx.x * X_THREADS + threadIdx.x; i0 < hi; i0 += stride * X_UNROLL) {
        float4 s[X_UNROLL];
#pragma unroll
        for (int u = 0; u < X_UNROLL; ++u) {
            const long long i = i0 + u * stride;
            if (i < hi && !(ADAM && a.mask && !a.mask[i >> 3])) s[u] = reduce_at(g, off + 4 * i);
        }
#pragma unroll
        for (int u = 0; u < X_UNROLL; ++u) {
            const long long i = i0 + u * stride;
            if (i >= hi) continue;
            if (ADAM) {
                if (a.mask && !a.mask[i >> 3]) continue;        // outside the frustum: parameters untouched
                const long long po = a.p_off + 4 * i;
                float4 p = *reinterpret_cast<const float4*>(g.param[g.rank] + po);
                float4 m = reinterpret_cast<float4*>(a.m)[i - lo], v = reinterpret_cast<float4*>(a.v)[i - lo];
                adam1(p.x, s[u].x, m.x, v.x, a.b1, a.b2, a.eps, step_size, bc2_sqrt);
                adam1(p.y, s[u].y, m.y, v.y, a.b1, a.b2, a.eps, step_size, bc2_sqrt);
                adam1(p.z, s[u].z, m.z, v.z, a.b1, a.b2, a.eps, step_size, bc2_sqrt);
                adam1(p.w, s[u].w, m.w, v.w, a.b1, a.b2, a.eps, step_size, bc2_sqrt);
                reinterpret_cast<float4*>(a.m)[i - lo] = m;
                reinterpret_cast<float4*>(a.v)[i - lo] = v;
                if (g.param_mc) mc_st(g.param_mc + po, p);
                else
                    for (int r = 0; r < g.world; ++r) peer_st(g.param[r] + po, p);
            } else {
                if (g.bucket_mc) mc_st(g.bucket_mc + off + 4 * i, s[u]);
                else
                    for (int r = 0; r < g.world; ++r) peer_st(g.bucket[r] + off + 4 * i, s[u]);
            }
        }
    }
    peer_barrier(g, (1 * X_MAX_BLOCKS + blockIdx.x) * NSB_MAX_PEERS);
    if (ADAM) {
        // the last CTA to finish publishes the new step count (every CTA has read the old one by then)
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&a.state2[1], 1) == nb - 1) {
                a.state2[1] = 0;
                a.state2[0] += 1;
            }
        }
    }
}

// Sparse two-shot all-reduce.  One mapping step touches 16 % (1000 rays) to 39 % (8000 rays) of the
// 128-byte voxel lines of the Replica bucket (scripts/bucket_sparsity.py), so most of a dense
// exchange moves zeros.  Unit = 32 lines (4 KB); rank o owns the units u = v * world + o (v < per_u): interleaved, because
// the touched lines of a step are a spatial subset of the grids (the camera frusta) and contiguous slices left some
// owners with most of the work while the others waited at the closing barrier.
//   A  every rank scans its local copy and publishes one bit per line (word per unit) in symmetric memory
//   -- barrier --
//   C  the owner ORs the words of all ranks and reduces / stores only lines some rank touched
//   -- barrier --
// Untouched lines are zero in every copy and stay so.  CTA b meets CTA b of the peers only, so the
// units CTA b scans in A (for every owner) are exactly the units CTA b of that owner handles in C.
constexpr int X_WARPS = X_THREADS / 32;
__device__ unsigned long long g_xprof[8];      // NSB_X_PROFILE_PRINT: globaltimer stamps of CTA 0 (start, A, barrier, C, barrier)
#define XSTAMP(k) if (blockIdx.x == 0 && threadIdx.x == 0) g_xprof[k] = globaltimer_ns();
__device__ __forceinline__ bool line_on(unsigned word, int j, int lane) { return (word >> (4 * j + (lane >> 3))) & 1u; }

__global__ void __launch_bounds__(X_THREADS) k_exchange_sparse(const __grid_constant__ PeersDev g, long long off,
                                                                long long n4, long long n_units, long long per_u) {
    const int lane = threadIdx.x & 31;
    const long long gw = (long long)blockIdx.x * X_WARPS + (threadIdx.x >> 5), nw = (long long)gridDim.x * X_WARPS;
    XSTAMP(0)
    // ---- A: touched bits of the local copy, pushed to the owner of each unit (posted peer stores: the
    //         owner later reads every rank's word from its OWN memory, no NVLink round trip)
    const float4* mine = reinterpret_cast<const float4*>(g.bucket[g.rank] + off);
    // the backward that filled the local copy marked the lines it touched (NsbPoints::line_marks): read ~4 bytes per
    // 4 KB unit instead of the unit, and clear the word for the next step.  Units the marks do not cover (the decoder
    // weight gradients) and backward paths that do not mark (flag word 0) are scanned.
    const bool marked = g.marks != nullptr && off == 0 && __ldcg(g.marks) != 0u;
    const long long covered = marked ? g.marks_units : 0;
    if (marked) {                          // 32 bytes per unit (lane = line) instead of the unit's 4 KB
        unsigned char* mb = reinterpret_cast<unsigned char*>(g.marks) + 128;
        for (int o = 0; o < g.world; ++o)
            for (long long v = gw; v < per_u; v += nw) {
                const long long u = v * g.world + o;
                if (u >= n_units || u >= covered) break;
                const unsigned char m = __ldcg(mb + u * 32 + lane);
                const unsigned word = __ballot_sync(FULL, m != 0);
                if (m) mb[u * 32 + lane] = 0;           // clean for the next step
                if (lane == 0) {
                    unsigned* dst = g.touched[o] + g.rank * per_u + v;
                    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(dst), "r"(word) : "memory");
                }
            }
    }
    for (int o = 0; o < g.world; ++o)
        for (long long v = gw; v < per_u; v += nw) {
            const long long u = v * g.world + o;
            if (u >= n_units) break;
            if (u < covered) continue;
            unsigned word = 0u;
            float4 x[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const long long i4 = u * 256 + j * 32 + lane;
                x[j] = i4 < n4 ? __ldcg(mine + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool nz = (x[j].x != 0.f) | (x[j].y != 0.f) | (x[j].z != 0.f) | (x[j].w != 0.f);
                const unsigned b = __ballot_sync(FULL, nz);
#pragma unroll
                for (int q = 0; q < 4; ++q) word |= ((b >> (8 * q)) & 0xffu) ? (1u << (4 * j + q)) : 0u;
            }
            if (lane == 0) {
                unsigned* dst = g.touched[o] + g.rank * per_u + v;
                asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(dst), "r"(word) : "memory");
            }
        }
    __syncthreads();
    XSTAMP(1)
    peer_barrier(g, (0 * X_MAX_BLOCKS + blockIdx.x) * NSB_MAX_PEERS);
    XSTAMP(2)
    // ---- C: reduce the lines of my units that some rank touched, two units per pass (16 requests in
    //         flight per lane)
    const unsigned* words = g.touched[g.rank];
    for (long long v0 = gw; v0 < per_u; v0 += 2 * nw) {
        const long long v[2] = {v0, v0 + nw};
        unsigned word[2];
        float4 s[2][8];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            unsigned w = 0u;
            const bool live = v[k] < per_u && v[k] * g.world + g.rank < n_units;
            if (live && lane < g.world) {
                const unsigned* p = words + lane * per_u + v[k];
                asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(w) : "l"(p) : "memory");
            }
            word[k] = __reduce_or_sync(FULL, w);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (!word[k]) continue;
            const long long u = v[k] * g.world + g.rank;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const long long i4 = u * 256 + j * 32 + lane;
                if (line_on(word[k], j, lane) && i4 < n4) s[k][j] = reduce_at(g, off + 4 * i4);
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (!word[k]) continue;
            const long long u = v[k] * g.world + g.rank;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const long long i4 = u * 256 + j * 32 + lane;
                if (!(line_on(word[k], j, lane) && i4 < n4)) continue;
                if (g.bucket_mc) mc_st(g.bucket_mc + off + 4 * i4, s[k][j]);
                else
                    for (int r = 0; r < g.world; ++r) peer_st(g.bucket[r] + off + 4 * i4, s[k][j]);
            }
        }
    }
    __syncthreads();
    XSTAMP(3)
    peer_barrier(g, (1 * X_MAX_BLOCKS + blockIdx.x) * NSB_MAX_PEERS);
    XSTAMP(4)
}

// CTAs of one exchange: NVLink round trips are several microseconds, so the wire is only full with
// ~10 MB of 16-byte requests in flight (CTAs x 512 threads x X_UNROLL x 16 B).  Every rank must pick
// the same count (CTA b of every rank meets CTA b of every other).  NSB_X_BLOCKS overrides (tuning).
// max(gt_depth) over the kept rays of ALL ranks (src/utils/Renderer.py:157,194 on the whole batch): the
// one batch-global scalar of the render.  One CTA: local max -> own symmetric slot -> barrier -> max of
// every rank's slot -> barrier (slots are free again).  NaN propagates like torch.max.
__device__ __forceinline__ float nanmax(float a, float b) { return (b > a || b != b) ? b : a; }
// Batch depth max over all ranks, push model: every rank posts (step number, local max) as ONE 8-byte store
// into its slot of every rank's scalar pad and spins on its own pad until every peer's value of this step has
// arrived -- one NVLink one-way latency, no barrier, no remote load.  Slots are double-buffered by the step's
// parity: a rank cannot be two steps ahead of a peer, it needs that peer's value of every step.
// pad (64 words per rank): u64 slot[2][NSB_MAX_PEERS] at words 0..31, the step counter at word 40.
__global__ void __launch_bounds__(1024) k_depth_max_peers(const __grid_constant__ PeersDev g, const float* __restrict__ gt,
                                                           const uint8_t* __restrict__ keep, long long n,
                                                           float* __restrict__ out) {
    static_assert(NSB_MAX_PEERS <= 8, "scalar pad layout");
    __shared__ float part[32];
    float m = -INFINITY;
    for (long long i = threadIdx.x; i < n; i += blockDim.x)
        if (!keep || keep[i]) m = nanmax(m, gt[i]);
    for (int s = 16; s > 0; s >>= 1) m = nanmax(m, __shfl_xor_sync(FULL, m, s));
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x >= 32) return;
    m = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : -INFINITY;
    for (int s = 16; s > 0; s >>= 1) m = nanmax(m, __shfl_xor_sync(FULL, m, s));
    unsigned* counter = reinterpret_cast<unsigned*>(g.scalar[g.rank]) + 40;
    unsigned seq = 0;
    if (threadIdx.x == 0) { seq = *counter + 1u; *counter = seq; }
    seq = __shfl_sync(FULL, seq, 0);
    const int t = threadIdx.x;
    float v = -INFINITY;
    if (t < g.world) {
        const unsigned long long word = ((unsigned long long)seq << 32) | __float_as_uint(m);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(g.scalar[t]) + (seq & 1u) * NSB_MAX_PEERS + g.rank;
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(dst), "l"(word) : "memory");
        const unsigned long long* src =
            reinterpret_cast<const unsigned long long*>(g.scalar[g.rank]) + (seq & 1u) * NSB_MAX_PEERS + t;
        const unsigned long long t0 = globaltimer_ns();
        unsigned long long got;
        for (;;) {
            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(src) : "memory");
            if ((unsigned)(got >> 32) == seq) break;
            if (globaltimer_ns() - t0 > 20000000000ull) { if (g.error) *g.error = 1; break; }
        }
        v = __uint_as_float((unsigned)got);
    }
    for (int s = 16; s > 0; s >>= 1) v = nanmax(v, __shfl_xor_sync(FULL, v, s));
    if (t == 0) out[0] = v;
}

static int blocks_for(long long per) {
    static int cap = 0;
    if (!cap) {
        const char* e = getenv("NSB_X_BLOCKS");
        int v = e ? atoi(e) : 0;
        cap = (v >= 1 && v <= X_MAX_BLOCKS) ? v : 128;
    }
    long long b = (per + (long long)X_THREADS * X_UNROLL - 1) / ((long long)X_THREADS * X_UNROLL);
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

static int to_dev(const NsbPeers* p, PeersDev& g, bool need_param) {
    if (!p) return set_error(NSB_EINVAL, "null peer group");
    if (p->world < 1 || p->world > NSB_MAX_PEERS || p->rank < 0 || p->rank >= p->world)
        return set_error(NSB_EINVAL, "peer group: rank %d of %d", p->rank, p->world);
    g.rank = p->rank; g.world = p->world;
    for (int r = 0; r < NSB_MAX_PEERS; ++r) {
        g.bucket[r] = r < p->world ? p->bucket[r] : nullptr;
        g.param[r] = r < p->world ? p->param[r] : nullptr;
        g.signal[r] = r < p->world ? p->signal[r] : nullptr;
        g.touched[r] = r < p->world ? p->touched[r] : nullptr;
        g.scalar[r] = r < p->world ? p->scalar[r] : nullptr;
        if (r < p->world && (!g.bucket[r] || !g.signal[r] || (need_param && !g.param[r])))
            return set_error(NSB_EINVAL, "peer group: missing mapping of rank %d", r);
    }
    g.bucket_mc = p->bucket_multicast;
    g.param_mc = p->param_multicast;
    g.error = p->error;
    g.marks = p->line_marks;
    g.marks_units = p->line_marks ? p->line_marks_floats / 1024 : 0;     // whole 4 KB units only
    return 0;
}

}  // namespace nsb

using namespace nsb;

extern "C" int64_t nsb_exchange_signal_words(void) { return 2ll * SIGNAL_WORDS; }

extern "C" int nsb_bucket_allreduce(const NsbPeers* peers, int64_t offset, int64_t n, void* stream) {
    PeersDev g;
    if (int e = to_dev(peers, g, false)) return e;
    if (n <= 0) return 0;
    if ((offset | n) & 3) return set_error(NSB_EINVAL, "exchange: offset and n must be multiples of 4 floats");
    for (int r = 0; r < g.world; ++r)
        if ((uintptr_t)g.bucket[r] & 15) return set_error(NSB_EINVAL, "exchange: buffers must be 16-byte aligned");
    const long long n4 = n / 4, per = (n4 + g.world - 1) / g.world;
    AdamDev a = {};
    timing_mark(3, 0, (cudaStream_t)stream);
    k_exchange<false><<<blocks_for(per), X_THREADS, 0, (cudaStream_t)stream>>>(g, (long long)offset, n4, per, a);
    timing_mark(3, 1, (cudaStream_t)stream);
    return check_launch("k_exchange");
}

extern "C" int64_t nsb_exchange_slice(int rank, int world, int64_t n, int64_t* lo) {
    if (world < 1 || rank < 0 || rank >= world || n < 0 || (n & 3)) return -1;
    const long long n4 = n / 4, per = (n4 + world - 1) / world;
    const long long a = (long long)rank * per, b = a + per < n4 ? a + per : n4;
    if (lo) *lo = 4 * (a < n4 ? a : n4);
    return b > a ? 4 * (b - a) : 0;
}

extern "C" int nsb_bucket_reduce_adam(const NsbPeers* peers, int64_t offset, int64_t n, int64_t param_offset, float* m_slice,
                                      float* v_slice, const uint8_t* voxel_mask, int32_t* state2, const float* lr,
                                      float beta1, float beta2, float eps, void* stream) {
    PeersDev g;
    if (int e = to_dev(peers, g, true)) return e;
    if (n <= 0) return 0;
    if (!m_slice || !v_slice || !state2 || !lr) return set_error(NSB_EINVAL, "null argument");
    if ((offset | n | param_offset) & 3) return set_error(NSB_EINVAL, "exchange: offsets and n must be multiples of 4 floats");
    if (voxel_mask && (n % 32)) return set_error(NSB_EINVAL, "masked Adam needs n to be a multiple of 32 (voxel lines)");
    for (int r = 0; r < g.world; ++r)
        if (((uintptr_t)g.bucket[r] | (uintptr_t)g.param[r]) & 15)
            return set_error(NSB_EINVAL, "exchange: buffers must be 16-byte aligned");
    if (((uintptr_t)m_slice | (uintptr_t)v_slice) & 15) return set_error(NSB_EINVAL, "exchange: moments must be 16-byte aligned");
    const long long n4 = n / 4, per = (n4 + g.world - 1) / g.world;
    AdamDev a;
    a.m = m_slice; a.v = v_slice; a.mask = voxel_mask; a.state2 = state2; a.lr = lr;
    a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.p_off = (long long)param_offset;
    timing_mark(3, 0, (cudaStream_t)stream);
    k_exchange<true><<<blocks_for(per), X_THREADS, 0, (cudaStream_t)stream>>>(g, (long long)offset, n4, per, a);
    timing_mark(3, 1, (cudaStream_t)stream);
    return check_launch("k_exchange_adam");
}

extern "C" int64_t nsb_exchange_touched_words(int64_t n) { return n <= 0 ? 0 : (n + 1023) / 1024 + NSB_MAX_PEERS; }

extern "C" int nsb_bucket_allreduce_sparse(const NsbPeers* peers, int64_t offset, int64_t n, void* stream) {
    return nsb_bucket_allreduce_sparse_ctas(peers, offset, n, 0, stream);
}

extern "C" int nsb_bucket_allreduce_sparse_ctas(const NsbPeers* peers, int64_t offset, int64_t n, int max_ctas, void* stream) {
    PeersDev g;
    if (int e = to_dev(peers, g, false)) return e;
    if (n <= 0) return 0;
    if ((offset & 31) || (n & 3)) return set_error(NSB_EINVAL, "sparse exchange: offset must be a multiple of 32 floats, n of 4");
    for (int r = 0; r < g.world; ++r) {
        if ((uintptr_t)g.bucket[r] & 127) return set_error(NSB_EINVAL, "sparse exchange: buffers must be 128-byte aligned");
        if (!g.touched[r]) return set_error(NSB_EINVAL, "sparse exchange: missing touched-line words of rank %d", r);
    }
    const long long n4 = n / 4, n_units = (n4 + 255) / 256, per_u = (n_units + g.world - 1) / g.world;
    // one warp per unit and pass; the scan covers world x per_u units
    long long b = (per_u + X_WARPS - 1) / X_WARPS;
    const int cap = blocks_for(1ll << 40);
    int nb = (int)(b < 1 ? 1 : (b > cap ? cap : b));
    if (max_ctas > 0 && nb > max_ctas) nb = max_ctas;
    timing_mark(3, 0, (cudaStream_t)stream);
    k_exchange_sparse<<<nb, X_THREADS, 0, (cudaStream_t)stream>>>(g, (long long)offset, n4, n_units, per_u);
    timing_mark(3, 1, (cudaStream_t)stream);
    if (getenv("NSB_X_PROFILE_PRINT")) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing((cudaStream_t)stream, &cs);
        if (cs == cudaStreamCaptureStatusNone) {
            cudaStreamSynchronize((cudaStream_t)stream);
            unsigned long long h[8];
            cudaMemcpyFromSymbol(h, g_xprof, sizeof(h));
            fprintf(stderr, "[xprof] rank %d ctas %d: scan/marks %.1f us, barrier %.1f us, reduce+store %.1f us, barrier %.1f us\n", g.rank, nb,
                    (h[1] - h[0]) * 1e-3, (h[2] - h[1]) * 1e-3, (h[3] - h[2]) * 1e-3, (h[4] - h[3]) * 1e-3);
        }
    }
    return check_launch("k_exchange_sparse");
}

extern "C" int nsb_depth_max_peers(const NsbPeers* peers, const float* gt_depth, const uint8_t* keep, int64_t N,
                                   float* out, void* stream) {
    PeersDev g;
    if (int e = to_dev(peers, g, false)) return e;
    if (!gt_depth || !out) return set_error(NSB_EINVAL, "null argument");
    for (int r = 0; r < g.world; ++r)
        if (!g.scalar[r]) return set_error(NSB_EINVAL, "depth max: missing scalar slot of rank %d", r);
    timing_mark(4, 0, (cudaStream_t)stream);
    k_depth_max_peers<<<1, 1024, 0, (cudaStream_t)stream>>>(g, gt_depth, keep, (long long)N, out);
    timing_mark(4, 1, (cudaStream_t)stream);
    return check_launch("k_depth_max_peers");
}
