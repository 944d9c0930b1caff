// Device-side helpers shared by the kernels of the render hot path (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/nice_slam_b200.h"
#include "nsb_layout.h"

namespace nsb {

constexpr int WARPS = 12;                // warps per CTA of the decoder kernels
constexpr int THREADS = WARPS * 32;
constexpr unsigned FULL = 0xffffffffu;

// ---- host-side error plumbing (nsb_api.cu) -------------------------------------------
int set_error(int code, const char* fmt, ...);
int check_launch(const char* what);
// Programmatic dependent launch: nsb_render_forward / nsb_render_backward know that the kernel just before a
// decoder kernel (k_sample_z, k_composite_*) does not produce what the decoder kernel's prologue reads (packed
// weights, tile counters), so they allow the next decoder launch to start while that kernel drains; the
// decoder kernels call pdl_wait() before they touch anything else.  Other callers get ordinary launches.
// nsb_render_backward in two calls (NsbRenderCall::dweight_phase): internal bits of nsb_decoders_backward's flags
constexpr unsigned NSB_FLAG_DEFER_DWEIGHT = 0x40000000u;     // decoder backward only
constexpr unsigned NSB_FLAG_ONLY_DWEIGHT = 0x20000000u;      // weight-gradient kernels only
void dweight_grid_cap(int ctas);     // 0: no cap; read (and reset) by the next nsb_decoders_backward of this thread
void pdl_allow_next(bool on);
bool pdl_take();
bool pdl_enabled();              // NSB_PDL=0 turns every programmatic launch into an ordinary one
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
#ifdef __CUDACC__
// Launch of a kernel that calls pdl_wait() before its first memory access: it may be scheduled while the kernel
// before it in the stream drains (that kernel calls pdl_launch_dependents() first thing), which hides the
// launch latency between the many short kernels of a step; the order of all memory accesses is the stream's.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif
void timing_mark(int which, int phase, cudaStream_t s);   // which: 0 decoder forward, 1 decoder backward, 2 weight gradients, 3 gradient exchange, 4 depth max over peers

// ---- points -------------------------------------------------------------------------
struct PointsDev {
    const float* ro;
    const float* rd;
    long long so, sd;
    const double* z;
    long long N;
    int S;
    const double* p64;
    const float* p32;
    long long P;
    unsigned* relu;      // NsbPoints::relu_masks
    float* pu;           // NsbPoints::point_cache: [P][8] = (p32 xyz, 0, u xyz, 0)
    unsigned* marks;     // NsbPoints::line_marks (word 0: valid flag), or NULL
    const float* marks_base;
};

inline PointsDev to_dev(const NsbPoints& p) {
    PointsDev d;
    d.ro = p.rays_o; d.rd = p.rays_d; d.so = p.rays_o_stride; d.sd = p.rays_d_stride;
    d.z = p.z; d.N = p.N; d.S = p.S; d.p64 = p.p64; d.p32 = p.p32;
    d.P = p.rays_o ? p.N * (long long)p.S : p.P;
    d.relu = p.relu_masks;
    d.pu = p.point_cache;
    d.marks = p.line_marks; d.marks_base = p.line_marks_base;
    return d;
}

struct BoundDev {
    double lo[3], hi[3];
};

inline BoundDev to_bound(const double* b6) {
    BoundDev b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = b6[2 * a]; b.hi[a] = b6[2 * a + 1]; }
    return b;
}

// p = rays_o + rays_d * z in fp64, un-fused, as torch evaluates
// src/utils/Renderer.py:227-228 (fp32 rays promoted to fp64 by the fp64 z).
__device__ __forceinline__ void ray_point(const PointsDev& q, long long i, double (&p)[3]) {
    long long r = i / q.S;
    double z = q.z[i];
#pragma unroll
    for (int a = 0; a < 3; ++a)
        p[a] = __dadd_rn((double)q.ro[r * q.so + a], __dmul_rn((double)q.rd[r * q.sd + a], z));
}

// strict in-bound test of Renderer.eval_points (Renderer.py:63-66)
__device__ __forceinline__ bool inbound64(const double (&p)[3], const BoundDev& b) {
    return p[0] < b.hi[0] && p[0] > b.lo[0] && p[1] < b.hi[1] && p[1] > b.lo[1] &&
           p[2] < b.hi[2] && p[2] > b.lo[2];
}
__device__ __forceinline__ bool inbound32(const float (&p)[3], const BoundDev& b) {
    return p[0] < (float)b.hi[0] && p[0] > (float)b.lo[0] && p[1] < (float)b.hi[1] &&
           p[1] > (float)b.lo[1] && p[2] < (float)b.hi[2] && p[2] > (float)b.lo[2];
}

__device__ __forceinline__ bool point_inbound(const PointsDev& q, long long i, const BoundDev& b) {
    if (q.ro) {
        double p[3];
        ray_point(q, i, p);
        return inbound64(p, b);
    }
    if (q.p64) {
        double p[3] = {q.p64[3 * i], q.p64[3 * i + 1], q.p64[3 * i + 2]};
        return inbound64(p, b);
    }
    float p[3] = {q.p32[3 * i], q.p32[3 * i + 1], q.p32[3 * i + 2]};
    return inbound32(p, b);
}

// ---- tensor-core primitives ---------------------------------------------------------------
// D += A(16x8, row) * B(8x8, col), tf32 inputs, fp32 accumulate.
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Exact two-term split x = hi + lo with hi representable in tf32 (low 13 mantissa bits
// cleared).  Three tf32 products hi*hi + lo*hi + hi*lo then carry ~2^-20 relative error
// per term, i.e. fp32-class accuracy on the tensor cores ("3xTF32").
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}

// acc[NT][4] += A[J slabs] * B, B in packed fragment order (nsb_layout.h), 3xTF32.
// a[j] is the A fragment of slab j: (row g, slot t), (row g+8, slot t), (row g, slot t+4),
// (row g+8, slot t+4).
template <int J, int NT>
__device__ __forceinline__ void gemm3(float (&acc)[NT][4], const float (&a)[J][4],
                                      const float* __restrict__ Bs, int lane) {
    static_assert(NT % 2 == 0, "n-tiles come in pairs");
    const float4* B4 = reinterpret_cast<const float4*>(Bs);
#pragma unroll
    for (int j = 0; j < J; ++j) {
        uint32_t ah[4], al[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_tf32(a[j][i], ah[i], al[i]);
#pragma unroll
        for (int h = 0; h < NT / 2; ++h) {
            float4 b = B4[(j * (NT / 2) + h) * 32 + lane];
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

// single n-tile variant: B stored as float2 per lane per slab (output layer)
template <int J>
__device__ __forceinline__ void gemm3_n1(float (&acc)[4], const float (&a)[J][4],
                                         const float* __restrict__ Bs, int lane) {
    const float2* B2 = reinterpret_cast<const float2*>(Bs);
#pragma unroll
    for (int j = 0; j < J; ++j) {
        uint32_t ah[4], al[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_tf32(a[j][i], ah[i], al[i]);
        float2 b = B2[j * 32 + lane];
        uint32_t bh[2], bl[2];
        split_tf32(b.x, bh[0], bl[0]);
        split_tf32(b.y, bh[1], bl[1]);
        mma_tf32(acc, al, bh[0], bh[1]);
        mma_tf32(acc, ah, bl[0], bl[1]);
        mma_tf32(acc, ah, bh[0], bh[1]);
    }
}

// accumulator (C fragment) of n-tile j -> A fragment of k-slab j under the "chain" slot map
template <int J>
__device__ __forceinline__ void chain(float (&a)[J][4], const float (&c)[J][4]) {
#pragma unroll
    for (int j = 0; j < J; ++j) {
        a[j][0] = c[j][0];
        a[j][1] = c[j][2];
        a[j][2] = c[j][1];
        a[j][3] = c[j][3];
    }
}

// ---- sine / cosine of the Fourier embedding ---------------------------------------------------
// sin(x), cos(x) to ~2 ulp (1.4e-7 absolute) for |x| < 1e5: x = k pi + r by a three-term
// Cody-Waite reduction, minimax polynomials on [-pi/2, pi/2] (degree 9 odd / degree 10 even,
// approximation errors 5e-9 / 3e-10) and the sign (-1)^k: 14-16 instructions instead of the ~27
// of the library fast path, whose inlined Payne-Hanek slow path also overflowed the instruction
// cache of the decoder kernels.  The embedding evaluates 96 sines (and, backward, 96 cosines)
// per point per decoder -- the largest block of issue slots in these kernels (profiles/r01).
// |p . B| stays below ~1e3 in any scene (B ~ N(0, 25^2), |p| of order 10 m); larger arguments
// take the library call, out of line.
static __device__ __noinline__ float sin_slow(float x) { return sinf(x); }
static __device__ __noinline__ float cos_slow(float x) { return cosf(x); }
__device__ __forceinline__ float reduce_pi(float x, float& k) {
    k = rintf(x * 0.318309886f);
    float r = fmaf(k, -3.14159202e+00f, x);
    r = fmaf(k, -6.27832946e-07f, r);
    return fmaf(k, -1.078060506e-14f, r);
}
__device__ __forceinline__ float sin_cw(float x) {
    if (fabsf(x) > 1.0e5f) return sin_slow(x);
    float k;
    const float r = reduce_pi(x, k), r2 = r * r;
    float q = fmaf(r2, 2.6000477646e-06f, -1.9806619093e-04f);
    q = fmaf(r2, q, 8.3330171183e-03f);
    q = fmaf(r2, q, -1.6666656733e-01f);
    const float v = fmaf(r * r2, q, r);
    return __uint_as_float(__float_as_uint(v) ^ ((uint32_t)(int)k << 31));
}
// Embedding sine of the tcgen05 forward: x = 2 pi k + r by a two-term Cody-Waite reduction (the FMA rounds once; exact to
// ~1e-7 for |x| < 1e5), then the special-function unit (MUFU.SIN: 2^-21.4 absolute error on [-pi, pi]).  9 instructions
// instead of sin_cw's 17 for the 13.4 M evaluations of a 1000-ray step: forward 104.6 -> 100.3 us.  Parity is unchanged
// within noise -- outputs against the fp32 torch oracle at the bench workload (scripts/accuracy_probe.py): depth / variance /
// colour 2.7e-6 / 7.6e-6 / 6.1e-6 with it, 2.9e-6 / 5.6e-6 / 7.7e-6 with sin_cw (bar 1e-4): the difference to torch is the
// summation order of the layers, not the sine.  -DNSB_EXACT_SIN (NSB_NVCC_EXTRA) restores sin_cw.
__device__ __forceinline__ float sin_embed(float x) {
#ifdef NSB_EXACT_SIN
    return sin_cw(x);
#else
    if (fabsf(x) > 1.0e5f) return sin_slow(x);
    const float k = rintf(x * 0.15915494309f);
    float r = fmaf(k, -6.2831855f, x);
    r = fmaf(k, 1.7484555e-07f, r);
    return __sinf(r);
#endif
}
__device__ __forceinline__ float cos_cw(float x) {
    if (fabsf(x) > 1.0e5f) return cos_slow(x);
    float k;
    const float r = reduce_pi(x, k), r2 = r * r;
    float q = fmaf(r2, -2.6076943982e-07f, 2.4761886380e-05f);
    q = fmaf(r2, q, -1.3888403773e-03f);
    q = fmaf(r2, q, 4.1666641831e-02f);
    q = fmaf(r2, q, -5.0e-01f);
    const float v = fmaf(r2, q, 1.0f);
    return __uint_as_float(__float_as_uint(v) ^ ((uint32_t)(int)k << 31));
}

// Adam on the 7 camera parameters + the tracker's best-candidate rule (Tracker.py:232-242,262-275), executed by one warp
// (lane i < 7 owns parameter i, g = its gradient); bias corrections in fp64 like torch's Python scalars.
__device__ __forceinline__ void cam_adam_warp(int i, float g, float* cam, float* m, float* v, int* step, float lr_q, float lr_t,
                                              float beta1, float beta2, float eps, const double* loss, double* best_loss,
                                              float* best_cam) {
    const int t = step[0] + 1;
    __syncwarp();
    if (i == 0) step[0] = t;
    float p = 0.f;
    if (i < 7) {
        const float mi = m[i] + (g - m[i]) * (1.f - beta1);              // exp_avg.lerp_(grad, 1 - beta1)
        const float vi = v[i] * beta2 + (1.f - beta2) * g * g;
        m[i] = mi;
        v[i] = vi;
        const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
        const float step_size = (float)((double)(i < 4 ? lr_q : lr_t) / bc1);
        const float denom = sqrtf(vi) / (float)sqrt(bc2) + eps;
        p = cam[i] - step_size * (mi / denom);
        cam[i] = p;
    }
    if (best_loss && loss[0] < best_loss[0]) {                            // uniform over the warp
        __syncwarp();
        if (i < 7) best_cam[i] = p;
        if (i == 0) best_loss[0] = loss[0];
    }
}

// cos(x) of the tcgen05 backward's embedding gradient d(e) cos(pB): sin_embed's reduction + MUFU.COS (gradients are
// compared at 1e-3; -DNSB_EXACT_SIN restores cos_cw as well).
__device__ __forceinline__ float cos_embed(float x) {
#ifdef NSB_EXACT_SIN
    return cos_cw(x);
#else
    if (fabsf(x) > 1.0e5f) return cos_slow(x);
    const float k = rintf(x * 0.15915494309f);
    float r = fmaf(k, -6.2831855f, x);
    r = fmaf(k, 1.7484555e-07f, r);
    return __cosf(r);
#endif
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};\n" ::"l"(addr), "f"(a), "f"(b),
                 "f"(c), "f"(d)
                 : "memory");
}

}  // namespace nsb
