// Callers of the hot path that run once per mapped frame, as kernels instead of the
// reference's numpy + cv2.remap pass over every voxel centre on the host:
//   nsb_frustum_mask      Mapper.get_mask_from_c2w            (src/Mapper.py:100-187)
//   nsb_keyframe_overlap  Mapper.keyframe_selection_overlap   (src/Mapper.py:189-293, the scores)
// Integer / boolean outputs: the arithmetic follows the reference operation by operation (float32
// rigid transform, float64 intrinsics product, OpenCV's 1/32-pixel fixed-point bilinear lookup with
// float tap weights, BORDER_CONSTANT 0), without FMA contraction, so the masks are bit-exact.
#include "nsb_common.cuh"

namespace nsb {

struct Proj {
    float w[12];            // rows 0..2 of w2c (float32, as np.linalg.inv(c2w) returns it)
    double fx, fy, cx, cy;
};

// Mapper.py:129-146 / 256-271: uv (float32) and camera z + 1e-5 (float64)
__device__ __forceinline__ void project(const Proj& P, float x, float y, float z, float& u, float& v, double& zz) {
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
        c[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, P.w[4 * k]), __fmul_rn(y, P.w[4 * k + 1])),
                                   __fmul_rn(z, P.w[4 * k + 2])), P.w[4 * k + 3]);
    const double cxd = (double)(-c[0]), cyd = (double)c[1], czd = (double)c[2];
    const double uh = __dadd_rn(__dadd_rn(__dmul_rn(P.fx, cxd), __dmul_rn(0.0, cyd)), __dmul_rn(P.cx, czd));
    const double vh = __dadd_rn(__dadd_rn(__dmul_rn(0.0, cxd), __dmul_rn(P.fy, cyd)), __dmul_rn(P.cy, czd));
    zz = __dadd_rn(czd, 1e-5);
    u = (float)__ddiv_rn(uh, zz);
    v = (float)__ddiv_rn(vh, zz);
}

// cvRound(x * INTER_TAB_SIZE) as cvtss2si does it: round half to even, INT_MIN for NaN / overflow
__device__ __forceinline__ int cv_fix(float x) {
    const float v = __fmul_rn(x, 32.f);
    if (!(fabsf(v) < 2147483648.f)) return (int)0x80000000;
    return __float2int_rn(v);
}

// cv2.remap(depth, u, v, INTER_LINEAR), float32 single channel, BORDER_CONSTANT 0
__device__ __forceinline__ float remap_bilinear(const float* __restrict__ img, int H, int W, float u, float v) {
    const int sx = cv_fix(u), sy = cv_fix(v);
    int ix = sx >> 5, iy = sy >> 5;
    ix = max(-32768, min(32767, ix));
    iy = max(-32768, min(32767, iy));
    const float fx = __fmul_rn((float)(sx & 31), 0.03125f), fy = __fmul_rn((float)(sy & 31), 0.03125f);
    const float wx0 = __fsub_rn(1.f, fx), wy0 = __fsub_rn(1.f, fy);
    auto tap = [&](int x, int y) -> float {
        return (x >= 0 && x < W && y >= 0 && y < H) ? img[(size_t)y * W + x] : 0.f;
    };
    float acc = __fmul_rn(tap(ix, iy), __fmul_rn(wy0, wx0));
    acc = __fadd_rn(acc, __fmul_rn(tap(ix + 1, iy), __fmul_rn(wy0, fx)));
    acc = __fadd_rn(acc, __fmul_rn(tap(ix, iy + 1), __fmul_rn(fy, wx0)));
    acc = __fadd_rn(acc, __fmul_rn(tap(ix + 1, iy + 1), __fmul_rn(fy, fx)));
    return acc;
}

// order-preserving float -> unsigned key (0 is below every float)
__device__ __forceinline__ unsigned f2key(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

struct FrustumParams {
    Proj P;
    float cam[3];
    const float *xs, *ys, *zs, *depth;
    int nx, ny, nz, H, W;
};

// pass 1: max over all voxels of the looked-up depth (Mapper.py:167: depths[zero_mask] = np.max(depths))
__global__ void __launch_bounds__(256) k_frustum_depth_max(FrustumParams f, unsigned* __restrict__ key_max) {
    const long long n = (long long)f.nx * f.ny * f.nz;
    unsigned best = 0u;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int iz = (int)(i % f.nz), iy = (int)((i / f.nz) % f.ny), ix = (int)(i / ((long long)f.nz * f.ny));
        float u, v;
        double zz;
        project(f.P, f.xs[ix], f.ys[iy], f.zs[iz], u, v, zz);
        best = max(best, f2key(remap_bilinear(f.depth, f.H, f.W, u, v)));
    }
    for (int s = 16; s > 0; s >>= 1) best = max(best, __shfl_xor_sync(FULL, best, s));
    if ((threadIdx.x & 31) == 0) atomicMax(key_max, best);
}

// pass 2: the mask, in the reference's [X,Y,Z] order and / or the grid's [Z,Y,X] voxel order
__global__ void __launch_bounds__(256) k_frustum_mask(FrustumParams f, const unsigned* __restrict__ key_max,
                                                       uint8_t* __restrict__ mask_xyz, uint8_t* __restrict__ mask_zyx) {
    const long long n = (long long)f.nx * f.ny * f.nz;
    const float dmax = key2f(key_max[0]);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int iz = (int)(i % f.nz), iy = (int)((i / f.nz) % f.ny), ix = (int)(i / ((long long)f.nz * f.ny));
        const float x = f.xs[ix], y = f.ys[iy], z = f.zs[iz];
        float u, v;
        double zz;
        project(f.P, x, y, z, u, v, zz);
        float d = remap_bilinear(f.depth, f.H, f.W, u, v);
        if (d == 0.f) d = dmax;
        bool m = (u < (float)f.W) && (u > 0.f) && (v < (float)f.H) && (v > 0.f);
        m = m && (0.0 <= -zz) && (-zz <= (double)__fadd_rn(d, 0.5f));
        const float dx = __fsub_rn(x, f.cam[0]), dy = __fsub_rn(y, f.cam[1]), dz = __fsub_rn(z, f.cam[2]);
        const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        m = m || (d2 < 0.25f);
        if (mask_xyz) mask_xyz[i] = m ? 1 : 0;
        if (mask_zyx) mask_zyx[((long long)iz * f.ny + iy) * f.nx + ix] = m ? 1 : 0;
    }
}

// one CTA per keyframe: how many of the N x N_samples points around the current frame's surface
// project inside the keyframe's image (edge margin) in front of the camera
__global__ void __launch_bounds__(256) k_keyframe_overlap(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                           const float* __restrict__ gt_depth, long long N, int Ns,
                                                           const float* __restrict__ t_vals,
                                                           const float* __restrict__ w2c, double fx, double fy, double cx,
                                                           double cy, int H, int W, int edge, int* __restrict__ counts) {
    __shared__ int s_count;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    Proj P;
#pragma unroll
    for (int k = 0; k < 12; ++k) P.w[k] = w2c[(size_t)blockIdx.x * 16 + k];
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    int local = 0;
    for (long long i = threadIdx.x; i < N * Ns; i += blockDim.x) {
        const long long r = i / Ns;
        const float t = t_vals[i - r * Ns];
        const float g = gt_depth[r];
        const float near = __fmul_rn(g, 0.8f), far = __fadd_rn(g, 0.5f);                    // Mapper.py:218-219
        const float zv = __fadd_rn(__fmul_rn(near, __fsub_rn(1.f, t)), __fmul_rn(far, t));
        const float x = __fadd_rn(rays_o[3 * r], __fmul_rn(rays_d[3 * r], zv));
        const float y = __fadd_rn(rays_o[3 * r + 1], __fmul_rn(rays_d[3 * r + 1], zv));
        const float z = __fadd_rn(rays_o[3 * r + 2], __fmul_rn(rays_d[3 * r + 2], zv));
        float u, v;
        double zz;
        project(P, x, y, z, u, v, zz);
        const bool m = (u < (float)(W - edge)) && (u > (float)edge) && (v < (float)(H - edge)) && (v > (float)edge) && (zz < 0.0);
        local += m ? 1 : 0;
    }
    for (int s = 16; s > 0; s >>= 1) local += __shfl_xor_sync(FULL, local, s);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_count, local);
    __syncthreads();
    if (threadIdx.x == 0) counts[blockIdx.x] = s_count;
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_frustum_mask(const float* w2c_host, const float* cam_pos_host, const float* xs, const float* ys,
                                const float* zs, int nx, int ny, int nz, const float* depth, int H, int W, double fx,
                                double fy, double cx, double cy, uint32_t* scratch, uint8_t* mask_xyz,
                                uint8_t* mask_zyx, void* stream) {
    if (!w2c_host || !cam_pos_host || !xs || !ys || !zs || !depth || !scratch) return set_error(NSB_EINVAL, "null argument");
    if (!mask_xyz && !mask_zyx) return set_error(NSB_EINVAL, "no output mask");
    if (nx < 1 || ny < 1 || nz < 1 || H < 1 || W < 1) return set_error(NSB_EINVAL, "bad extents");
    FrustumParams f;
    for (int k = 0; k < 12; ++k) f.P.w[k] = w2c_host[k];
    f.P.fx = fx; f.P.fy = fy; f.P.cx = cx; f.P.cy = cy;
    for (int k = 0; k < 3; ++k) f.cam[k] = cam_pos_host[k];
    f.xs = xs; f.ys = ys; f.zs = zs; f.depth = depth;
    f.nx = nx; f.ny = ny; f.nz = nz; f.H = H; f.W = W;
    const long long n = (long long)nx * ny * nz;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t), (cudaStream_t)stream);
    if (e != cudaSuccess) return set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
    k_frustum_depth_max<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(f, scratch);
    if (int rc = check_launch("k_frustum_depth_max")) return rc;
    k_frustum_mask<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(f, scratch, mask_xyz, mask_zyx);
    return check_launch("k_frustum_mask");
}

extern "C" int nsb_keyframe_overlap(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t N,
                                    int N_samples, const float* t_vals, const float* w2c, int K, int H, int W,
                                    double fx, double fy, double cx, double cy, int edge, int32_t* counts, void* stream) {
    if (!rays_o || !rays_d || !gt_depth || !t_vals || !w2c || !counts) return set_error(NSB_EINVAL, "null argument");
    if (K <= 0 || N <= 0) return 0;
    if (N_samples < 1) return set_error(NSB_EINVAL, "N_samples must be positive");
    k_keyframe_overlap<<<K, 256, 0, (cudaStream_t)stream>>>(rays_o, rays_d, gt_depth, N, N_samples, t_vals, w2c, fx, fy, cx,
                                                           cy, H, W, edge, counts);
    return check_launch("k_keyframe_overlap");
}
