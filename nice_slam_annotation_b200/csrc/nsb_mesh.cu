// Mesh depth rendering for the 2-D reconstruction metric: the reference rasterises the
// ground-truth and the reconstructed mesh through an Open3D GL window and reads the depth buffer back
// (src/tools/eval_recon.py:131-210, mesh_show_back_face = True, z far 20) -- unavailable on this box.  Here
// the same image is ray-cast headless: pixel (u, v) of a pinhole view (extrinsic = world -> camera, camera
// looks down +z, y down, as o3d.camera.PinholeCameraParameters) sees the nearest triangle along
// dir = ((u - cx) / fx, (v - cy) / fy, 1), so the ray parameter is the z-depth the GL buffer holds; pixels
// that see nothing are 0.
//
// One thread per (triangle, view): camera-space vertices, conservative pixel box of the projection (the whole
// part in front of the near plane for triangles that cross it), a Moeller-Trumbore test per covered pixel and an
// atomicMin on the depth bits (positive floats order like their bit patterns).  Triangles whose box is larger
// than INLINE_PIXELS go to a list that a second kernel walks with one block per triangle.  Reconstructed
// meshes are marching-cubes output (a few pixels per triangle), so the work is ~T x K pixel tests, not
// T x K x H x W.  The arithmetic of the test is fixed operation by operation (fp32, no FMA contraction):
// oracle/mesh_oracle.py restates it in numpy float32 and the images agree bit for bit.
#include "nsb_common.cuh"

namespace nsb {

constexpr int INLINE_PIXELS = 256;
constexpr float BARY_EPS = 1e-4f;      // neighbouring triangles overlap by this much: no cracks along shared edges

struct MeshView {
    const float* w2c;       // [K,16] row-major, device
    int H, W;
    float fx, fy, cx, cy, z_near, z_far;
};

struct Tri {
    float a[3], e1[3], e2[3];      // camera-space vertex 0 and the two edges
    int u0, u1, v0, v1;            // pixel box, inclusive; empty when u0 > u1
};

__device__ __forceinline__ void to_cam(const float* __restrict__ M, const float* __restrict__ p, float* c) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
        c[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[4 * k], p[0]), __fmul_rn(M[4 * k + 1], p[1])),
                                   __fmul_rn(M[4 * k + 2], p[2])), M[4 * k + 3]);
}

__device__ __forceinline__ bool load_tri(const MeshView& V, int view, const float* __restrict__ vert,
                                         const int32_t* __restrict__ tri, long long t, Tri& T) {
    const float* M = V.w2c + 16 * (size_t)view;
    float c[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) to_cam(M, vert + 3 * (size_t)tri[3 * t + i], c[i]);
    const float zmax = fmaxf(c[0][2], fmaxf(c[1][2], c[2][2])), zmin = fminf(c[0][2], fminf(c[1][2], c[2][2]));
    if (!(zmax >= V.z_near) || !(zmin <= V.z_far)) return false;      // NaN vertices fall out here too
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        T.a[k] = c[0][k];
        T.e1[k] = __fsub_rn(c[1][k], c[0][k]);
        T.e2[k] = __fsub_rn(c[2][k], c[0][k]);
    }
    // pixel box of the part of the triangle at z >= z_near (hits nearer than that are rejected anyway): the
    // vertices on that side and the points where edges cross the plane.  Triangles that straddle the camera
    // plane -- a ring of them around every camera inside a room -- project to a bounded, mostly off-screen box
    float ulo = INFINITY, uhi = -INFINITY, vlo = INFINITY, vhi = -INFINITY;
    auto add = [&](float x, float y, float z) {
        const float u = V.fx * x / z + V.cx, v = V.fy * y / z + V.cy;
        ulo = fminf(ulo, u); uhi = fmaxf(uhi, u); vlo = fminf(vlo, v); vhi = fmaxf(vhi, v);
    };
    const float zc = V.z_near;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3;
        const bool in_i = c[i][2] >= zc, in_j = c[j][2] >= zc;
        if (in_i) add(c[i][0], c[i][1], c[i][2]);
        if (in_i != in_j) {
            const float w = (zc - c[i][2]) / (c[j][2] - c[i][2]);
            add(c[i][0] + w * (c[j][0] - c[i][0]), c[i][1] + w * (c[j][1] - c[i][1]), zc);
        }
    }
    // slack: one pixel for the rounding of the projection, plus what BARY_EPS of the edge lengths can move
    // a hit at the nearest admissible depth
    const float len = fabsf(T.e1[0]) + fabsf(T.e1[1]) + fabsf(T.e1[2]) + fabsf(T.e2[0]) + fabsf(T.e2[1]) + fabsf(T.e2[2]);
    const float pad = 1.f + BARY_EPS * len * fmaxf(V.fx, V.fy) / fmaxf(zmin, zc);
    const float W1 = (float)(V.W - 1), H1 = (float)(V.H - 1);
    if (!(ulo <= uhi) || !(vlo <= vhi)) return false;
    T.u0 = (int)fminf(fmaxf(floorf(ulo - pad), 0.f), W1 + 1.f);
    T.u1 = (int)fmaxf(fminf(ceilf(uhi + pad), W1), -1.f);
    T.v0 = (int)fminf(fmaxf(floorf(vlo - pad), 0.f), H1 + 1.f);
    T.v1 = (int)fmaxf(fminf(ceilf(vhi + pad), H1), -1.f);
    return T.u0 <= T.u1 && T.v0 <= T.v1;
}

// the test oracle/mesh_oracle.py:ray_triangle restates; returns the depth or a negative number
__device__ __forceinline__ float ray_triangle(const MeshView& V, const Tri& T, int u, int v) {
    const float dx = __fdiv_rn(__fsub_rn((float)u, V.cx), V.fx), dy = __fdiv_rn(__fsub_rn((float)v, V.cy), V.fy);
    const float px = __fsub_rn(__fmul_rn(dy, T.e2[2]), T.e2[1]);
    const float py = __fsub_rn(T.e2[0], __fmul_rn(dx, T.e2[2]));
    const float pz = __fsub_rn(__fmul_rn(dx, T.e2[1]), __fmul_rn(dy, T.e2[0]));
    const float det = __fadd_rn(__fadd_rn(__fmul_rn(T.e1[0], px), __fmul_rn(T.e1[1], py)), __fmul_rn(T.e1[2], pz));
    if (!(fabsf(det) > 1e-20f)) return -1.f;
    const float inv = __fdiv_rn(1.f, det);
    const float tx = -T.a[0], ty = -T.a[1], tz = -T.a[2];
    const float bu = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(tx, px), __fmul_rn(ty, py)), __fmul_rn(tz, pz)), inv);
    if (!(bu >= -BARY_EPS) || !(bu <= 1.f + BARY_EPS)) return -1.f;
    const float qx = __fsub_rn(__fmul_rn(ty, T.e1[2]), __fmul_rn(tz, T.e1[1]));
    const float qy = __fsub_rn(__fmul_rn(tz, T.e1[0]), __fmul_rn(tx, T.e1[2]));
    const float qz = __fsub_rn(__fmul_rn(tx, T.e1[1]), __fmul_rn(ty, T.e1[0]));
    const float bv = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, qx), __fmul_rn(dy, qy)), qz), inv);
    if (!(bv >= -BARY_EPS) || !(__fadd_rn(bu, bv) <= 1.f + BARY_EPS)) return -1.f;
    const float t = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(T.e2[0], qx), __fmul_rn(T.e2[1], qy)),
                                        __fmul_rn(T.e2[2], qz)), inv);
    return (t >= V.z_near && t <= V.z_far) ? t : -1.f;
}

__global__ void k_mesh_clear(uint32_t* __restrict__ zbuf, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        zbuf[i] = 0x7f800000u;      // +inf
}

__global__ void k_mesh_raster(MeshView V, const float* __restrict__ vert, const int32_t* __restrict__ tri,
                              long long n_tri, uint32_t* __restrict__ zbuf, int2* __restrict__ big,
                              unsigned* __restrict__ n_big, unsigned big_cap) {
    const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int view = blockIdx.y;
    if (t >= n_tri) return;
    Tri T;
    if (!load_tri(V, view, vert, tri, t, T)) return;
    const long long area = (long long)(T.u1 - T.u0 + 1) * (T.v1 - T.v0 + 1);
    if (area > INLINE_PIXELS && big) {
        const unsigned slot = atomicAdd(n_big, 1u);
        if (slot < big_cap) {
            big[slot] = make_int2(view, (int)t);
            return;
        }                            // list full: fall through and do it here
    }
    uint32_t* img = zbuf + (size_t)view * V.H * V.W;
    for (int v = T.v0; v <= T.v1; ++v)
        for (int u = T.u0; u <= T.u1; ++u) {
            const float d = ray_triangle(V, T, u, v);
            if (d >= 0.f) atomicMin(img + (size_t)v * V.W + u, __float_as_uint(d));
        }
}

__global__ void k_mesh_raster_big(MeshView V, const float* __restrict__ vert, const int32_t* __restrict__ tri,
                                  uint32_t* __restrict__ zbuf, const int2* __restrict__ big,
                                  const unsigned* __restrict__ n_big, unsigned big_cap) {
    const unsigned n = min(*n_big, big_cap);
    for (unsigned e = blockIdx.x; e < n; e += gridDim.x) {
        const int2 it = big[e];
        Tri T;
        if (!load_tri(V, it.x, vert, tri, it.y, T)) continue;
        uint32_t* img = zbuf + (size_t)it.x * V.H * V.W;
        const int bw = T.u1 - T.u0 + 1;
        const long long area = (long long)bw * (T.v1 - T.v0 + 1);
        for (long long i = threadIdx.x; i < area; i += blockDim.x) {
            const int u = T.u0 + (int)(i % bw), v = T.v0 + (int)(i / bw);
            const float d = ray_triangle(V, T, u, v);
            if (d >= 0.f) atomicMin(img + (size_t)v * V.W + u, __float_as_uint(d));
        }
    }
}

__global__ void k_mesh_resolve(uint32_t* __restrict__ zbuf, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        if (zbuf[i] == 0x7f800000u) zbuf[i] = 0u;      // nothing seen: 0.0f like the GL read-back
}

}  // namespace nsb

using namespace nsb;

extern "C" int64_t nsb_mesh_depth_workspace_bytes(int64_t n_triangles, int K) {
    // counter (256 B) + list of (view, triangle) pairs whose pixel box is large
    int64_t cap = n_triangles * (int64_t)K;
    if (cap > (1ll << 22)) cap = 1ll << 22;
    if (cap < 1024) cap = 1024;
    return 256 + cap * 8;
}

extern "C" int nsb_mesh_depth(const float* vertices, int64_t n_vertices, const int32_t* triangles,
                              int64_t n_triangles, const float* w2c, int K, int H, int W, float fx, float fy,
                              float cx, float cy, float z_near, float z_far, void* workspace,
                              int64_t workspace_bytes, float* depth, void* stream) {
    if (!depth || !w2c || (n_triangles > 0 && (!vertices || !triangles)))
        return set_error(NSB_EINVAL, "null argument");
    if (K <= 0 || H <= 0 || W <= 0 || n_triangles < 0 || n_vertices < 0 || K > 65535)
        return set_error(NSB_EINVAL, "bad mesh depth extents (K %d, H %d, W %d)", K, H, W);
    if (!(z_near > 0.f) || !(z_far >= z_near)) return set_error(NSB_EINVAL, "need 0 < z_near <= z_far");
    if (n_triangles > 0x7fffffffll) return set_error(NSB_EINVAL, "too many triangles");
    cudaStream_t s = (cudaStream_t)stream;
    MeshView V;
    V.w2c = w2c; V.H = H; V.W = W; V.fx = fx; V.fy = fy; V.cx = cx; V.cy = cy; V.z_near = z_near; V.z_far = z_far;
    const long long n_pix = (long long)K * H * W;
    uint32_t* zbuf = (uint32_t*)depth;
    const int fill_blocks = (int)((n_pix + 1023) / 1024 < 148 * 8 ? (n_pix + 1023) / 1024 : 148 * 8);
    k_mesh_clear<<<fill_blocks, 256, 0, s>>>(zbuf, n_pix);
    if (int rc = check_launch("k_mesh_clear")) return rc;
    if (n_triangles > 0) {
        unsigned* n_big = nullptr;
        int2* big = nullptr;
        unsigned cap = 0;
        if (workspace && workspace_bytes >= 256 + 8) {
            n_big = (unsigned*)workspace;
            big = (int2*)((char*)workspace + 256);
            const int64_t c = (workspace_bytes - 256) / 8;
            cap = (unsigned)(c > 0x7fffffffll ? 0x7fffffffll : c);
            cudaError_t e = cudaMemsetAsync(n_big, 0, 256, s);
            if (e != cudaSuccess) return set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
        }
        dim3 grid((unsigned)((n_triangles + 127) / 128), (unsigned)K);
        k_mesh_raster<<<grid, 128, 0, s>>>(V, vertices, triangles, n_triangles, zbuf, big, n_big, cap);
        if (int rc = check_launch("k_mesh_raster")) return rc;
        if (big) {
            k_mesh_raster_big<<<148 * 4, 256, 0, s>>>(V, vertices, triangles, zbuf, big, n_big, cap);
            if (int rc = check_launch("k_mesh_raster_big")) return rc;
        }
    }
    k_mesh_resolve<<<fill_blocks, 256, 0, s>>>(zbuf, n_pix);
    return check_launch("k_mesh_resolve");
}
