// K1: pixel -> ray, bound pre-filter, batch depth max, sample depths along the rays and the
// iMAP importance resampling.  Replaces get_samples / get_rays (src/common.py:74-134,
// 259-277), the sampling half of Renderer.render_batch_ray (src/utils/Renderer.py:116-224)
// and sample_pdf (src/common.py:19-63).  Depths are fp64 like the reference (its scene bound
// is an fp64 tensor, NICE_SLAM.py:196, which promotes everything downstream).
#include "nsb_common.cuh"

namespace nsb {

constexpr int SAMPLE_Z_INLINE_MAX = 4096;   // up to here k_sample_z reduces the batch depth max in every block
constexpr int MAX_LIST = 128;   // per-ray limit of N_samples, N_surface, N_importance each

struct CamDev {
    float fx, fy, cx, cy;
};

// dirs = [(i-cx)/fx, -(j-cy)/fy, -1]; rays_d = sum(dirs * R, -1) evaluated without fusing, in
// the order torch evaluates common.py:82-87
__device__ __forceinline__ void pixel_ray(float i, float j, const CamDev& cam,
                                          const float* __restrict__ c2w, float (&d)[3]) {
    const float x = __fdiv_rn(__fsub_rn(i, cam.cx), cam.fx);
    const float y = -__fdiv_rn(__fsub_rn(j, cam.cy), cam.fy);
    const float z = -1.f;
#pragma unroll
    for (int a = 0; a < 3; ++a)
        d[a] = __fadd_rn(__fadd_rn(__fmul_rn(x, c2w[4 * a]), __fmul_rn(y, c2w[4 * a + 1])),
                         __fmul_rn(z, c2w[4 * a + 2]));
}

// NaN-propagating max/min like torch.max / torch.min
__device__ __forceinline__ double tmax(double a, double b) { return (a > b || a != a) ? a : b; }
__device__ __forceinline__ double tmin(double a, double b) { return (a < b || a != a) ? a : b; }

// slab exit of the scene box: min_axis max(lo,hi)((bound - o)/d), fp64 (Renderer.py:139-150)
__device__ __forceinline__ double box_exit(const float (&o)[3], const float (&d)[3], const BoundDev& b) {
    double far = 0.0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        double tl = __ddiv_rn(__dsub_rn(b.lo[a], (double)o[a]), (double)d[a]);
        double th = __ddiv_rn(__dsub_rn(b.hi[a], (double)o[a]), (double)d[a]);
        double m = tmax(tl, th);
        far = a == 0 ? m : tmin(far, m);
    }
    return far;
}

// [quat, t] -> 3x4 pose in fp32 with the operation order of quad2rotation (common.py:137-160)
__device__ __forceinline__ void cam7_to_c2w(const float* __restrict__ q7, float (&m)[12]) {
    const float qr = q7[0], qi = q7[1], qj = q7[2], qk = q7[3];
    const float nn = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(qr, qr), __fmul_rn(qi, qi)), __fmul_rn(qj, qj)), __fmul_rn(qk, qk));
    const float s = __fdiv_rn(2.0f, nn);
    auto sq = [](float a) { return __fmul_rn(a, a); };
    auto mm = [](float a, float b) { return __fmul_rn(a, b); };
    m[0] = __fsub_rn(1.f, mm(s, __fadd_rn(sq(qj), sq(qk))));
    m[1] = mm(s, __fsub_rn(mm(qi, qj), mm(qk, qr)));
    m[2] = mm(s, __fadd_rn(mm(qi, qk), mm(qj, qr)));
    m[4] = mm(s, __fadd_rn(mm(qi, qj), mm(qk, qr)));
    m[5] = __fsub_rn(1.f, mm(s, __fadd_rn(sq(qi), sq(qk))));
    m[6] = mm(s, __fsub_rn(mm(qj, qk), mm(qi, qr)));
    m[8] = mm(s, __fsub_rn(mm(qi, qk), mm(qj, qr)));
    m[9] = mm(s, __fadd_rn(mm(qj, qk), mm(qi, qr)));
    m[10] = __fsub_rn(1.f, mm(s, __fadd_rn(sq(qi), sq(qj))));
    m[3] = q7[4]; m[7] = q7[5]; m[11] = q7[6];
}

__global__ void k_sample_rays(const int64_t* __restrict__ indices, long long n, int H0, int W0,
                              int hc, int wc, int W_img, CamDev cam, const float* __restrict__ c2w_in,
                              const float* __restrict__ cam7,
                              const float* __restrict__ depth_img, const void* __restrict__ color_img,
                              int color_is_f64, BoundDev bound, int have_bound,
                              float* __restrict__ rays_o, float* __restrict__ rays_d,
                              float* __restrict__ gt_depth, float* __restrict__ gt_color,
                              uint8_t* __restrict__ keep) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= n) return;
    long long idx = indices[r];
    const long long total = (long long)hc * wc;
    idx = idx < 0 ? 0 : (idx >= total ? total - 1 : idx);
    const int pi = W0 + (int)(idx % wc), pj = H0 + (int)(idx / wc);
    float d[3], o[3];
    float c2w[12];
    if (cam7) {
        cam7_to_c2w(cam7, c2w);
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) c2w[k] = c2w_in[k];
    }
    pixel_ray((float)pi, (float)pj, cam, c2w, d);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        o[a] = c2w[4 * a + 3];
        rays_o[3 * r + a] = o[a];
        rays_d[3 * r + a] = d[a];
    }
    const long long pix = (long long)pj * W_img + pi;
    const float gd = depth_img[pix];
    if (gt_depth) gt_depth[r] = gd;
    if (gt_color) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
            gt_color[3 * r + c] = color_is_f64 ? (float)reinterpret_cast<const double*>(color_img)[3 * pix + c]
                                               : reinterpret_cast<const float*>(color_img)[3 * pix + c];
    }
    if (keep) keep[r] = have_bound ? (box_exit(o, d, bound) >= (double)gd ? 1 : 0) : 1;
}

__global__ void k_all_rays(int H, int W, CamDev cam, const float* __restrict__ c2w,
                           float* __restrict__ rays_o, float* __restrict__ rays_d) {
    long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= (long long)H * W) return;
    float d[3];
    pixel_ray((float)(r % W), (float)(r / W), cam, c2w, d);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        rays_o[3 * r + a] = c2w[4 * a + 3];
        rays_d[3 * r + a] = d[a];
    }
}

__global__ void k_depth_max(const float* __restrict__ gt, const uint8_t* __restrict__ keep, long long n,
                            float* __restrict__ out) {
    __shared__ float part[32];
    float m = -INFINITY;
    bool nan = false;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) {
        if (keep && !keep[i]) continue;
        float v = gt[i];
        nan |= (v != v);
        m = fmaxf(m, v);
    }
    if (nan) m = NAN;
    for (int s = 16; s > 0; s >>= 1) {
        float o = __shfl_xor_sync(FULL, m, s);
        m = (o > m || o != o) ? o : m;
    }
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : -INFINITY;
        for (int s = 16; s > 0; s >>= 1) {
            float o = __shfl_xor_sync(FULL, m, s);
            m = (o > m || o != o) ? o : m;
        }
        if (threadIdx.x == 0) out[0] = m;
    }
}

__device__ __forceinline__ bool is_sorted(const double* v, int n) {
    bool ok = true;
    for (int i = 1; i < n; ++i) ok &= (v[i - 1] <= v[i]);
    return ok;
}

__device__ __forceinline__ void insertion_sort(double* v, int n) {
    for (int i = 1; i < n; ++i) {
        double x = v[i];
        int j = i - 1;
        // NaNs sort last, as torch.sort does
        while (j >= 0 && (v[j] > x || (v[j] != v[j] && x == x))) { v[j + 1] = v[j]; --j; }
        v[j + 1] = x;
    }
}

// ascending merge of two ascending lists (what torch.sort(torch.cat(...)) returns for them)
__device__ __forceinline__ void merge_sorted(const double* a, int na, const double* b, int nb,
                                             double* __restrict__ out) {
    int i = 0, j = 0, k = 0;
    while (i < na && j < nb) out[k++] = (b[j] < a[i]) ? b[j++] : a[i++];
    while (i < na) out[k++] = a[i++];
    while (j < nb) out[k++] = b[j++];
}

// One thread per (ray, sample): the thread evaluates its own depth (uniform or surface list,
// jitter from its two neighbours), the block ranks the S values of each ray against each other
// in shared memory (NaNs last, ties by position: a stable ascending sort, i.e. what
// torch.sort(torch.cat([z_uniform, z_surface])) returns) and writes z[rank].
__global__ void __launch_bounds__(256) k_sample_z(PointsDev q, const float* __restrict__ gt_depth, BoundDev bound,
                           int Ns, int Nf, int lindisp, const float* __restrict__ t_uniform,
                           const double* __restrict__ t_surface, const float* __restrict__ u,
                           const float* __restrict__ depth_max, double* __restrict__ z_out, int rays_per_block,
                           unsigned* __restrict__ zero_word, const uint8_t* __restrict__ keep) {
    extern __shared__ double sh_z[];
    __shared__ float s_part[8];
    __shared__ float s_dmax;
    pdl_launch_dependents();     // a decoder kernel launched with the PDL attribute may set itself up meanwhile
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const int S = Ns + Nf;
    if (gt_depth && !depth_max) {
        // small batches: every block reduces the batch depth max itself (k_depth_max's rule: kept rays,
        // NaN propagates) instead of waiting for a separate launch
        float m = -INFINITY;
        for (long long i = threadIdx.x; i < q.N; i += blockDim.x)
            if (!keep || keep[i]) { const float v = gt_depth[i]; m = (v > m || v != v) ? v : m; }
        for (int sft = 16; sft > 0; sft >>= 1) { const float o = __shfl_xor_sync(FULL, m, sft); m = (o > m || o != o) ? o : m; }
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { const float o = s_part[w]; m = (o > m || o != o) ? o : m; }
            s_dmax = m;
        }
        __syncthreads();
    } else if (threadIdx.x == 0) {
        s_dmax = depth_max ? depth_max[0] : 0.f;
    }
    if (!(gt_depth && !depth_max)) __syncthreads();
    const float dmax = s_dmax;
    const int local = threadIdx.x / S, k = threadIdx.x - local * S;
    const long long r = blockIdx.x * (long long)rays_per_block + local;
    if (zero_word && blockIdx.x == 0 && threadIdx.x == 0) *zero_word = 0u;
    const bool live = local < rays_per_block && r < q.N;
    double v = 0.0;
    if (live) {
        if (k < Ns) {
            float o[3], d[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) { o[a] = q.ro[r * q.so + a]; d[a] = q.rd[r * q.sd + a]; }
            const double far_bb = __dadd_rn(box_exit(o, d, bound), 0.01);
            float near;
            double far;
            if (gt_depth) {
                near = __fmul_rn(gt_depth[r], 0.01f);
                const double cap = (double)__fmul_rn(dmax, 1.2f);             // max(gt_depth * 1.2)
                far = tmin(tmax(far_bb, 0.0), cap);                            // torch.clamp(far_bb, 0, cap)
                if (far_bb != far_bb) far = far_bb;
            } else {
                near = 0.01f;
                far = far_bb;
            }
            auto uniform_z = [&](int kk) -> double {
                const float t = t_uniform[kk];
                const float omt = __fsub_rn(1.f, t);
                if (!lindisp) return __dadd_rn((double)__fmul_rn(near, omt), __dmul_rn(far, (double)t));
                const float inv_near = gt_depth ? __fdiv_rn(1.f, near) : 100.f;
                const double s = __dadd_rn((double)__fmul_rn(inv_near, omt), __dmul_rn(__ddiv_rn(1.0, far), (double)t));
                return __ddiv_rn(1.0, s);
            };
            v = uniform_z(k);
            if (u) {   // stratified jitter (Renderer.py:212-219)
                const double lower = k > 0 ? __dmul_rn(0.5, __dadd_rn(v, uniform_z(k - 1))) : v;
                const double upper = (k + 1 < Ns) ? __dmul_rn(0.5, __dadd_rn(uniform_z(k + 1), v)) : v;
                v = __dadd_rn(lower, __dmul_rn(__dsub_rn(upper, lower), (double)u[r * Ns + k]));
            }
        } else {
            const double ts = t_surface[k - Ns];
            const float g = gt_depth[r];
            if (g > 0.f) {       // 0.95 d .. 1.05 d (Renderer.py:177-185)
                const double a = (double)__fmul_rn(0.95f, g), b = (double)__fmul_rn(1.05f, g);
                v = __dadd_rn(__dmul_rn(a, __dsub_rn(1.0, ts)), __dmul_rn(b, ts));
            } else {             // no sensor depth: 0.001 .. max depth of the batch (Renderer.py:191-200)
                v = __dadd_rn(__dmul_rn(0.001, __dsub_rn(1.0, ts)), __dmul_rn((double)dmax, ts));
            }
        }
        sh_z[threadIdx.x] = v;
    }
    __syncthreads();
    if (!live) return;
    int rank = k;
    if (Nf > 0) {                 // the uniform list alone is written in place, unsorted, like the reference
        rank = 0;
        const double* row = sh_z + local * S;
        const bool vnan = v != v;
        for (int j = 0; j < S; ++j) {
            const double w = row[j];
            const bool wnan = w != w;
            const bool w_less = (w < v) || (!wnan && vnan);
            const bool v_less = (v < w) || (!vnan && wnan);
            rank += (w_less || (!v_less && j < k)) ? 1 : 0;
        }
    }
    z_out[r * (long long)S + rank] = v;
    if (q.pu) {
        // point cache for the decoder kernels (NsbPoints::point_cache): p = o + d z in fp64, un-fused, as torch evaluates
        // Renderer.py:227-228, its fp32 value (the embedding input, decoder.py:204) and the normalised grid coordinate
        // (common.py:280-295) -- the expressions of ray_point / tc_load_point, evaluated once per point here instead of
        // once per point and decoder there
        float* c = q.pu + 8 * (r * (long long)S + rank);
        float p32[3], un[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const double p = __dadd_rn((double)q.ro[r * q.so + a], __dmul_rn((double)q.rd[r * q.sd + a], v));
            p32[a] = (float)p;
            un[a] = (float)__dsub_rn(__dmul_rn(__ddiv_rn(__dsub_rn(p, bound.lo[a]), __dsub_rn(bound.hi[a], bound.lo[a])), 2.0), 1.0);
        }
        reinterpret_cast<float4*>(c)[0] = make_float4(p32[0], p32[1], p32[2], 0.f);
        reinterpret_cast<float4*>(c)[1] = make_float4(un[0], un[1], un[2], 0.f);
    }
}

// sample_pdf (common.py:19-63) on bins = mid(z), weights = w[1:-1]; then the sort-merge of
// Renderer.py:249
__global__ void k_importance_z(const double* __restrict__ z, const float* __restrict__ w,
                               long long N, int S, int Ni, const float* __restrict__ u,
                               long long u_stride, double* __restrict__ z_samples,
                               double* __restrict__ z_out) {
    long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= N) return;
    const double* zr = z + r * (long long)S;
    const float* wr = w + r * (long long)S;
    const int M = S - 1;          // bins
    const int K = S - 2;          // weights
    float cdf[MAX_LIST];
    double zs[MAX_LIST];
    float sum = 0.f;
    for (int k = 0; k < K; ++k) sum = __fadd_rn(sum, __fadd_rn(wr[k + 1], 1e-5f));
    cdf[0] = 0.f;
    float run = 0.f;
    for (int k = 0; k < K; ++k) {
        run = __fadd_rn(run, __fdiv_rn(__fadd_rn(wr[k + 1], 1e-5f), sum));
        cdf[k + 1] = run;
    }
    const int L = K + 1;          // cdf length == number of bins
    for (int k = 0; k < Ni; ++k) {
        const float uu = u[r * u_stride + k];
        int lo = 0, hi = L;       // searchsorted(right=True): first index with cdf > u
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (cdf[mid] > uu) hi = mid; else lo = mid + 1;
        }
        const int below = lo - 1 > 0 ? lo - 1 : 0;
        const int above = lo < L - 1 ? lo : L - 1;
        float denom = __fsub_rn(cdf[above], cdf[below]);
        if (denom < 1e-5f) denom = 1.f;
        const float t = __fdiv_rn(__fsub_rn(uu, cdf[below]), denom);
        const double bb = __dmul_rn(0.5, __dadd_rn(zr[below + 1], zr[below]));
        const double ba = __dmul_rn(0.5, __dadd_rn(zr[above + 1], zr[above]));
        zs[k] = __dadd_rn(bb, __dmul_rn((double)t, __dsub_rn(ba, bb)));
        if (z_samples) z_samples[r * (long long)Ni + k] = zs[k];
    }
    (void)M;
    if (!is_sorted(zs, Ni)) insertion_sort(zs, Ni);
    double* out = z_out + r * (long long)(S + Ni);
    if (is_sorted(zr, S)) {
        merge_sorted(zr, S, zs, Ni, out);
    } else {
        for (int k = 0; k < S; ++k) out[k] = zr[k];
        for (int k = 0; k < Ni; ++k) out[S + k] = zs[k];
        insertion_sort(out, S + Ni);
    }
}

// sample_pdf (common.py:19-63) as a function of explicit bins [N,M] and weights [N,M-1]
__global__ void k_sample_pdf(const double* __restrict__ bins, const float* __restrict__ w, long long N,
                             int M, int Ni, const float* __restrict__ u, long long u_stride,
                             double* __restrict__ samples) {
    long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= N) return;
    const double* br = bins + r * (long long)M;
    const float* wr = w + r * (long long)(M - 1);
    const int K = M - 1;
    float cdf[MAX_LIST];
    float sum = 0.f;
    for (int k = 0; k < K; ++k) sum = __fadd_rn(sum, __fadd_rn(wr[k], 1e-5f));
    cdf[0] = 0.f;
    float run = 0.f;
    for (int k = 0; k < K; ++k) {
        run = __fadd_rn(run, __fdiv_rn(__fadd_rn(wr[k], 1e-5f), sum));
        cdf[k + 1] = run;
    }
    const int L = K + 1;
    for (int k = 0; k < Ni; ++k) {
        const float uu = u[r * u_stride + k];
        int lo = 0, hi = L;
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (cdf[mid] > uu) hi = mid; else lo = mid + 1;
        }
        const int below = lo - 1 > 0 ? lo - 1 : 0;
        const int above = lo < L - 1 ? lo : L - 1;
        float denom = __fsub_rn(cdf[above], cdf[below]);
        if (denom < 1e-5f) denom = 1.f;
        const float t = __fdiv_rn(__fsub_rn(uu, cdf[below]), denom);
        samples[r * (long long)Ni + k] = __dadd_rn(br[below], __dmul_rn((double)t, __dsub_rn(br[above], br[below])));
    }
}

}  // namespace nsb

using namespace nsb;

static inline unsigned blocks_for(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

extern "C" int nsb_sample_rays(const int64_t* indices, int64_t n, int H0, int H1, int W0, int W1,
                               int W_img, float fx, float fy, float cx, float cy, const float* c2w,
                               const float* cam7, const float* depth_img, const void* color_img,
                               int color_is_f64, const double* bound6_host, float* rays_o, float* rays_d,
                               float* gt_depth, float* gt_color, uint8_t* keep, void* stream) {
    if (!indices || (!c2w && !cam7) || !depth_img || !rays_o || !rays_d) return set_error(NSB_EINVAL, "null argument");
    if (gt_color && !color_img) return set_error(NSB_EINVAL, "gt_color requested without an image");
    if (H1 <= H0 || W1 <= W0) return set_error(NSB_EINVAL, "empty crop");
    if (n <= 0) return 0;
    BoundDev b{};
    if (bound6_host) b = to_bound(bound6_host);
    CamDev cam{fx, fy, cx, cy};
    launch_pdl(k_sample_rays, dim3(blocks_for(n, 128)), dim3(128), 0, (cudaStream_t)stream,
        indices, n, H0, W0, H1 - H0, W1 - W0, W_img, cam, c2w, cam7, depth_img, color_img, color_is_f64, b,
        bound6_host != nullptr, rays_o, rays_d, gt_depth, gt_color, keep);
    return check_launch("k_sample_rays");
}

// Box pre-filter of explicit rays (Tracker.py:100-107, Mapper.py:587-594): keep[r] = (exit depth of ray r from
// the scene box, fp64 like the reference's .to(float64) bound) >= gt_depth[r].  Same box_exit as k_sample_rays.
__global__ void k_box_prefilter(const float* __restrict__ ro, const float* __restrict__ rd,
                                const float* __restrict__ gt_depth, long long n, BoundDev bound,
                                uint8_t* __restrict__ keep) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= n) return;
    float o[3], d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) { o[a] = ro[3 * r + a]; d[a] = rd[3 * r + a]; }
    keep[r] = box_exit(o, d, bound) >= (double)gt_depth[r] ? 1 : 0;
}

extern "C" int nsb_box_prefilter(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t n,
                                 const double* bound6_host, uint8_t* keep, void* stream) {
    if (n <= 0) return 0;
    if (!rays_o || !rays_d || !gt_depth || !bound6_host || !keep) return set_error(NSB_EINVAL, "null argument");
    k_box_prefilter<<<(unsigned)blocks_for(n, 128), 128, 0, (cudaStream_t)stream>>>(rays_o, rays_d, gt_depth, n,
                                                                                    to_bound(bound6_host), keep);
    return check_launch("k_box_prefilter");
}

extern "C" int nsb_all_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w,
                            float* rays_o, float* rays_d, void* stream) {
    if (!c2w || !rays_o || !rays_d) return set_error(NSB_EINVAL, "null argument");
    if (H <= 0 || W <= 0) return 0;
    CamDev cam{fx, fy, cx, cy};
    k_all_rays<<<blocks_for((long long)H * W, 256), 256, 0, (cudaStream_t)stream>>>(H, W, cam, c2w, rays_o, rays_d);
    return check_launch("k_all_rays");
}

namespace nsb {
// one batch into the fixed-shape input buffers of an engine step: n <= N rays are copied, the rest is
// masked off (keep = 0) behind harmless finite rays
__global__ void k_load_batch(const float* __restrict__ ro, const float* __restrict__ rd, const float* __restrict__ gd,
                             const float* __restrict__ gc, const uint8_t* __restrict__ keep, long long n, long long N,
                             float* __restrict__ d_ro, float* __restrict__ d_rd, float* __restrict__ d_gd,
                             float* __restrict__ d_gc, uint8_t* __restrict__ d_keep) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= N) return;
    const bool live = i < n;
    if (!live || ro) {          // ro == NULL: the live entries are already in place (copied by the caller)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            d_ro[3 * i + a] = live ? ro[3 * i + a] : 0.f;
            d_rd[3 * i + a] = live ? rd[3 * i + a] : 1.f;
            d_gc[3 * i + a] = live ? gc[3 * i + a] : 0.f;
        }
        d_gd[i] = live ? gd[i] : 0.f;
    }
    d_keep[i] = live ? (keep ? keep[i] : (uint8_t)1) : (uint8_t)0;
}

// the same from one packed [n,10] array (rays_o 3, rays_d 3, gt_depth 1, gt_color 3 per ray): what a host
// producer hands over with a single H2D copy
__global__ void k_load_batch_packed(const float* __restrict__ src, const uint8_t* __restrict__ keep, long long n,
                                    long long N, float* __restrict__ d_ro, float* __restrict__ d_rd,
                                    float* __restrict__ d_gd, float* __restrict__ d_gc, uint8_t* __restrict__ d_keep) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= N) return;
    const bool live = i < n;
    const float* r = src + 10 * i;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        d_ro[3 * i + a] = live ? r[a] : 0.f;
        d_rd[3 * i + a] = live ? r[3 + a] : 1.f;
        d_gc[3 * i + a] = live ? r[7 + a] : 0.f;
    }
    d_gd[i] = live ? r[6] : 0.f;
    d_keep[i] = live ? (keep ? keep[i] : (uint8_t)1) : (uint8_t)0;
}

int launch_sample_z(const PointsDev& q, const float* gt_depth, const BoundDev& bound, int Ns, int Nf, int lindisp,
                    const float* t_uniform, const double* t_surface, const float* u, const float* depth_max,
                    double* z_out, unsigned* zero_word, const uint8_t* keep, void* stream) {
    const int S = Ns + Nf;
    const int rpb = S >= 256 ? 1 : 256 / S;
    const int threads = (rpb * S + 31) / 32 * 32;             // whole warps (the in-block depth max shuffles)
    launch_pdl(k_sample_z, dim3(blocks_for(q.N, rpb)), dim3(threads), threads * sizeof(double), (cudaStream_t)stream,
        q, gt_depth, bound, Ns, Nf, lindisp, t_uniform, t_surface, u, depth_max, z_out, rpb, zero_word, keep);
    return check_launch("k_sample_z");
}
}  // namespace nsb

extern "C" int nsb_depth_max(const float* gt_depth, const uint8_t* keep, int64_t N, float* out, void* stream) {
    if (!gt_depth || !out) return set_error(NSB_EINVAL, "null argument");
    k_depth_max<<<1, 1024, 0, (cudaStream_t)stream>>>(gt_depth, keep, N, out);
    return check_launch("k_depth_max");
}

namespace nsb {
int sample_z_impl(const NsbPoints* rays, const float* gt_depth, const double* bound6_host, int N_samples,
                  int N_surface, int lindisp, const float* t_uniform, const double* t_surface, const float* u,
                  const float* depth_max, double* z_out, unsigned* zero_word, const uint8_t* keep, void* stream) {
    if (!rays || !rays->rays_o || !rays->rays_d || !bound6_host || !t_uniform || !z_out)
        return set_error(NSB_EINVAL, "null argument");
    if (N_samples < 1 || N_samples > MAX_LIST || N_surface < 0 || N_surface > MAX_LIST)
        return set_error(NSB_EINVAL, "N_samples / N_surface out of range (1..%d)", MAX_LIST);
    if (!gt_depth) N_surface = 0;
    if (gt_depth && !depth_max && rays->N > SAMPLE_Z_INLINE_MAX)
        return set_error(NSB_EINVAL, "depth_max is required with gt_depth (batches above %d rays)", SAMPLE_Z_INLINE_MAX);
    if (N_surface > 0 && !t_surface) return set_error(NSB_EINVAL, "t_surface is NULL");
    if (rays->N <= 0) return 0;
    PointsDev q = to_dev(*rays);
    return launch_sample_z(q, gt_depth, to_bound(bound6_host), N_samples, N_surface, lindisp, t_uniform, t_surface,
                           u, depth_max, z_out, zero_word, keep, stream);
}
}  // namespace nsb

extern "C" int nsb_sample_z(const NsbPoints* rays, const float* gt_depth, const double* bound6_host,
                            int N_samples, int N_surface, int lindisp, const float* t_uniform,
                            const double* t_surface, const float* u, const float* depth_max,
                            double* z_out, void* stream) {
    if (gt_depth && !depth_max) return set_error(NSB_EINVAL, "depth_max is required with gt_depth");
    return sample_z_impl(rays, gt_depth, bound6_host, N_samples, N_surface, lindisp, t_uniform, t_surface, u,
                         depth_max, z_out, nullptr, nullptr, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// iMAP free-space regulariser points (Renderer.py:343-409): S stratified, always-jittered fp32 depths in
// [0, 0.85 d] of each ray and the points o + d z they name, one thread per (ray, sample).  Each thread evaluates
// its own depth and its two neighbours' (three products, no shared memory); every product and sum is the un-fused
// fp32 operation torch evaluates, so points and depths equal the eager expressions bit for bit.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_regulation_points(const float* __restrict__ ro, const float* __restrict__ rd,
                                                           const float* __restrict__ gt_depth, long long N, int S,
                                                           const float* __restrict__ t_uniform,
                                                           const float* __restrict__ u, float* __restrict__ pts,
                                                           float* __restrict__ z_out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= N * S) return;
    const long long r = i / S;
    const int k = (int)(i - r * S);
    const float far = __fmul_rn(gt_depth[r], 0.85f);
    auto zk = [&](int kk) -> float {                       // near * (1 - t) + far * t with near = 0
        const float t = t_uniform[kk];
        return __fadd_rn(__fmul_rn(0.f, __fsub_rn(1.f, t)), __fmul_rn(far, t));
    };
    const float zc = zk(k);
    const float lower = k > 0 ? __fmul_rn(0.5f, __fadd_rn(zc, zk(k - 1))) : zc;
    const float upper = k + 1 < S ? __fmul_rn(0.5f, __fadd_rn(zk(k + 1), zc)) : zc;
    const float z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), u[i]));
    if (z_out) z_out[i] = z;
#pragma unroll
    for (int a = 0; a < 3; ++a) pts[i * 3 + a] = __fadd_rn(ro[r * 3 + a], __fmul_rn(rd[r * 3 + a], z));
}

// adjoint of p = o + d z (z carries no gradient: the sensor depth is data): d(o) = sum_s d(p), d(d) = sum_s z d(p);
// one warp per ray, lanes stride the samples, fixed-order butterfly sum.
__global__ void __launch_bounds__(256) k_regulation_points_bwd(const float* __restrict__ d_pts,
                                                               const float* __restrict__ z, long long N, int S,
                                                               float* __restrict__ d_ro, float* __restrict__ d_rd) {
    const long long r = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= N) return;
    float go[3] = {0.f, 0.f, 0.f}, gd[3] = {0.f, 0.f, 0.f};
    for (int k = lane; k < S; k += 32) {
        const float zz = z[r * S + k];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float g = d_pts[(r * S + k) * 3 + a];
            go[a] += g;
            gd[a] = fmaf(g, zz, gd[a]);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
        for (int sft = 16; sft > 0; sft >>= 1) {
            go[a] += __shfl_xor_sync(FULL, go[a], sft);
            gd[a] += __shfl_xor_sync(FULL, gd[a], sft);
        }
    if (lane == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (d_ro) d_ro[r * 3 + a] = go[a];
            if (d_rd) d_rd[r * 3 + a] = gd[a];
        }
    }
}

extern "C" int nsb_regulation_points(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t N,
                                     int N_samples, const float* t_uniform, const float* u, float* pts,
                                     float* z_out, void* stream) {
    if (N < 0 || N_samples < 1) return set_error(NSB_EINVAL, "regulation_points: N >= 0 and N_samples >= 1");
    if (N == 0) return 0;
    if (!rays_o || !rays_d || !gt_depth || !t_uniform || !u || !pts)
        return set_error(NSB_EINVAL, "regulation_points: null pointer");
    const long long total = (long long)N * N_samples;
    k_regulation_points<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        rays_o, rays_d, gt_depth, N, N_samples, t_uniform, u, pts, z_out);
    return check_launch("k_regulation_points");
}

extern "C" int nsb_regulation_points_backward(const float* d_pts, const float* z, int64_t N, int N_samples,
                                              float* d_rays_o, float* d_rays_d, void* stream) {
    if (N < 0 || N_samples < 1) return set_error(NSB_EINVAL, "regulation_points_backward: N >= 0 and N_samples >= 1");
    if (N == 0) return 0;
    if (!d_pts || !z) return set_error(NSB_EINVAL, "regulation_points_backward: null pointer");
    k_regulation_points_bwd<<<(unsigned)((N * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        d_pts, z, N, N_samples, d_rays_o, d_rays_d);
    return check_launch("k_regulation_points_bwd");
}

extern "C" int nsb_importance_z(const double* z, const float* weights, int64_t N, int S, int N_imp,
                                const float* u, int64_t u_row_stride, double* z_samples,
                                double* z_out, void* stream) {
    if (!z || !weights || !u || !z_out) return set_error(NSB_EINVAL, "null argument");
    if (S < 3 || S > MAX_LIST || N_imp < 1 || N_imp > MAX_LIST)
        return set_error(NSB_EINVAL, "S / N_importance out of range");
    if (N <= 0) return 0;
    k_importance_z<<<blocks_for(N, 64), 64, 0, (cudaStream_t)stream>>>(z, weights, N, S, N_imp, u,
                                                                      u_row_stride, z_samples, z_out);
    return check_launch("k_importance_z");
}

extern "C" int nsb_sample_pdf(const double* bins, const float* weights, int64_t N, int M, int N_samples,
                              const float* u, int64_t u_row_stride, double* samples, void* stream) {
    if (!bins || !weights || !u || !samples) return set_error(NSB_EINVAL, "null argument");
    if (M < 2 || M > MAX_LIST || N_samples < 1) return set_error(NSB_EINVAL, "bins / N_samples out of range");
    if (N <= 0) return 0;
    k_sample_pdf<<<blocks_for(N, 64), 64, 0, (cudaStream_t)stream>>>(bins, weights, N, M, N_samples, u,
                                                                    u_row_stride, samples);
    return check_launch("k_sample_pdf");
}

extern "C" int nsb_load_batch(const float* rays_o, const float* rays_d, const float* gt_depth, const float* gt_color,
                              const uint8_t* keep, int64_t n, int64_t N, float* dst_rays_o, float* dst_rays_d,
                              float* dst_gt_depth, float* dst_gt_color, uint8_t* dst_keep, void* stream) {
    if (!dst_rays_o || !dst_rays_d || !dst_gt_depth || !dst_gt_color || !dst_keep) return set_error(NSB_EINVAL, "null argument");
    if (n < 0 || n > N) return set_error(NSB_EINVAL, "batch of %lld rays exceeds the capacity %lld", (long long)n, (long long)N);
    if (rays_o && (!rays_d || !gt_depth || !gt_color)) return set_error(NSB_EINVAL, "null argument");
    if (N <= 0) return 0;
    launch_pdl(k_load_batch, dim3(blocks_for(N, 256)), dim3(256), 0, (cudaStream_t)stream, rays_o, rays_d, gt_depth, gt_color, keep, n, N,
                                                                      dst_rays_o, dst_rays_d, dst_gt_depth, dst_gt_color,
                                                                      dst_keep);
    return check_launch("k_load_batch");
}

extern "C" int nsb_load_batch_packed(const float* packed, const uint8_t* keep, int64_t n, int64_t N, float* dst_rays_o,
                                     float* dst_rays_d, float* dst_gt_depth, float* dst_gt_color, uint8_t* dst_keep,
                                     void* stream) {
    if (!dst_rays_o || !dst_rays_d || !dst_gt_depth || !dst_gt_color || !dst_keep) return set_error(NSB_EINVAL, "null argument");
    if (n < 0 || n > N) return set_error(NSB_EINVAL, "batch of %lld rays exceeds the capacity %lld", (long long)n, (long long)N);
    if (n > 0 && !packed) return set_error(NSB_EINVAL, "null argument");
    if (N <= 0) return 0;
    launch_pdl(k_load_batch_packed, dim3(blocks_for(N, 256)), dim3(256), 0, (cudaStream_t)stream, packed, keep, n, N, dst_rays_o, dst_rays_d,
                                                                             dst_gt_depth, dst_gt_color, dst_keep);
    return check_launch("k_load_batch_packed");
}
