// One-call forward / backward of Renderer.render_batch_ray (src/utils/Renderer.py:95-262) for
// the common case without importance resampling, and the two fused losses of the SLAM steps
// (src/Mapper.py:604-612, src/Tracker.py:117-130).  These only chain the kernels of the other
// files on one stream with a caller-provided workspace, so that the host issues two calls per
// iteration instead of a dozen.
#include "nsb_common.cuh"

namespace nsb {

struct Carve {
    char* p;
    explicit Carve(void* base) : p((char*)(((uintptr_t)base + 255) & ~(uintptr_t)255)) {}
    template <typename T>
    T* take(int64_t n) {
        T* r = (T*)p;
        p += ((n * (int64_t)sizeof(T) + 255) / 256) * 256;
        return r;
    }
};

struct Ws {
    float* depth_max;
    double* z;
    float *occ0, *occ1, *rgb, *d_raw, *d_point, *d_dnorm;
    uint32_t* relu;     // [3][P][8] ReLU sign words, forward -> backward
    float* pu;          // [P][8] point cache (fp32 position, normalised coordinate), forward -> backward
    uint32_t* ticket;   // fused mapping loss: arrival counter (zeroed by k_sample_z) ...
    double* partials;   // ... and per-block partial sums
};

int sample_z_impl(const NsbPoints* rays, const float* gt_depth, const double* bound6_host, int N_samples,
                  int N_surface, int lindisp, const float* t_uniform, const double* t_surface, const float* u,
                  const float* depth_max, double* z_out, unsigned* zero_word, const uint8_t* keep, void* stream);

static Ws carve_ws(void* base, int64_t N, int S) {
    Carve c(base);
    Ws w;
    const int64_t P = N * S;
    w.depth_max = c.take<float>(64);
    w.z = c.take<double>(P);
    w.occ0 = c.take<float>(P);
    w.occ1 = c.take<float>(P);
    w.rgb = c.take<float>(4 * P);
    w.d_raw = c.take<float>(4 * P);
    w.d_point = c.take<float>(9 * P);
    w.d_dnorm = c.take<float>(3 * N);
    w.relu = c.take<uint32_t>(24 * P);
    w.pu = c.take<float>(8 * P);
    w.ticket = c.take<uint32_t>(64);
    w.partials = c.take<double>(N / 4 + 1);
    return w;
}

static NsbPoints ray_pts(const NsbRenderCall* r, const double* z, int S) {
    NsbPoints p{};
    p.rays_o = r->rays_o; p.rays_d = r->rays_d;
    p.rays_o_stride = r->rays_o_stride; p.rays_d_stride = r->rays_d_stride;
    p.z = z; p.N = r->N; p.S = S; p.P = r->N * S;
    if (r->deposit_dweight_inputs && (r->flags & NSB_GRAD_WEIGHT) && r->dweight_mask && r->scratch) {
        p.x_scratch = r->scratch; p.x_scratch_bytes = r->scratch_bytes; p.x_mask = r->dweight_mask;
    }
    return p;
}

static void stage_planes(int stage, const Ws& w, float*& o0, float*& o1, float*& rgb) {
    o0 = w.occ0;
    o1 = (stage == NSB_STAGE_FINE || stage == NSB_STAGE_COLOR) ? w.occ1 : nullptr;
    rgb = stage == NSB_STAGE_COLOR ? w.rgb : nullptr;
}

// ---- losses ----------------------------------------------------------------------------------
// Mapping: sum_{gt>0} |gt - d| + w * sum |gt_c - c| (Mapper.py:604-612) and its gradient.
__global__ void k_mapping_loss(const double* __restrict__ depth, const float* __restrict__ color,
                               const float* __restrict__ gd, const float* __restrict__ gc,
                               const uint8_t* __restrict__ keep, long long N, float w_color, int use_color,
                               double* __restrict__ loss, double* __restrict__ g_depth,
                               float* __restrict__ g_color) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    __shared__ double part[32];
    double acc = 0.0;
    for (long long i = threadIdx.x; i < N; i += blockDim.x) {
        const bool live = !keep || keep[i];
        double gdv = 0.0;
        if (live && gd[i] > 0.f) {
            const double diff = (double)gd[i] - depth[i];
            acc += fabs(diff);
            gdv = diff > 0 ? -1.0 : (diff < 0 ? 1.0 : 0.0);
        }
        g_depth[i] = gdv;
        for (int c = 0; c < 3; ++c) {
            float g = 0.f;
            if (live && use_color) {
                const float diff = gc[3 * i + c] - color[3 * i + c];
                acc += (double)(w_color * fabsf(diff));
                g = diff > 0 ? -w_color : (diff < 0 ? w_color : 0.f);
            }
            g_color[3 * i + c] = g;
        }
    }
    for (int s = 16; s > 0; s >>= 1) {
        int lo = __shfl_xor_sync(FULL, __double2loint(acc), s), hi = __shfl_xor_sync(FULL, __double2hiint(acc), s);
        acc += __hiloint2double(hi, lo);
    }
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += part[k];
        loss[0] = t;
    }
}

// Median of the live depth residuals |gt - depth| / sqrt(var + 1e-10) (Tracker.py:118-121, torch.median = the
// lower middle of the sorted values, NaN if any value is NaN) by rank counting over many blocks: every block
// holds all N residuals in shared memory, every warp ranks one of them (lanes stride the others, ties by
// position), and the one whose rank is (n_live - 1) / 2 writes out[0].  1000 rays: 32 blocks, 32 compares per
// lane, where the single-block bitonic sort below needs 55 barrier-separated stages.
__global__ void k_residual_median(const double* __restrict__ depth, const double* __restrict__ var,
                                  const float* __restrict__ gd, const uint8_t* __restrict__ keep, int N,
                                  double* __restrict__ out) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    extern __shared__ double sh[];      // N doubles
    __shared__ int s_live;
    if (threadIdx.x == 0) s_live = 0;
    __syncthreads();
    int live_local = 0, nan_local = 0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        double v = INFINITY;
        if (!keep || keep[i]) {
            v = fabs((double)gd[i] - depth[i]) / sqrt(var[i] + 1e-10);
            ++live_local;
            nan_local |= v != v;
        }
        sh[i] = v;
    }
    if (live_local) atomicAdd(&s_live, live_local);
    const int any_nan = __syncthreads_or(nan_local);
    const int n_live = s_live;
    if (n_live == 0 || any_nan) {        // torch.median: NaN as soon as one value is NaN
        if (blockIdx.x == 0 && threadIdx.x == 0)
            out[0] = any_nan ? __longlong_as_double(0x7ff8000000000000ll) : INFINITY;
        return;
    }
    const int want = (n_live - 1) / 2;
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    const double a = sh[i];
    int rank = 0;
    for (int j = lane; j < N; j += 32) {
        const double b = sh[j];
        rank += (b < a || (b == a && j < i)) ? 1 : 0;
    }
    for (int s = 16; s > 0; s >>= 1) rank += __shfl_xor_sync(FULL, rank, s);
    if (lane == 0 && rank == want) out[0] = a;
}

// Tracking (Tracker.py:117-130): tmp = |gt - d| / sqrt(var + 1e-10); mask = tmp < 10 median(tmp)
// & gt > 0 (handle_dynamic) or gt > 0; loss = sum_mask tmp + w sum_mask |gt_c - c|.  The median
// is torch's lower median of the live rays, found by a bitonic sort in shared memory.
__global__ void k_tracking_loss(const double* __restrict__ depth, const double* __restrict__ var,
                                const float* __restrict__ color, const float* __restrict__ gd,
                                const float* __restrict__ gc, const uint8_t* __restrict__ keep, long long N,
                                int npow2, float w_color, int use_color, int handle_dynamic,
                                const double* median, double* __restrict__ loss,
                                double* __restrict__ g_depth, float* __restrict__ g_color) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    extern __shared__ double sh[];      // npow2 doubles
    __shared__ double part[32];
    __shared__ int s_live;
    if (threadIdx.x == 0) s_live = 0;
    __syncthreads();
    int live_local = 0;
    for (long long i = threadIdx.x; i < npow2; i += blockDim.x) {
        double v = INFINITY;
        if (i < N && (!keep || keep[i])) {
            v = fabs((double)gd[i] - depth[i]) / sqrt(var[i] + 1e-10);
            ++live_local;
        }
        sh[i] = v;
    }
    atomicAdd(&s_live, live_local);
    __syncthreads();
    double thr = INFINITY;
    if (handle_dynamic && median) {
        thr = 10.0 * median[0];          // k_residual_median ran before this kernel
    } else if (handle_dynamic) {
        for (int k = 2; k <= npow2; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const bool up = (i & k) == 0;
                        const double a = sh[i], b = sh[ixj];
                        // NaN sorts last like torch
                        const bool gt = (a > b) || (a != a && b == b);
                        if (gt == up) { sh[i] = b; sh[ixj] = a; }
                    }
                }
                __syncthreads();
            }
        const int n_live = s_live;
        // torch.median returns NaN as soon as one value is NaN (they sort last)
        thr = n_live > 0 ? 10.0 * (sh[npow2 - 1] != sh[npow2 - 1] ? sh[npow2 - 1] : sh[(n_live - 1) / 2]) : INFINITY;
    }
    __syncthreads();
    double acc = 0.0;
    for (long long i = threadIdx.x; i < N; i += blockDim.x) {
        const bool live = !keep || keep[i];
        double gdv = 0.0;
        float gcv[3] = {0.f, 0.f, 0.f};
        if (live) {
            const double sd = sqrt(var[i] + 1e-10);
            const double diff = (double)gd[i] - depth[i];
            const double tmp = fabs(diff) / sd;
            const bool m = (handle_dynamic ? tmp < thr : true) && gd[i] > 0.f;
            if (m) {
                acc += tmp;
                gdv = (diff > 0 ? -1.0 : (diff < 0 ? 1.0 : 0.0)) / sd;
                if (use_color)
                    for (int c = 0; c < 3; ++c) {
                        const float dc = gc[3 * i + c] - color[3 * i + c];
                        acc += (double)w_color * (double)fabsf(dc);
                        gcv[c] = dc > 0 ? -w_color : (dc < 0 ? w_color : 0.f);
                    }
            }
        }
        g_depth[i] = gdv;
        for (int c = 0; c < 3; ++c) g_color[3 * i + c] = gcv[c];
    }
    for (int s = 16; s > 0; s >>= 1) {
        int lo = __shfl_xor_sync(FULL, __double2loint(acc), s), hi = __shfl_xor_sync(FULL, __double2hiint(acc), s);
        acc += __hiloint2double(hi, lo);
    }
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += part[k];
        loss[0] = t;
    }
}

}  // namespace nsb

using namespace nsb;

extern "C" int64_t nsb_render_workspace_bytes(int64_t N, int S) {
    const int64_t P = N * S;
    // depth_max, z, occ0, occ1, rgb, d_raw, d_point (3 planes), d_dnorm, each rounded to 256 B
    // ... relu sign words, ticket, loss partials
    return 256 * 14 + 256 + P * 8 + P * 4 * 2 + P * 16 * 2 + P * 36 + N * 12 + P * 96 + P * 32 + 256 + (N / 4 + 1) * 8;
}

extern "C" int nsb_render_forward(const NsbRenderCall* r, void* stream) {
    if (!r || !r->rays_o || !r->rays_d || !r->workspace || !r->depth || !r->var || !r->color)
        return set_error(NSB_EINVAL, "null argument");
    const int Nf = r->gt_depth ? r->N_surface : 0;
    const int S = r->N_samples + Nf;
    if (r->workspace_bytes < nsb_render_workspace_bytes(r->N, S)) return set_error(NSB_EINVAL, "workspace too small");
    if (r->N <= 0) return 0;
    Ws w = carve_ws(r->workspace, r->N, S);
    const double* b6 = &r->scene.bound[0][0];
    const float* dmax = r->depth_max;
    if (r->gt_depth && !dmax && r->N > 4096) {      // small batches: k_sample_z reduces the max in every block
        if (int rc = nsb_depth_max(r->gt_depth, r->keep, r->N, w.depth_max, stream)) return rc;
        dmax = w.depth_max;
    }
    NsbPoints rays = ray_pts(r, nullptr, 0);
    rays.point_cache = r->stage == NSB_STAGE_COARSE ? nullptr : w.pu;      // the coarse decoder has its own (enlarged) bound
    const bool fused_loss = r->loss != nullptr;
    if (fused_loss && (!r->gt_depth || !r->gt_color)) return set_error(NSB_EINVAL, "fused mapping loss needs gt_depth and gt_color");
    if (int rc = sample_z_impl(&rays, r->gt_depth, b6, r->N_samples, Nf, r->lindisp, r->t_uniform, r->t_surface,
                               r->u, dmax, w.z, fused_loss ? w.ticket : nullptr, r->keep, stream)) return rc;
    NsbPoints pts = ray_pts(r, w.z, S);
    pts.relu_masks = w.relu;
    pts.point_cache = rays.point_cache;
    float *o0, *o1, *rgb;
    stage_planes(r->stage, w, o0, o1, rgb);
    pdl_allow_next(true);         // the decoder kernel's prologue may overlap k_sample_z
    if (int rc = nsb_decoders_forward(&pts, &r->scene, r->stage, o0, o1, rgb, stream)) return rc;
    if (fused_loss) return 0;     // compositing, loss and adjoint run as one kernel in nsb_render_backward
    return nsb_composite_forward(&pts, b6, r->stage, r->occupancy, o0, o1, rgb, r->depth, r->var, r->color,
                                 nullptr, stream);
}

extern "C" int nsb_render_backward(const NsbRenderCall* r, void* stream) {
    if (!r || !r->rays_o || !r->rays_d || !r->workspace || !r->depth) return set_error(NSB_EINVAL, "null argument");
    const int Nf = r->gt_depth ? r->N_surface : 0;
    const int S = r->N_samples + Nf;
    if (r->N <= 0) return 0;
    Ws w = carve_ws(r->workspace, r->N, S);
    const double* b6 = &r->scene.bound[0][0];
    NsbPoints pts = ray_pts(r, w.z, S);
    pts.relu_masks = w.relu;             // written by nsb_render_forward of this call
    pts.point_cache = r->stage == NSB_STAGE_COARSE ? nullptr : w.pu;
    pts.line_marks = r->line_marks; pts.line_marks_base = r->line_marks_base;
    float *o0, *o1, *rgb;
    stage_planes(r->stage, w, o0, o1, rgb);
    const bool want_rays = (r->flags & NSB_GRAD_POINT) != 0;
    float* dn = (!r->occupancy && want_rays) ? w.d_dnorm : nullptr;
    if (r->dweight_phase < 0 || r->dweight_phase > 2) return set_error(NSB_EINVAL, "dweight_phase %d", r->dweight_phase);
    if (r->dweight_phase == 2) {      // second call: only the weight-gradient kernels of the phase-1 call
        if (!(r->flags & NSB_GRAD_WEIGHT) || !r->dweight_mask) return 0;
        dweight_grid_cap(r->dweight_max_ctas);
        return nsb_decoders_backward(&pts, &r->scene, r->stage, w.d_raw, r->flags | NSB_FLAG_ONLY_DWEIGHT, r->dweight_mask,
                                     r->d_weight, w.d_point, r->scratch, r->scratch_bytes, stream);
    }
    const unsigned phase_bit = r->dweight_phase == 1 ? NSB_FLAG_DEFER_DWEIGHT : 0u;
    if (r->dweight_phase == 0) dweight_grid_cap(r->dweight_max_ctas);
    if (r->loss) {
        if (dn) return set_error(NSB_EINVAL, "fused mapping loss: density-mode ray gradients are not supported");
        if (!r->var || !r->color) return set_error(NSB_EINVAL, "null argument");
        if (int rc = nsb_composite_mapping(&pts, b6, r->stage, r->occupancy, o0, o1, rgb, r->gt_depth, r->gt_color,
                                           r->keep, r->w_color, r->use_color, r->depth, r->var, r->color, w.d_raw,
                                           w.partials, w.ticket, r->loss, stream)) return rc;
    } else if (int rc = nsb_composite_backward(&pts, b6, r->stage, r->occupancy, o0, o1, rgb, r->depth, r->g_depth,
                                               r->g_var, r->g_color, w.d_raw, dn, stream)) return rc;
    pdl_allow_next(true);         // ... and the backward kernel's weight staging the compositing kernel
    if (r->flags)
        if (int rc = nsb_decoders_backward(&pts, &r->scene, r->stage, w.d_raw, r->flags | phase_bit, r->dweight_mask,
                                           r->d_weight, w.d_point, r->scratch, r->scratch_bytes, stream)) return rc;
    pdl_allow_next(false);
    if (want_rays) {
        if (!r->g_rays_o || !r->g_rays_d) return set_error(NSB_EINVAL, "ray gradient outputs are NULL");
        const int ndec = r->stage == NSB_STAGE_COLOR ? 3 : (r->stage == NSB_STAGE_FINE ? 2 : 1);
        return nsb_ray_grad_reduce(&pts, w.d_point, ndec, dn, r->g_rays_o, r->g_rays_d, stream);
    }
    return 0;
}

__global__ void k_live_tiles(const float* __restrict__ d_raw, long long P, int tile, long long* counts) {
    const long long nt = (P + tile - 1) / tile;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < nt; t += (long long)gridDim.x * blockDim.x) {
        bool occ = false, rgb = false;
        const long long hi = (t + 1) * tile < P ? (t + 1) * tile : P;
        for (long long i = t * tile; i < hi; ++i) {
            const float4 v = reinterpret_cast<const float4*>(d_raw)[i];
            occ |= v.w != 0.f;
            rgb |= v.x != 0.f || v.y != 0.f || v.z != 0.f;
        }
        if (occ) atomicAdd((unsigned long long*)(counts + 1), 1ull);
        if (rgb) atomicAdd((unsigned long long*)(counts + 2), 1ull);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[0] = nt;
}

extern "C" int nsb_render_live_tiles(const NsbRenderCall* r, int tile_points, int64_t* counts, void* stream) {
    if (!r || !r->workspace || !counts || tile_points <= 0) return set_error(NSB_EINVAL, "bad argument");
    const int S = r->N_samples + (r->gt_depth ? r->N_surface : 0);
    if (r->N <= 0) return 0;
    Ws w = carve_ws(r->workspace, r->N, S);
    cudaMemsetAsync(counts, 0, 3 * sizeof(int64_t), (cudaStream_t)stream);
    const long long P = (long long)r->N * S, nt = (P + tile_points - 1) / tile_points;
    k_live_tiles<<<(unsigned)((nt + 255) / 256 > 1184 ? 1184 : (nt + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        w.d_raw, P, tile_points, (long long*)counts);
    return check_launch("k_live_tiles");
}

extern "C" int nsb_mapping_loss(const double* depth, const float* color, const float* gt_depth,
                                const float* gt_color, const uint8_t* keep, int64_t N, float w_color,
                                int use_color, double* loss, double* g_depth, float* g_color, void* stream) {
    if (!depth || !color || !gt_depth || !gt_color || !loss || !g_depth || !g_color)
        return set_error(NSB_EINVAL, "null argument");
    launch_pdl(k_mapping_loss, dim3(1), dim3(1024), 0, (cudaStream_t)stream, depth, color, gt_depth, gt_color, keep, N, w_color,
                                                       use_color, loss, g_depth, g_color);
    return check_launch("k_mapping_loss");
}

// Adam on the 7 camera parameters + the tracker's best-candidate rule (Tracker.py:232-242,262-275).
// One warp; bias corrections in fp64 like torch's Python scalars.
__global__ void k_cam_adam(float* cam, const float* d_cam, float* m, float* v, int* step, float lr_q, float lr_t,
                           float beta1, float beta2, float eps, const double* loss, double* best_loss,
                           float* best_cam) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const int i = threadIdx.x;
    cam_adam_warp(i, i < 7 ? d_cam[i] : 0.f, cam, m, v, step, lr_q, lr_t, beta1, beta2, eps, loss, best_loss, best_cam);
}

extern "C" int nsb_cam_adam(float* cam, const float* d_cam, float* m, float* v, int32_t* step, float lr_q, float lr_t,
                            float beta1, float beta2, float eps, const double* loss, double* best_loss,
                            float* best_cam, void* stream) {
    if (!cam || !d_cam || !m || !v || !step) return set_error(NSB_EINVAL, "null argument");
    if (best_loss && (!loss || !best_cam)) return set_error(NSB_EINVAL, "best_loss needs loss and best_cam");
    launch_pdl(k_cam_adam, dim3(1), dim3(32), 0, (cudaStream_t)stream, cam, d_cam, m, v, step, lr_q, lr_t, beta1, beta2, eps, loss,
                                                   best_loss, best_cam);
    return check_launch("k_cam_adam");
}

// Adam over a flat tensor, optionally restricted to masked voxel lines (Mapper.py:462-482 on the
// compact val[mask] copies + the write-back of :633-642, as one in-place pass).
__global__ void __launch_bounds__(256) k_adam(float* __restrict__ p, const float* __restrict__ g,
                                               float* __restrict__ m, float* __restrict__ v, long long n,
                                               const unsigned char* __restrict__ mask, int* state2,
                                               const float* __restrict__ lr_dev, float b1, float b2, float eps,
                                               int vec) {
    const int t = state2[0] + 1;
    const double bc1 = 1.0 - pow((double)b1, (double)t), bc2 = 1.0 - pow((double)b2, (double)t);
    const float step_size = (float)((double)lr_dev[0] / bc1);
    const float bc2_sqrt = (float)sqrt(bc2);
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {
        const long long n4 = n >> 2;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
            if (mask && !mask[i >> 3]) continue;                 // 8 float4 per 128-byte voxel line
            const float4 gg = reinterpret_cast<const float4*>(g)[i];
            float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i],
                   vv = reinterpret_cast<float4*>(v)[i];
            mm.x += (gg.x - mm.x) * (1.f - b1); mm.y += (gg.y - mm.y) * (1.f - b1);
            mm.z += (gg.z - mm.z) * (1.f - b1); mm.w += (gg.w - mm.w) * (1.f - b1);
            vv.x = vv.x * b2 + (1.f - b2) * gg.x * gg.x; vv.y = vv.y * b2 + (1.f - b2) * gg.y * gg.y;
            vv.z = vv.z * b2 + (1.f - b2) * gg.z * gg.z; vv.w = vv.w * b2 + (1.f - b2) * gg.w * gg.w;
            pp.x -= step_size * (mm.x / (sqrtf(vv.x) / bc2_sqrt + eps));
            pp.y -= step_size * (mm.y / (sqrtf(vv.y) / bc2_sqrt + eps));
            pp.z -= step_size * (mm.z / (sqrtf(vv.z) / bc2_sqrt + eps));
            pp.w -= step_size * (mm.w / (sqrtf(vv.w) / bc2_sqrt + eps));
            reinterpret_cast<float4*>(p)[i] = pp;
            reinterpret_cast<float4*>(m)[i] = mm;
            reinterpret_cast<float4*>(v)[i] = vv;
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            if (mask && !mask[i >> 5]) continue;
            const float gi = g[i];
            const float mi = m[i] + (gi - m[i]) * (1.f - b1);
            const float vi = v[i] * b2 + (1.f - b2) * gi * gi;
            m[i] = mi;
            v[i] = vi;
            p[i] -= step_size * (mi / (sqrtf(vi) / bc2_sqrt + eps));
        }
    }
    // the last block to finish publishes the new step count (every block has read the old one by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&state2[1], 1) == (int)gridDim.x - 1) {
            state2[1] = 0;
            state2[0] = t;
        }
    }
}

extern "C" int nsb_adam(float* p, const float* g, float* m, float* v, int64_t n, const uint8_t* voxel_mask,
                        int32_t* state2, const float* lr, float beta1, float beta2, float eps, void* stream) {
    if (!p || !g || !m || !v || !state2 || !lr) return set_error(NSB_EINVAL, "null argument");
    if (n <= 0) return 0;
    if (voxel_mask && (n % 32)) return set_error(NSB_EINVAL, "masked Adam needs n to be a multiple of 32 (voxel lines)");
    const bool vec = (n % 4 == 0) && !(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15);
    long long work = vec ? n / 4 : n;
    long long blocks = (work + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_adam<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (long long)n, voxel_mask, state2, lr, beta1, beta2,
                                                          eps, vec ? 1 : 0);
    return check_launch("k_adam");
}

extern "C" int nsb_tracking_loss(const double* depth, const double* var, const float* color,
                                 const float* gt_depth, const float* gt_color, const uint8_t* keep, int64_t N,
                                 float w_color, int use_color, int handle_dynamic, double* loss,
                                 double* g_depth, float* g_color, void* stream) {
    if (!depth || !var || !color || !gt_depth || !gt_color || !loss || !g_depth || !g_color)
        return set_error(NSB_EINVAL, "null argument");
    int npow2 = 1;
    while (npow2 < N) npow2 <<= 1;
    if (npow2 > 16384) return set_error(NSB_EINVAL, "tracking loss supports up to 16384 rays per batch");
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(k_tracking_loss, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8);
        if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        attr = true;
    }
    const double* median = nullptr;
    if (handle_dynamic && N > 0 && N <= 4096) {
        // loss[0] carries the median between the two kernels; the second overwrites it with the loss
        const int warps = 32;
        launch_pdl(k_residual_median, dim3((unsigned)((N + warps - 1) / warps)), dim3(warps * 32), (size_t)N * 8, (cudaStream_t)stream,
            depth, var, gt_depth, keep, (int)N, loss);
        if (int rc = check_launch("k_residual_median")) return rc;
        median = loss;
    }
    launch_pdl(k_tracking_loss, dim3(1), dim3(1024), (size_t)(median ? 8 : npow2 * 8), (cudaStream_t)stream,
        depth, var, color, gt_depth, gt_color, keep, N, median ? 1 : npow2, w_color, use_color, handle_dynamic,
        median, loss, g_depth, g_color);
    return check_launch("k_tracking_loss");
}
