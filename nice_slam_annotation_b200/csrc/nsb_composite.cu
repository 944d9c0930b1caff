// K4 (+ its part of K5): alpha compositing along the ray as a warp-level scan, its adjoint,
// the per-ray reduction of point gradients and the camera (quaternion + translation)
// gradient.  Replaces raw2outputs_nerf_color (src/common.py:215-256), the out-of-bound rule of
// Renderer.eval_points (src/utils/Renderer.py:63-66,87), the stage mixing tail of NICE.forward
// (decoder.py:344-369) and the autograd of get_rays_from_uv / get_camera_from_tensor
// (src/common.py:74-89,137-176).
//
// One warp per ray.  Lane l owns the CH = ceil(S/32) consecutive samples [l*CH, (l+1)*CH):
// transmittance is a lane-local running product plus an exclusive warp scan of the lane
// products; the adjoint needs the mirrored suffix scan of G_k w_k.
#include "nsb_common.cuh"

namespace nsb {

constexpr int MAX_CH = 8;   // S <= 256

struct CompParams {
    PointsDev q;
    BoundDev bound;
    int stage, occupancy;
    const float* occ0;
    const float* occ1;
    const float* rgb;     // [P,4]
};

__device__ __forceinline__ double shfl_xor_d(double v, int m) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_xor_sync(FULL, lo, m);
    hi = __shfl_xor_sync(FULL, hi, m);
    return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double warp_sum_d(double v) {
    for (int m = 16; m > 0; m >>= 1) v += shfl_xor_d(v, m);
    return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(FULL, v, m);
    return v;
}

// raw occupancy / density of sample i with the stage mixing and the out-of-bound rule
__device__ __forceinline__ float raw_sigma(const CompParams& c, long long i, bool& inb) {
    inb = point_inbound(c.q, i, c.bound);
    float v = c.occ0 ? c.occ0[i] : 0.f;
    if (c.occ1) v = c.occ1[i] + v;                     // decoder.py:369  fine_occ + middle_occ
    if (c.stage == NSB_STAGE_COLOR && !c.occ1 && !c.occ0) v = c.rgb[4 * i + 3];   // iMAP single MLP
    return inb ? v : 100.f;                            // Renderer.py:87
}

struct RaySamples {
    float alpha[MAX_CH], trans[MAX_CH], sig[MAX_CH], dist[MAX_CH];
    bool inb[MAX_CH];
    int n;        // samples owned by this lane
    int s0;       // first sample
};

// alpha and transmittance of the lane's samples (common.py:233-250)
__device__ __forceinline__ void ray_alpha(const CompParams& c, long long ray, int lane, int CH,
                                          float dnorm, RaySamples& R) {
    const int S = c.q.S;
    R.s0 = lane * CH;
    R.n = max(0, min(CH, S - R.s0));
    float prod = 1.f;
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k) {
        if (k < R.n) {
            const long long i = ray * S + R.s0 + k;
            const float s = raw_sigma(c, i, R.inb[k]);
            R.sig[k] = s;
            float a;
            if (c.occupancy) {
                a = 1.f / (1.f + expf(-10.f * s));
                R.dist[k] = 0.f;
            } else {
                float dz = (R.s0 + k + 1 < S) ? (float)__dsub_rn(c.q.z[i + 1], c.q.z[i]) : 1e10f;
                dz = __fmul_rn(dz, dnorm);
                R.dist[k] = dz;
                a = 1.f - expf(-fmaxf(s, 0.f) * dz);
            }
            R.alpha[k] = a;
            R.trans[k] = prod;                                     // lane-local exclusive product
            prod = __fmul_rn(prod, __fadd_rn(__fsub_rn(1.f, a), 1e-10f));
        }
    }
    // exclusive scan of the lane products
    float incl = prod;
    for (int d = 1; d < 32; d <<= 1) {
        float o = __shfl_up_sync(FULL, incl, d);
        if (lane >= d) incl *= o;
    }
    float excl = __shfl_up_sync(FULL, incl, 1);
    if (lane == 0) excl = 1.f;
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k)
        if (k < R.n) R.trans[k] *= excl;
}

__device__ __forceinline__ float ray_dnorm(const PointsDev& q, long long ray) {
    const float x = q.rd[ray * q.sd], y = q.rd[ray * q.sd + 1], z = q.rd[ray * q.sd + 2];
    return sqrtf(x * x + y * y + z * z);
}

__global__ void k_composite_fwd(CompParams c, double* __restrict__ depth, double* __restrict__ var,
                                float* __restrict__ rgb_out, float* __restrict__ weights) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const int lane = threadIdx.x & 31;
    const long long ray = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    if (ray >= c.q.N) return;
    const int S = c.q.S, CH = (S + 31) / 32;
    const float dnorm = c.occupancy ? 0.f : ray_dnorm(c.q, ray);
    RaySamples R;
    ray_alpha(c, ray, lane, CH, dnorm, R);
    double dsum = 0.0;
    float col[3] = {0.f, 0.f, 0.f};
    float w[MAX_CH];
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k) {
        if (k < R.n) {
            const long long i = ray * S + R.s0 + k;
            w[k] = R.alpha[k] * R.trans[k];
            if (weights) weights[i] = w[k];
            dsum += (double)w[k] * c.q.z[i];
            if (c.rgb && c.stage == NSB_STAGE_COLOR) {
                const float4 v = *reinterpret_cast<const float4*>(c.rgb + 4 * i);
                col[0] = fmaf(w[k], v.x, col[0]);
                col[1] = fmaf(w[k], v.y, col[1]);
                col[2] = fmaf(w[k], v.z, col[2]);
            }
        }
    }
    const double D = warp_sum_d(dsum);
    double vsum = 0.0;
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k) {
        if (k < R.n) {
            const double t = c.q.z[ray * S + R.s0 + k] - D;
            vsum += (double)w[k] * t * t;
        }
    }
    const double V = warp_sum_d(vsum);
#pragma unroll
    for (int a = 0; a < 3; ++a) col[a] = warp_sum_f(col[a]);
    if (lane == 0) {
        depth[ray] = D;
        var[ray] = V;
        rgb_out[3 * ray] = col[0];
        rgb_out[3 * ray + 1] = col[1];
        rgb_out[3 * ray + 2] = col[2];
    }
}

__global__ void k_composite_bwd(CompParams c, const double* __restrict__ depth,
                                const double* __restrict__ g_depth, const double* __restrict__ g_var,
                                const float* __restrict__ g_rgb, float* __restrict__ d_raw,
                                float* __restrict__ d_dnorm) {
    pdl_launch_dependents();     // the backward decoder kernel may stage its weights meanwhile
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const int lane = threadIdx.x & 31;
    const long long ray = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    if (ray >= c.q.N) return;
    const int S = c.q.S, CH = (S + 31) / 32;
    const float dnorm = c.occupancy ? 0.f : ray_dnorm(c.q, ray);
    RaySamples R;
    ray_alpha(c, ray, lane, CH, dnorm, R);
    const bool color = c.rgb && c.stage == NSB_STAGE_COLOR;
    const double D = depth[ray];
    const double gd = g_depth ? g_depth[ray] : 0.0;
    const double gv = g_var ? g_var[ray] : 0.0;
    float gc[3] = {0.f, 0.f, 0.f};
    if (g_rgb && color) { gc[0] = g_rgb[3 * ray]; gc[1] = g_rgb[3 * ray + 1]; gc[2] = g_rgb[3 * ray + 2]; }
    // W = sum of weights (needed by the variance adjoint through depth)
    float wl = 0.f;
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k)
        if (k < R.n) wl += R.alpha[k] * R.trans[k];
    const double W = (double)warp_sum_f(wl);
    const double gd_tot = gd - 2.0 * gv * D * (1.0 - W);
    // G_k = dL/dw_k ; lane-local suffix sums of G_k w_k
    float G[MAX_CH], suf[MAX_CH];
    float lane_total = 0.f;
#pragma unroll
    for (int k = MAX_CH - 1; k >= 0; --k) {
        if (k < R.n) {
            const long long i = ray * S + R.s0 + k;
            const double z = c.q.z[i];
            double g = gd_tot * z + gv * (z - D) * (z - D);
            float w = R.alpha[k] * R.trans[k];
            float rgbdot = 0.f;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (color) {
                v = *reinterpret_cast<const float4*>(c.rgb + 4 * i);
                rgbdot = gc[0] * v.x + gc[1] * v.y + gc[2] * v.z;
            }
            G[k] = (float)g + rgbdot;
            suf[k] = lane_total;                 // sum over later samples of this lane
            lane_total += G[k] * w;
            // d(rgb raw) = w g_rgb
            d_raw[4 * i] = w * gc[0];
            d_raw[4 * i + 1] = w * gc[1];
            d_raw[4 * i + 2] = w * gc[2];
        }
    }
    // exclusive suffix scan over lanes
    float incl = lane_total;
    for (int d = 1; d < 32; d <<= 1) {
        float o = __shfl_down_sync(FULL, incl, d);
        if (lane + d < 32) incl += o;
    }
    float excl = __shfl_down_sync(FULL, incl, 1);
    if (lane == 31) excl = 0.f;
    float ddn = 0.f;
#pragma unroll
    for (int k = 0; k < MAX_CH; ++k) {
        if (k < R.n) {
            const long long i = ray * S + R.s0 + k;
            const float a = R.alpha[k];
            const float later = suf[k] + excl;
            const float dalpha = G[k] * R.trans[k] - later / ((1.f - a) + 1e-10f);
            float ds;
            if (c.occupancy) {
                ds = dalpha * (1.f - a) * a * 10.f;
            } else {
                const float ex = 1.f - a;                      // exp(-relu(s) dist)
                ds = R.sig[k] > 0.f ? dalpha * ex * R.dist[k] : 0.f;
                // d(dist) = dalpha * ex * relu(s); dist = dz * |d|
                if (R.s0 + k + 1 < S) {
                    const float dz = (float)__dsub_rn(c.q.z[i + 1], c.q.z[i]);
                    ddn += dalpha * ex * fmaxf(R.sig[k], 0.f) * dz;
                } else {
                    ddn += dalpha * ex * fmaxf(R.sig[k], 0.f) * 1e10f;
                }
            }
            d_raw[4 * i + 3] = R.inb[k] ? ds : 0.f;            // Renderer.py:87 cuts the gradient
        }
    }
    if (d_dnorm) {
        ddn = warp_sum_f(ddn);
        if (lane == 0) {
            const float inv = dnorm > 0.f ? 1.f / dnorm : 0.f;
#pragma unroll
            for (int a = 0; a < 3; ++a)
                d_dnorm[3 * ray + a] = c.occupancy ? 0.f : ddn * c.q.rd[ray * c.q.sd + a] * inv;
        }
    }
}

// Mapping iteration middle section as ONE kernel: compositing (k_composite_fwd), the mapping loss
// sum_{gt>0} |gt - depth| + w sum |gt_c - color| with its gradient (Mapper.py:604-612; no cross-ray
// term, unlike the tracker's median mask) and the compositing adjoint (k_composite_bwd with
// g_var = 0), without re-reading the decoder planes or round-tripping depth / g_depth through HBM.
// Loss: per-block partial sums, the last block to arrive adds them in block order (deterministic)
// and re-arms the ticket.  ticket[0] must be 0 on entry (k_sample_z of the same call zeroes it).
__global__ void __launch_bounds__(128) k_composite_map(CompParams c, const float* __restrict__ gt_depth,
                                                        const float* __restrict__ gt_color,
                                                        const uint8_t* __restrict__ keep, float w_color, int use_color,
                                                        double* __restrict__ depth, double* __restrict__ var,
                                                        float* __restrict__ rgb_out, float* __restrict__ d_raw,
                                                        double* __restrict__ partials, unsigned* __restrict__ ticket,
                                                        double* __restrict__ loss) {
    pdl_launch_dependents();     // the backward decoder kernel may stage its weights meanwhile
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    __shared__ double s_part[4];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long ray = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int S = c.q.S, CH = (S + 31) / 32;
    const bool color = c.rgb && c.stage == NSB_STAGE_COLOR;
    double ray_loss = 0.0;
    if (ray < c.q.N) {
        const float dnorm = c.occupancy ? 0.f : ray_dnorm(c.q, ray);
        RaySamples R;
        ray_alpha(c, ray, lane, CH, dnorm, R);
        double dsum = 0.0;
        float col[3] = {0.f, 0.f, 0.f};
        float w[MAX_CH];
        double zz[MAX_CH];
        float4 cv[MAX_CH];
#pragma unroll
        for (int k = 0; k < MAX_CH; ++k) {
            if (k < R.n) {
                const long long i = ray * S + R.s0 + k;
                w[k] = R.alpha[k] * R.trans[k];
                zz[k] = c.q.z[i];
                dsum += (double)w[k] * zz[k];
                cv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (color) {
                    cv[k] = *reinterpret_cast<const float4*>(c.rgb + 4 * i);
                    col[0] = fmaf(w[k], cv[k].x, col[0]);
                    col[1] = fmaf(w[k], cv[k].y, col[1]);
                    col[2] = fmaf(w[k], cv[k].z, col[2]);
                }
            }
        }
        const double D = warp_sum_d(dsum);
        double vsum = 0.0;
#pragma unroll
        for (int k = 0; k < MAX_CH; ++k)
            if (k < R.n) { const double t = zz[k] - D; vsum += (double)w[k] * t * t; }
        const double V = warp_sum_d(vsum);
#pragma unroll
        for (int a = 0; a < 3; ++a) col[a] = warp_sum_f(col[a]);
        if (lane == 0) {
            depth[ray] = D;
            var[ray] = V;
            rgb_out[3 * ray] = col[0];
            rgb_out[3 * ray + 1] = col[1];
            rgb_out[3 * ray + 2] = col[2];
        }
        // loss of this ray and its gradient (every lane holds D and col)
        const bool live = !keep || keep[ray];
        double gd = 0.0;
        float gc[3] = {0.f, 0.f, 0.f};
        const float gdv = gt_depth[ray];
        if (live && gdv > 0.f) {
            const double diff = (double)gdv - D;
            ray_loss += fabs(diff);
            gd = diff > 0 ? -1.0 : (diff < 0 ? 1.0 : 0.0);
        }
        if (live && use_color)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const float diff = gt_color[3 * ray + a] - col[a];
                ray_loss += (double)(w_color * fabsf(diff));
                gc[a] = diff > 0 ? -w_color : (diff < 0 ? w_color : 0.f);
            }
        if (!color) gc[0] = gc[1] = gc[2] = 0.f;
        // adjoint (k_composite_bwd with g_var = 0)
        float G[MAX_CH], suf[MAX_CH];
        float lane_total = 0.f;
#pragma unroll
        for (int k = MAX_CH - 1; k >= 0; --k) {
            if (k < R.n) {
                const long long i = ray * S + R.s0 + k;
                const double g = gd * zz[k];
                const float rgbdot = color ? gc[0] * cv[k].x + gc[1] * cv[k].y + gc[2] * cv[k].z : 0.f;
                G[k] = (float)g + rgbdot;
                suf[k] = lane_total;
                lane_total += G[k] * w[k];
                d_raw[4 * i] = w[k] * gc[0];
                d_raw[4 * i + 1] = w[k] * gc[1];
                d_raw[4 * i + 2] = w[k] * gc[2];
            }
        }
        float incl = lane_total;
        for (int d = 1; d < 32; d <<= 1) {
            float o = __shfl_down_sync(FULL, incl, d);
            if (lane + d < 32) incl += o;
        }
        float excl = __shfl_down_sync(FULL, incl, 1);
        if (lane == 31) excl = 0.f;
#pragma unroll
        for (int k = 0; k < MAX_CH; ++k) {
            if (k < R.n) {
                const long long i = ray * S + R.s0 + k;
                const float a = R.alpha[k];
                const float later = suf[k] + excl;
                const float dalpha = G[k] * R.trans[k] - later / ((1.f - a) + 1e-10f);
                float ds;
                if (c.occupancy) ds = dalpha * (1.f - a) * a * 10.f;
                else ds = R.sig[k] > 0.f ? dalpha * (1.f - a) * R.dist[k] : 0.f;
                d_raw[4 * i + 3] = R.inb[k] ? ds : 0.f;
            }
        }
    }
    // ---- loss: warp 0 lane 0 of each warp holds ray_loss (identical on all lanes) ----
    if (lane == 0) s_part[warp] = ray_loss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += s_part[k];
        partials[blockIdx.x] = t;
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && warp == 0) {
        __threadfence();
        double t = 0.0;
        for (unsigned b = lane; b < gridDim.x; b += 32) t += __ldcg(partials + b);
        t = warp_sum_d(t);
        if (lane == 0) { loss[0] = t; *ticket = 0u; }
    }
}

__global__ void k_assemble_raw(CompParams c, float* __restrict__ raw, uint8_t* __restrict__ inbound) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= c.q.P) return;
    bool inb;
    const float s = raw_sigma(c, i, inb);
    float4 o = make_float4(0.f, 0.f, 0.f, s);
    if (c.rgb && c.stage == NSB_STAGE_COLOR) {
        const float4 v = *reinterpret_cast<const float4*>(c.rgb + 4 * i);
        o.x = v.x; o.y = v.y; o.z = v.z;
    }
    *reinterpret_cast<float4*>(raw + 4 * i) = o;
    if (inbound) inbound[i] = inb ? 1 : 0;
}

// g_o[n] = sum_s dp ; g_d[n] = sum_s z dp (+ density-mode |d| term)
__global__ void k_ray_grad_reduce(PointsDev q, const float* __restrict__ d_point, int ndec,
                                  const float* __restrict__ d_dnorm, float* __restrict__ g_o,
                                  float* __restrict__ g_d) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    const int lane = threadIdx.x & 31;
    const long long ray = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    if (ray >= q.N) return;
    const int S = q.S;
    double so[3] = {0, 0, 0}, sd[3] = {0, 0, 0};
    for (int s = lane; s < S; s += 32) {
        const long long i = ray * S + s;
        const double z = q.z[i];
        for (int k = 0; k < ndec; ++k) {
            const float* dp = d_point + ((size_t)k * q.P + i) * 3;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                so[a] += (double)dp[a];
                sd[a] += z * (double)dp[a];
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { so[a] = warp_sum_d(so[a]); sd[a] = warp_sum_d(sd[a]); }
    if (lane == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            g_o[3 * ray + a] = (float)so[a];
            g_d[3 * ray + a] = (float)sd[a] + (d_dnorm ? d_dnorm[3 * ray + a] : 0.f);
        }
    }
}

struct CamAdam {        // optional Adam step on the camera tensor at the end of k_pose_grad (cam == NULL: none)
    float* cam; float* m; float* v; int* step;
    float lr_q, lr_t, beta1, beta2, eps;
    const double* loss; double* best_loss; float* best_cam;
};

// d(camera_tensor): dt = sum g_o ; dR = sum g_d (x) dirs ; dq = J(quad2rotation)^T dR
__global__ void k_pose_grad(const float* __restrict__ g_o, const float* __restrict__ g_d,
                            const int64_t* __restrict__ indices, const uint8_t* __restrict__ keep,
                            long long n, int H0, int W0, int wc, float fx, float fy, float cx,
                            float cy, const float* __restrict__ cam7, float* __restrict__ d_cam7, CamAdam A) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    __shared__ double red[12][32];
    __shared__ float s_dc[8];
    double acc[12];
    for (int k = 0; k < 12; ++k) acc[k] = 0.0;
    for (long long r = threadIdx.x; r < n; r += blockDim.x) {
        if (keep && !keep[r]) continue;
        const long long idx = indices[r];
        const float pi = (float)(W0 + (int)(idx % wc)), pj = (float)(H0 + (int)(idx / wc));
        const double dir[3] = {(double)__fdiv_rn(__fsub_rn(pi, cx), fx), (double)(-__fdiv_rn(__fsub_rn(pj, cy), fy)), -1.0};
        for (int a = 0; a < 3; ++a) {
            acc[9 + a] += (double)g_o[3 * r + a];
            for (int b = 0; b < 3; ++b) acc[3 * a + b] += (double)g_d[3 * r + a] * dir[b];
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = 0; k < 12; ++k) {
        double v = warp_sum_d(acc[k]);
        if (lane == 0) red[k][warp] = v;
    }
    __syncthreads();
    if (warp != 0) return;
    const int nw = blockDim.x >> 5;
    for (int k = 0; k < 12; ++k) acc[k] = warp_sum_d(lane < nw ? red[k][lane] : 0.0);
    if (threadIdx.x == 0) {
        double dR[3][3], dt[3];
        for (int k = 0; k < 12; ++k) {
            if (k < 9) dR[k / 3][k % 3] = acc[k]; else dt[k - 9] = acc[k];
        }
        const double qr = cam7[0], qi = cam7[1], qj = cam7[2], qk = cam7[3];
        const double nrm = qr * qr + qi * qi + qj * qj + qk * qk;
        const double s = 2.0 / nrm;
        // R = I + s P(q)
        const double P[3][3] = {{-(qj * qj + qk * qk), qi * qj - qk * qr, qi * qk + qj * qr},
                                {qi * qj + qk * qr, -(qi * qi + qk * qk), qj * qk - qi * qr},
                                {qi * qk - qj * qr, qj * qk + qi * qr, -(qi * qi + qj * qj)}};
        double dP = 0.0;
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) dP += dR[a][b] * P[a][b];
        // dP_ab/dq_m
        const double gqr = dR[0][1] * (-qk) + dR[0][2] * qj + dR[1][0] * qk + dR[1][2] * (-qi) +
                           dR[2][0] * (-qj) + dR[2][1] * qi;
        const double gqi = dR[0][1] * qj + dR[0][2] * qk + dR[1][0] * qj + dR[1][1] * (-2 * qi) +
                           dR[1][2] * (-qr) + dR[2][0] * qk + dR[2][1] * qr + dR[2][2] * (-2 * qi);
        const double gqj = dR[0][0] * (-2 * qj) + dR[0][1] * qi + dR[0][2] * qr + dR[1][0] * qi +
                           dR[1][2] * qk + dR[2][0] * (-qr) + dR[2][1] * qk + dR[2][2] * (-2 * qj);
        const double gqk = dR[0][0] * (-2 * qk) + dR[0][1] * (-qr) + dR[0][2] * qi + dR[1][0] * qr +
                           dR[1][1] * (-2 * qk) + dR[1][2] * qj + dR[2][0] * qi + dR[2][1] * qj;
        // s = 2/n, n = |q|^2: d(s)/d(q_m) = -2/n^2 * 2 q_m = -(s*s) q_m
        s_dc[0] = (float)(s * gqr - s * s * qr * dP);
        s_dc[1] = (float)(s * gqi - s * s * qi * dP);
        s_dc[2] = (float)(s * gqj - s * s * qj * dP);
        s_dc[3] = (float)(s * gqk - s * s * qk * dP);
        s_dc[4] = (float)dt[0];
        s_dc[5] = (float)dt[1];
        s_dc[6] = (float)dt[2];
    }
    __syncwarp();
    const float g = lane < 7 ? s_dc[lane] : 0.f;
    if (lane < 7) d_cam7[lane] = g;
    // the camera optimiser's step in the same launch (nsb_pose_grad_adam): cam7 == A.cam is read above, written here
    if (A.cam)
        cam_adam_warp(lane, g, A.cam, A.m, A.v, A.step, A.lr_q, A.lr_t, A.beta1, A.beta2, A.eps, A.loss, A.best_loss, A.best_cam);
}

static int comp_params(const NsbPoints* rays, const double* b6, int stage, int occupancy,
                       const float* occ0, const float* occ1, const float* rgb, bool need_rays,
                       CompParams& c) {
    if (!rays || !b6) return set_error(NSB_EINVAL, "null argument");
    c.q = to_dev(*rays);
    if (need_rays && (!c.q.ro || !c.q.rd || !c.q.z)) return set_error(NSB_EINVAL, "ray mode required");
    if (need_rays && c.q.S > 32 * MAX_CH) return set_error(NSB_EINVAL, "too many samples per ray (max %d)", 32 * MAX_CH);
    if (stage < 0 || stage > 3) return set_error(NSB_EINVAL, "bad stage %d", stage);
    c.bound = to_bound(b6);
    c.stage = stage;
    c.occupancy = occupancy;
    c.occ0 = occ0; c.occ1 = occ1; c.rgb = rgb;
    if (!occ0 && !occ1 && !rgb) return set_error(NSB_EINVAL, "no decoder output planes");
    return 0;
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_composite_forward(const NsbPoints* rays, const double* bound6_host, int stage,
                                     int occupancy, const float* out_occ0, const float* out_occ1,
                                     const float* out_rgb, double* depth, double* var, float* rgb,
                                     float* weights, void* stream) {
    CompParams c;
    if (int rc = comp_params(rays, bound6_host, stage, occupancy, out_occ0, out_occ1, out_rgb, true, c)) return rc;
    if (!depth || !var || !rgb) return set_error(NSB_EINVAL, "null output");
    if (c.q.N <= 0) return 0;
    const int bs = 128;
    launch_pdl(k_composite_fwd, dim3((unsigned)((c.q.N * 32 + bs - 1) / bs)), dim3(bs), 0, (cudaStream_t)stream, c, depth, var, rgb, weights);
    return check_launch("k_composite_fwd");
}

extern "C" int nsb_composite_backward(const NsbPoints* rays, const double* bound6_host, int stage,
                                      int occupancy, const float* out_occ0, const float* out_occ1,
                                      const float* out_rgb, const double* depth, const double* g_depth,
                                      const double* g_var, const float* g_rgb, float* d_raw,
                                      float* d_rays_d_norm, void* stream) {
    CompParams c;
    if (int rc = comp_params(rays, bound6_host, stage, occupancy, out_occ0, out_occ1, out_rgb, true, c)) return rc;
    if (!depth || !d_raw) return set_error(NSB_EINVAL, "null argument");
    if (c.q.N <= 0) return 0;
    const int bs = 128;
    launch_pdl(k_composite_bwd, dim3((unsigned)((c.q.N * 32 + bs - 1) / bs)), dim3(bs), 0, (cudaStream_t)stream, 
        c, depth, g_depth, g_var, g_rgb, d_raw, d_rays_d_norm);
    return check_launch("k_composite_bwd");
}

extern "C" int nsb_assemble_raw(const NsbPoints* pts, const double* bound6_host, int stage,
                                const float* out_occ0, const float* out_occ1, const float* out_rgb,
                                float* raw, uint8_t* inbound, void* stream) {
    CompParams c;
    if (int rc = comp_params(pts, bound6_host, stage, 1, out_occ0, out_occ1, out_rgb, false, c)) return rc;
    if (!raw) return set_error(NSB_EINVAL, "null output");
    if (c.q.P <= 0) return 0;
    k_assemble_raw<<<(unsigned)((c.q.P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(c, raw, inbound);
    return check_launch("k_assemble_raw");
}

extern "C" int nsb_ray_grad_reduce(const NsbPoints* rays, const float* d_point, int ndec,
                                   const float* d_rays_d_norm, float* g_rays_o, float* g_rays_d,
                                   void* stream) {
    if (!rays || !d_point || !g_rays_o || !g_rays_d) return set_error(NSB_EINVAL, "null argument");
    PointsDev q = to_dev(*rays);
    if (!q.z) return set_error(NSB_EINVAL, "ray mode required");
    if (q.N <= 0) return 0;
    const int bs = 128;
    launch_pdl(k_ray_grad_reduce, dim3((unsigned)((q.N * 32 + bs - 1) / bs)), dim3(bs), 0, (cudaStream_t)stream, 
        q, d_point, ndec, d_rays_d_norm, g_rays_o, g_rays_d);
    return check_launch("k_ray_grad_reduce");
}

extern "C" int nsb_pose_grad(const float* g_rays_o, const float* g_rays_d, const int64_t* indices,
                             const uint8_t* keep, int64_t n, int H0, int W0, int Wc, float fx, float fy,
                             float cx, float cy, const float* cam7, float* d_cam7, void* stream) {
    if (!g_rays_o || !g_rays_d || !indices || !cam7 || !d_cam7) return set_error(NSB_EINVAL, "null argument");
    launch_pdl(k_pose_grad, dim3(1), dim3(1024), 0, (cudaStream_t)stream, g_rays_o, g_rays_d, indices, keep, n, H0, W0, Wc, fx, fy,
                                                     cx, cy, cam7, d_cam7, CamAdam{});
    return check_launch("k_pose_grad");
}

extern "C" int nsb_pose_grad_adam(const float* g_rays_o, const float* g_rays_d, const int64_t* indices,
                                  const uint8_t* keep, int64_t n, int H0, int W0, int Wc, float fx, float fy,
                                  float cx, float cy, float* cam7, float* d_cam7, float* m, float* v, int32_t* step,
                                  float lr_q, float lr_t, float beta1, float beta2, float eps, const double* loss,
                                  double* best_loss, float* best_cam, void* stream) {
    if (!g_rays_o || !g_rays_d || !indices || !cam7 || !d_cam7 || !m || !v || !step)
        return set_error(NSB_EINVAL, "null argument");
    if (best_loss && (!loss || !best_cam)) return set_error(NSB_EINVAL, "best_loss needs loss and best_cam");
    CamAdam A{cam7, m, v, step, lr_q, lr_t, beta1, beta2, eps, loss, best_loss, best_cam};
    launch_pdl(k_pose_grad, dim3(1), dim3(1024), 0, (cudaStream_t)stream, g_rays_o, g_rays_d, indices, keep, n, H0, W0, Wc, fx, fy,
                                                     cx, cy, (const float*)cam7, d_cam7, A);
    return check_launch("k_pose_grad");
}

extern "C" int64_t nsb_composite_mapping_partials(int64_t N) { return (N * 32 + 127) / 128; }

extern "C" int nsb_composite_mapping(const NsbPoints* rays, const double* bound6_host, int stage, int occupancy,
                                     const float* out_occ0, const float* out_occ1, const float* out_rgb,
                                     const float* gt_depth, const float* gt_color, const uint8_t* keep, float w_color,
                                     int use_color, double* depth, double* var, float* rgb, float* d_raw,
                                     double* partials, uint32_t* ticket, double* loss, void* stream) {
    CompParams c;
    if (int rc = comp_params(rays, bound6_host, stage, occupancy, out_occ0, out_occ1, out_rgb, true, c)) return rc;
    if (!gt_depth || !gt_color || !depth || !var || !rgb || !d_raw || !partials || !ticket || !loss)
        return set_error(NSB_EINVAL, "null argument");
    if (c.q.N <= 0) return 0;
    const int bs = 128;
    launch_pdl(k_composite_map, dim3((unsigned)((c.q.N * 32 + bs - 1) / bs)), dim3(bs), 0, (cudaStream_t)stream, 
        c, gt_depth, gt_color, keep, w_color, use_color, depth, var, rgb, d_raw, partials, ticket, loss);
    return check_launch("k_composite_map");
}

// ---- allocation-free variant of nsb_composite_mapping ------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel, same
// inputs and outputs; its scratch (per-block loss partials + the arrival ticket) lives in ONE file-scope device
// array of a fixed, modest size that every GPU loading the library holds until exit:
//     [0, 256)                      the 32-bit ticket (rest of the 256-byte slot unused)
//     [256, 256 + 8 * MAX_PARTIALS) one fp64 partial sum per 128-thread block (4 rays per block)
// = 128.25 KiB for NSB_STATIC_MAX_RAYS = 65536 rays per call.
// In HOST code `g_cm_static_scratch` is only a stand-in for the symbol, not a device address: the address is taken
// with cudaGetSymbolAddress, once per device, with that device current, and that pointer is what the kernel gets.
// The array is shared by every call and keeps the last call's contents: the ticket must start at zero, so it is
// cleared on the caller's stream before every launch (the partials need no clearing: every block writes its slot
// before the last block reads them).  Calls that overlap on the GPU (two streams, two host threads) RACE on it.
constexpr long long CM_STATIC_MAX_RAYS = NSB_STATIC_MAX_RAYS;
constexpr long long CM_STATIC_MAX_PARTIALS = (CM_STATIC_MAX_RAYS * 32 + 127) / 128;
constexpr size_t CM_STATIC_BYTES = 256 + 8 * (size_t)CM_STATIC_MAX_PARTIALS;
static __device__ __align__(256) unsigned char g_cm_static_scratch[CM_STATIC_BYTES];

constexpr int CM_STATIC_MAX_DEVICES = 64;
static unsigned char* g_cm_static_ptr[CM_STATIC_MAX_DEVICES] = {};      // device address of the array, per device ordinal

// device address of the static scratch on the CURRENT device (set-up once per device; no allocation)
static int cm_static_scratch(unsigned char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error((int)e, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= CM_STATIC_MAX_DEVICES)
        return set_error(NSB_EINVAL, "nsb_composite_mapping_static: device ordinal %d out of range (max %d)", dev,
                         CM_STATIC_MAX_DEVICES - 1);
    if (!g_cm_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_cm_static_scratch);
        if (e != cudaSuccess || !p)
            return set_error(e != cudaSuccess ? (int)e : NSB_EINVAL, "cudaGetSymbolAddress(static scratch) on device %d: %s",
                             dev, cudaGetErrorString(e));
        g_cm_static_ptr[dev] = (unsigned char*)p;
    }
    *out = g_cm_static_ptr[dev];
    return 0;
}

extern "C" int nsb_composite_mapping_static(const NsbPoints* rays, const double* bound6_host, int stage, int occupancy,
                                            const float* out_occ0, const float* out_occ1, const float* out_rgb,
                                            const float* gt_depth, const float* gt_color, const uint8_t* keep,
                                            float w_color, int use_color, double* depth, double* var, float* rgb,
                                            float* d_raw, double* loss, void* stream) {
    CompParams c;
    if (int rc = comp_params(rays, bound6_host, stage, occupancy, out_occ0, out_occ1, out_rgb, true, c)) return rc;
    if (!gt_depth || !gt_color || !depth || !var || !rgb || !d_raw || !loss) return set_error(NSB_EINVAL, "null argument");
    if (c.q.N > CM_STATIC_MAX_RAYS)
        return set_error(NSB_EINVAL,
                         "nsb_composite_mapping_static: N = %lld rays exceeds the static scratch limit of %lld rays per call "
                         "(NSB_STATIC_MAX_RAYS); use nsb_composite_mapping with caller-provided scratch",
                         (long long)c.q.N, CM_STATIC_MAX_RAYS);
    if (c.q.N <= 0) return 0;
    unsigned char* base = nullptr;
    if (int rc = cm_static_scratch(&base)) return rc;
    unsigned* ticket = reinterpret_cast<unsigned*>(base);
    double* partials = reinterpret_cast<double*>(base + 256);
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned), (cudaStream_t)stream);      // must start at zero
    if (e != cudaSuccess) return set_error((int)e, "cudaMemsetAsync(ticket): %s", cudaGetErrorString(e));
    const int bs = 128;
    // an ordinary launch: the kernel's griddepcontrol instructions are no-ops without the programmatic attribute
    k_composite_map<<<(unsigned)((c.q.N * 32 + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(
        c, gt_depth, gt_color, keep, w_color, use_color, depth, var, rgb, d_raw, partials, ticket, loss);
    return check_launch("k_composite_map (static scratch)");
}
