/*
 * nice_slam_b200.h -- C ABI of the B200 (sm_100a) render hot path of NICE-SLAM.
 *
 * The reference (WohimLee/NICE-SLAM-annotation, pure Python/PyTorch) has no FFI: its hot
 * path is the Python API Renderer.render_batch_ray / eval_points + the conv_onet decoders
 * (SURVEY.md section 8b).  This header is what a native binding for that path binds: every
 * entry point names the reference code it replaces (paths relative to the reference root).
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends
 *    in _host; nothing is allocated, retained or freed by the library; all work is queued
 *    on `stream` (a cudaStream_t passed as void*) and the call returns without syncing.
 *  - every function returns 0 on success, a negative NSB_E* code for a bad argument, or a
 *    positive cudaError_t; nsb_last_error() returns a thread-local description.
 *  - feature grids are fp32, logical shape [1, 32, Z, Y, X] as in the reference
 *    (src/NICE_SLAM.py:243-320) but CHANNELS-LAST in memory: element (c, z, y, x) lives at
 *    ((z*Y + y)*X + x)*32 + c, i.e. one 128-byte line per voxel
 *    (torch.channels_last_3d).  Gradients use the same layout.
 *  - points are addressed either as rays (p = rays_o + rays_d * z, evaluated in fp64 like
 *    src/utils/Renderer.py:227-231) or as an explicit [P,3] array (fp64 or fp32).
 */
#ifndef NICE_SLAM_B200_H
#define NICE_SLAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NSB_VERSION 1

/* error codes (negative); positive values are cudaError_t */
#define NSB_EINVAL (-22)
#define NSB_ENOSYS (-38)

/* decoder kinds: src/conv_onet/models/decoder.py:316-334 (NICE sub-decoders) */
#define NSB_DEC_COARSE 0 /* MLP_no_xyz, c_dim 32, 1 output  (decoder.py:222-297) */
#define NSB_DEC_MIDDLE 1 /* MLP, fourier 93, c_dim 32, 1 output (decoder.py:93-219) */
#define NSB_DEC_FINE 2   /* MLP, fourier 93, c_dim 64 (fine ++ middle feature), 1 output */
#define NSB_DEC_COLOR 3  /* MLP, fourier 93, c_dim 32, 4 outputs */
#define NSB_NUM_DEC 4

/* stages of NICE.forward (decoder.py:336-370) */
#define NSB_STAGE_COARSE 0
#define NSB_STAGE_MIDDLE 1
#define NSB_STAGE_FINE 2
#define NSB_STAGE_COLOR 3
/* NSB_STAGE_SINGLE + NSB_DEC_x: run decoder x alone (MLP.forward / MLP_no_xyz.forward,
 * decoder.py:190-219,285-297); its plane is out_occ0 (coarse, middle), out_occ1 (fine) or
 * out_rgb (color).  Accepted by nsb_decoders_forward / nsb_decoders_backward only. */
#define NSB_STAGE_SINGLE 16

/* backward selection flags */
#define NSB_GRAD_GRID 1u   /* scatter-add d(grid) for the levels that carry gradient */
#define NSB_GRAD_POINT 2u  /* d(point) -> per-ray d(rays_o), d(rays_d) */
#define NSB_GRAD_WEIGHT 4u /* d(decoder parameters), bitmask of decoders in dweight_mask */

typedef struct NsbGrid {
    const float* data; /* channels-last fp32, may be NULL when the stage does not use it */
    float* grad;       /* same layout, accumulated into (caller zeroes), may be NULL */
    int32_t X, Y, Z;   /* grid_sample convention: X <-> p[:,0] <-> LAST logical dim */
} NsbGrid;

/* Point set of one call.  Ray mode: rays_o/rays_d [N,3] fp32 (row strides in floats, so a
 * stride-0 rays_o as produced by src/common.py:88 is accepted) and z [N,S] fp64.
 * Explicit mode (rays_o == NULL): p32 or p64 [P,3]. */
typedef struct NsbPoints {
    const float* rays_o;
    const float* rays_d;
    int64_t rays_o_stride; /* floats between consecutive rays (0 or 3) */
    int64_t rays_d_stride;
    const double* z;       /* [N,S] */
    int64_t N;
    int32_t S;
    const double* p64;     /* [P,3] explicit points (fp64), or NULL */
    const float* p32;      /* [P,3] explicit points (fp32), or NULL */
    int64_t P;             /* number of points (N*S in ray mode) */
    uint32_t* relu_masks;  /* optional [n decoders of the stage][P][8] words (opaque: the ReLU sign bits of the
                            * five layers of every point).  nsb_decoders_forward writes them,
                            * nsb_decoders_backward of the SAME points / stage / weights reads them and skips
                            * the forward recomputation for decoders without a weight gradient.  NULL: recompute. */
    /* optional: the scratch / size / dweight_mask that nsb_decoders_backward of the SAME points will be
     * given.  nsb_decoders_forward then deposits the layer inputs (embedding, features, hidden
     * activations) of the trained decoders into the weight-gradient records, and the backward skips their
     * recomputation too.  Needs relu_masks and a scratch that holds the whole batch
     * (nsb_dweight_scratch_bytes); otherwise, or with NULL, nothing changes. */
    void* x_scratch;
    int64_t x_scratch_bytes;
    uint32_t x_mask;
    /* optional [P][8] floats (opaque): every point's fp32 position and normalised grid coordinate, evaluated in
     * fp64 from rays and depths.  As the rays argument of nsb_sample_z's render-call variant (nsb_render_forward) the
     * sampler fills it for the depths it writes; nsb_decoders_forward / nsb_decoders_backward of those points then
     * read it instead of evaluating the points per decoder.  Scene bound only (not the coarse decoder's).  NULL: evaluate. */
    float* point_cache;
    /* optional (ray-sharded mapping): one BYTE per 128-byte line of the flat gradient bucket that starts at
     * line_marks_base, at byte 128 + (byte offset >> 7) of line_marks; its first word = "marks are valid".
     * The tcgen05 backward writes 1 for the voxel lines it scatters into (plain stores) and sets the first word;
     * any other backward path clears the first word.  nsb_bucket_allreduce_sparse (NsbPeers.line_marks) then reads
     * 32 bytes per 4 KB unit instead of scanning the unit, and clears them.  Local device memory of
     * 128 + bucket bytes / 128 bytes, zero-initialised by the caller once. */
    uint32_t* line_marks;
    const float* line_marks_base;
} NsbPoints;

/* Scene of one call: bound (src/NICE_SLAM.py:188-208), grids, packed decoders. */
typedef struct NsbScene {
    double bound[3][2];        /* lo/hi per axis, the Renderer's (non-enlarged) bound */
    double coarse_bound[3][2]; /* bound * coarse_bound_enlarge (NICE_SLAM.py:208) */
    NsbGrid grid[4];           /* indexed by NSB_DEC_* : coarse, middle, fine, color */
    const float* packed[4];    /* nsb_pack_weights output per decoder, or NULL */
} NsbScene;

int nsb_version(void);
const char* nsb_last_error(void);

/* ---- instrumentation (bench.py) -------------------------------------------------------
 * nsb_launch_count: kernels launched by this library in this process (optionally reset).
 * nsb_timing(1) brackets every decoder forward / backward kernel with CUDA events on the
 * launching stream; nsb_timing_read sums their durations (ms) and counts since nsb_timing(). */
int64_t nsb_launch_count(int reset);
int nsb_timing(int enable);
int nsb_timing_read(double* ms_forward, int* n_forward, double* ms_backward, int* n_backward);
/* which: 0 k_decoders_forward, 1 k_decoders_backward, 2 k_dweight */
int nsb_timing_read_kernel(int which, double* ms, int* n);

/* ---- decoder parameter packing ------------------------------------------------------
 * Canonical parameter order of one decoder (reference state_dict names, decoder.py):
 *   pts_linears.{0..4}.weight, pts_linears.{0..4}.bias, fc_c.{0..4}.weight,
 *   fc_c.{0..4}.bias (absent for COARSE), output_linear.weight, output_linear.bias,
 *   embedder._B (absent for COARSE).
 * nsb_param_count / nsb_param_numel describe it; nsb_packed_numel is the size of the
 * packed blob; nsb_pack_table_host fills a host int32 table (one entry per packed float:
 * (param_index << 20) | element_offset, or -1 for a zero pad).  nsb_pack_weights gathers the
 * blob on the device from the live parameter storages (dev array of pointers).
 * nsb_unpack_weight_grads is not needed: weight gradients are produced in canonical flat
 * order (concatenation of the parameters above). */
int nsb_param_count(int kind);
int64_t nsb_param_numel(int kind, int index);
int64_t nsb_flat_numel(int kind);   /* sum of nsb_param_numel */
int64_t nsb_packed_numel(int kind);
int nsb_pack_table_host(int kind, int32_t* table_host);
/* test hook: 32 ints describing the blob layout (offsets of B, biases and every packed GEMM
 * block, in the order listed in csrc/nsb_api.cu:layout_offsets) */
int nsb_layout_offsets(int kind, int32_t* out32_host);
int nsb_pack_weights(int kind, const float* const* param_ptrs, const int32_t* table,
                     float* packed, void* stream);

/* ---- K1: rays and sample depths -------------------------------------------------------
 * nsb_sample_rays: get_samples / get_rays_from_uv (src/common.py:74-134).  indices are flat
 * indices into the crop [H0:H1, W0:W1] (row-major), as drawn by torch.randint
 * (common.py:99).  c2w is 12 floats (3x4 row-major); alternatively cam7 = [quat(4), t(3)]
 * (get_camera_from_tensor, common.py:163-176, evaluated in the kernel).  color may be fp32 or fp64
 * (color_is_f64), the sampled colour is written as fp32.  prefilter (optional, may be
 * NULL): keep[n] = (bbox exit >= gt depth), src/Tracker.py:100-107, src/Mapper.py:580-591. */
int nsb_sample_rays(const int64_t* indices, int64_t n, int H0, int H1, int W0, int W1,
                    int W_img, float fx, float fy, float cx, float cy, const float* c2w,
                    const float* cam7, const float* depth_img, const void* color_img, int color_is_f64,
                    const double* bound6_host, float* rays_o, float* rays_d, float* gt_depth,
                    float* gt_color, uint8_t* keep, void* stream);

/* all-pixel rays, get_rays (src/common.py:259-277) */
int nsb_all_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w,
                 float* rays_o, float* rays_d, void* stream);

/* max over the batch of gt_depth (Renderer.py:157,194): out[0] = max(gt), fp32.  keep (optional,
 * [N] bytes) restricts the max to the rays the box pre-filter kept (the reference compacts the
 * batch before rendering; a fixed-shape caller passes the mask instead). */
int nsb_depth_max(const float* gt_depth, const uint8_t* keep, int64_t N, float* out, void* stream);

/* sorted sample depths z[N, N_samples + N_surface] in fp64 (Renderer.py:116-224).
 * gt_depth NULL => near 0.01, far = box exit + 0.01, no surface samples.
 * t_uniform [N_samples] fp32 and t_surface [N_surface] fp64 are torch.linspace(0,1,.)
 * values supplied by the host; u (optional) [N,N_samples] fp32 uniforms for perturb > 0. */
int nsb_sample_z(const NsbPoints* rays, const float* gt_depth, const double* bound6_host,
                 int N_samples, int N_surface, int lindisp, const float* t_uniform,
                 const double* t_surface, const float* u, const float* depth_max,
                 double* z_out, void* stream);

/* box pre-filter of explicit rays (src/Tracker.py:100-107, src/Mapper.py:587-594): keep[r] = 1 when ray r leaves
 * the scene box (bound6_host: x0,x1,y0,y1,z0,z1, fp64) at or beyond its sensor depth.  The reference compacts the
 * batch with this mask; the render entry points take it as NsbRenderCall.keep instead. */
int nsb_box_prefilter(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t n,
                      const double* bound6_host, uint8_t* keep, void* stream);

/* iMAP free-space regulariser points (src/utils/Renderer.py:360-394): N_samples stratified, always-jittered fp32
 * depths in [0, 0.85 gt_depth] per ray and pts [N*N_samples,3] = o + d z; the densities are then one
 * nsb_imap_* / eval_points call on pts (Renderer.py:395-409).  t_uniform [N_samples] = torch.linspace(0,1,.),
 * u [N,N_samples] fp32 uniforms (the reference draws them from the CPU generator, Renderer.py:385).
 * z_out (optional) [N,N_samples] keeps the depths for the backward. */
int nsb_regulation_points(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t N,
                          int N_samples, const float* t_uniform, const float* u, float* pts, float* z_out,
                          void* stream);
/* adjoint of the above towards the rays (bundle adjustment, Mapper.py:571-584 makes them depend on the camera
 * tensors): d_rays_o = sum_s d_pts, d_rays_d = sum_s z d_pts.  Either output may be NULL. */
int nsb_regulation_points_backward(const float* d_pts, const float* z, int64_t N, int N_samples,
                                   float* d_rays_o, float* d_rays_d, void* stream);

/* importance resampling: sample_pdf (src/common.py:19-63) + sort-merge with the old depths
 * (Renderer.py:244-249).  weights [N,S] fp32 of the first pass, u [N,N_imp] fp32
 * (torch.linspace for the deterministic case, uniforms otherwise). */
int nsb_importance_z(const double* z, const float* weights, int64_t N, int S, int N_imp,
                     const float* u, int64_t u_row_stride, double* z_samples, double* z_out,
                     void* stream);

/* sample_pdf itself (src/common.py:19-63): bins [N,M] fp64, weights [N,M-1] fp32 ->
 * samples [N,N_samples] fp64 (unsorted, as the reference returns them). */
int nsb_sample_pdf(const double* bins, const float* weights, int64_t N, int M, int N_samples,
                   const float* u, int64_t u_row_stride, double* samples, void* stream);

/* ---- K2 + K3: grid gather + decoders --------------------------------------------------
 * Forward of NICE.forward at `stage` (decoder.py:336-370) for all points.  Outputs are
 * the raw decoder outputs per point, one plane per decoder that the stage runs:
 *   out_occ0 [P]   coarse (stage coarse) or middle decoder output
 *   out_occ1 [P]   fine decoder output (stages fine, color)
 *   out_rgb  [P,4] colour decoder output (stage color)
 * Planes a stage does not run may be NULL. */
int nsb_decoders_forward(const NsbPoints* pts, const NsbScene* scene, int stage,
                         float* out_occ0, float* out_occ1, float* out_rgb, void* stream);

/* raw [P,4] as returned by Renderer.eval_points (Renderer.py:36-93): rgb from out_rgb (or 0),
 * occupancy = occ1 + occ0 (decoder.py:369), 100 for points outside the bound
 * (Renderer.py:63-66,87).  inbound (optional) receives the mask. */
int nsb_assemble_raw(const NsbPoints* pts, const double* bound6_host, int stage,
                     const float* out_occ0, const float* out_occ1, const float* out_rgb,
                     float* raw, uint8_t* inbound, void* stream);

/* Backward of the above.  d_raw [P,4] is dL/d(raw) as produced by nsb_composite_backward
 * (channel 3 already zero for out-of-bound points).  Selected by `flags`:
 *   NSB_GRAD_GRID   -> scene->grid[k].grad accumulated (levels with gradient at this stage)
 *   NSB_GRAD_POINT  -> d_point [ndec, P, 3] fp32, one plane per decoder that ran
 *   NSB_GRAD_WEIGHT -> d_weight[k] (flat canonical order, accumulated) for k in dweight_mask;
 *                      scratch must hold nsb_dweight_scratch_bytes(P) bytes per trained decoder: the records of
 *                      the whole batch up to 4 Mi points (47.6 KB per 16 points), above that one 512 Ki-point
 *                      chunk (the backward then runs in chunks and the forward cannot deposit, see x_scratch)
 * Points whose incoming gradient is exactly zero are skipped. */
int64_t nsb_dweight_scratch_bytes(int kind, int64_t P);
int nsb_decoders_backward(const NsbPoints* pts, const NsbScene* scene, int stage,
                          const float* d_raw, unsigned flags, unsigned dweight_mask,
                          float* const* d_weight_host, float* d_point, void* scratch,
                          int64_t scratch_bytes, void* stream);

/* ---- K4: volume rendering -------------------------------------------------------------
 * raw2outputs_nerf_color (src/common.py:215-256) on the decoder planes of a stage, with the
 * out-of-bound rule of eval_points folded in.  occupancy != 0: alpha = sigmoid(10 raw3);
 * else alpha = 1 - exp(-relu(raw3) * dz * |rays_d|).  Outputs depth/var fp64, rgb fp32,
 * weights [N,S] fp32 (optional). */
int nsb_composite_forward(const NsbPoints* rays, const double* bound6_host, int stage,
                          int occupancy, const float* out_occ0, const float* out_occ1,
                          const float* out_rgb, double* depth, double* var, float* rgb,
                          float* weights, void* stream);

/* adjoint: (g_depth, g_var fp64; g_rgb fp32) -> d_raw [P,4]; in density mode also
 * d_rays_d_norm [N,3] (gradient through |rays_d|), may be NULL otherwise. */
int nsb_composite_backward(const NsbPoints* rays, const double* bound6_host, int stage,
                           int occupancy, const float* out_occ0, const float* out_occ1,
                           const float* out_rgb, const double* depth, const double* g_depth,
                           const double* g_var, const float* g_rgb, float* d_raw,
                           float* d_rays_d_norm, void* stream);

/* per-ray reduction of point gradients: g_o[n] = sum_s dp, g_d[n] = sum_s z dp
 * (+ d_rays_d_norm).  d_point [ndec, P, 3]. */
int nsb_ray_grad_reduce(const NsbPoints* rays, const float* d_point, int ndec,
                        const float* d_rays_d_norm, float* g_rays_o, float* g_rays_d,
                        void* stream);

/* camera gradient (src/common.py:137-176 adjoint): from per-ray
 * gradients and pixel directions to d(camera_tensor[7]) = [dquat(4), dt(3)].
 * dirs are recomputed from the pixel indices of nsb_sample_rays. */
int nsb_pose_grad(const float* g_rays_o, const float* g_rays_d, const int64_t* indices,
                  const uint8_t* keep, int64_t n, int H0, int W0, int Wc, float fx, float fy,
                  float cx, float cy, const float* cam7, float* d_cam7, void* stream);
/* the same followed, in the same launch, by the camera optimiser's step of nsb_cam_adam (src/Tracker.py:232-242,
 * 262-275): cam7 is read for the gradient and then updated in place. */
int nsb_pose_grad_adam(const float* g_rays_o, const float* g_rays_d, const int64_t* indices,
                       const uint8_t* keep, int64_t n, int H0, int W0, int Wc, float fx, float fy,
                       float cx, float cy, float* cam7, float* d_cam7, float* m, float* v, int32_t* step,
                       float lr_q, float lr_t, float beta1, float beta2, float eps, const double* loss,
                       double* best_loss, float* best_cam, void* stream);

/* test hook: D[128,N] = A[128,K] W[N,K]^T on the tcgen05 tensor cores (kind::tf32).  mode bit 0:
 * A operand from tensor memory instead of shared memory; bit 1: 3xTF32 error compensation. */
int nsb_tc_selftest(const float* A, const float* W, float* D, int K, int N, int mode, void* stream);

/* micro-benchmark: SM cycles for `count` tf32 MMAs (M 128, width N, K 8) issued back to back into
 * `nacc` accumulators in rotation; out2[0] = issue cycles, out2[1] = cycles until all completed.
 * mode 0: A in tensor memory, B K-major; 1: A and B K-major in shared memory; 2: both MN-major
 * (N a multiple of 32); 3: A K-major, B MN-major; 4: A MN-major, B K-major. */
int nsb_tc_mma_latency(long long* out2, int N, int count, int nacc, int mode, void* stream);

/* ---- whole render call ---------------------------------------------------------------------
 * Renderer.render_batch_ray (src/utils/Renderer.py:95-262) without importance resampling, and
 * its backward, as ONE host call each: depth max -> sample depths -> decoders -> compositing
 * (forward), compositing adjoint -> decoders backward (+ weight gradients) -> per-ray reduction
 * (backward), all queued on `stream`.  The forward keeps z and the decoder planes in
 * `workspace` (nsb_render_workspace_bytes), which the backward of the same call reads, so the
 * caller must leave it untouched in between.  `keep` (optional) marks the rays of a fixed-shape
 * batch that survived the box pre-filter: dropped rays are excluded from the batch depth max;
 * the caller gives them zero incoming gradient. */
typedef struct NsbRenderCall {
    const float* rays_o;
    const float* rays_d;
    int64_t rays_o_stride, rays_d_stride;
    int64_t N;
    const float* gt_depth;  /* [N] or NULL (coarse stage / no sensor depth) */
    const uint8_t* keep;    /* [N] or NULL */
    const float* depth_max; /* [1] or NULL: batch max of gt_depth computed by the caller (a rank that
                             * renders a shard of the batch passes the max over the WHOLE batch) */
    int32_t N_samples, N_surface, lindisp, stage, occupancy;
    const float* t_uniform;  /* [N_samples] torch.linspace(0,1) fp32 */
    const double* t_surface; /* [N_surface] fp64 */
    const float* u;          /* [N,N_samples] jitter uniforms or NULL */
    NsbScene scene;          /* grid[k].grad are the scatter targets of the backward */
    void* workspace;
    int64_t workspace_bytes;
    /* forward outputs */
    double* depth; /* [N] */
    double* var;   /* [N] */
    float* color;  /* [N,3] */
    /* backward inputs (NULL = zero) and outputs */
    const double* g_depth;
    const double* g_var;
    const float* g_color;
    uint32_t flags, dweight_mask;
    float* d_weight[4]; /* flat canonical gradients per decoder kind (accumulated) */
    void* scratch;      /* weight-gradient records, nsb_dweight_scratch_bytes */
    int64_t scratch_bytes;
    float* g_rays_o; /* [N,3] with NSB_GRAD_POINT */
    float* g_rays_d;
    /* fused mapping loss (src/Mapper.py:604-612), optional: with loss != NULL nsb_render_forward stops
     * after the decoders and nsb_render_backward runs compositing + loss + the adjoint of both as ONE
     * kernel (nsb_composite_mapping): depth / var / color and loss[0] are then written by
     * nsb_render_backward and g_depth / g_var / g_color are ignored.  Needs gt_depth. */
    const float* gt_color; /* [N,3] */
    float w_color;
    int32_t use_color;
    double* loss; /* [1] */
    /* opt-in: nsb_render_forward deposits the layer inputs of the trained decoders (dweight_mask) into the
     * weight-gradient records in `scratch`, and nsb_render_backward skips their recomputation.  flags,
     * dweight_mask, scratch and this field must then be the same in the forward and the backward call of
     * a step (they are when one struct serves both), and scratch must hold the whole batch. */
    int32_t deposit_dweight_inputs;
    /* optional, ray-sharded mapping: see NsbPoints.line_marks / line_marks_base */
    uint32_t* line_marks;
    const float* line_marks_base;
    /* nsb_render_backward in two calls, so that a caller can start the exchange of the grid gradients between them
     * (they are complete after phase 1): 0 = everything (default), 1 = compositing adjoint + decoder backward only,
     * 2 = only the weight-gradient kernels of a phase-1 call with the same struct (needs deposit_dweight_inputs).
     * dweight_max_ctas > 0 caps their grid (leaves SMs to a concurrent exchange kernel). */
    int32_t dweight_phase;
    int32_t dweight_max_ctas;
} NsbRenderCall;

/* One ray batch (contiguous [n,3] / [n] device tensors, n <= N) into the fixed-shape input buffers of a
 * render call that is replayed as a CUDA graph: rays n..N-1 are masked off (keep = 0).  keep may be NULL
 * (all n rays kept; it may alias dst_keep).  rays_o == NULL: the n live entries are already in place (a
 * caller that copied them from host memory) and only keep and the tail are written.  One launch instead
 * of a copy per tensor. */
int nsb_load_batch(const float* rays_o, const float* rays_d, const float* gt_depth, const float* gt_color,
                   const uint8_t* keep, int64_t n, int64_t N, float* dst_rays_o, float* dst_rays_d,
                   float* dst_gt_depth, float* dst_gt_color, uint8_t* dst_keep, void* stream);
/* The same from ONE packed [n,10] fp32 array (rays_o 3, rays_d 3, gt_depth 1, gt_color 3 per ray). */
int nsb_load_batch_packed(const float* packed, const uint8_t* keep, int64_t n, int64_t N, float* dst_rays_o,
                          float* dst_rays_d, float* dst_gt_depth, float* dst_gt_color, uint8_t* dst_keep,
                          void* stream);
int64_t nsb_render_workspace_bytes(int64_t N, int S);
int nsb_render_forward(const NsbRenderCall* call, void* stream);
int nsb_render_backward(const NsbRenderCall* call, void* stream);
/* Diagnostic for benchmarks (no reference counterpart): after nsb_render_backward of `call`, count the
 * tiles of `tile_points` consecutive points whose incoming gradient d(raw) is not exactly zero -- the
 * decoder backward skips the others, so this is the share of the backward's work a fixture really asks
 * for.  counts (device, int64 [3], zeroed by the callee): tiles, tiles with d(occupancy) != 0, tiles with
 * d(rgb) != 0. */
int nsb_render_live_tiles(const NsbRenderCall* call, int tile_points, int64_t* counts, void* stream);

/* ---- losses of the two SLAM steps, value and gradient in one kernel --------------------------
 * mapping (src/Mapper.py:604-612): loss = sum_{gt>0} |gt - depth| + w_color sum |gt_c - color|
 * (colour term when use_color).  tracking (src/Tracker.py:113-130): tmp = |gt - depth| /
 * sqrt(var + 1e-10); mask = gt > 0 (& tmp < 10 median(tmp) when handle_dynamic); loss =
 * sum_mask tmp + w_color sum_mask |gt_c - color|.  Outputs: loss[0] fp64, g_depth [N] fp64,
 * g_color [N,3] fp32 (d loss / d depth, d loss / d color; var is detached as in the reference).
 * keep (optional) drops rays of a fixed-shape batch. */
int nsb_mapping_loss(const double* depth, const float* color, const float* gt_depth,
                     const float* gt_color, const uint8_t* keep, int64_t N, float w_color,
                     int use_color, double* loss, double* g_depth, float* g_color, void* stream);
/* Compositing (src/common.py:215-256) + the mapping loss + the adjoint of both in one kernel, one warp
 * per ray: writes depth/var/rgb [N], d_raw [P,4] (the decoders' incoming gradient) and loss[0].
 * partials: nsb_composite_mapping_partials(N) doubles of scratch; ticket: one 32-bit word that is 0 on
 * entry (left 0 on exit).  The loss is summed in a fixed order (deterministic). */
int64_t nsb_composite_mapping_partials(int64_t N);
int nsb_composite_mapping(const NsbPoints* rays, const double* bound6_host, int stage, int occupancy,
                          const float* out_occ0, const float* out_occ1, const float* out_rgb,
                          const float* gt_depth, const float* gt_color, const uint8_t* keep, float w_color,
                          int use_color, double* depth, double* var, float* rgb, float* d_raw,
                          double* partials, uint32_t* ticket, double* loss, void* stream);
/* nsb_composite_mapping without caller scratch and without ANY allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernel, inputs, outputs and results as nsb_composite_mapping; the
 * per-block loss partials and the ticket live in one file-scope `static __device__` array of the library (128.25 KiB,
 * held by every GPU that loads the library until exit), whose device address is taken with cudaGetSymbolAddress once
 * per device on the first call made with that device current (make that first call outside a stream capture).
 * Limit: N <= NSB_STATIC_MAX_RAYS rays per call; a larger batch is rejected with NSB_EINVAL and a message naming the
 * limit.  The callee clears the ticket on `stream` before every launch (the array keeps the last call's contents).
 * Every call on a device shares the array: calls that OVERLAP on the GPU (two streams, two host threads) race --
 * serialise them on one stream.  Calls no cudaMalloc / cudaFree / allocating library routine. */
#define NSB_STATIC_MAX_RAYS 65536
int nsb_composite_mapping_static(const NsbPoints* rays, const double* bound6_host, int stage, int occupancy,
                                 const float* out_occ0, const float* out_occ1, const float* out_rgb,
                                 const float* gt_depth, const float* gt_color, const uint8_t* keep, float w_color,
                                 int use_color, double* depth, double* var, float* rgb, float* d_raw, double* loss,
                                 void* stream);
int nsb_tracking_loss(const double* depth, const double* var, const float* color,
                      const float* gt_depth, const float* gt_color, const uint8_t* keep, int64_t N,
                      float w_color, int use_color, int handle_dynamic, double* loss,
                      double* g_depth, float* g_color, void* stream);

/* ---- camera optimiser step of the tracker (src/Tracker.py:232-242, 262-275) --------------------
 * torch.optim.Adam (no weight decay, no amsgrad) on the 7-dof camera tensor [quat(4), t(3)] with
 * the reference's two learning rates (quaternion: lr_q, translation: lr_t), state m, v [7] and
 * the step counter step[0] on the device (all zero before the first call).  Then the
 * reference's candidate rule: if loss[0] < best_loss[0], best_loss = loss and best_cam = the
 * UPDATED camera tensor (Tracker.py:270-273 clones camera_tensor after optimizer.step()).
 * Everything stays on the device so a whole tracking iteration can be replayed as a CUDA graph. */
int nsb_cam_adam(float* cam, const float* d_cam, float* m, float* v, int32_t* step, float lr_q,
                 float lr_t, float beta1, float beta2, float eps, const double* loss,
                 double* best_loss, float* best_cam, void* stream);

/* ---- optimiser step of the mapper (src/Mapper.py:393-404, 462-482, 625, 633-642) ---------------
 * torch.optim.Adam (no weight decay, no amsgrad) over a flat fp32 tensor p[n] with gradient g[n] and
 * state m, v [n]; state2[0] = step counter (device, zero before the first call), state2[1] is a
 * ticket the kernel uses and leaves zero.  lr is read from the device (lr[0]) so that a captured
 * CUDA graph follows the stage schedule.  voxel_mask (optional): one byte per 32-float voxel line of a
 * channels-last grid; lines whose byte is 0 are left untouched -- the reference optimises the
 * compact copy val[mask] and writes it back every iteration (frustum_feature_selection), which
 * is the same elementwise update restricted to the mask.  n must be a multiple of 32 with a mask. */
int nsb_adam(float* p, const float* g, float* m, float* v, int64_t n, const uint8_t* voxel_mask,
             int32_t* state2, const float* lr, float beta1, float beta2, float eps, void* stream);

/* ---- once per mapped frame: frustum feature selection and keyframe overlap ----------
 * nsb_frustum_mask = Mapper.get_mask_from_c2w (src/Mapper.py:100-187): voxel centres of a grid
 * (xs [nx], ys [ny], zs [nz]: the torch.linspace values of the bound, device) are projected into the
 * current frame (w2c = np.linalg.inv(c2w) float32 row-major [16] and cam_pos = c2w[:3,3], both HOST),
 * the depth image (device [H,W] fp32) is looked up as cv2.remap INTER_LINEAR does (1/32-pixel fixed
 * point, BORDER_CONSTANT 0; zero look-ups take the maximum look-up), and a voxel is selected when it
 * projects inside the image with 0 <= -z <= depth + 0.5, or lies within 0.5 m of the camera.  Outputs
 * (either may be NULL): mask_xyz [nx,ny,nz] (the reference's array) and mask_zyx [nz,ny,nx] (one byte per
 * 128-byte voxel line of a channels-last grid: what nsb_adam / nsb_bucket_reduce_adam take).  scratch:
 * one device word.  Bit-exact with the reference (tests/golden/frustum.npz).
 * nsb_keyframe_overlap = the scores of Mapper.keyframe_selection_overlap (src/Mapper.py:189-293): for
 * every keyframe k (w2c [K,16] fp32 device) the number of the N x N_samples points o + d z,
 * z = 0.8 gt (1 - t) + (gt + 0.5) t, that project inside its image shrunk by `edge` pixels with
 * camera z < 0; percent_inside = counts[k] / (N N_samples). */
int nsb_frustum_mask(const float* w2c_host, const float* cam_pos_host, const float* xs, const float* ys,
                     const float* zs, int nx, int ny, int nz, const float* depth, int H, int W, double fx,
                     double fy, double cx, double cy, uint32_t* scratch, uint8_t* mask_xyz, uint8_t* mask_zyx,
                     void* stream);
int nsb_keyframe_overlap(const float* rays_o, const float* rays_d, const float* gt_depth, int64_t N,
                         int N_samples, const float* t_vals, const float* w2c, int K, int H, int W, double fx,
                         double fy, double cx, double cy, int edge, int32_t* counts, void* stream);

/* ---- mesh depth images for the 2-D reconstruction metric -----------------------------
 * Replaces the Open3D GL window of calc_2d_metric (src/tools/eval_recon.py:131-210: add_geometry,
 * convert_from_pinhole_camera_parameters, capture_depth_float_buffer(True), back faces shown) by a headless
 * ray-cast: depth [K,H,W] fp32 = z-depth of the nearest triangle of the mesh (vertices [V,3] fp32,
 * triangles [T,3] int32, device) seen through pixel (u, v) of view k (w2c [K,16] fp32 row-major device = the
 * extrinsic np.linalg.inv(c2w); camera looks down +z, y down), 0 where nothing is hit within
 * [z_near, z_far].  workspace: nsb_mesh_depth_workspace_bytes(T, K) bytes (may be NULL: slower on meshes
 * with triangles that cover many pixels).  The intersection arithmetic is fixed operation by operation
 * (oracle/mesh_oracle.py), the images agree with the oracle bit for bit. */
int64_t nsb_mesh_depth_workspace_bytes(int64_t n_triangles, int K);
int nsb_mesh_depth(const float* vertices, int64_t n_vertices, const int32_t* triangles, int64_t n_triangles,
                   const float* w2c, int K, int H, int W, float fx, float fy, float cx, float cy, float z_near,
                   float z_far, void* workspace, int64_t workspace_bytes, float* depth, void* stream);

/* ---- gradient exchange of the sharded mapping step over NVLink peer memory -------
 * The reference maps on one GPU; its loss is a plain sum over rays (src/Mapper.py:605-612), so ranks
 * that render disjoint ray slices hold partial gradients that add.  The gradient bucket of every rank
 * (and, for the fused optimiser, the parameter buffer) lives in SYMMETRIC memory: the same allocation
 * size on every rank, every rank's copy mapped into this process (peer pointers), and -- when the
 * NVSwitch fabric offers NVLS -- one multicast mapping of all copies (NULL otherwise).  The host side
 * obtains the mappings from torch.distributed._symmetric_memory (plumbing only).
 * signal: per-rank pad of nsb_exchange_signal_words() zero-initialised 32-bit words, symmetric too.
 * error (optional, local): set to 1 if a peer did not arrive within ~20 s.
 * Every rank must issue the same sequence of exchange calls with the same offsets and sizes. */
#define NSB_MAX_PEERS 8
typedef struct NsbPeers {
    int32_t rank, world;
    float* bucket[NSB_MAX_PEERS];   /* gradient bucket of rank i as mapped here; bucket[rank] is the local copy */
    float* bucket_multicast;        /* multicast mapping of the bucket, or NULL */
    float* param[NSB_MAX_PEERS];    /* parameter buffer of rank i (nsb_bucket_reduce_adam only) */
    float* param_multicast;
    uint32_t* signal[NSB_MAX_PEERS];
    int32_t* error;
    float* scalar[NSB_MAX_PEERS];     /* nsb_depth_max_peers only: a symmetric, zero-initialised pad of 64 words per rank */
    uint32_t* touched[NSB_MAX_PEERS]; /* nsb_bucket_allreduce_sparse only: per-rank symmetric scratch of
                                       * nsb_exchange_touched_words(bucket floats) words, mapped like bucket */
    uint32_t* line_marks;             /* optional, LOCAL: NsbPoints.line_marks of the backward that filled bucket[rank] */
    int64_t line_marks_floats;        /* leading floats of the bucket the marks cover (the grid segments); the rest
                                       * (decoder weight gradients) is scanned */
} NsbPeers;
int64_t nsb_exchange_signal_words(void);

/* SUM all-reduce of bucket[offset, offset + n) over the ranks, in place in every copy, as ONE kernel:
 * rank r reduces the r-th slice of the range (multimem.ld_reduce through the switch, or peer loads in
 * rank order) and stores the sum into every copy (multimem.st / peer stores); block-wise flag barriers
 * before and after.  Replaces dist.all_reduce on the flat gradient bucket.  offset, n: multiples of 4. */
int nsb_bucket_allreduce(const NsbPeers* peers, int64_t offset, int64_t n, void* stream);
/* The same sum, moving only the 128-byte lines some rank touched: every rank scans its copy and
 * publishes one bit per line (peers->touched), owners reduce the union.  A mapping step touches 16-39 % of
 * the lines of a Replica-scale bucket, so this is the exchange of the sharded step; lines that are zero
 * on every rank are left alone (they stay zero).  offset: multiple of 32 floats; buffers 128-byte aligned. */
int64_t nsb_exchange_touched_words(int64_t n);
/* nsb_depth_max over the kept rays of ALL ranks (the render's one batch-global scalar, Renderer.py:157),
 * through peers->scalar: what a rank that renders a shard passes as NsbRenderCall.depth_max.  Every rank posts
 * (call number, local max) into its slot of every rank's pad and waits for the peers' values of the same call:
 * all ranks must make the same sequence of calls. */
int nsb_depth_max_peers(const NsbPeers* peers, const float* gt_depth, const uint8_t* keep, int64_t N,
                        float* out, void* stream);
int nsb_bucket_allreduce_sparse(const NsbPeers* peers, int64_t offset, int64_t n, void* stream);
/* The same with at most max_ctas CTAs (0: the default of nsb_bucket_allreduce_sparse): an exchange that runs beside
 * another kernel and should leave it most of the SMs.  Every rank must pass the same value. */
int nsb_bucket_allreduce_sparse_ctas(const NsbPeers* peers, int64_t offset, int64_t n, int max_ctas, void* stream);

/* The same reduction fused with the mapper's Adam step (src/Mapper.py:462-482, 633-642; semantics of
 * nsb_adam): rank r sums the r-th slice of bucket[offset, offset + n), updates
 * param[param_offset + slice] with ITS moments m_slice / v_slice (nsb_exchange_slice() floats each --
 * the optimiser state is sharded over the ranks) and stores the new parameters into every rank's
 * copy.  voxel_mask: one byte per 32 floats of the WHOLE range (identical on every rank), or NULL. */
int64_t nsb_exchange_slice(int rank, int world, int64_t n, int64_t* lo);
int nsb_bucket_reduce_adam(const NsbPeers* peers, int64_t offset, int64_t n, int64_t param_offset,
                           float* m_slice, float* v_slice, const uint8_t* voxel_mask, int32_t* state2,
                           const float* lr, float beta1, float beta2, float eps, void* stream);

/* ---- iMAP* decoder: the single MLP of configs/imap.yaml ----------------------------------------
 * src/conv_onet/config.py:33-37 builds MLP(dim 3, c_dim 0, color, hidden_size 256, skips [],
 * n_blocks 4); forward = src/conv_onet/models/decoder.py:190-219 without grid features:
 * e = sin(p.float() @ B); h = relu(W_i h + b_i) four times; raw = Wo h + bo ([rgb(3), density]).
 * Parameters are the LIVE fp32 tensors of the nn.Module in their torch layout (weight [out, in],
 * row-major); nothing is packed.  hidden: a multiple of 16 up to 256.
 * The layers run as chained 3xTF32 GEMMs on the tcgen05 tensor cores (csrc/nsb_gemm.cu). */
typedef struct NsbImapDecoder {
    const float* B;    /* embedder._B [3, 93] */
    const float* W[4]; /* pts_linears.i.weight: [hidden, 93], then [hidden, hidden] */
    const float* b[4]; /* pts_linears.i.bias [hidden] */
    const float* Wo;   /* output_linear.weight [4, hidden] */
    const float* bo;   /* output_linear.bias [4] */
    int32_t hidden;
} NsbImapDecoder;

/* gradient targets, same shapes, ACCUMULATED into (the caller zeroes them) */
typedef struct NsbImapGrads {
    float* B;
    float* W[4];
    float* b[4];
    float* Wo;
    float* bo;
} NsbImapGrads;

/* forward of all points -> raw [P, 4].  workspace: nsb_imap_workspace_floats(P, hidden, save)
 * floats.  save != 0 keeps the embedding and the four activations of every point for
 * nsb_imap_backward; save == 0 runs in a fixed-size workspace (points are processed in chunks). */
int64_t nsb_imap_workspace_floats(int64_t P, int hidden, int save);
int nsb_imap_forward(const NsbPoints* pts, const NsbImapDecoder* dec, float* raw, float* workspace,
                     int save, void* stream);

/* backward: d_raw [P, 4] -> parameter gradients (NSB_GRAD_WEIGHT, into *grads) and / or
 * d_point [P, 3] (NSB_GRAD_POINT, overwritten).  workspace: the one a save != 0 forward of the SAME
 * points and parameters filled; scratch: nsb_imap_backward_scratch_floats(P, hidden) floats. */
int64_t nsb_imap_backward_scratch_floats(int64_t P, int hidden);
int nsb_imap_backward(const NsbPoints* pts, const NsbImapDecoder* dec, const float* workspace,
                      const float* d_raw, unsigned flags, const NsbImapGrads* grads, float* d_point,
                      float* scratch, void* stream);

/* test hook: one fp32-accurate GEMM through the chained tcgen05 kernel.
 * C[m, n] = epi(sum_k A(m, k) B(n, k)), m < M, n < N_valid <= 256.  a_rc / b_rc = 0: element (r, k)
 * at p[r * ld + k]; 1: at p[k * ld + r].  epi: 0 store, 1 +bias & ReLU, 2 +bias, 3 zero where
 * mask[m, n] <= 0, 4 atomic accumulate (required for n_split > 1).  colsum (optional, a_rc = 1):
 * colsum[m] += sum_k A(m, k).  b_scratch (optional, nsb_gemm3x_pack_floats(N_valid, K) floats, 16-byte
 * aligned): B is split once into its operand images there and fetched by bulk copies -- the route the
 * iMAP* chains take for their weight matrices. */
int64_t nsb_gemm3x_pack_floats(int N_valid, int64_t K);
int nsb_gemm3x(const float* A, int64_t lda, int a_rc, const float* B, int64_t ldb, int b_rc, float* C,
               int64_t ldc, int64_t M, int N_valid, int64_t K, int epi, const float* bias,
               const float* mask, int64_t ldmask, float* colsum, int n_split, float* b_scratch,
               void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NICE_SLAM_B200_H */
