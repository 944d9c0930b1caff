// iMAP* decoder (the single MLP of configs/imap.yaml: src/conv_onet/config.py:33-37 builds
// MLP(dim 3, c_dim 0, color, hidden_size 256, skips [], n_blocks 4); forward in
// src/conv_onet/models/decoder.py:190-219): Gaussian-Fourier embedding + four ReLU layers + a
// 4-wide linear head, forward and backward, as chained 3xTF32 GEMMs on tcgen05 (nsb_gemm.cu).
//
//   e = sin(p32 @ B)                     k_imap_embed          [P, 96] (93 + 3 zero columns)
//   h1 = relu(e W0^T + b0) ... h4        one chained launch    [P, hidden] each
//   raw = h4 Wo^T + bo                   (same launch)         [P, 4]
//
// backward: g4 = (d_raw Wo) . [h4 > 0], g3 = (g4 W3) . [h3 > 0], ..., de = g1 W0 (one chained
// launch); dW_l = g_l^T h_{l-1}, db_l = colsum(g_l) (one split-K launch, fp32 atomics);
// d(arg) = de . cos(p B), dB = p^T d(arg), dp = d(arg) B^T (k_imap_embed_bwd).
#include "nsb_common.cuh"
#include "nsb_gemm.cuh"

namespace nsb {
namespace {

constexpr int IM_EMB = 93, IM_EPAD = 96;
constexpr long long IM_CHUNK = 262144;      // points per forward-only chain launch (fixed workspace)

__device__ __forceinline__ void point32(const PointsDev& q, long long i, float (&p)[3]) {
    if (q.ro) {
        double d[3];
        ray_point(q, i, d);
        p[0] = (float)d[0]; p[1] = (float)d[1]; p[2] = (float)d[2];
    } else if (q.p64) {
        p[0] = (float)q.p64[3 * i]; p[1] = (float)q.p64[3 * i + 1]; p[2] = (float)q.p64[3 * i + 2];
    } else {
        p[0] = q.p32[3 * i]; p[1] = q.p32[3 * i + 1]; p[2] = q.p32[3 * i + 2];
    }
}

__device__ __forceinline__ float phase96(const float (&p)[3], const float* Bs, int f) {
    return fmaf(p[2], Bs[2 * IM_EPAD + f], fmaf(p[1], Bs[IM_EPAD + f], p[0] * Bs[f]));
}

__device__ __forceinline__ void load_B(const float* __restrict__ B, float* Bs) {
    for (int i = threadIdx.x + threadIdx.y * blockDim.x; i < 3 * IM_EPAD; i += blockDim.x * blockDim.y) {
        const int a = i / IM_EPAD, f = i % IM_EPAD;
        Bs[i] = f < IM_EMB ? B[a * IM_EMB + f] : 0.f;
    }
    __syncthreads();
}

// E[p - i0, 4 jq .. 4 jq + 3] = sin(p . B[:, f]); one thread per (point, feature quad)
__global__ void __launch_bounds__(256) k_imap_embed(PointsDev q, long long i0, long long n, const float* __restrict__ B,
                                                    float* __restrict__ E) {
    __shared__ float Bs[3 * IM_EPAD];
    load_B(B, Bs);
    const long long total = n * (IM_EPAD / 4);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long p = t / (IM_EPAD / 4);
        const int jq = (int)(t % (IM_EPAD / 4));
        float pf[3];
        point32(q, i0 + p, pf);
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int f = 4 * jq + j;
            v[j] = f < IM_EMB ? sin_cw(phase96(pf, Bs, f)) : 0.f;
        }
        *reinterpret_cast<float4*>(E + p * IM_EPAD + 4 * jq) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// block (96, 4): x = embedding feature, y = point lane
__global__ void __launch_bounds__(384) k_imap_embed_bwd(PointsDev q, long long P, const float* __restrict__ B,
                                                        const float* __restrict__ dE, float* dB, float* d_point) {
    __shared__ float Bs[3 * IM_EPAD];
    load_B(B, Bs);
    const int f = threadIdx.x;
    float acc[3] = {0.f, 0.f, 0.f};
    for (long long p = blockIdx.x * (long long)blockDim.y + threadIdx.y; p < P; p += (long long)gridDim.x * blockDim.y) {
        float pf[3];
        point32(q, p, pf);
        float g = 0.f;
        if (f < IM_EMB) g = dE[p * IM_EPAD + f] * cos_cw(phase96(pf, Bs, f));
#pragma unroll
        for (int a = 0; a < 3; ++a) acc[a] = fmaf(pf[a], g, acc[a]);
        if (d_point) {
            float v[3] = {g * Bs[f], g * Bs[IM_EPAD + f], g * Bs[2 * IM_EPAD + f]};
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                v[0] += __shfl_xor_sync(FULL, v[0], o);
                v[1] += __shfl_xor_sync(FULL, v[1], o);
                v[2] += __shfl_xor_sync(FULL, v[2], o);
            }
            if ((threadIdx.x & 31) == 0) {
                atomicAdd(d_point + 3 * p, v[0]);
                atomicAdd(d_point + 3 * p + 1, v[1]);
                atomicAdd(d_point + 3 * p + 2, v[2]);
            }
        }
    }
    if (dB && f < IM_EMB) {
        atomicAdd(dB + f, acc[0]);
        atomicAdd(dB + IM_EMB + f, acc[1]);
        atomicAdd(dB + 2 * IM_EMB + f, acc[2]);
    }
}

// floats of the B operand images of one five-layer chain (k_gemm_pack_b): the forward chain (W0 .. W3, head) and
// the input-gradient chain (Wo^T, W3^T .. W0^T) need the same amount
long long pack_floats(int h) {
    return gemm3x_pack_floats(h, IM_EMB) + 3 * gemm3x_pack_floats(h, h) + gemm3x_pack_floats(16, h) +
           gemm3x_pack_floats(h, 4) + gemm3x_pack_floats(IM_EPAD, h);
}
// hands every layer of a chain its slice of `pack`
void assign_pack(GemmBatch& gb, float* pack) {
    for (int l = 0; l < gb.n_layers; ++l) {
        gb.L[l].b_packed = pack;
        pack += gemm3x_pack_floats(gb.L[l].N, gb.L[l].K);
    }
}

int check_decoder(const NsbImapDecoder* d) {
    if (!d) return set_error(NSB_EINVAL, "imap: null decoder");
    if (d->hidden < 16 || d->hidden > 256 || (d->hidden & 15)) return set_error(NSB_EINVAL, "imap: hidden size %d (multiple of 16 up to 256)", d->hidden);
    if (!d->B || !d->Wo || !d->bo) return set_error(NSB_EINVAL, "imap: null parameter");
    for (int i = 0; i < 4; ++i)
        if (!d->W[i] || !d->b[i]) return set_error(NSB_EINVAL, "imap: null parameter of layer %d", i);
    return 0;
}

// the five forward layers of n points: E [n, 96] -> H[0..3] [n, hidden] -> raw [n, 4]
void forward_chain(GemmBatch& gb, const NsbImapDecoder& d, long long n, const float* E, float* const (&H)[4], float* raw) {
    const int h = d.hidden;
    gb = GemmBatch{};
    gb.n_layers = 5; gb.chain = 1; gb.M = n;
    for (int l = 0; l < 4; ++l) {
        GemmLayer& L = gb.L[l];
        L.A = l ? H[l - 1] : E; L.lda = l ? h : IM_EPAD; L.a_rc = 0;
        L.B = d.W[l]; L.ldb = l ? h : IM_EMB; L.b_rc = 0;
        L.K = l ? h : IM_EMB;
        L.C = H[l]; L.ldc = h; L.N = L.N_valid = h;
        L.bias = d.b[l]; L.epi = GEMM_EPI_BIAS_RELU;
    }
    GemmLayer& L = gb.L[4];
    L.A = H[3]; L.lda = h; L.B = d.Wo; L.ldb = h; L.K = h;
    L.C = raw; L.ldc = 4; L.N = 16; L.N_valid = 4; L.bias = d.bo; L.epi = GEMM_EPI_BIAS;
}

int embed_launch(const PointsDev& q, long long i0, long long n, const float* B, float* E, cudaStream_t s) {
    const long long total = n * (IM_EPAD / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_imap_embed<<<(int)blocks, 256, 0, s>>>(q, i0, n, B, E);
    return check_launch("k_imap_embed");
}

}  // namespace
}  // namespace nsb

using namespace nsb;

extern "C" int64_t nsb_imap_workspace_floats(int64_t P, int hidden, int save) {
    if (P < 1) P = 1;
    if (save) return P * (IM_EPAD + 4LL * hidden) + pack_floats(hidden);
    const long long n = P < IM_CHUNK ? P : IM_CHUNK;
    return n * (IM_EPAD + 2LL * hidden) + pack_floats(hidden);
}

extern "C" int64_t nsb_imap_backward_scratch_floats(int64_t P, int hidden) {
    if (P < 1) P = 1;
    return P * (IM_EPAD + 4LL * hidden) + pack_floats(hidden);
}

extern "C" int nsb_imap_forward(const NsbPoints* pts, const NsbImapDecoder* dec, float* raw, float* workspace,
                                int save, void* stream) {
    if (!pts || !raw || !workspace) return set_error(NSB_EINVAL, "imap: null argument");
    if (int rc = check_decoder(dec)) return rc;
    const PointsDev q = to_dev(*pts);
    if (q.P <= 0) return 0;
    const int h = dec->hidden;
    cudaStream_t s = (cudaStream_t)stream;
    GemmBatch gb;
    if (save) {
        float* E = workspace;
        float* H[4];
        for (int l = 0; l < 4; ++l) H[l] = workspace + q.P * IM_EPAD + (long long)l * q.P * h;
        if (int rc = embed_launch(q, 0, q.P, dec->B, E, s)) return rc;
        forward_chain(gb, *dec, q.P, E, H, raw);
        assign_pack(gb, workspace + q.P * (IM_EPAD + 4LL * h));      // behind E and H[0..3] (all multiples of 16 floats)
        return gemm3x_launch(gb, s);
    }
    const long long ch = q.P < IM_CHUNK ? q.P : IM_CHUNK;
    float* E = workspace;
    float* bufA = workspace + ch * IM_EPAD;
    float* bufB = bufA + ch * h;
    float* H[4] = {bufA, bufB, bufA, bufB};
    for (long long i0 = 0; i0 < q.P; i0 += ch) {
        const long long n = q.P - i0 < ch ? q.P - i0 : ch;
        if (int rc = embed_launch(q, i0, n, dec->B, E, s)) return rc;
        forward_chain(gb, *dec, n, E, H, raw + 4 * i0);
        assign_pack(gb, bufB + ch * h);
        if (int rc = gemm3x_launch(gb, s)) return rc;
    }
    return 0;
}

extern "C" int nsb_imap_backward(const NsbPoints* pts, const NsbImapDecoder* dec, const float* workspace,
                                 const float* d_raw, unsigned flags, const NsbImapGrads* grads, float* d_point,
                                 float* scratch, void* stream) {
    if (!pts || !workspace || !d_raw || !scratch) return set_error(NSB_EINVAL, "imap: null argument");
    if (int rc = check_decoder(dec)) return rc;
    const bool want_w = flags & NSB_GRAD_WEIGHT, want_p = flags & NSB_GRAD_POINT;
    if (want_w) {
        if (!grads || !grads->B || !grads->Wo || !grads->bo) return set_error(NSB_EINVAL, "imap: null gradient target");
        for (int i = 0; i < 4; ++i)
            if (!grads->W[i] || !grads->b[i]) return set_error(NSB_EINVAL, "imap: null gradient target of layer %d", i);
    }
    if (want_p && !d_point) return set_error(NSB_EINVAL, "imap: d_point is NULL");
    if (!want_w && !want_p) return 0;
    const PointsDev q = to_dev(*pts);
    const long long P = q.P;
    if (P <= 0) return 0;
    const int h = dec->hidden;
    cudaStream_t s = (cudaStream_t)stream;
    const float* E = workspace;
    const float* H[4];
    for (int l = 0; l < 4; ++l) H[l] = workspace + P * IM_EPAD + (long long)l * P * h;
    float* dE = scratch;
    float* G[4];
    for (int l = 0; l < 4; ++l) G[l] = scratch + P * IM_EPAD + (long long)l * P * h;

    // ---- input-gradient chain
    GemmBatch gb{};
    gb.n_layers = 5; gb.chain = 1; gb.M = P;
    {
        GemmLayer& L = gb.L[0];                       // g4 = (d_raw Wo) . [h4 > 0]
        L.A = d_raw; L.lda = 4; L.a_rc = 0; L.K = 4;
        L.B = dec->Wo; L.ldb = h; L.b_rc = 1;         // B(n, k) = Wo[k, n]
        L.C = G[3]; L.ldc = h; L.N = L.N_valid = h;
        L.mask = H[3]; L.ldmask = h; L.epi = GEMM_EPI_MASK;
    }
    for (int l = 3; l >= 1; --l) {                    // g_{l-1} = (g_l W_l) . [h_{l-1} > 0]
        GemmLayer& L = gb.L[4 - l];
        L.A = G[l]; L.lda = h; L.a_rc = 0; L.K = h;
        L.B = dec->W[l]; L.ldb = h; L.b_rc = 1;
        L.C = G[l - 1]; L.ldc = h; L.N = L.N_valid = h;
        L.mask = H[l - 1]; L.ldmask = h; L.epi = GEMM_EPI_MASK;
    }
    {
        GemmLayer& L = gb.L[4];                       // de = g_0 W_0
        L.A = G[0]; L.lda = h; L.a_rc = 0; L.K = h;
        L.B = dec->W[0]; L.ldb = IM_EMB; L.b_rc = 1;
        L.C = dE; L.ldc = IM_EPAD; L.N = IM_EPAD; L.N_valid = IM_EMB;
        L.epi = GEMM_EPI_STORE;
    }
    assign_pack(gb, scratch + P * (IM_EPAD + 4LL * h));
    if (int rc = gemm3x_launch(gb, s)) return rc;

    // ---- weight gradients: dW_l = g_l^T h_{l-1}, db_l = column sums of g_l
    if (want_w) {
        long long split = P / 1024;
        const int tiles = 1 + 4 * ((h + 127) / 128);
        const long long cap = (4LL * 148 + tiles - 1) / tiles;
        if (split > cap) split = cap;
        if (split < 1) split = 1;
        GemmBatch gw{};
        gw.n_layers = 5; gw.chain = 0;
        {
            GemmLayer& L = gw.L[0];                   // dWo [4, h] = d_raw^T h4
            L.A = d_raw; L.lda = 4; L.a_rc = 1; L.M = 4;
            L.B = H[3]; L.ldb = h; L.b_rc = 1; L.N = L.N_valid = h;
            L.K = P; L.C = grads->Wo; L.ldc = h; L.colsum = grads->bo;
            L.epi = GEMM_EPI_ATOMIC; L.n_split = (int)split;
        }
        for (int l = 3; l >= 0; --l) {
            GemmLayer& L = gw.L[4 - l];
            L.A = G[l]; L.lda = h; L.a_rc = 1; L.M = h;
            L.B = l ? H[l - 1] : E; L.ldb = l ? h : IM_EPAD; L.b_rc = 1;
            L.N = l ? h : IM_EPAD; L.N_valid = l ? h : IM_EMB;
            L.K = P; L.C = grads->W[l]; L.ldc = l ? h : IM_EMB; L.colsum = grads->b[l];
            L.epi = GEMM_EPI_ATOMIC; L.n_split = (int)split;
        }
        if (int rc = gemm3x_launch(gw, s)) return rc;
    }

    // ---- embedding adjoint
    if (want_p) {
        cudaError_t e = cudaMemsetAsync(d_point, 0, (size_t)P * 3 * sizeof(float), s);
        if (e != cudaSuccess) return set_error((int)e, "imap: memset d_point: %s", cudaGetErrorString(e));
    }
    long long blocks = (P + 3) / 4;
    if (blocks > 148 * 4) blocks = 148 * 4;
    k_imap_embed_bwd<<<(int)blocks, dim3(IM_EPAD, 4), 0, s>>>(q, P, dec->B, dE, want_w ? grads->B : nullptr,
                                                              want_p ? d_point : nullptr);
    return check_launch("k_imap_embed_bwd");
}
