// Batched / chained fp32-accurate GEMM on the tcgen05 tensor cores (3xTF32), the dense contraction
// behind the iMAP* decoder (src/conv_onet/models/decoder.py:93-219 with hidden_size 256, c_dim 0).
//
//   C[m, n] = epilogue( sum_k A(m, k) * B(n, k) ),   m < M (points or output features), n < N <= 256
//
// Operands are plain fp32 arrays in global memory, addressed either with k contiguous ("KC":
// element (r, k) at p[r * ld + k]) or with the row index contiguous ("RC": p[k * ld + r]), so the
// forward layers (X W^T), the input-gradient layers (G W) and the weight gradients (G^T X) all read
// the live torch tensors without a packing pass.  Producer warps split every value into tf32 hi/lo
// and write operand images to shared memory; one thread issues three MMAs per K step
// (hi*hi + lo*hi + hi*lo, fp32 accumulate in tensor memory); eight epilogue warps read the
// accumulator back (tcgen05.ld) and apply bias / ReLU / ReLU-mask / atomic accumulation.
// Two shortcuts keep the producers off the critical path:
//  * a B operand that is a WEIGHT matrix (the same for every row tile) can be split once per launch into
//    its per-chunk hi/lo images in global scratch (GemmLayer::b_packed, k_gemm_pack_b); the kernel then
//    fetches a chunk with one bulk asynchronous copy (cp.async.bulk -> mbarrier) and no ALU work;
//  * when BOTH operands are row-contiguous (the weight gradients G^T X: K = points) the images are
//    MN-major (nsb_tc.cuh), so the producers move float4s along the rows instead of transposing scalars.
//
// A batch is either a CHAIN (all layers share the row dimension; a CTA carries a pair of 128-row
// tiles through every layer, layer l+1 of a tile reading what layer l of the same CTA wrote, so the
// activations of a tile stay in L1/L2 and the whole MLP is one launch) or a list of INDEPENDENT
// GEMMs with split-K (the weight gradients: K = points), whose partial sums are accumulated with
// fp32 atomics.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nsb {

enum GemmEpi : int {
    GEMM_EPI_STORE = 0,      // C = acc
    GEMM_EPI_BIAS_RELU = 1,  // C = max(acc + bias[n], 0)
    GEMM_EPI_BIAS = 2,       // C = acc + bias[n]
    GEMM_EPI_MASK = 3,       // C = mask[m, n] > 0 ? acc : 0        (ReLU adjoint)
    GEMM_EPI_ATOMIC = 4,     // C += acc                            (split-K partial sums)
};

struct GemmLayer {
    const float* A;
    long long lda;
    const float* B;
    long long ldb;
    float* C;
    long long ldc;
    const float* bias;
    const float* mask;
    long long ldmask;
    float* colsum;      // optional (RC-addressed A only): colsum[m] += sum_k A(m, k)
    float* b_packed;    // optional scratch of gemm3x_pack_floats(N, K) floats (16-byte aligned): B's operand images,
                        // written by gemm3x_launch before the GEMM; all layers of a batch or none
    long long M;        // valid rows of A / C (independent mode; a chain uses GemmBatch::M)
    long long K;        // valid reduction length
    long long k_per_split;   // multiple of 16 (independent mode)
    int a_rc, b_rc;     // 0: k contiguous, 1: row contiguous
    int N, N_valid;     // MMA width (multiple of 16, <= 256) and valid columns
    int epi;
    int m_tiles, n_split;    // independent mode
};

constexpr int GEMM_MAX_LAYERS = 6;

struct GemmBatch {
    int n_layers;
    int chain;
    long long M;        // chain: rows shared by all layers
    int m_tiles;        // chain: ceil(M / 128)
    int packed;         // set by gemm3x_launch: every layer's B comes from its b_packed images
    int debug;          // set by gemm3x_launch from NSB_GEMM_DEBUG (timing experiments, results are wrong):
                        // 1 no epilogue stores, 2 no A production, 8 no bulk copies of B
    GemmLayer L[GEMM_MAX_LAYERS];
};

// floats of GemmLayer::b_packed for an operand of N (padded) rows and reduction length K
inline long long gemm3x_pack_floats(int N, long long K) { return (K + 15) / 16 * (long long)N * 32; }

// validates the batch, fills m_tiles, launches (k_gemm_pack_b and) k_gemm3x on `stream`
int gemm3x_launch(GemmBatch& gb, cudaStream_t stream);

}  // namespace nsb
