// Known-answer test of the tcgen05 building blocks (nsb_tc.cuh): one 128 x N x K tf32 GEMM
// D = A W^T with the A operand either in shared memory (mode 0) or in tensor memory (mode 1),
// optionally as the 3xTF32 sum (mode | 2).  Used by tests/test_tc_selftest.py to pin the
// descriptor encodings and operand layouts the decoder kernels rely on.
#include "nsb_common.cuh"
#include "nsb_tc.cuh"

namespace nsb {

constexpr int ST_MAXK = 96, ST_MAXN = 64;

__global__ void __launch_bounds__(128, 1)
k_tc_selftest(const float* __restrict__ A, const float* __restrict__ W, float* __restrict__ D, int K, int N, int mode) {
    extern __shared__ __align__(128) float sm[];
    float* Whi = sm;                              // N x K, K-major canonical
    float* Wlo = Whi + ST_MAXN * ST_MAXK;
    float* Ahi = Wlo + ST_MAXN * ST_MAXK;         // 128 x K (mode SS)
    float* Alo = Ahi + 128 * ST_MAXK;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const bool ts = mode & 1, x3 = mode & 2;
    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::mbar_init_fence(); }
    for (int i = tid; i < N * K; i += 128) {
        const int n = i / K, k = i % K;
        uint32_t hi, lo;
        split_tf32(W[i], hi, lo);
        Whi[tc::kmajor_index(n, k, N)] = __uint_as_float(hi);
        Wlo[tc::kmajor_index(n, k, N)] = __uint_as_float(lo);
    }
    for (int i = tid; i < 128 * K; i += 128) {
        const int m = i / K, k = i % K;
        uint32_t hi, lo;
        split_tf32(A[i], hi, lo);
        Ahi[tc::kmajor_index(m, k, 128)] = __uint_as_float(hi);
        Alo[tc::kmajor_index(m, k, 128)] = __uint_as_float(lo);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tbase = tmem_slot;
    const int C_D = 0, C_AHI = 128, C_ALO = 256;           // column offsets
    if (ts) {     // each thread stores ITS row of A (hi and lo) to tensor memory
        const uint32_t lane_base = tc::tmem_addr(tbase, 32 * warp, 0);
        for (int k0 = 0; k0 < K; k0 += 16) {
            float h[16], l[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                uint32_t hi = 0, lo = 0;
                if (k0 + j < K) split_tf32(A[tid * K + k0 + j], hi, lo);
                h[j] = __uint_as_float(hi); l[j] = __uint_as_float(lo);
            }
            tc::tmem_st16(lane_base + C_AHI + k0, h);
            tc::tmem_st16(lane_base + C_ALO + k0, l);
        }
        tc::tmem_st_wait();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (tid == 0) {
        tc::fence_after_sync();
        const uint32_t idesc = tc::idesc_tf32(128, N);
        const uint32_t whi = tc::smem_u32(Whi), wlo = tc::smem_u32(Wlo), ahi = tc::smem_u32(Ahi), alo = tc::smem_u32(Alo);
        bool acc = false;
        for (int k0 = 0; k0 < K; k0 += 8) {
            const uint64_t bh = tc::kmajor_desc(whi, N, k0), bl = tc::kmajor_desc(wlo, N, k0);
            if (ts) {
                if (x3) {
                    tc::mma_ts(tbase + C_D, tbase + C_ALO + k0, bh, idesc, acc); acc = true;
                    tc::mma_ts(tbase + C_D, tbase + C_AHI + k0, bl, idesc, acc);
                }
                tc::mma_ts(tbase + C_D, tbase + C_AHI + k0, bh, idesc, acc); acc = true;
            } else {
                if (x3) {
                    tc::mma_ss(tbase + C_D, tc::kmajor_desc(alo, 128, k0), bh, idesc, acc); acc = true;
                    tc::mma_ss(tbase + C_D, tc::kmajor_desc(ahi, 128, k0), bl, idesc, acc);
                }
                tc::mma_ss(tbase + C_D, tc::kmajor_desc(ahi, 128, k0), bh, idesc, acc); acc = true;
            }
        }
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    {
        const uint32_t lane_base = tc::tmem_addr(tbase, 32 * warp, C_D);
        for (int n0 = 0; n0 < N; n0 += 16) {
            float v[16];
            tc::tmem_ld16(lane_base + n0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) D[tid * N + n0 + j] = v[j];
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tbase);
}

// Micro-benchmark behind the kernel design: cycles for `count` tf32 MMAs of width N
// (K = 8 each, A from tensor memory) issued back to back by one thread, accumulating into
// `nacc` different accumulators in rotation (nacc = 1: one dependent chain).
// mode 0: A from tensor memory, B K-major (the decoder kernels' layer chain); 1: A and B K-major in shared memory
// (nsb_gemm.cu chains); 2: A and B MN-major (weight gradients); 3: A K-major, B MN-major; 4: A MN-major, B K-major.
__global__ void __launch_bounds__(128, 1) k_tc_mma_latency(long long* out, int N, int count, int nacc, int mode) {
    extern __shared__ __align__(1024) float sm[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
    if (tid == 0) { tc::mbar_init(&bar, 1); tc::mbar_init_fence(); }
    for (int i = tid; i < 3 * 2048; i += 128) sm[i] = 0.f;       // A: 8 KB, B: 16 KB (MN-major atoms are 2 KB apart)
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tbase = tmem_slot;
    {
        float z[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = 0.f;
        tc::tmem_st16(tc::tmem_addr(tbase, 32 * warp, 0), z);
        tc::tmem_st_wait();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0 && tc::elect_one()) {
        tc::fence_after_sync();
        const bool a_mn = mode == 2 || mode == 4, b_mn = mode == 2 || mode == 3;
        const uint32_t idesc = tc::idesc_tf32(128, N) | (a_mn ? tc::IDESC_A_MN : 0u) | (b_mn ? tc::IDESC_B_MN : 0u);
        const uint32_t sa = tc::smem_u32(sm), sb = sa + 8192;
        const uint64_t a = a_mn ? tc::mnmajor_desc(sa) : tc::kmajor_desc(sa, 128, 0);
        const uint64_t b = b_mn ? tc::mnmajor_desc(sb) : tc::kmajor_desc(sb, N, 0);
        const long long t0 = clock64();
        if (mode == 0) {
            for (int i = 0; i < count; ++i)
                tc::mma_ts(tbase + 16 + (uint32_t)((i % nacc) * N), tbase, b, idesc, i >= nacc);
        } else {
            for (int i = 0; i < count; ++i)
                tc::mma_ss(tbase + 16 + (uint32_t)((i % nacc) * N), a, b, idesc, i >= nacc);
        }
        tc::mma_commit(&bar);
        const long long t1 = clock64();
        tc::mbar_wait(&bar, 0);
        const long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    __syncthreads();
    tc::mbar_wait(&bar, 0);
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tbase);
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_tc_mma_latency(long long* out2, int N, int count, int nacc, int mode, void* stream) {
    if (!out2 || N < 16 || N > 256 || N % 16 || nacc < 1 || 16 + nacc * N > 512 || mode < 0 || mode > 4 ||
        (mode >= 2 && N % 32))
        return set_error(NSB_EINVAL, "bad argument");
    k_tc_mma_latency<<<1, 128, 3 * 2048 * 4, (cudaStream_t)stream>>>(out2, N, count, nacc, mode);
    return check_launch("k_tc_mma_latency");
}

extern "C" int nsb_tc_selftest(const float* A, const float* W, float* D, int K, int N, int mode, void* stream) {
    if (!A || !W || !D) return set_error(NSB_EINVAL, "null argument");
    if (K < 8 || K > ST_MAXK || K % 8 || N < 16 || N > ST_MAXN || N % 16) return set_error(NSB_EINVAL, "bad K or N");
    const int smem = (2 * ST_MAXN * ST_MAXK + 2 * 128 * ST_MAXK) * 4;
    cudaError_t e = cudaFuncSetAttribute(k_tc_selftest, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    k_tc_selftest<<<1, 128, smem, (cudaStream_t)stream>>>(A, W, D, K, N, mode);
    return check_launch("k_tc_selftest");
}
