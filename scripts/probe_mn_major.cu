// Developer probe (not part of the library): tcgen05.mma kind::tf32 with MN-major shared-memory operands.
//   D[m][n] = sum_k X[k][m] G[k][n],  k = 16 points, m < 128, n < N
// X and G are stored the way the weight-gradient records store them: per 32 channels an atom [16 points][32 channels]
// (128 B per point) whose 32-byte chunks are XOR-swizzled with (point & 3) -- the only layout tcgen05 accepts for
// MN-major tf32 operands (SWIZZLE_128B_BASE32B; measured with probe_mn_decode.cu): SBO (next 4 points) = 512 B,
// LBO (next 32 channels) = 2048 B, one MMA (K = 8 points) starts 1024 B after the previous one.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I nice_slam_annotation_b200/csrc scripts/probe_mn_major.cu -o build_prof/probe_mn
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>
#include "nsb_tc.cuh"

using namespace nsb;

__host__ __device__ constexpr uint32_t idesc_mn(int M, int N) { return tc::idesc_tf32(M, N) | (1u << 15) | (1u << 16); }

template <int N>
__global__ void k_probe(const float* X, const float* G, float* D) {
    __shared__ __align__(1024) float sx[128 * 16];
    __shared__ __align__(1024) float sg[256 * 16];
    __shared__ uint32_t slot;
    __shared__ __align__(8) uint64_t bar;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 128 * 16; i += blockDim.x) sx[i] = X[i];
    for (int i = threadIdx.x; i < N * 16; i += blockDim.x) sg[i] = G[i];
    if (warp == 0) tc::tmem_alloc<256>(&slot);
    if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::mbar_init_fence(); }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tm = slot;
    if (warp == 0 && tc::elect_one()) {
        for (int ks = 0; ks < 2; ++ks) {
            const uint64_t a = tc::smem_desc(tc::smem_u32(sx) + ks * 1024, 2048, 512) | (1ull << 61);
            const uint64_t b = tc::smem_desc(tc::smem_u32(sg) + ks * 1024, 2048, 512) | (1ull << 61);
            tc::mma_ss(tm, a, b, idesc_mn(128, N), ks > 0);
        }
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    if (warp < 4) {
        const int row = 32 * warp + (threadIdx.x & 31);
        for (int c0 = 0; c0 < N; c0 += 8) {
            float v[8];
            tc::tmem_ld8(tc::tmem_addr(tm, 32 * warp, c0), v);
            tc::tmem_ld_wait();
            for (int j = 0; j < 8; ++j) D[row * N + c0 + j] = v[j];
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<256>(tm);
}

static float tf32_trunc(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

template <int N>
static int run() {
    std::vector<float> X(128 * 16), G(N * 16), D(128 * N), x(16 * 128), g(16 * N);
    srand(1 + N);
    for (auto& v : x) v = (float)(rand() % 2001 - 1000) / 64.f;      // exactly representable in tf32? no: up to 11 bits + sign
    for (auto& v : g) v = (float)(rand() % 2001 - 1000) / 64.f;
    for (auto& v : x) v = tf32_trunc(v);
    for (auto& v : g) v = tf32_trunc(v);
    for (int k = 0; k < 16; ++k) {
        for (int m = 0; m < 128; ++m) X[(m / 32) * 512 + k * 32 + ((m % 32) ^ ((k & 3) << 3))] = x[k * 128 + m];
        for (int n = 0; n < N; ++n) G[(n / 32) * 512 + k * 32 + ((n % 32) ^ ((k & 3) << 3))] = g[k * N + n];
    }
    float *dX, *dG, *dD;
    cudaMalloc(&dX, X.size() * 4); cudaMalloc(&dG, G.size() * 4); cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dG, G.data(), G.size() * 4, cudaMemcpyHostToDevice);
    k_probe<N><<<1, 128>>>(dX, dG, dD);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d: %s\n", N, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double worst = 0, scale = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < 16; ++k) s += (double)x[k * 128 + m] * g[k * N + n];
            worst = fmax(worst, fabs(s - D[m * N + n]));
            scale = fmax(scale, fabs(s));
        }
    printf("N=%d: max |err| %.3g of max |ref| %.3g -> %s\n", N, worst, scale, worst <= 1e-5 * scale ? "OK" : "MISMATCH");
    return worst <= 1e-5 * scale ? 0 : 1;
}

int main() {
    int bad = 0;
    bad += run<32>();
    bad += run<128>();
    bad += run<224>();
    return bad;
}
