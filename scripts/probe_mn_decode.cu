// Developer probe: decode how tcgen05 (kind::tf32) addresses an MN-major, no-swizzle A operand.
// B is a K-major selector (known-good layout): g[n][k] = (n == k), so D[m][n] = A(k = n, m).
// The smem image of A holds its own float index, so D shows which smem float the hardware reads for (k, m).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "nsb_tc.cuh"
using namespace nsb;
__host__ __device__ constexpr uint32_t idesc_a_mn(int M, int N) { return tc::idesc_tf32(M, N) | (1u << 15); }

__global__ void k_probe(float* D, uint32_t lbo, uint32_t sbo, int mn, int layout) {
    __shared__ __align__(1024) float sx[8192];
    __shared__ __align__(1024) float sg[16 * 8];
    __shared__ uint32_t slot;
    __shared__ __align__(8) uint64_t bar;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sx[i] = (float)i;
    for (int i = threadIdx.x; i < 128; i += blockDim.x) sg[i] = 0.f;
    __syncthreads();
    if (threadIdx.x < 8) sg[tc::kmajor_index(threadIdx.x, threadIdx.x, 16)] = 1.f;     // n == k, R = 16 rows
    if (warp == 0) tc::tmem_alloc<32>(&slot);
    if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::mbar_init_fence(); }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tm = slot;
    if (warp == 0 && tc::elect_one()) {
        const uint64_t a = mn ? (tc::smem_desc(tc::smem_u32(sx), lbo, sbo) | ((uint64_t)layout << 61)) : tc::kmajor_desc(tc::smem_u32(sx), 128, 0);
        const uint64_t b = tc::kmajor_desc(tc::smem_u32(sg), 16, 0);
        tc::mma_ss(tm, a, b, mn ? idesc_a_mn(128, 16) : tc::idesc_tf32(128, 16), false);
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    if (warp < 4) {
        const int row = 32 * warp + (threadIdx.x & 31);
        float v[8];
        tc::tmem_ld8(tc::tmem_addr(tm, 32 * warp, 0), v);
        tc::tmem_ld_wait();
        for (int j = 0; j < 8; ++j) D[row * 8 + j] = v[j];
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<32>(tm);
}

int main() {
    float* dD;
    cudaMalloc(&dD, 128 * 8 * 4);
    const uint32_t cfg[5][3] = {{0, 0, 0}, {2048, 4096, 1}, {4096, 2048, 1}, {512, 1024, 1}, {1024, 512, 1}};
    for (auto& c : cfg) {
        k_probe<<<1, 128>>>(dD, c[0], c[1], c[0] != 0, (int)c[2]);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
        std::vector<float> D(128 * 8);
        cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
        printf("LBO %u SBO %u: smem float index read for (k, m):\n", c[0], c[1]);
        for (int m : {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 12, 13, 15, 16, 17, 20, 24, 28, 31, 32, 33, 36, 63, 64, 65, 96, 127}) {
            printf("  m=%3d:", m);
            for (int k = 0; k < 8; ++k) printf(" %7.1f", D[m * 8 + k]);
            printf("\n");
        }
    }
    return 0;
}
