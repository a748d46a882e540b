// Validation of the tcgen05.mma.kind::i8 building blocks the base-conversion contraction needs, in isolation:
// one CTA, D[128 x 256] (s32, TMEM) = A[128 x 128] (u8, K-major, no swizzle) x B[256 x 128]^T (u8, K-major, no swizzle), four
// K = 32 instructions, completion through tcgen05.commit -> mbarrier, accumulators read back with tcgen05.ld.32x32b.x16 by four
// warps, compared with the host.  Layout of both operands in shared memory (canonical K-major INTERLEAVE, units of 16 bytes:
// ((8, n), 2) : ((1, SBO), LBO)):   byte (row r, k) at  (k / 16) * (rows * 16) + r * 16 + k % 16 ,  SBO = 128 B, LBO = rows * 16 B.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_i8_test umma_i8_test.cu ; run: ./umma_i8_test
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 256, K = 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__host__ __device__ constexpr uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32) |
           ((uint64_t)1 << 46);                       // version = 1 (Blackwell), base_offset 0, layout_type 0 = SWIZZLE_NONE
}
// instruction descriptor: c_format S32 (2) bits [4,6), a/b_format UINT8 (0), K-major both, n_dim = N >> 3 bits [17,23), m_dim = M >> 4 bits [24,29)
constexpr uint32_t kIdesc = (2u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);

__global__ void __launch_bounds__(128) k_test(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int32_t* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t sm[];
    uint8_t* sA = sm;                       // 128 x 128 = 16 KB
    uint8_t* sB = sm + M * K;               // 256 x 128 = 32 KB
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < M * K; i += 128) { const int r = i / K, k = i % K; sA[(k / 16) * (M * 16) + r * 16 + k % 16] = A[i]; }
    for (int i = tid; i < N * K; i += 128) { const int r = i / K, k = i % K; sB[(k / 16) * (N * 16) + r * 16 + k % 16] = B[i]; }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy smem writes -> visible to the tensor core (async proxy)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    if (tid == 0) {
        for (int s = 0; s < K / 32; s++) {
            const uint64_t ad = make_desc(smem_u32(sA) + s * 2 * (M * 16), M * 16, 128);
            const uint64_t bd = make_desc(smem_u32(sB) + s * 2 * (N * 16), N * 16, 128);
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tm), "l"(ad), "l"(bd), "r"(kIdesc), "r"(s > 0 ? 1 : 0)
                         : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // everyone waits for the accumulators
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(&bar))
                 : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        const uint32_t ta = tm + ((uint32_t)(warp * 32) << 16) + c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(ta));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 16; j++) D[tid * N + c0 + j] = (int32_t)v[j];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tm) : "memory");
}

int main() {
    std::vector<uint8_t> A(M * K), B(N * K);
    srand(1);
    for (auto& x : A) x = rand() & 255;
    for (auto& x : B) x = rand() & 255;
    uint8_t *dA, *dB;
    int32_t* dD;
    cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, M * N * 4);
    const int smem = M * K + N * K;
    cudaFuncSetAttribute(k_test, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k_test<<<1, 128, smem>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<int32_t> D(M * N);
    cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int i = 0; i < M; i++)
        for (int n = 0; n < N; n++) {
            int32_t s = 0;
            for (int k = 0; k < K; k++) s += (int32_t)A[i * K + k] * (int32_t)B[n * K + k];
            if (s != D[i * N + n]) { if (bad < 5) printf("mismatch D[%d][%d] = %d, want %d\n", i, n, D[i * N + n], s); bad++; }
        }
    printf("tcgen05.mma.kind::i8 u8 x u8 -> s32, M=128 N=256 K=128: %ld mismatches of %d\n", bad, M * N);
    return bad ? 2 : 0;
}
