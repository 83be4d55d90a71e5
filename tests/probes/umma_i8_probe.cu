// umma_i8_probe.cu -- bring-up probe (not part of the product): one CTA, one 128 x N x 128 int8 GEMM through
// tcgen05.mma kind::i8 with operands written to shared memory by ordinary threads in the canonical
// K-major SWIZZLE_128B layout.  Validates: smem descriptor, instruction descriptor, swizzle, TMEM load.
//   D[m][n] = sum_k A[m][k] * B[n][k],  A u8 [128][K], B s8 [N][K], K = 128 * KB
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr)
{
    // K-major, SWIZZLE_128B: rows 128 B apart, 8-row groups SBO = 1024 B, LBO unused (1), version 1
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // start address [0,14)
    d |= (uint64_t)1 << 16;                           // LBO (ignored for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                 // SBO
    d |= (uint64_t)1 << 46;                           // version = 1 (Blackwell)
    d |= (uint64_t)2 << 61;                           // layout type SWIZZLE_128B
    return d;
}

extern "C" __global__ void __launch_bounds__(128) umma_probe(const uint8_t *A, const int8_t *B, int32_t *D, int N, int KB)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    uint8_t *sA = smem;                       // KB tiles of 128 rows x 128 B
    uint8_t *sB = smem + KB * 16384;          // KB tiles of N rows x 128 B (N multiple of 8)
    const int tid = threadIdx.x, warp = tid >> 5;
    const int K = 128 * KB;
    // stage A and B with the swizzle: byte (row r, col c) of K-block kb -> tile + (r/8)*1024 + (r%8)*128 + (((c/16) ^ (r%8))*16) + c%16
    for (int i = tid; i < 128 * K / 4; i += 128) {
        const int r = i / (K / 4), c4 = (i % (K / 4)) * 4, kb = c4 / 128, c = c4 % 128;
        const uint32_t v = *reinterpret_cast<const uint32_t *>(A + (size_t)r * K + c4);
        *reinterpret_cast<uint32_t *>(sA + kb * 16384 + (r / 8) * 1024 + (r % 8) * 128 + (((c / 16) ^ (r % 8)) * 16) + c % 16) = v;
    }
    for (int i = tid; i < N * K / 4; i += 128) {
        const int r = i / (K / 4), c4 = (i % (K / 4)) * 4, kb = c4 / 128, c = c4 % 128;
        const uint32_t v = *reinterpret_cast<const uint32_t *>(B + (size_t)r * K + c4);
        *reinterpret_cast<uint32_t *>(sB + kb * (N * 128) + (r / 8) * 1024 + (r % 8) * 128 + (((c / 16) ^ (r % 8)) * 16) + c % 16) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "n"(256));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tm = tmem_base;
    if (tid == 0) {
        // instruction descriptor: c=S32 (2<<4), a=U8 (0<<7), b=S8 (1<<10), K-major both, N>>3 at [17,23), M>>4 at [24,29)
        const uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        for (int kb = 0; kb < KB; kb++)
            for (int ks = 0; ks < 4; ks++) {
                const uint64_t da = make_desc_sw128(smem_u32(sA + kb * 16384) + ks * 32);
                const uint64_t db = make_desc_sw128(smem_u32(sB + kb * (N * 128)) + ks * 32);
                const uint32_t acc = (kb | ks) ? 1u : 0u;
                asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                             ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc));
            }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
    }
    // everyone waits for the MMA to complete
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0));
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    // epilogue: warp w reads TMEM lanes 32w..32w+31, 8 columns at a time
    for (int c0 = 0; c0 < N; c0 += 8) {
        uint32_t v[8];
        const uint32_t addr = tm + ((uint32_t)(warp * 32) << 16) + c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        for (int i = 0; i < 8; i++) D[(size_t)tid * N + c0 + i] = (int32_t)v[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "n"(256));
}

int main()
{
    const int N = 48, KB = 4, K = 128 * KB;
    uint8_t *hA = new uint8_t[128 * K];
    int8_t *hB = new int8_t[N * K];
    int32_t *hD = new int32_t[128 * N], *ref = new int32_t[128 * N];
    uint32_t seed = 12345;
    auto rnd = [&]() { seed = seed * 1664525u + 1013904223u; return seed >> 24; };
    for (int i = 0; i < 128 * K; i++) hA[i] = (uint8_t)rnd();
    for (int i = 0; i < N * K; i++) hB[i] = (int8_t)rnd();
    for (int m = 0; m < 128; m++)
        for (int n = 0; n < N; n++) {
            int32_t s = 0;
            for (int k = 0; k < K; k++) s += (int32_t)hA[m * K + k] * (int32_t)hB[n * K + k];
            ref[m * N + n] = s;
        }
    uint8_t *dA; int8_t *dB; int32_t *dD;
    cudaMalloc(&dA, 128 * K); cudaMalloc(&dB, N * K); cudaMalloc(&dD, 128 * N * 4);
    cudaMemcpy(dA, hA, 128 * K, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, N * K, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, 128 * N * 4);
    const int smem = KB * 16384 + KB * N * 128 + 1024;
    cudaFuncSetAttribute(umma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    umma_probe<<<1, 128, smem>>>(dA, dB, dD, N, KB);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    cudaMemcpy(hD, dD, 128 * N * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < 128 * N; i++) if (hD[i] != ref[i]) { if (bad < 10) printf("mismatch at m=%d n=%d: got %d want %d\n", i / N, i % N, hD[i], ref[i]); bad++; }
    printf("mismatches: %d of %d\n", bad, 128 * N);
    return bad != 0;
}
