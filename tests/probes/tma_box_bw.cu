// tma_box_bw.cu -- probe: HBM read bandwidth of TMA box shapes over S stream rows (pitch 4 MiB), no compute.
// tensor view: dim0 = 32 samples (128 B), dim1 = time blocks of 128 B, dim2 = stream rows.
// box (32, T, R): one TMA instruction fetches R rows x T*128 contiguous bytes.  Not part of the product.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}

// each warp owns `nslot` slots of slot_bytes; walks its (row group, time segment) issuing one box per slot
__global__ void __launch_bounds__(512, 1) tma_walk(const __grid_constant__ CUtensorMap map, int warps_per_cta, int nslot, int slot_bytes,
                                                   int T, int R, int n_groups, int n_seg, int tb_per_seg, unsigned *sink)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[16 * 16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *sbase = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);
    const uint32_t ring = smem_u32(sbase) + warp * nslot * slot_bytes;
    const uint32_t bar0 = smem_u32(&bars[warp * 16]);
    if (lane == 0) {
        for (int s = 0; s < nslot; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int item = blockIdx.x * warps_per_cta + warp;
    if (item >= n_groups * n_seg) return;
    const int g = item % n_groups, seg = item / n_groups;
    const int n_steps = tb_per_seg / T;
    auto issue = [&](int step, int slot) {
        const uint32_t bar = bar0 + 8 * slot;
        asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(bar), "r"(slot_bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(ring + slot * slot_bytes), "l"(&map), "r"(0), "r"(seg * tb_per_seg + step * T), "r"(g * R), "r"(bar) : "memory");
    };
    if (lane == 0)
        for (int s = 0; s < nslot && s < n_steps; s++) issue(s, s);
    unsigned acc = 0;
    int slot = 0;
    uint32_t phase = 0;
    for (int step = 0; step < n_steps; step++) {
        mbar_wait(bar0 + 8 * slot, phase);
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(ring + slot * slot_bytes + lane * 16));
        acc ^= v;
        __syncwarp();
        if (lane == 0 && step + nslot < n_steps) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(step + nslot, slot);
        }
        if (++slot == nslot) { slot = 0; phase ^= 1; }
    }
    if (acc == 0x12345678u) *sink = acc;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main()
{
    const int S = 1024;
    const size_t row_bytes = 4u << 20;
    void *f = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)f;
    uint8_t *d; cudaMalloc(&d, row_bytes * S); cudaMemset(d, 1, row_bytes * S);
    unsigned *sink; cudaMalloc(&sink, 4);
    struct V { int T, R, warps, nslot; CUtensorMapL2promotion l2; const char *name; };
    const V vs[] = {
        {1, 32, 16, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T1 R32 w16 s2 (v2 now)"},
        {1, 32, 8, 4, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T1 R32 w8 s4"},
        {1, 32, 8, 5, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "T1 R32 w8 s5 l2_128"},
        {1, 32, 8, 5, CU_TENSOR_MAP_L2_PROMOTION_NONE, "T1 R32 w8 s5 l2_none"},
        {2, 32, 8, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R32 w8 s2"},
        {2, 32, 8, 2, CU_TENSOR_MAP_L2_PROMOTION_NONE, "T2 R32 w8 s2 l2_none"},
        {2, 32, 10, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R32 w10 s2"},
        {2, 32, 7, 3, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R32 w7 s3"},
        {4, 32, 5, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R32 w5 s2"},
        {4, 32, 4, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R32 w4 s2"},
        {8, 8, 8, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T8 R8 w8 s2"},
        {8, 8, 16, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T8 R8 w16 s2"},
        {4, 8, 16, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R8 w16 s2"},
        {4, 8, 16, 3, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R8 w16 s3"},
        {2, 8, 16, 4, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R8 w16 s4"},
        {4, 16, 8, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R16 w8 s2"},
        {4, 16, 10, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T4 R16 w10 s2"},
        {2, 16, 16, 2, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R16 w16 s2"},
        {2, 16, 16, 3, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "T2 R16 w16 s3"},
    };
    for (const V &v : vs) {
        CUtensorMap m;
        const cuuint64_t dims[3] = {32, row_bytes / 128, (cuuint64_t)S};
        const cuuint64_t strides[2] = {128, row_bytes};
        const cuuint32_t box[3] = {32, (cuuint32_t)v.T, (cuuint32_t)v.R};
        const cuuint32_t es[3] = {1, 1, 1};
        CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, v.l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", v.name, (int)r); continue; }
        const int slot_bytes = v.T * v.R * 128;
        const int n_groups = S / v.R;
        int n_seg = (148 * v.warps) / n_groups;
        if (n_seg < 1) n_seg = 1;
        const int tb_total = (int)(row_bytes / 128);
        while (tb_total % (n_seg * v.T)) n_seg--;
        const int items = n_groups * n_seg;
        const size_t smem = (size_t)v.warps * v.nslot * slot_bytes + 1024;
        if (smem > 227 * 1024) { printf("%s: smem %zu too large\n", v.name, smem); continue; }
        cudaFuncSetAttribute(tma_walk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float best = 1e9;
        for (int it = 0; it < 4; it++) {
            cudaEventRecord(e0);
            tma_walk<<<(items + v.warps - 1) / v.warps, v.warps * 32, smem>>>(m, v.warps, v.nslot, slot_bytes, v.T, v.R, n_groups, n_seg, tb_total / n_seg, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (it > 0 && ms < best) best = ms;
        }
        printf("%-26s slot %5d B  in-flight/SM %3d KB  items %5d ctas %4d  %.3f ms  %.0f GB/s  (%s)\n", v.name, slot_bytes,
               v.warps * (v.nslot) * slot_bytes / 1024, items, (items + v.warps - 1) / v.warps, best, (double)S * row_bytes / best / 1e6,
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
