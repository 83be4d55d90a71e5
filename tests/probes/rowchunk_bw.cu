// rowchunk_bw.cu -- probe: HBM read bandwidth when each warp walks R stream rows in pieces of B contiguous
// bytes (the access pattern of the fused bank kernel: lane = stream).  Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o rowchunk_bw rowchunk_bw.cu && ./rowchunk_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

// grid: one warp per (row group, time segment).  Each step the warp reads R rows x B bytes = R*B bytes with
// 16-byte lanes; U steps are kept in flight.
template <int R, int B>
__global__ void __launch_bounds__(512, 1) walk(const uint4 *base, size_t pitch_b, int n_groups, int n_seg, size_t row_bytes, unsigned *sink)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_groups * n_seg) return;
    const int g = warp % n_groups, seg = warp / n_groups;
    const size_t seg_bytes = row_bytes / n_seg;
    constexpr int LPR = B / 16;                 // lanes per row piece
    constexpr int RPI = 32 / LPR > 0 ? 32 / LPR : 1;  // rows per load instruction (B <= 512)
    constexpr int IPS = (R * B) / 512;          // load instructions per step
    unsigned acc = 0;
    for (size_t off = 0; off < seg_bytes; off += B) {
        uint4 v[IPS];
#pragma unroll
        for (int i = 0; i < IPS; i++) {
            int row, col;
            if (B <= 512) { row = i * RPI + lane / LPR; col = (lane % LPR) * 16; }
            else { row = i / (B / 512); col = (i % (B / 512)) * 512 + lane * 16; }
            const char *p = (const char *)base + (size_t)(g * R + row) * pitch_b + seg * seg_bytes + off + col;
            v[i] = __ldg((const uint4 *)p);
        }
#pragma unroll
        for (int i = 0; i < IPS; i++) acc += v[i].x ^ v[i].y ^ v[i].z ^ v[i].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int R, int B>
void run(const uint4 *d, size_t pitch_b, int rows, size_t row_bytes, unsigned *sink, const char *tag)
{
    int sms = 148, wps = 16;
    int n_groups = rows / R;
    int n_seg = (sms * wps) / n_groups;
    if (n_seg < 1) n_seg = 1;
    while (row_bytes % ((size_t)n_seg * B)) n_seg--;
    int warps = n_groups * n_seg;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    for (int it = 0; it < 5; it++) {
        cudaEventRecord(e0);
        walk<R, B><<<(warps * 32 + 511) / 512, 512>>>(d, pitch_b, n_groups, n_seg, row_bytes, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it > 0 && ms < best) best = ms;
    }
    printf("%-22s R=%2d B=%4d pitch=%zu warps=%d  %.3f ms  %.0f GB/s\n", tag, R, B, pitch_b, warps, best, (double)rows * row_bytes / best / 1e6);
}

int main()
{
    const int rows = 1024;
    const size_t row_bytes = 4u << 20;
    unsigned *sink; cudaMalloc(&sink, 4);
    for (size_t pad : {(size_t)0, (size_t)2048, (size_t)(1 << 20) + 4096}) {
        size_t pitch = row_bytes + pad;
        uint4 *d; cudaMalloc(&d, pitch * rows); cudaMemset(d, 1, pitch * rows);
        run<32, 128>(d, pitch, rows, row_bytes, sink, "lane=stream 128B");
        run<32, 256>(d, pitch, rows, row_bytes, sink, "32 rows x 256B");
        run<32, 512>(d, pitch, rows, row_bytes, sink, "32 rows x 512B");
        run<16, 256>(d, pitch, rows, row_bytes, sink, "16 rows x 256B");
        run<8, 512>(d, pitch, rows, row_bytes, sink, "8 rows x 512B");
        run<8, 1024>(d, pitch, rows, row_bytes, sink, "8 rows x 1KB");
        run<4, 1024>(d, pitch, rows, row_bytes, sink, "4 rows x 1KB");
        run<4, 2048>(d, pitch, rows, row_bytes, sink, "4 rows x 2KB");
        run<1, 4096>(d, pitch, rows, row_bytes, sink, "1 row x 4KB");
        cudaFree(d);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
