// scatter_probe.cu -- what does one small scattered write per session cost on this GPU, whatever kernel issues it?
// The aggregator's append step (agg_kernels.cuh: k_agg_append) reads a push of 2^20 quanta linearly (0.54 GB) and
// writes ~86 bytes into each of 2^20 sessions' staged rows (2.7 KB apart).  This probe does the same traffic with
// the simplest possible kernels, to tell the kernel's share of its time from the memory system's.
//   build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/scatter_probe tools/scatter_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

// one warp per session: `bytes` bytes at byte offset `skew` of row perm(s); optionally reads 512 B of input first
__global__ void k_probe(const uint4 *__restrict__ in, unsigned char *__restrict__ rows, uint32_t n, uint32_t row_bytes,
                        uint32_t bytes, uint32_t skew, int read_in, uint32_t mul)
{
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= n) return;
    uint32_t acc = 0;
    if (read_in) { const uint4 v = in[(size_t)w * 32u + lane]; acc = v.x ^ v.y ^ v.z ^ v.w; }
    const uint32_t s = (uint32_t)(((unsigned long long)w * mul) % n);          // a fixed permutation (mul odd, n a power of two)
    unsigned char *dst = rows + (size_t)s * row_bytes + skew;
    if ((bytes & 3u) == 0u && (skew & 3u) == 0u) {
        for (uint32_t b = 4u * lane; b < bytes; b += 128u) *reinterpret_cast<uint32_t *>(dst + b) = acc + b;
    } else {
        for (uint32_t b = 2u * lane; b < bytes; b += 64u) *reinterpret_cast<uint16_t *>(dst + b) = (uint16_t)(acc + b);
    }
}

int main()
{
    const uint32_t n = 1u << 20, row_bytes = 2752;
    uint4 *in; unsigned char *rows;
    cudaMalloc(&in, (size_t)n * 512); cudaMalloc(&rows, (size_t)n * row_bytes);
    cudaMemset(in, 1, (size_t)n * 512); cudaMemset(rows, 0, (size_t)n * row_bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    struct Case { const char *name; uint32_t bytes, skew; int read_in; uint32_t mul; } cases[] = {
        {"read 512 B + write 86 B at odd offset, scattered rows", 86, 946, 1, 2654435761u},
        {"read 512 B + write 96 B sector-aligned, scattered rows", 96, 960, 1, 2654435761u},
        {"write 86 B at odd offset only, scattered rows", 86, 946, 0, 2654435761u},
        {"write 96 B sector-aligned only, scattered rows", 96, 960, 0, 2654435761u},
        {"read 512 B + write 86 B, rows in push order", 86, 946, 1, 1u},
        {"read 512 B only (bytes = 0)", 0, 0, 1, 1u},
        {"read 512 B + write 512 B aligned, scattered rows (the float ring)", 512, 1024, 1, 2654435761u},
    };
    for (const Case &c : cases) {
        float best = 1e9f;
        for (int rep = 0; rep < 8; rep++) {
            cudaEventRecord(e0);
            k_probe<<<n / 8, 256>>>(in, rows, n, row_bytes, c.bytes, c.skew, c.read_in, c.mul);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep >= 2 && ms < best) best = ms;
        }
        printf("{\"case\": \"%s\", \"sessions\": %u, \"us\": %.1f}\n", c.name, n, best * 1000.f);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
