// PCIe copy rates relevant to the host-buffer path (tools only): linear vs 2-D (odd width -> padded pitch) H2D, D2H.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
int main() {
    const int W = 1241, H = 376, B = 512, P = 1280;
    uint8_t *h, *d, *h2;
    cudaMallocHost(&h, (size_t)W * H * B); cudaMalloc(&d, (size_t)P * H * B); cudaMallocHost(&h2, (size_t)64 << 20);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(a, s); cudaMemcpyAsync(d, h, (size_t)W * H * B, cudaMemcpyHostToDevice, s); cudaEventRecord(b, s);
        cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("H2D linear 239MB: %.2f ms  %.1f GB/s\n", ms, W * (double)H * B / ms / 1e6);
        cudaEventRecord(a, s); cudaMemcpy2DAsync(d, P, h, W, W, (size_t)H * B, cudaMemcpyHostToDevice, s); cudaEventRecord(b, s);
        cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("H2D 2D 1241->1280 one call: %.2f ms  %.1f GB/s\n", ms, W * (double)H * B / ms / 1e6);
        cudaEventRecord(a, s);
        for (int k = 0; k < 8; k++) cudaMemcpy2DAsync(d + (size_t)k * 64 * P * H, P, h + (size_t)k * 64 * W * H, W, W, (size_t)H * 64, cudaMemcpyHostToDevice, s);
        cudaEventRecord(b, s); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("H2D 2D 8 chunks: %.2f ms  %.1f GB/s\n", ms, W * (double)H * B / ms / 1e6);
        cudaEventRecord(a, s); cudaMemcpyAsync(h2, d, (size_t)62 << 20, cudaMemcpyDeviceToHost, s); cudaEventRecord(b, s);
        cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("D2H linear 62MiB: %.2f ms  %.1f GB/s\n", ms, (62 << 20) / ms / 1e6);
    }
    return 0;
}
