// Probe: can SMs pull pinned host memory over PCIe as fast as the copy engine does?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/zc zero_copy_probe.cu && /tmp/zc
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int UNROLL>
__global__ void pull(const uint4 *__restrict__ src, uint4 *__restrict__ dst, uint64_t n16)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < n16; i += UNROLL * stride) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) dst[i + u * stride] = v[u];
    }
    for (; i < n16; i += stride) dst[i] = __ldcs(src + i);
}

int main()
{
    const uint64_t total = 294497180ull / 16 * 16, piece = 1495000 / 16 * 16;
    uint8_t *h, *d;
    CK(cudaHostAlloc((void **)&h, total, cudaHostAllocDefault));
    for (uint64_t i = 0; i < total; i += 4096) h[i] = (uint8_t)i;
    CK(cudaMalloc((void **)&d, total));
    cudaStream_t s[2];
    CK(cudaStreamCreate(&s[0])); CK(cudaStreamCreate(&s[1]));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0, s[0]));
        CK(cudaMemcpyAsync(d, h, total, cudaMemcpyHostToDevice, s[0]));
        CK(cudaEventRecord(e1, s[0])); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("copy engine, one piece: %.3f ms %.1f GB/s\n", ms, total / ms / 1e6);
    }
    const int grids[] = { 4, 8, 16, 32, 64, 148 };
    for (int g : grids) for (int thr : { 256, 1024 }) {
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0, s[0]));
            pull<4><<<g, thr, 0, s[0]>>>((const uint4 *)h, (uint4 *)d, total / 16);
            CK(cudaEventRecord(e1, s[0])); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        }
        printf("SM pull whole buffer, %3d CTAs x %4d thr x4: %.3f ms %.1f GB/s\n", g, thr, ms, total / ms / 1e6);
    }
    // per-split launches, one stream and two alternating streams
    for (int ns = 1; ns <= 2; ++ns) for (int g : { 8, 16, 32 }) {
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, s[0]));
            if (ns == 2) { CK(cudaStreamWaitEvent(s[1], e0, 0)); }
            int k = 0;
            for (uint64_t off = 0; off < total; off += piece, ++k) {
                uint64_t n = total - off < piece ? total - off : piece;
                pull<4><<<g, 512, 0, s[k % ns]>>>((const uint4 *)(h + off), (uint4 *)(d + off), n / 16);
            }
            if (ns == 2) { CK(cudaEventRecord(e1, s[1])); CK(cudaStreamWaitEvent(s[0], e1, 0)); }
            CK(cudaEventRecord(e1, s[0])); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        }
        printf("SM pull per split (%d launches), %d stream(s), %2d CTAs x 512: %.3f ms %.1f GB/s\n",
               (int)((total + piece - 1) / piece), ns, g, ms, total / ms / 1e6);
    }
    CK(cudaMemcpy(h, d, 4096, cudaMemcpyDeviceToHost));
    return 0;
}
