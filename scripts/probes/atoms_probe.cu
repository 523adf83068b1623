// atoms_probe.cu -- what one SM can do per cycle of the operations the map kernel is made of
// (shared-memory atomics, shared loads, global probes + reductions), measured the way the kernel
// runs them: 148 CTAs x 1024 threads, every lane its own pseudo-random address.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o atoms_probe atoms_probe.cu && ./atoms_probe
// Prints cycles per warp-instruction per SM (SM-wide: elapsed cycles x 1 / warp-instructions issued
// on that SM) and the same per lane.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t h)
{
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
    return h;
}

// index distributions over `slots` (power of two): 0 uniform, 1 Zipf-like head (half of the draws fall
// on 16 slots, a quarter on the next 256), 2 all lanes the same slot
__device__ __forceinline__ uint32_t draw(uint32_t r, uint32_t slots, int dist)
{
    if (dist == 2) return 5;
    if (dist == 1) {
        const uint32_t c = r >> 30;
        if (c >= 2) return (r >> 8) & 15u;
        if (c == 1) return 16u + ((r >> 8) & 255u);
    }
    return (r >> 4) & (slots - 1);
}

constexpr int SLOTS = 4096;
constexpr int ITERS = 2048;

template <int MODE>
__global__ void __launch_bounds__(1024, 1) probe(unsigned long long *gtab, uint32_t gmask, int dist, unsigned long long *out,
                                                 long long *cycles)
{
    __shared__ unsigned int cnt[SLOTS];
    __shared__ __align__(16) unsigned long long keys[SLOTS];
    for (int i = threadIdx.x; i < SLOTS; i += blockDim.x) { cnt[i] = 0; keys[i] = mix(i) | 1ull << 40; }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    uint32_t r = mix(blockIdx.x * 1024 + threadIdx.x + 12345);
    unsigned long long acc = 0;
    const long long t0 = clock64();
#pragma unroll 4
    for (int it = 0; it < ITERS; ++it) {
        r = r * 1664525u + 1013904223u;
        const uint32_t s = draw(mix(r), SLOTS, dist);
        if (MODE == 0) {                       // atomicAdd(.., 1): ATOMS.POPC.INC
            atomicAdd(&cnt[s], 1u);
        } else if (MODE == 1) {                // atomicAdd(.., v): ATOMS.ADD
            atomicAdd(&cnt[s], 2u + (r & 1u));
        } else if (MODE == 2) {                // match_any on the slot, leader adds the popcount
            const uint32_t m = __match_any_sync(0xFFFFFFFFu, s);
            if ((uint32_t)(__ffs((int)m) - 1) == lane) atomicAdd(&cnt[s], (unsigned)__popc(m));
        } else if (MODE == 3) {                // 16-byte shared load of a bucket (two keys)
            const ulonglong2 q = *(const ulonglong2 *)&keys[s & ~1u];
            acc += q.x ^ q.y;
        } else if (MODE == 4) {                // 4-byte shared load
            acc += cnt[s];
        } else if (MODE == 5) {                // global: 64-bit reduction into a 32-byte slot, spread over the table
            const uint32_t g = draw(mix(r), gmask + 1, dist == 1 ? 1 : 0);
            atomicAdd(&gtab[(size_t)g * 4 + 2], 1ull);
        } else if (MODE == 6) {                // global: 16-byte probe load + reduction (the insertion path)
            const uint32_t g = draw(mix(r), gmask + 1, 0);
            const ulonglong2 v = __ldcg((const ulonglong2 *)&gtab[(size_t)g * 4]);
            if (v.y != 0x1234567ull) atomicAdd(&gtab[(size_t)g * 4 + 2], 1ull);
            acc += v.x;
        } else if (MODE == 7) {                // shared load + compare + POPC.INC: the CTA-table hit path
            const ulonglong2 q = *(const ulonglong2 *)&keys[s & ~1u];
            const unsigned long long want = mix(s & ~1u) | 1ull << 40;
            const int way = q.x == want ? 0 : 1;
            atomicAdd(&cnt[(s & ~1u) + way], 1u);
        } else if (MODE == 8) {                // plain (racy) read-modify-write: what a non-atomic update would cost
            cnt[s] = cnt[s] + 1;
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    if (acc == 0x123456789abcdefull) out[0] = acc;
    if (MODE <= 2 || MODE >= 7) { __syncthreads(); if (threadIdx.x < 32) atomicAdd(&out[1], (unsigned long long)cnt[threadIdx.x]); }
}

template <int MODE>
void run(const char *name, int dist, unsigned long long *gtab, uint32_t gmask, unsigned long long *out, long long *cyc, int sms)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    probe<MODE><<<sms, 1024>>>(gtab, gmask, dist, out, cyc);
    CK(cudaEventRecord(a));
    probe<MODE><<<sms, 1024>>>(gtab, gmask, dist, out, cyc);
    CK(cudaEventRecord(b));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    long long *h = (long long *)malloc(sizeof(long long) * sms);
    CK(cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < sms; ++i) avg += (double)h[i];
    avg /= sms;
    const double winst = 32.0 * ITERS;         // warp-instructions of the probed kind per SM
    static const char *dn[] = { "uniform", "zipf-head", "one-slot" };
    printf("%-44s %-9s  %8.3f ms  %8.2f cyc/warp-op/SM  %6.3f cyc/lane-op/SM\n", name, dn[dist], ms, avg / winst, avg / winst / 32.0);
    free(h);
}

int main()
{
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    const uint32_t gslots = 1u << 21;          // 64 MB of 32-byte slots, like the job's table
    unsigned long long *gtab, *out;
    long long *cyc;
    CK(cudaMalloc(&gtab, (size_t)gslots * 32));
    CK(cudaMemset(gtab, 0, (size_t)gslots * 32));
    CK(cudaMalloc(&out, 64));
    CK(cudaMemset(out, 0, 64));
    CK(cudaMalloc(&cyc, sizeof(long long) * sms));
    printf("%s, %d SMs, %d iterations x 32 warps per SM\n", p.name, sms, ITERS);
    for (int d = 0; d < 3; ++d) {
        run<0>("shared atomicAdd(+1)  [ATOMS.POPC.INC]", d, gtab, gslots - 1, out, cyc, sms);
        run<1>("shared atomicAdd(+v)  [ATOMS.ADD]", d, gtab, gslots - 1, out, cyc, sms);
        run<2>("match_any + leader atomicAdd(popc)", d, gtab, gslots - 1, out, cyc, sms);
        run<7>("LDS.128 + compare + atomicAdd(+1)", d, gtab, gslots - 1, out, cyc, sms);
        run<8>("plain shared read-modify-write (racy)", d, gtab, gslots - 1, out, cyc, sms);
    }
    run<3>("LDS.128 of a random bucket", 0, gtab, gslots - 1, out, cyc, sms);
    run<4>("LDS.32 of a random word", 0, gtab, gslots - 1, out, cyc, sms);
    run<5>("global RED.64 into a random 32-byte slot", 0, gtab, gslots - 1, out, cyc, sms);
    run<5>("global RED.64 into a random 32-byte slot", 1, gtab, gslots - 1, out, cyc, sms);
    run<6>("global LDG.128 probe + RED.64", 0, gtab, gslots - 1, out, cyc, sms);
    return 0;
}
