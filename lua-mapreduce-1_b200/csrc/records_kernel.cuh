// records_kernel.cuh -- map + combine for jobs whose map output is already (number key, number
// value) records: group-by-sum on uint32 keys (BASELINE config 5).
//
// In the reference such a job's mapfn emits (key, value) numbers, emit groups them per key
// (job.lua:83-97) and the combiner / reducer sums them (reducefn.lua:1-5).  Here one thread
// takes one record: heavy-hitter keys are summed in the CTA's shared-memory table (4-way
// buckets, 64-bit sums), everything else goes to the global table with the same 128-bit
// compare-and-swap insertion as the text path.  A key travels as a 4-byte inline key holding
// its big-endian bytes, so the byte-wise key order of the reduce side IS numeric order
// (Lua `<` on numbers, utils.lua:214) and nothing downstream needs to know about numbers.
#pragma once
#include "common.cuh"

namespace mr {

struct RecordsParams {
    const uint32_t *keys, *vals;
    uint64_t n;
    GEntry *gtab;
    uint32_t gmask;
    unsigned long long *records; // running record count
    unsigned int *err;
    uint32_t op;                 // ReduceOp
};

__device__ __forceinline__ Key key_of_u32(uint32_t k)
{
    Key key;
    key.k0 = __byte_perm(k, 0, 0x0123); // big-endian bytes: byte 0 = most significant
    key.k1 = 4ull << 56;
    return key;
}

__device__ __forceinline__ uint32_t u32_of_key(const Key &key) { return __byte_perm((uint32_t)key.k0, 0, 0x0123); }

constexpr unsigned long long REC_TAG = 1ull << 40; // marks an occupied shared-table slot (key 0 is valid)

template <int LOG2_SLOTS>
__global__ void __launch_bounds__(1024, 1) map_records_kernel(RecordsParams p)
{
    constexpr uint32_t SLOTS = 1u << LOG2_SLOTS;
    extern __shared__ __align__(16) uint8_t smem[];
    unsigned long long *tkeys = (unsigned long long *)smem;
    unsigned long long *tsum = tkeys + SLOTS;
    for (uint32_t i = threadIdx.x; i < SLOTS; i += blockDim.x) { tkeys[i] = 0; tsum[i] = 0; }
    __syncthreads();
    unsigned long long mine = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += stride) {
        const uint32_t k = __ldg(p.keys + i), v = __ldg(p.vals + i);
        ++mine;
        const unsigned long long sk = (unsigned long long)k | REC_TAG;
        const uint32_t bucket = (k * 0x9E3779B1u) >> (32 - (LOG2_SLOTS - 2));
        unsigned long long *bk = tkeys + bucket * 4;
        int way = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned long long cur = *(volatile unsigned long long *)&bk[q];
            if (cur == 0) cur = atomicCAS(&bk[q], 0ull, sk) == 0 ? sk : *(volatile unsigned long long *)&bk[q];
            if (cur == sk) { way = q; break; }
        }
        const unsigned long long x = op_encode(p.op, v);
        if (way >= 0) op_apply(&tsum[bucket * 4 + way], x, p.op);
        else gtable_add(p.gtab, p.gmask, nullptr, key_of_u32(k), x, p.err, p.op);
    }
    for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(p.records, mine);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < SLOTS; i += blockDim.x) {
        const unsigned long long sk = tkeys[i];
        if (sk) gtable_add(p.gtab, p.gmask, nullptr, key_of_u32((uint32_t)sk), tsum[i], p.err, p.op);
    }
}

} // namespace mr
