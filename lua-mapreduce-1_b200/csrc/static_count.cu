// static_count.cu -- mr_count_tokens_static: token counting for callers that must not
// allocate device memory.
//
// Who needs this: a host that has already handed every byte of device memory to its own
// pool (so cudaMalloc would fail), or one that sits in the middle of a latency-critical
// pipeline and cannot afford the device-wide synchronisation that cudaMalloc / cudaFree
// imply.  The context-based API (mr_ctx_create ...) allocates its arena, hash table and
// sort buffers; this entry point allocates nothing at all: the caller owns the input and
// the output buffers, and the one piece of scratch the kernel needs -- the hash table the
// tokens are combined in -- lives in the fixed array below, which every GPU that loads this
// library carries until the process exits (4 MiB + a small control block).
//
// It is a variant of the tokenise + hash-combine kernel (map_kernel.cuh): same whitespace
// rule, same 128-bit inline keys, same 128-bit compare-and-swap insertion (common.cuh), but
// a plain one-thread-per-token-start loop instead of the staged warp pipeline, because the
// inputs it accepts are small.
//
// Limits, all checked: input at most MR_STATIC_MAX_BYTES (so the table can never fill: it
// has twice as many slots as the input can have tokens), tokens at most 15 bytes (the inline
// key), out_capacity at least the number of distinct tokens.
//
// Sharing: all calls on one device use the same array and it keeps whatever the previous
// call left, so every call clears it first; two calls whose GPU work overlaps (different
// streams, or different host threads) race on it and must be serialised by the caller.
// The address of the array is looked up once per device, on the first call made with that
// device current; make that first call outside any stream capture.
#include "mrgpu.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "common.cuh"

using namespace mr;

namespace {

constexpr uint32_t STATIC_SLOTS = 1u << 17;                   // 32-byte slots: 4 MiB
constexpr uint32_t STATIC_CTL_WORDS = 32;                     // control block after the table
constexpr size_t STATIC_WORDS = (size_t)STATIC_SLOTS * 4 + STATIC_CTL_WORDS;
static_assert(MR_STATIC_MAX_BYTES <= STATIC_SLOTS, "an input of N bytes has at most N/2+1 tokens; keep load <= 1/2");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress,
// never an address to hand to a kernel.
static __device__ __align__(256) unsigned long long g_static_scratch[STATIC_WORDS];

struct StaticCtl {          // lives in the control block
    unsigned int err;
    unsigned int pad;
    unsigned long long tokens;
};

constexpr int MAX_DEVICES = 64;
void *g_scratch_ptr[MAX_DEVICES];
std::mutex g_scratch_mutex;

__device__ __forceinline__ bool is_space_byte(uint32_t c) { return c == 0x20u || (c - 9u) < 5u; }

__global__ void __launch_bounds__(256) count_tokens_static_kernel(const uint8_t *text, uint64_t n, GEntry *table,
                                                                  StaticCtl *ctl)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long mine = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (is_space_byte(text[i])) continue;
        if (i > 0 && !is_space_byte(text[i - 1])) continue; // not the first byte of its token
        Key key;
        key.k0 = 0; key.k1 = 0;
        uint32_t len = 0;
        while (i + len < n && len <= INLINE_MAX) {
            const uint32_t c = text[i + len];
            if (is_space_byte(c)) break;
            if (len < 8) key.k0 |= (unsigned long long)c << (8 * len);
            else if (len < INLINE_MAX) key.k1 |= (unsigned long long)c << (8 * (len - 8));
            ++len;
        }
        if (len > INLINE_MAX) { atomicOr(&ctl->err, ERR_TOKEN_TOO_LONG); continue; }
        key.k1 |= (unsigned long long)len << 56;
        gtable_add(table, STATIC_SLOTS - 1, text, key, 1, &ctl->err);
        ++mine;
    }
    if (mine) atomicAdd(&ctl->tokens, mine);
}

// occupied slots -> caller's records {key[15], len, count}, in table order
__global__ void __launch_bounds__(256) emit_static_kernel(const GEntry *table, unsigned long long *out,
                                                          unsigned long long cap, unsigned long long *n_out,
                                                          StaticCtl *ctl)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= STATIC_SLOTS) return;
    const GEntry e = table[slot];
    if (e.k1 == 0) return;
    const unsigned long long idx = atomicAdd(n_out, 1ull);
    if (idx >= cap) { atomicOr(&ctl->err, ERR_KEYS_FULL); return; }
    out[3 * idx] = e.k0;       // key bytes 0..7
    out[3 * idx + 1] = e.k1;   // key bytes 8..14, then the length byte
    out[3 * idx + 2] = e.cnt;
}

int fail_static(char *err_text, uint64_t err_cap, int code, const char *msg)
{
    if (err_text && err_cap) snprintf(err_text, (size_t)err_cap, "%s", msg);
    return code;
}

} // namespace

int mr_count_tokens_static(const void *dev_text, uint64_t n_bytes, mr_token_count *dev_out, uint64_t out_capacity,
                           uint64_t *dev_n_out, uint64_t *n_tokens, void *stream, char *err_text, uint64_t err_cap)
{
    static_assert(sizeof(mr_token_count) == 24, "record layout");
    char msg[256];
    if (!dev_n_out || (n_bytes && !dev_text) || (out_capacity && !dev_out))
        return fail_static(err_text, err_cap, MR_ERR_INVALID, "mr_count_tokens_static: null pointer");
    if (n_bytes > MR_STATIC_MAX_BYTES) {
        snprintf(msg, sizeof msg, "mr_count_tokens_static: input of %llu bytes exceeds MR_STATIC_MAX_BYTES (%u); "
                 "use the context API for larger inputs", (unsigned long long)n_bytes, (unsigned)MR_STATIC_MAX_BYTES);
        return fail_static(err_text, err_cap, MR_ERR_INVALID, msg);
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        snprintf(msg, sizeof msg, "mr_count_tokens_static: no usable current device (%s)", cudaGetErrorString(e));
        return fail_static(err_text, err_cap, MR_ERR_CUDA, msg);
    }
    void *scratch = nullptr;
    {
        // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lock(g_scratch_mutex);
        if (!g_scratch_ptr[dev]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || !p) {
                snprintf(msg, sizeof msg, "mr_count_tokens_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
                return fail_static(err_text, err_cap, MR_ERR_CUDA, msg);
            }
            g_scratch_ptr[dev] = p;
        }
        scratch = g_scratch_ptr[dev];
    }
    GEntry *table = (GEntry *)scratch;
    StaticCtl *ctl = (StaticCtl *)((unsigned long long *)scratch + (size_t)STATIC_SLOTS * 4);
    cudaStream_t s = (cudaStream_t)stream;
    // the array still holds the previous call's table and flags
    e = cudaMemsetAsync(scratch, 0, STATIC_WORDS * sizeof(unsigned long long), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(dev_n_out, 0, sizeof(uint64_t), s);
    if (e == cudaSuccess && n_bytes) {
        const unsigned grid = (unsigned)((n_bytes + 255) / 256);
        count_tokens_static_kernel<<<grid, 256, 0, s>>>((const uint8_t *)dev_text, n_bytes, table, ctl);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) {
        emit_static_kernel<<<STATIC_SLOTS / 256, 256, 0, s>>>(table, (unsigned long long *)dev_out, out_capacity,
                                                             (unsigned long long *)dev_n_out, ctl);
        e = cudaGetLastError();
    }
    StaticCtl h;
    h.err = 0; h.pad = 0; h.tokens = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, ctl, sizeof h, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "mr_count_tokens_static: CUDA error: %s", cudaGetErrorString(e));
        return fail_static(err_text, err_cap, MR_ERR_CUDA, msg);
    }
    if (n_tokens) *n_tokens = h.tokens;
    if (h.err & ERR_TOKEN_TOO_LONG)
        return fail_static(err_text, err_cap, MR_ERR_OVERFLOW,
                           "mr_count_tokens_static: a token is longer than 15 bytes; use the context API");
    if (h.err & ERR_KEYS_FULL)
        return fail_static(err_text, err_cap, MR_ERR_NOMEM,
                           "mr_count_tokens_static: more distinct tokens than out_capacity");
    if (h.err & ERR_TABLE_FULL)
        return fail_static(err_text, err_cap, MR_ERR_OVERFLOW, "mr_count_tokens_static: static table full");
    return MR_OK;
}
