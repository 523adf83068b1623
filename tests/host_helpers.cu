// host_helpers.cu -- the device helpers of csrc/common.cuh compiled FOR THE HOST (they are
// __host__ __device__), behind a tiny C interface for the CPU test suite: the partitioner
// (FnvLua52 / partition_of_key), the sort words (sort_chunk) and the owner map (owner_of) are the
// very source the kernels run.  Test infrastructure; built by tests/test_device_helpers_on_host.py.
#include <cstdint>
#include <cstring>
#include <vector>

#include "common.cuh"

using namespace mr;

namespace {

// a key as the map kernels build it: inline up to 15 bytes, else (offset, length) into a byte space
struct HostKey {
    Key key;
    std::vector<uint8_t> space; // byte space of a long key, with the slack load8 may read
};

HostKey make_key(const uint8_t *bytes, uint64_t len)
{
    HostKey h;
    h.key.k0 = 0;
    h.key.k1 = 0;
    if (len <= (uint64_t)INLINE_MAX) {
        for (uint64_t i = 0; i < len && i < 8; ++i) h.key.k0 |= (unsigned long long)bytes[i] << (8 * i);
        for (uint64_t i = 8; i < len; ++i) h.key.k1 |= (unsigned long long)bytes[i] << (8 * (i - 8));
        h.key.k1 |= (unsigned long long)len << 56;
    } else {
        h.space.assign(len + 64, 0);
        memcpy(h.space.data() + 16, bytes, len); // not at offset 0: exercise unaligned load8
        h.key = make_long_key(16, (uint32_t)len, 0x12345678u);
    }
    return h;
}

} // namespace

extern "C" {

uint32_t th_fnv(const uint8_t *bytes, uint64_t len)
{
    FnvLua52 f;
    for (uint64_t i = 0; i < len; ++i) f.byte(bytes[i]);
    return f.h;
}

uint32_t th_partition_of(const uint8_t *bytes, uint64_t len, uint32_t num_reducers)
{
    HostKey h = make_key(bytes, len);
    return partition_of_key(h.key, h.space.data(), num_reducers);
}

uint64_t th_sort_word(const uint8_t *bytes, uint64_t len, uint32_t round)
{
    HostKey h = make_key(bytes, len);
    return sort_chunk(h.key, h.space.data(), round);
}

uint32_t th_owner_of(uint32_t n, const uint32_t *sp_part, const unsigned long long *sp_word, uint32_t world,
                     uint32_t part, uint64_t word)
{
    OwnerMap m;
    m.part = sp_part; m.word = sp_word; m.n = n; m.world = world;
    return owner_of(m, part, word);
}

} // extern "C"
