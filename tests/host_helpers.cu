// host_helpers.cu -- the device helpers of csrc/common.cuh compiled FOR THE HOST (they are
// __host__ __device__), behind a tiny C interface for the CPU test suite: the partitioner
// (FnvLua52 / partition_of_key), the sort words (sort_chunk) and the owner map (owner_of) are the
// very source the kernels run.  Test infrastructure; built by tests/test_device_helpers_on_host.py.
#include <cstdint>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "reduce_kernels.cuh"
#include "msd_kernels.cuh"

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

// whitespace classification of the tokeniser: 0x80 in every byte lane of x that is C-locale isspace
uint32_t th_space_flags(uint32_t x) { return space_flags(x); }
uint32_t th_nonspace_nibble(uint32_t x) { return nonspace_nibble(x); }

// the radix sort's digit plan for the varying bits (vlo, vhi): returns the number of passes and
// writes, for every pass, its runs as (shift, width, dst) triples into out[pass][run][3]
uint32_t th_digit_plan(uint64_t vlo, uint32_t vhi, uint8_t *out /* [MAX_PASSES][8][3] */, uint32_t *nruns_out /* [MAX_PASSES] */)
{
    DigitRun runs[MAX_PASSES][8];
    unsigned int nruns[MAX_PASSES];
    memset(runs, 0, sizeof runs);
    memset(nruns, 0, sizeof nruns);
    const uint32_t np = build_digit_plan(vlo, vhi, runs, nruns);
    for (uint32_t p = 0; p < (uint32_t)MAX_PASSES; ++p) {
        nruns_out[p] = nruns[p];
        for (int r = 0; r < 8; ++r) {
            out[(p * 8 + r) * 3 + 0] = runs[p][r].shift;
            out[(p * 8 + r) * 3 + 1] = runs[p][r].width;
            out[(p * 8 + r) * 3 + 2] = runs[p][r].dst;
        }
    }
    return np;
}

// digit of pass `pass` of the key (hi, lo) under the plan of (vlo, vhi)
uint32_t th_digit(uint64_t vlo, uint32_t vhi, uint32_t pass, uint32_t hi, uint64_t lo)
{
    DigitRun runs[MAX_PASSES][8];
    unsigned int nruns[MAX_PASSES];
    memset(runs, 0, sizeof runs);
    memset(nruns, 0, sizeof nruns);
    const uint32_t np = build_digit_plan(vlo, vhi, runs, nruns);
    return pass < np ? extract_digit(runs[pass], nruns[pass], hi, lo) : 0;
}

// token bounds of the map kernel: out = {starts, s_long, s_med, s_short}
void th_token_masks(uint32_t m, uint32_t nxt, uint32_t carry, uint32_t *out)
{
    const TokenMasks t = token_masks(m, nxt, carry);
    out[0] = t.starts; out[1] = t.s_long; out[2] = t.s_med; out[3] = t.s_short;
}
uint32_t th_token_len(uint32_t m, uint32_t nxt, uint32_t bit)
{
    return token_len((unsigned long long)m | ((unsigned long long)nxt << 32), bit);
}

// inline keys from staged words: words = the aligned 32-bit words of the staging buffer, pos = byte
// position of the token in it, len = its length; out = {k0, k1} (a short key has k1 = 0 here: its
// length travels in the top byte of k0, as in the CTA table)
void th_inline_key(const uint32_t *words, uint32_t pos, uint32_t len, unsigned long long *out)
{
    const uint32_t wi = pos >> 2, bs = (pos & 3) * 8;
    if (len <= 7) {
        const ShortKey k = short_key(words[wi], words[wi + 1], words[wi + 2], bs, len);
        out[0] = (unsigned long long)k.w0 | ((unsigned long long)k.w1 << 32);
        out[1] = 0;
    } else {
        const Key k = medium_key(words[wi], words[wi + 1], words[wi + 2], words[wi + 3], words[wi + 4], bs, len);
        out[0] = k.k0;
        out[1] = k.k1;
    }
}

// the long-token scanner (tokens of 16+ bytes): length up to the next whitespace byte, 0 if longer
// than max_len; *hash = the 32-bit hash a long key carries
uint32_t th_scan_long_token(const uint8_t *bytes, uint64_t off, uint32_t max_len, uint32_t *hash)
{
    return scan_long_token(bytes, off, max_len, hash);
}

// equality of two long keys that live at offsets a and b of one byte space (hash and length equal)
int th_same_long_key(const uint8_t *space, uint64_t a, uint64_t b, uint32_t la, uint32_t lb, uint32_t ha, uint32_t hb)
{
    return same_key(make_long_key(a, la, ha), make_long_key(b, lb, hb), space) ? 1 : 0;
}

// reducers: the encoded running value (identity 0) and back
unsigned long long th_op_roundtrip(uint32_t op, unsigned long long a, unsigned long long b)
{
    // what the tables do with two values of one key: encode both, merge (sum or max), decode
    const unsigned long long ea = op_encode(op, a), eb = op_encode(op, b);
    const unsigned long long m = (op == OP_SUM || op == OP_COUNT) ? ea + eb : (ea > eb ? ea : eb);
    return op_decode(op, m);
}

// the most-significant-digit path's plan for a job of n items whose key bytes take the values in
// occ (7 x 256-bit masks): out = {nbins, k, e, low_bits, fallback}; returns the bin of the sort word
// (hi, lo), *low = its packed low word
uint32_t th_msd_extract(const uint32_t *occ /* [7][8] */, uint32_t num_reducers, uint64_t n, uint32_t hi, uint64_t lo,
                        uint64_t *low, uint32_t *out)
{
    MsdPlan p;
    memset(&p, 0, sizeof p);
    msd_build_plan((const unsigned int (*)[8])occ, 0, num_reducers - 1, n, &p);
    out[0] = p.nbins; out[1] = p.k; out[2] = p.e; out[3] = p.low_bits; out[4] = p.fallback;
    static unsigned char lut[7][256];
    for (int f = 0; f < 7; ++f)
        for (uint32_t v = 0; v < 256; ++v) lut[f][v] = (unsigned char)msd_rank_of(occ + 8 * f, v);
    uint32_t d = 0;
    unsigned long long l = 0;
    if (!p.fallback) msd_extract(p, lut, hi, lo, d, l);
    *low = l;
    return d;
}

// digit of pass `pass` of a unit's shared-memory sort: bits [8 pass, 8 pass + 8) of (local bin : low)
uint32_t th_msd_pass_digit(uint64_t low, uint32_t lbin, uint32_t low_bits, uint32_t pass)
{
    return msd_pass_digit(low, lbin, low_bits, pass);
}

// order of two inline keys that agree in bytes 0..6 (msd_kernels.cuh: the tie groups' one-word compare),
// and the same decision made the long way: sort_chunk round by round as key_less does
int th_inline_tail_less(uint64_t a0, uint64_t a1, uint64_t b0, uint64_t b1, int a_first_if_equal)
{
    Key a, b; a.k0 = a0; a.k1 = a1; b.k0 = b0; b.k1 = b1;
    return inline_tail_less(a, b, a_first_if_equal != 0) ? 1 : 0;
}
int th_rounds_less(uint64_t a0, uint64_t a1, uint64_t b0, uint64_t b1, int a_first_if_equal)
{
    Key a, b; a.k0 = a0; a.k1 = a1; b.k0 = b0; b.k1 = b1;
    for (uint32_t r = 1;; ++r) {
        const uint64_t ca = sort_chunk(a, nullptr, r), cb = sort_chunk(b, nullptr, r);
        if (ca != cb) return ca < cb ? 1 : 0;
        if ((ca & 0xFF) < 8) break;
    }
    return a_first_if_equal ? 1 : 0;
}

} // extern "C"
