// Host-side construction of the constant tables the kernels use.
//
//  * sq[]        per-byte squared offsets of the UC8 magnitude table
//                (reference dump1090/dsp/helpers/tables.c:24-29)
//  * bit_synd[]  CRC-24 (generator 0xFFF409, dump1090/crc.c:28) syndrome of each
//                single message bit (crc.c:57-62)
//  * synd_*      the 1-/2-bit repair tables of crc.c:182-381, rebuilt with the same
//                rules (drop every ambiguous syndrome; with two-bit repair also drop
//                syndromes aliased by 3- and 4-bit errors) and stored as hash tables
//  * DF bitsets  demod_2400.c:80-93 and mode_s.c:260-265
#include "adsbcu_internal.h"

#include <algorithm>
#include <cstring>
#include <vector>

namespace adsbcu {
namespace {

uint32_t crc_bitwise(const uint8_t *m, int nbytes)
{
    // remainder of the first nbytes-3 bytes, xor the trailing parity field
    uint32_t rem = 0;
    for (int i = 0; i < nbytes - 3; ++i) {
        rem ^= uint32_t(m[i]) << 16;
        for (int b = 0; b < 8; ++b)
            rem = (rem & 0x800000) ? ((rem << 1) ^ 0xFFF409u) & 0xFFFFFF : (rem << 1) & 0xFFFFFF;
    }
    return rem ^ (uint32_t(m[nbytes - 3]) << 16) ^ (uint32_t(m[nbytes - 2]) << 8) ^ m[nbytes - 1];
}

struct Entry {
    uint32_t syndrome;
    int nerr, b0, b1;
};

uint32_t damage_set(unsigned df, int dist)
{
    uint32_t s = 1u << df;
    if (dist > 0)
        for (int b = 0; b < 5; ++b) s |= damage_set(df ^ (1u << b), dist - 1);
    return s;
}

void build_synd(const uint32_t *bit_synd, int bits, int max_correct, int max_detect, uint64_t *slots, uint32_t nslots)
{
    for (uint32_t i = 0; i < nslots; ++i) slots[i] = kSyndEmpty;
    if (max_correct <= 0) return;
    const int off = 112 - bits;
    std::vector<Entry> all;
    for (int a = 0; a < bits; ++a) {
        all.push_back({bit_synd[a + off], 1, a, 0});
        if (max_correct >= 2)
            for (int b = a + 1; b < bits; ++b) all.push_back({bit_synd[a + off] ^ bit_synd[b + off], 2, a, b});
    }
    std::sort(all.begin(), all.end(), [](const Entry &x, const Entry &y) { return x.syndrome < y.syndrome; });
    std::vector<Entry> uniq;
    for (size_t i = 0; i < all.size();) {
        size_t k = i + 1;
        while (k < all.size() && all[k].syndrome == all[i].syndrome) ++k;
        if (k == i + 1) uniq.push_back(all[i]);
        i = k;
    }
    if (max_detect > max_correct) {
        // remove syndromes that a heavier error pattern (max_correct < weight <= max_detect) also produces
        std::vector<uint32_t> keys(uniq.size());
        for (size_t i = 0; i < uniq.size(); ++i) keys[i] = uniq[i].syndrome;
        std::vector<char> dead(uniq.size(), 0);
        auto mark = [&](uint32_t s) {
            auto it = std::lower_bound(keys.begin(), keys.end(), s);
            if (it != keys.end() && *it == s) dead[size_t(it - keys.begin())] = 1;
        };
        for (int a = 0; a < bits; ++a)
            for (int b = a + 1; b < bits; ++b) {
                const uint32_t sab = bit_synd[a + off] ^ bit_synd[b + off];
                for (int c = b + 1; c < bits; ++c) {
                    const uint32_t sabc = sab ^ bit_synd[c + off];
                    if (max_correct < 3 && max_detect >= 3) mark(sabc);
                    if (max_detect >= 4)
                        for (int d = c + 1; d < bits; ++d) mark(sabc ^ bit_synd[d + off]);
                }
            }
        std::vector<Entry> kept;
        for (size_t i = 0; i < uniq.size(); ++i)
            if (!dead[i]) kept.push_back(uniq[i]);
        uniq.swap(kept);
    }
    for (const Entry &e : uniq) {
        uint32_t h = synd_hash(e.syndrome, nslots);
        while (slots[h] != kSyndEmpty) h = (h + 1) & (nslots - 1);
        slots[h] = uint64_t(e.syndrome) | (uint64_t(e.nerr) << 24) | (uint64_t(e.b0) << 26) | (uint64_t(e.b1) << 33);
    }
}

}  // namespace

void build_tables(DevTables &t, int nfix_crc, int fix_df)
{
    std::memset(&t, 0, sizeof(t));
    for (int i = 0; i < 256; ++i) {
        volatile float f = float((i - 127.4) / 128);   // double expression narrowed once, as the reference does
        volatile float s = f * f;
        t.sq[i] = s * 17179869184.0f;                  // 2^34: exact; sqrt then yields 2^17 * sqrtf(I^2+Q^2)
    }
    uint8_t z[14];
    for (int k = 0; k < 112; ++k) {
        std::memset(z, 0, sizeof(z));
        z[k >> 3] = uint8_t(0x80 >> (k & 7));
        t.bit_synd[k] = crc_bitwise(z, 14);
    }
    for (int b = 0; b < 11; ++b)
        for (int v = 0; v < 256; ++v) {
            uint32_t x = 0;
            for (int k = 0; k < 8; ++k)
                if (v & (0x80 >> k)) x ^= t.bit_synd[8 * b + k];
            t.crc_byte[b][v] = x;
        }
    static const uint32_t fix_short[3] = {0x00000800u, 0x08008e08u, 0x08008e08u};
    static const uint32_t fix_long[3] = {0x00060000u, 0x066f0006u, 0x6fff066fu};
    const int nfix = nfix_crc < 0 ? 0 : nfix_crc > 2 ? 2 : nfix_crc;
    const int sel = fix_df ? nfix : 0;
    t.nfix = nfix;
    t.fixable_short = fix_short[sel];
    t.fixable_long = fix_long[sel];
    t.valid_short = (1u << 0) | (1u << 4) | (1u << 5) | (1u << 11);
    t.valid_long = (1u << 16) | (1u << 17) | (1u << 18) | (1u << 20) | (1u << 21);
    if (fix_df && nfix) {
        t.valid_short |= damage_set(11, 1);
        t.valid_long |= damage_set(17, nfix) | damage_set(18, nfix);
    }
    build_synd(t.bit_synd, 112, nfix, nfix == 2 ? 4 : nfix, t.synd_long, kSyndSlotsLong);
    build_synd(t.bit_synd, 56, nfix, nfix == 2 ? 4 : nfix, t.synd_short, kSyndSlotsShort);
    std::memset(t.occ_long, 0, sizeof(t.occ_long));
    std::memset(t.occ_short, 0, sizeof(t.occ_short));
    for (int h = 0; h < kSyndSlotsLong; ++h) if (t.synd_long[h] != kSyndEmpty) t.occ_long[h >> 5] |= 1u << (h & 31);
    for (int h = 0; h < kSyndSlotsShort; ++h) if (t.synd_short[h] != kSyndEmpty) t.occ_short[h >> 5] |= 1u << (h & 31);
}

}  // namespace adsbcu
