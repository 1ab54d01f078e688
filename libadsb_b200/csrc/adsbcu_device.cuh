// Device-side building blocks shared by the scan and finalize kernels.
#pragma once

#include "adsbcu_internal.h"

namespace adsbcu {

// ---- magnitude (reference dump1090/dsp/helpers/tables.c:24-35) ------------------------
// The reference looks up round(sqrtf(fI*fI + fQ*fQ) * 65536) clamped to 65535 in a 64K-entry
// table.  Here the same value is computed in registers: sq[] holds fI*fI scaled by 2^34 (an
// exact power-of-two scaling, so every float rounding lands on the same mantissa), the IEEE
// square root then equals 2^17*sqrtf(...), and floor(v+0.5) == (floor(2v)+1)>>1.
__device__ __forceinline__ uint32_t mag_from_sq(float x)
{
    const float s = __fsqrt_rn(x);
    const uint32_t w = __float2uint_rz(s);
    return min((w + 1u) >> 1, 65535u);
}

__device__ __forceinline__ uint32_t mag_iq(const float *sq, uint32_t i, uint32_t q)
{
    return mag_from_sq(__fadd_rn(sq[i], sq[q]));
}

// ---- bit slicer (demod_2400.c:45-59): correlator for sub-phase `sub` at p[0..3] --------
//   sub 0: 5,-3,-2   1: 4,-1,-3   2: 3,1,-4   3: 2,3,-5   4: 1,5,-5,-1
__device__ __forceinline__ int slice_corr(int m0, int m1, int m2, int m3, int sub)
{
    return (5 - sub) * m0 + (2 * sub - 3) * m1 - min(2 + sub, 5) * m2 - (sub == 4 ? m3 : 0);
}

// A sliced frame: bits 0..55 in the low 56 bits of h0 (bit 0 = MSB), bits 56..111 in h1.
struct Frame {
    uint64_t h0, h1;
};
__device__ __forceinline__ uint32_t frame_bit(const Frame &f, int k)   // k: 0-based message bit
{
    return k < 56 ? uint32_t(f.h0 >> (55 - k)) & 1u : uint32_t(f.h1 >> (111 - k)) & 1u;
}
__device__ __forceinline__ void frame_flip(Frame &f, int k)
{
    if (k < 56) f.h0 ^= 1ull << (55 - k); else f.h1 ^= 1ull << (111 - k);
}
__device__ __forceinline__ uint32_t frame_byte(const Frame &f, int b)
{
    return b < 7 ? uint32_t(f.h0 >> (48 - 8 * b)) & 0xFFu : uint32_t(f.h1 >> (48 - 8 * (b - 7))) & 0xFFu;
}
__device__ __forceinline__ uint32_t frame_df(const Frame &f) { return uint32_t(f.h0 >> 51) & 31u; }
__device__ __forceinline__ uint32_t frame_aa(const Frame &f) { return uint32_t(f.h0 >> 24) & 0xFFFFFFu; }

// Slice `nbits` symbols of try_phase `ph` (4..8).  mag(i) returns the magnitude i samples
// after the first data sample m[j+19].  Also accumulates the two CRC-24 syndromes as the XOR
// of per-bit syndromes (CRC is linear; equals crc.c:65-81 on the sliced bytes).
template <class MagFn>
__device__ __forceinline__ void slice_bits(MagFn mag, int ph, int first, int last, const uint32_t *bit_synd,
                                           Frame &f, uint32_t &xs, uint32_t &xl)
{
    int u = ph + 12 * first;
    int s = u / 5, sub = u - 5 * s;
    for (int k = first; k < last; ++k) {
        const int c = slice_corr(mag(s), mag(s + 1), mag(s + 2), mag(s + 3), sub);
        if (c > 0) {
            frame_flip(f, k);
            xl ^= bit_synd[k];
            if (k < 56) xs ^= bit_synd[k + 56];
        }
        s += 2; sub += 2;
        if (sub >= 5) { sub -= 5; s += 1; }
    }
}

// ---- syndrome lookup (crc.c:386-405 as a hash probe) -----------------------------------
struct Diag {
    int nerr;       // -1: not correctable
    int b0, b1;
};
__device__ __forceinline__ Diag diagnose(uint32_t syn, const uint64_t *slots, uint32_t nslots)
{
    Diag d{-1, 0, 0};
    if (syn == 0) { d.nerr = 0; return d; }
    uint32_t h = synd_hash(syn, nslots);
    for (;;) {
        const uint64_t e = slots[h];
        if (e == kSyndEmpty) return d;
        if (uint32_t(e & 0xFFFFFF) == syn) {
            d.nerr = int(e >> 24) & 3; d.b0 = int(e >> 26) & 127; d.b1 = int(e >> 33) & 127;
            return d;
        }
        h = (h + 1) & (nslots - 1);
    }
}
__device__ __forceinline__ void apply_diag(Frame &f, const Diag &d)
{
    if (d.nerr >= 1) frame_flip(f, d.b0);
    if (d.nerr >= 2) frame_flip(f, d.b1);
}

// DF18: does decodeExtendedSquitter flag the address as non-ICAO (mode_s.c:1447-1548; IMF bit
// positions :869 :991 :1031 :1126 :1137 :1333)?  The flag is part of the key later given to
// icaoFilterAdd (mode_s.c:786-788).
__device__ __forceinline__ bool df18_non_icao(const Frame &f)
{
    const uint32_t cf = frame_byte(f, 0) & 7u;
    auto me = [&](int n) { return frame_bit(f, 31 + n); };   // ME bit n (1-based)
    const uint32_t metype = (me(1) << 4) | (me(2) << 3) | (me(3) << 2) | (me(4) << 1) | me(5);
    const uint32_t sub = (me(6) << 2) | (me(7) << 1) | me(8);
    bool imf = false;
    switch (cf) {
    case 0: return false;
    case 1: case 5: return true;
    case 2: case 6: imf = true; break;
    case 3: return me(1) != 0;
    default: return true;
    }
    if (!imf) return false;
    switch (metype) {
    case 19: return (sub >= 1 && sub <= 4) ? me(9) != 0 : false;
    case 5: case 6: case 7: case 8: return me(21) != 0;
    case 0: case 9: case 10: case 11: case 12: case 13: case 14: case 15: case 16: case 17: case 18:
    case 20: case 21: case 22: return me(8) != 0;
    case 28: return sub == 1 ? me(56) != 0 : false;
    case 29: return me(51) != 0;
    case 31: return me(56) != 0;
    default: return false;
    }
}

// ---- correctMessage + the state-independent half of scoreModesMessage ------------------
// (mode_s.c:252-346 and :350-466).  `xs`/`xl` are the syndromes of the sliced bytes.
struct Classified {
    bool     keep;         // can reach SR_ACCEPT_THRESHOLD under some filter state
    bool     needs_known;  // ... only if the filter knows test_key
    uint32_t kind;
    int      corrections;
    Frame    fixed;
    uint32_t crc;          // mm.crc of decodeModesMessage (mode_s.c:541-549)
    uint32_t test_key;     // argument of icaoFilterTest
};

__device__ __forceinline__ Classified classify(const Frame &in, uint32_t xs, uint32_t xl, const DevTables *t)
{
    Classified r;
    r.keep = false; r.needs_known = true; r.kind = K_AP; r.corrections = -1; r.fixed = in; r.crc = 0; r.test_key = 0;
    if (in.h0 == 0) return r;                                   // SR_ALL_ZEROS

    const uint32_t df = frame_df(in), dfbit = 1u << df;
    uint32_t ss = kUnchecked, ls = kUnchecked;
    Diag le{-1, 0, 0}, se{-1, 0, 0};
    bool done = false;
    int nc = -1;
    Frame out = in;

    if (t->fixable_long & dfbit) {
        ls = xl;
        if ((df == 17 || df == 18) && ls == 0) { nc = 0; done = true; }
        else le = diagnose(ls, t->synd_long, kSyndSlotsLong);
    }
    if (!done && (t->fixable_short & dfbit)) {
        ss = xs;
        if (df == 11 && (ss & 0xFFFF80u) == 0) { nc = 0; done = true; }
        else se = diagnose(ss, t->synd_short, kSyndSlotsShort);
    }
    if (!done) {
        const unsigned lerr = le.nerr >= 0 ? unsigned(le.nerr) : 999u;
        const unsigned serr = se.nerr >= 0 ? unsigned(se.nerr) : 999u;
        if (le.nerr >= 0 && lerr <= serr) {
            Frame c = in; apply_diag(c, le);
            const uint32_t d2 = frame_df(c);
            if (d2 == 17 || d2 == 18) { out = c; nc = int(lerr); done = true; }
        }
        if (!done && se.nerr >= 0 && serr == 1) {
            Frame c = in; apply_diag(c, se);
            if (frame_df(c) == 11) { out = c; nc = 1; done = true; }
        }
        if (!done && le.nerr >= 0 && lerr > serr) {
            Frame c = in; apply_diag(c, le);
            const uint32_t d2 = frame_df(c);
            if (d2 == 17 || d2 == 18) { out = c; nc = int(lerr); done = true; }
        }
    }
    r.fixed = out;
    r.corrections = nc;
    const uint32_t d2 = frame_df(out);
    const uint32_t s_short = ss == kUnchecked ? xs : ss;
    const uint32_t s_long = ls == kUnchecked ? xl : ls;
    r.crc = (d2 & 16) ? s_long : s_short;
    switch (d2) {
    case 0: case 4: case 5:
        r.keep = true; r.kind = K_AP; r.test_key = s_short; break;
    case 16: case 20: case 21:
        r.keep = true; r.kind = K_AP; r.test_key = s_long; break;
    case 11: {
        const bool acq = (s_short & 0x7Fu) == 0;
        r.test_key = frame_aa(out);
        if (nc == 0) { r.keep = true; r.kind = acq ? K_DF11_ACQ_0 : K_DF11_IID_0; r.needs_known = !acq; }
        else if (nc == 1) { r.keep = true; r.kind = acq ? K_DF11_ACQ_1 : K_DF11_IID_1; }
        break;
    }
    case 17:
        r.test_key = frame_aa(out);
        if (nc >= 0 && nc <= 2) { r.keep = true; r.kind = K_DF17_0 + uint32_t(nc); r.needs_known = nc == 2; }
        break;
    case 18:
        r.test_key = frame_aa(out) | kNtFlag;
        if (nc >= 0 && nc <= 2) { r.keep = true; r.kind = K_DF18_0 + uint32_t(nc); r.needs_known = nc == 2; }
        break;
    default:
        break;                                                  // unknown DF / DF24+ (enable_df24 off)
    }
    return r;
}

// bitmap of addresses that may ever be in the filter: bit index = 24-bit address (+2^24 for the NT table)
__device__ __forceinline__ uint32_t bitmap_index(uint32_t key) { return (key & 0xFFFFFFu) | ((key & kNtFlag) ? (1u << 24) : 0u); }

}  // namespace adsbcu
