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

// ---- fast slicer: uniform code for every phase ---------------------------------------------
// Symbol k of try_phase p starts u = p + 12k fifths-of-a-sample after m[j+19].  Re-index with a
// "virtual" symbol v = k + v0 where v0 = 3p mod 5: then u = 5c + 12v with c = (p - 12 v0)/5, so the
// correlator sub-phase is 2v mod 5 for EVERY phase and the pattern repeats every 5 symbols / 12
// samples:  v mod 5 = 0..4 -> sample offset 0,2,4,7,9 with sub-phase 0,2,4,1,3.  One "period" is
// therefore 12 consecutive samples -> 5 symbols with fixed weights (demod_2400.c:45-59):
//     r0 = 5m0-3m1-2m2   r1 = 3m2+m3-4m4   r2 = m4+5m5-5m6-m7   r3 = 4m7-m8-3m9   r4 = 2m9+3m10-5m11
// A lane only differs in where its sample 0 is (c) and in how many leading virtual symbols are
// garbage (v0).  Samples are consumed as packed uint16 pairs straight from the shared-memory tile
// with IDP.2A (u16 x s8 dot products); the two possible alignments of sample 0 inside a 32-bit
// word are folded into per-lane weight registers, so there is no divergence on parity either.
#ifdef ADSBCU_EMU      // (u16 x s8) two-way dot products, as the PTX instruction defines them
__device__ __forceinline__ int dp2a_lo(uint32_t a, uint32_t b, int c)
{
    return c + int(a & 0xFFFFu) * int(int8_t(b & 0xFFu)) + int(a >> 16) * int(int8_t((b >> 8) & 0xFFu));
}
__device__ __forceinline__ int dp2a_hi(uint32_t a, uint32_t b, int c)
{
    return c + int(a & 0xFFFFu) * int(int8_t((b >> 16) & 0xFFu)) + int(a >> 16) * int(int8_t(b >> 24));
}
#else
__device__ __forceinline__ int dp2a_lo(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.lo.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.hi.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
#endif

// weights are stored NEGATED so that (corr > 0) is the sign bit of the accumulated value
__host__ __device__ constexpr uint32_t wpair(int lo, int hi) { return uint32_t(uint8_t(-lo)) | (uint32_t(uint8_t(-hi)) << 8); }
__host__ __device__ constexpr uint32_t wreg(uint32_t pa, uint32_t pb) { return pa | (pb << 16); }
// 13 (word, weight-pair) terms per period; register k holds terms 2k (dp2a.lo) and 2k+1 (dp2a.hi)
//   term: 0:r0/W0 1:r0/W1 2:r1/W1 3:r1/W2 4:r2/W2 5:r2/W3 6:r2/W4 7:r3/W3 8:r3/W4 9:r3/W5 10:r4/W4 11:r4/W5 12:r4/W6
__device__ __forceinline__ void slicer_weights(bool odd, uint32_t (&w)[7])
{
    // sample 0 in the low half of word 0 (even) or in the high half (odd)
    constexpr uint32_t E[7] = {wreg(wpair(5, -3), wpair(-2, 0)), wreg(wpair(3, 1), wpair(-4, 0)), wreg(wpair(1, 5), wpair(-5, -1)),
                               wreg(wpair(0, 0), wpair(0, 4)), wreg(wpair(-1, -3), wpair(0, 0)), wreg(wpair(0, 2), wpair(3, -5)),
                               wreg(wpair(0, 0), 0)};
    constexpr uint32_t O[7] = {wreg(wpair(0, 5), wpair(-3, -2)), wreg(wpair(0, 3), wpair(1, -4)), wreg(wpair(0, 1), wpair(5, -5)),
                               wreg(wpair(-1, 0), wpair(0, 0)), wreg(wpair(4, -1), wpair(-3, 0)), wreg(wpair(0, 0), wpair(2, 3)),
                               wreg(wpair(-5, 0), 0)};
#pragma unroll
    for (int k = 0; k < 7; ++k) w[k] = odd ? O[k] : E[k];
}

// Slice NPER periods (5*NPER virtual symbols) starting at 32-bit word pointer wp (shared memory).
// acc[i] receives virtual symbols 32i..32i+31, first symbol in the MSB once the word is full.
template <int NPER>
__device__ __forceinline__ void slice_periods(const uint32_t *wp, const uint32_t (&w)[7], uint32_t (&acc)[4])
{
    uint32_t W0 = wp[0];
#pragma unroll
    for (int q = 0; q < NPER; ++q) {
        const uint32_t W1 = wp[6 * q + 1], W2 = wp[6 * q + 2], W3 = wp[6 * q + 3], W4 = wp[6 * q + 4],
                       W5 = wp[6 * q + 5], W6 = wp[6 * q + 6];
        int c[5];
        c[0] = dp2a_hi(W1, w[0], dp2a_lo(W0, w[0], 0));
        c[1] = dp2a_hi(W2, w[1], dp2a_lo(W1, w[1], 0));
        c[2] = dp2a_lo(W4, w[3], dp2a_hi(W3, w[2], dp2a_lo(W2, w[2], 0)));
        c[3] = dp2a_hi(W5, w[4], dp2a_lo(W4, w[4], dp2a_hi(W3, w[3], 0)));
        c[4] = dp2a_lo(W6, w[6], dp2a_hi(W5, w[5], dp2a_lo(W4, w[5], 0)));
#pragma unroll
        for (int r = 0; r < 5; ++r) {
            const int v = 5 * q + r;
            acc[v >> 5] = __funnelshift_l(uint32_t(c[r]), acc[v >> 5], 1);   // (acc << 1) | (corr > 0)
        }
        W0 = W6;
    }
}

// per-phase constants of the re-indexing above
__device__ __forceinline__ int phase_v0(int ph) { return (3 * ph) % 5; }
__device__ __forceinline__ int phase_c(int ph) { return (ph - 12 * phase_v0(ph)) / 5; }

// Align the virtual-symbol accumulators to message bits: returns M[0..3], bit 0 of the message in the
// MSB of M[0].  NBITS = number of virtual symbols accumulated (5*NPER).
template <int NBITS>
__device__ __forceinline__ void align_message(const uint32_t (&acc)[4], int v0, uint32_t (&M)[4])
{
    uint32_t a[5] = {acc[0], acc[1], acc[2], acc[3], 0u};
    constexpr int last = (NBITS - 1) >> 5, rem = NBITS & 31;
    if (rem) a[last] <<= (32 - rem);                    // left-align the partially filled word
#pragma unroll
    for (int i = last + 1; i < 5; ++i) a[i] = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i) M[i] = __funnelshift_l(a[i + 1], a[i], v0);
}

// CRC-24 syndromes of the sliced bytes via per-byte contribution tables (linear code):
// long: bytes 0..10 through tables 0..10, xor bytes 11..13; short: bytes 0..3 through tables 7..10, xor bytes 4..6
__device__ __forceinline__ uint32_t msg_byte(const uint32_t (&M)[4], int b) { return (M[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu; }
__device__ __forceinline__ uint32_t crc_long(const uint32_t (&M)[4], const uint32_t *crc_byte)
{
    uint32_t x = (M[2] & 0xFFu) << 16 | (M[3] >> 16);
#pragma unroll
    for (int b = 0; b < 11; ++b) x ^= crc_byte[b * 256 + msg_byte(M, b)];
    return x;
}
__device__ __forceinline__ uint32_t crc_short(const uint32_t (&M)[4], const uint32_t *crc_byte)
{
    uint32_t x = M[1] >> 8;
#pragma unroll
    for (int b = 0; b < 4; ++b) x ^= crc_byte[(7 + b) * 256 + msg_byte(M, b)];
    return x;
}
__device__ __forceinline__ Frame frame_from_words(const uint32_t (&M)[4], bool is_long)
{
    Frame f;
    f.h0 = (uint64_t(M[0]) << 24) | (M[1] >> 8);
    f.h1 = is_long ? (uint64_t(M[1] & 0xFFu) << 48) | (uint64_t(M[2]) << 16) | (M[3] >> 16) : 0ull;
    return f;
}

// ---- syndrome lookup (crc.c:386-405 as a hash probe) -----------------------------------
struct Diag {
    int nerr;       // -1: not correctable
    int b0, b1;
};
__device__ __forceinline__ Diag diagnose(uint32_t syn, const uint64_t *slots, uint32_t nslots, const uint32_t *occ)
{
    Diag d{-1, 0, 0};
    if (syn == 0) { d.nerr = 0; return d; }
    uint32_t h = synd_hash(syn, nslots);
    for (;;) {
        if (occ && !((occ[h >> 5] >> (h & 31)) & 1u)) return d;    // empty slot, known without touching the table
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

// occ_long / occ_short: shared-memory copies of DevTables::occ_* (or nullptr)
__device__ __forceinline__ Classified classify(const Frame &in, uint32_t xs, uint32_t xl, const DevTables *t,
                                               const uint32_t *occ_long = nullptr, const uint32_t *occ_short = nullptr)
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
        else le = diagnose(ls, t->synd_long, kSyndSlotsLong, occ_long);
    }
    if (!done && (t->fixable_short & dfbit)) {
        ss = xs;
        if (df == 11 && (ss & 0xFFFF80u) == 0) { nc = 0; done = true; }
        else se = diagnose(ss, t->synd_short, kSyndSlotsShort, occ_short);
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

// ---- symbol planes -> message bits (front kernel layout: planes[c][run] holds 10 decisions, MSB first) ----------
// The message whose first data symbol is at time t0 = 5 * (staged sample) + phase occupies the bits n0 .. n0+111 of
// plane c = t0 mod 12 (n0 = t0 div 12).  `pf` points at the field holding bit n0, `r` = n0 mod 10.  The 10-bit fields
// from there on are repacked into a dense bit stream (every field is < 1024, so the shifted fields add up without
// overlap) and shifted to the message's first bit.  Also returns both CRC-24 syndromes.
__device__ __forceinline__ void slice_planes(const uint16_t *pf, uint32_t r, bool is_long, const uint32_t *crcb, uint32_t fixable_short,
                                             uint32_t (&M)[4], uint32_t &xs, uint32_t &xl)
{
    // 32-bit loads of field pairs from the aligned address (half as many memory instructions as one load per field);
    // an odd first field shifts every pair by 16 bits.  c[j] = fields 2j, 2j+1 as one 20-bit piece of the bit stream.
    const uint32_t odd = uint32_t(reinterpret_cast<uintptr_t>(pf) >> 1) & 1u, sh = 16u * odd;
    const uint32_t *pw = reinterpret_cast<const uint32_t *>(pf - odd);
    auto piece = [&](uint32_t lo, uint32_t hi) {
        const uint32_t a = __funnelshift_r(lo, hi, sh);          // {field 2j, field 2j+1}, each below 1024
        return ((a & 0xFFFFu) << 10) | (a >> 16);
    };
    const uint32_t q0 = pw[0], q1 = pw[1], q2 = pw[2], q3 = pw[3];
    const uint32_t c0 = piece(q0, q1), c1 = piece(q1, q2), c2 = piece(q2, q3);
    xs = 0; xl = 0;
    if (is_long) {
        const uint32_t q4 = pw[4], q5 = pw[5], q6 = pw[6];
        const uint32_t c3 = piece(q3, q4), c4 = piece(q4, q5), c5 = piece(q5, q6), c6 = piece(q6, 0u);
        const uint32_t w0 = (c0 << 12) | (c1 >> 8), w1 = (c1 << 24) | (c2 << 4) | (c3 >> 16), w2 = (c3 << 16) | (c4 >> 4),
                       w3 = (c4 << 28) | (c5 << 8) | (c6 >> 12), w4 = c6 << 20;
        M[0] = __funnelshift_l(w1, w0, r); M[1] = __funnelshift_l(w2, w1, r);
        M[2] = __funnelshift_l(w3, w2, r);
        M[3] = __funnelshift_l(w4, w3, r) & 0xFFFF0000u;
        xl = crc_long(M, crcb);
        // the 56-bit syndrome is also needed when a short DF was sliced long (DF0/DF5 are within two bit flips of
        // DF17/18, so --fix-2bit slices them to 14 bytes but scores them as 56-bit Address/Parity replies)
        if (!(M[0] >> 31) || (fixable_short & (1u << (M[0] >> 27)))) xs = crc_short(M, crcb);
    } else {
        const uint32_t c3 = piece(q3, 0u);
        const uint32_t w0 = (c0 << 12) | (c1 >> 8), w1 = (c1 << 24) | (c2 << 4) | (c3 >> 16), w2 = c3 << 16;
        M[0] = __funnelshift_l(w1, w0, r);
        M[1] = __funnelshift_l(w2, w1, r) & 0xFFFFFF00u;
        M[2] = 0u; M[3] = 0u;
        xs = crc_short(M, crcb);
    }
}

// The evaluated record of one kept (position, phase): what decodeModesMessage fills before the field decode
// (mode_s.c:520-626) plus the keys of the two filter calls.  signal_level is filled in by the caller.
__device__ __forceinline__ void fill_rec(FullRec &q, uint32_t pos, uint32_t ph, const Frame &f, const Classified &c)
{
    q.pos = pos;
    q.phase = uint8_t(ph);
    q.kind = uint8_t(c.keep ? c.kind : K_COUNT);
    q.corrections = int8_t(c.corrections < 0 ? 0 : c.corrections);
    q.msgtype = uint8_t(frame_df(c.fixed));
#pragma unroll
    for (int b = 0; b < 14; ++b) {
        q.verbatim[b] = uint8_t(frame_byte(f, b));
        q.msg[b] = uint8_t(frame_byte(c.fixed, b));
    }
    q.crc = c.crc;
    q.test_key = c.test_key;
    q.add_key = kUnchecked;
    const uint32_t aa = frame_aa(c.fixed);
    switch (q.msgtype) {
    case 11: q.addr = aa; if (c.kind == K_DF11_ACQ_0) q.add_key = aa; break;
    case 17: q.addr = aa; if (c.kind == K_DF17_0) q.add_key = aa; break;
    case 18:
        q.addr = aa | (df18_non_icao(c.fixed) ? kNonIcao : 0u);
        if (c.kind == K_DF18_0) q.add_key = q.addr | kNtFlag;
        break;
    default: q.addr = c.crc; break;
    }
    q.signal_level = 0.0;
}

// mean squared magnitude over the signal window of the message at start position `pos` (demod_2400.c:382-403):
// IEEE double divisions, the same on the device as on the host.  One warp per record.
__device__ __forceinline__ double signal_level_warp(const ChunkArgs &a, uint32_t pos, uint32_t df_uncorrected, int lane)
{
    const int siglen = (df_uncorrected & 16u) ? 268 : 134;
    const int64_t xb = int64_t(pos) - kOverlap + 19;               // feed sample of m[j+19]
    const int64_t n = a.nsamples;
    unsigned long long acc = 0;
    if (!a.mag_in && xb >= 0 && xb + 288 <= n) {
        // the whole window lies inside the feed: nine independent sample loads per lane, then nine LUT gathers
        const uint16_t *s16 = reinterpret_cast<const uint16_t *>(a.iq) + xb;
        uint32_t s[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) s[k] = lane + 32 * k < siglen ? uint32_t(s16[lane + 32 * k]) : 0u;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const uint32_t v = lane + 32 * k < siglen ? uint32_t(a.tab->mag_lut_full[lutf_index(s[k])]) : 0u;
            acc += uint64_t(v * v);
        }
    } else {
        for (int i = lane; i < siglen; i += 32) {
            const int64_t x = xb + i;
            uint32_t v = 0;
            if (x < 0) v = (x + kOverlap >= 0) ? a.carry_mag[x + kOverlap] : 0u;
            else if (x < n) v = a.mag_in ? uint32_t(a.mag_in[x])
                                         : uint32_t(a.tab->mag_lut_full[lutf_index(uint32_t(a.iq[2 * x]) | (uint32_t(a.iq[2 * x + 1]) << 8))]);
            acc += uint64_t(v * v);
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, d);
    return double(acc) / 65535.0 / 65535.0 / double(siglen);
}

// ---- the common case of classify(), branch-light --------------------------------------------------------------
// Almost every sliced candidate is noise: its syndromes are neither zero nor in the repair tables.  For those
// correctMessage changes nothing and scoreModesMessage only looks at the DF: Address/Parity formats (DF0/4/5 with the
// 56-bit syndrome as address, DF16/20/21 with the 112-bit one) can be accepted for a known address, everything else is
// SR_UNCORRECTABLE / SR_UNKNOWN_DF.  Returns 0: dropped, 1: Address/Parity candidate with filter key `key`,
// 2: anything else (a syndrome that is zero, DF11-clean or present in a repair table) -> run classify().
// `occ_*`: occupancy bitmaps of the syndrome hash tables; a syndrome whose home slot is empty is in no table.
__device__ __forceinline__ int classify_fast(uint32_t m0, bool zero56, uint32_t xs, uint32_t xl, uint32_t fixable_long, uint32_t fixable_short,
                                             const uint32_t *occ_long, const uint32_t *occ_short, uint32_t &key)
{
    if (zero56) return 0;                                           // SR_ALL_ZEROS
    const uint32_t df = m0 >> 27, dfbit = 1u << df;
    if (fixable_long & dfbit) {
        const uint32_t h = synd_hash(xl, kSyndSlotsLong);
        if (xl == 0u || ((occ_long[h >> 5] >> (h & 31)) & 1u)) return 2;
    }
    if (fixable_short & dfbit) {
        const uint32_t h = synd_hash(xs, kSyndSlotsShort);
        if ((xs & 0xFFFF80u) == 0u || ((occ_short[h >> 5] >> (h & 31)) & 1u)) return 2;
    }
    // no correction applies: corrected == uncorrected, corrections == -1
    const uint32_t ap_short = (1u << 0) | (1u << 4) | (1u << 5), ap_long = (1u << 16) | (1u << 20) | (1u << 21);
    if (dfbit & ap_short) { key = xs; return 1; }
    if (dfbit & ap_long) { key = xl; return 1; }
    return 0;
}

// bitmap of addresses that may ever be in the filter: bit index = 24-bit address (+2^24 for the NT table)
__device__ __forceinline__ uint32_t bitmap_index(uint32_t key) { return (key & 0xFFFFFFu) | ((key & kNtFlag) ? (1u << 24) : 0u); }

}  // namespace adsbcu
