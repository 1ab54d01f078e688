/*
 * TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path
 * (libadsb_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it.
 *
 * oracle1090b: plain CPU restatement of the 1090 MHz path libadsb itself compiles
 * (reference ADSB1090.cpp, the antirez 2.0 MS/s demodulator), with the reference's
 * fixed configuration (ADSB1090.cpp:111-122: fixErrors = true, checkCRC = true,
 * aggressive = false).  One call == one ADSB1090Handler::HandleData(span) call: the
 * reference keeps no sample state between calls (ADSB1090.cpp:158-175, :772); only the
 * cache of recently seen ICAO addresses persists (ADSB1090.cpp:195-207).
 *
 * Parity status: PINNED at function and counter level.  oracle/ref_harness_b.cpp
 * compiles the reference's own ADSB1090.cpp where it lies (unmodified, -D__ANDROID__
 * selects the branch of RTLSDR.hpp that compiles; stub rtl-sdr.h) and
 * tests/test_oracle_b.py checks against it: the magnitude LUT, ModesChecksum,
 * FixSingleBitErrors and DecodeModesMessage on explicit frames (all fields), and, per
 * window of a synthetic capture, the statistics counters (ADSB1090.cpp:235-243), the
 * number and address/altitude/callsign content of IListener::OnChanged callbacks and the
 * set of cached ICAO addresses.
 */
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <unordered_map>
#include <vector>

namespace {

constexpr int kPreambleUs = 8, kLongBits = 112, kShortBits = 56;
constexpr int kFullLength = kPreambleUs + kLongBits;   /* ADSB1090.cpp:106 */

#pragma pack(push, 1)
struct Msg {              /* one UseModesMessage() call that passed checkCRC (ADSB1090.cpp:968-971) */
    uint32_t pos;         /* j: sample index of the preamble inside this buffer */
    uint32_t crc;         /* mm.crc */
    uint32_t addr;        /* aa1<<16 | aa2<<8 | aa3 (recovered address for AP formats) */
    uint8_t  msg[14];     /* mm.msg (after single-bit repair) */
    uint8_t  msgbits;     /* 56 or 112 */
    uint8_t  msgtype;     /* DF as sliced (not updated by a repair, ADSB1090.cpp:501,513-519) */
    int8_t   errorbit;    /* repaired bit or -1 */
    uint8_t  corrected;   /* 1: decoded on the phase-correction retry (ADSB1090.cpp:817-826) */
    uint16_t pad;
};
#pragma pack(pop)
static_assert(sizeof(Msg) == 32, "record layout");

struct Stats { uint64_t valid_preamble, demodulated, goodcrc, badcrc, fixed, single_bit_fix, out_of_phase; };

/* ADSB1090.cpp:266-275 is the table of CRC-24 remainders of single bits; rebuilt from the
 * generator 0xFFF409 (its entry 87), last 24 entries zero. */
uint32_t g_table[112];
void init_table()
{
    static bool done = false;
    if (done) return;
    for (int k = 0; k < 88; ++k) {
        /* remainder of x^(111-k) ... computed by clocking a single 1 bit followed by zeros through the register */
        uint32_t rem = 0;
        for (int b = k; b < 88; ++b) {
            const uint32_t in = (b == k) ? 1u : 0u;
            const uint32_t top = (rem >> 23) & 1u;
            rem = (rem << 1) & 0xFFFFFFu;
            if (top ^ in) rem ^= 0xFFF409u;
        }
        g_table[k] = rem;
    }
    for (int k = 88; k < 112; ++k) g_table[k] = 0;
    done = true;
}

uint32_t checksum(const uint8_t *msg, int bits)   /* ADSB1090.cpp:277-291 */
{
    uint32_t crc = 0;
    const int offset = bits == 112 ? 0 : 112 - 56;
    for (int j = 0; j < bits; ++j)
        if (msg[j / 8] & (1 << (7 - j % 8))) crc ^= g_table[j + offset];
    return crc;
}

int len_by_type(int type)   /* ADSB1090.cpp:295-299 */
{
    return (type == 16 || type == 17 || type == 19 || type == 20 || type == 21) ? kLongBits : kShortBits;
}

int fix_single(uint8_t *msg, int bits)   /* ADSB1090.cpp:304-332 */
{
    uint8_t aux[14];
    for (int j = 0; j < bits; ++j) {
        std::memcpy(aux, msg, 14);
        aux[j / 8] ^= uint8_t(1u << (7 - j % 8));
        const uint32_t crc1 = (uint32_t(aux[bits / 8 - 3]) << 16) | (uint32_t(aux[bits / 8 - 2]) << 8) | aux[bits / 8 - 1];
        if (crc1 == checksum(aux, bits)) { std::memcpy(msg, aux, 14); return j; }
    }
    return -1;
}

struct Decoded { uint8_t msg[14]; int msgbits, msgtype, crcok, errorbit; uint32_t crc, addr; };

}  // namespace

struct oracle1090b {
    std::vector<uint16_t> lut;          /* ADSB1090.cpp:131-142 */
    std::vector<uint16_t> mag;
    std::unordered_map<uint32_t, std::chrono::system_clock::time_point> icao;   /* ADSB1090.cpp:195-207,223 */
    Stats st{};

    bool recently_seen(uint32_t addr)
    {
        auto it = icao.find(addr);
        return it != icao.end() && (std::chrono::system_clock::now() - it->second) <= std::chrono::seconds{60};
    }

    Decoded decode(const uint8_t *in)   /* ADSB1090.cpp:491-596 (the parity-relevant prefix) */
    {
        Decoded d{};
        std::memcpy(d.msg, in, 14);
        d.msgtype = d.msg[0] >> 3;
        d.msgbits = len_by_type(d.msgtype);
        const int nb = d.msgbits / 8;
        d.crc = (uint32_t(d.msg[nb - 3]) << 16) | (uint32_t(d.msg[nb - 2]) << 8) | d.msg[nb - 1];
        const uint32_t crc2 = checksum(d.msg, d.msgbits);
        d.errorbit = -1;
        d.crcok = d.crc == crc2;
        if (!d.crcok && (d.msgtype == 11 || d.msgtype == 17)) {
            if ((d.errorbit = fix_single(d.msg, d.msgbits)) != -1) {
                d.crc = checksum(d.msg, d.msgbits);
                d.crcok = 1;
            }
        }
        d.addr = (uint32_t(d.msg[1]) << 16) | (uint32_t(d.msg[2]) << 8) | d.msg[3];
        if (d.msgtype != 11 && d.msgtype != 17) {
            d.crcok = 0;                    /* ADSB1090.cpp:569-583 + BruteForceAp :396-435 */
            const int t = d.msgtype;
            if (t == 0 || t == 4 || t == 5 || t == 16 || t == 20 || t == 21 || t == 24) {
                const uint32_t c = checksum(d.msg, d.msgbits);
                const uint32_t field = (uint32_t(d.msg[nb - 3]) << 16) | (uint32_t(d.msg[nb - 2]) << 8) | d.msg[nb - 1];
                const uint32_t a = field ^ c;
                if (recently_seen(a)) { d.addr = a; d.crcok = 1; }
            }
        } else if (d.crcok && d.errorbit == -1) {
            icao[d.addr] = std::chrono::system_clock::now();
        }
        return d;
    }

    size_t handle_data(const uint8_t *p, size_t nbytes, Msg *out, size_t cap)
    {
        const size_t n = nbytes / 2;
        if (mag.size() < n + 16) mag.resize(n + 16, 0xffff);
        uint16_t *m = mag.data();
        for (size_t j = 0; j < nbytes; j += 2) {            /* ADSB1090.cpp:158-174 */
            int i = p[j] - 127, q = p[j + 1] - 127;
            if (i < 0) i = -i;
            if (q < 0) q = -q;
            m[j / 2] = lut[size_t(i * 129 + q)];
        }
        size_t nout = 0;
        if (n < size_t(kFullLength * 2)) return 0;           /* the reference's unsigned loop bound would wrap */
        detect(m, uint32_t(n), out, cap, nout);
        return nout;
    }

    void detect(uint16_t *m, uint32_t mlen, Msg *out, size_t cap, size_t &nout)   /* ADSB1090.cpp:741-959 */
    {
        uint8_t bits[kLongBits], msg[14];
        uint16_t aux[kLongBits * 2];
        bool use_correction = false;
        for (uint32_t j = 0; j < mlen - kFullLength * 2; j++) {
            int low, high, delta, errors, good = 0;
            if (!use_correction) {
                if (m[j] <= m[j + 1] || m[j + 1] >= m[j + 2] || m[j + 2] <= m[j + 3] || m[j + 3] >= m[j] || m[j + 4] >= m[j]
                    || m[j + 5] >= m[j] || m[j + 6] >= m[j] || m[j + 7] <= m[j + 8] || m[j + 8] >= m[j + 9] || m[j + 9] <= m[j + 6])
                    continue;
                high = (m[j] + m[j + 2] + m[j + 7] + m[j + 9]) / 6;
                if (m[j + 4] >= high || m[j + 5] >= high) continue;
                if (m[j + 11] >= high || m[j + 12] >= high || m[j + 13] >= high || m[j + 14] >= high) continue;
                st.valid_preamble++;
            } else {
                std::memcpy(aux, m + j + kPreambleUs * 2, sizeof(aux));
                if (j && out_of_phase(m + j)) { phase_correct(m + j); st.out_of_phase++; }
            }
            errors = 0;
            for (int i = 0; i < kLongBits * 2; i += 2) {
                low = m[j + i + kPreambleUs * 2];
                high = m[j + i + kPreambleUs * 2 + 1];
                delta = low - high;
                if (delta < 0) delta = -delta;
                if (i > 0 && delta < 256) bits[i / 2] = bits[i / 2 - 1];
                else if (low == high) { bits[i / 2] = 2; if (i < kShortBits * 2) errors++; }
                else if (low > high) bits[i / 2] = 1;
                else bits[i / 2] = 0;
            }
            if (use_correction) std::memcpy(m + j + kPreambleUs * 2, aux, sizeof(aux));
            for (int i = 0; i < kLongBits; i += 8)
                msg[i / 8] = uint8_t(bits[i] << 7 | bits[i + 1] << 6 | bits[i + 2] << 5 | bits[i + 3] << 4 | bits[i + 4] << 3
                                     | bits[i + 5] << 2 | bits[i + 6] << 1 | bits[i + 7]);
            const int msgtype = msg[0] >> 3;
            const uint32_t msglen = uint32_t(len_by_type(msgtype)) / 8;
            delta = 0;
            for (uint32_t i = 0; i < msglen * 8 * 2; i += 2)
                delta += std::abs(int(m[j + i + kPreambleUs * 2]) - int(m[j + i + kPreambleUs * 2 + 1]));
            delta = int(uint32_t(delta) / (msglen * 4));
            if (delta < 10 * 255) { use_correction = false; continue; }
            if (errors == 0) {
                const Decoded mm = decode(msg);
                if (mm.crcok || use_correction) {
                    st.demodulated++;
                    if (mm.errorbit == -1) { if (mm.crcok) st.goodcrc++; else st.badcrc++; }
                    else { st.badcrc++; st.fixed++; st.single_bit_fix++; }
                }
                if (mm.crcok) {
                    if (nout < cap) {                          /* UseModesMessage: checkCRC lets only crcok through */
                        Msg &o = out[nout];
                        std::memset(&o, 0, sizeof(o));
                        o.pos = j; o.crc = mm.crc; o.addr = mm.addr;
                        std::memcpy(o.msg, mm.msg, 14);
                        o.msgbits = uint8_t(mm.msgbits); o.msgtype = uint8_t(mm.msgtype);
                        o.errorbit = int8_t(mm.errorbit); o.corrected = use_correction;
                    }
                    ++nout;
                    j += (kPreambleUs + msglen * 8) * 2;
                    good = 1;
                }
            }
            if (!good && !use_correction) { j--; use_correction = true; }
            else use_correction = false;
        }
    }

    static int out_of_phase(const uint16_t *m)   /* ADSB1090.cpp:683-690 */
    {
        if (m[3] > m[2] / 3) return 1;
        if (m[10] > m[9] / 3) return 1;
        if (m[6] > m[7] / 3) return -1;
        if (m[-1] > m[1] / 3) return -1;
        return 0;
    }
    static void phase_correct(uint16_t *m)       /* ADSB1090.cpp:720-736 */
    {
        m += 16;
        for (int j = 0; j < (kLongBits - 1) * 2; j += 2) {
            if (m[j] > m[j + 1]) m[j + 2] = uint16_t((m[j + 2] * 5) / 4);
            else m[j + 2] = uint16_t((m[j + 2] * 4) / 5);
        }
    }
};

extern "C" {

oracle1090b *oracle1090b_create(void)
{
    init_table();
    oracle1090b *o = new oracle1090b;
    o->lut.assign(129u * 129u * 2u, 0);
    for (int i = 0; i <= 128; i++)
        for (int q = 0; q <= 128; q++)
            o->lut[size_t(i * 129 + q)] = uint16_t(std::round(std::sqrt(double(i * i + q * q)) * 360));
    return o;
}
void oracle1090b_destroy(oracle1090b *o) { delete o; }
size_t oracle1090b_handle_data(oracle1090b *o, const uint8_t *iq, size_t nbytes, void *out, size_t cap)
{
    return o->handle_data(iq, nbytes, static_cast<Msg *>(out), cap);
}
void oracle1090b_stats(const oracle1090b *o, uint64_t out[7]) { std::memcpy(out, &o->st, sizeof(Stats)); }
void oracle1090b_lut(const oracle1090b *o, uint16_t out[129 * 129]) { std::memcpy(out, o->lut.data(), 129 * 129 * 2); }
uint32_t oracle1090b_checksum(const uint8_t *msg, int bits) { init_table(); return checksum(msg, bits); }
int oracle1090b_fix_single(uint8_t *msg, int bits) { init_table(); return fix_single(msg, bits); }
/* DecodeModesMessage on an explicit frame: out = {msgbits, msgtype, crcok, errorbit, crc, addr}, msg updated in place */
void oracle1090b_decode(oracle1090b *o, uint8_t *msg, int32_t out[6])
{
    const Decoded d = o->decode(msg);
    std::memcpy(msg, d.msg, 14);
    out[0] = d.msgbits; out[1] = d.msgtype; out[2] = d.crcok; out[3] = d.errorbit; out[4] = int32_t(d.crc); out[5] = int32_t(d.addr);
}
size_t oracle1090b_cache(const oracle1090b *o, uint32_t *out, size_t cap)
{
    size_t k = 0;
    for (auto &kv : o->icao) { if (k < cap) out[k] = kv.first; ++k; }
    return k;
}

}  // extern "C"
