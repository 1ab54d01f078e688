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

/* ---- field decode (ADSB1090.cpp:525-672) and CPR (ADSB1090.cpp:986-1121): SURVEY 8f rank 2 ---- */
namespace {
struct Fields {     /* == adsbcu_fields_b */
    int32_t ca, metype, mesub, heading, fflag, raw_latitude, raw_longitude;
    char flight[8];
    int32_t ew_dir, ew_velocity, ns_dir, ns_velocity, velocity, fs, dr, um, identity, altitude, unit;
};
static_assert(sizeof(Fields) == 80, "Fields layout");

/* 13-bit altitude code of DF0/4/16/20 (ADSB1090.cpp:440-466), 12-bit of DF17 (:470-486) */
int altitude13(const uint8_t *m, int32_t &unit)
{
    if (m[3] & 0x40) { unit = 1; return 0; }
    unit = 0;
    if (!(m[3] & 0x10)) return 0;
    return ((((m[2] & 31) << 6) | ((m[3] & 0x80) >> 2) | ((m[3] & 0x20) >> 1) | (m[3] & 15)) * 25) - 1000;
}
int altitude12(const uint8_t *m, int32_t &unit)
{
    if (!(m[5] & 1)) return 0;
    unit = 0;
    return ((((m[5] >> 1) << 4) | (m[6] >> 4)) * 25) - 1000;
}

/* m = the frame after the CRC stage's single-bit repair, df = mm.msgtype, which the repair does not refresh (:501,513-519) */
Fields fields_of(const uint8_t *m, int df)
{
    static const char ais[] = "?ABCDEFGHIJKLMNOPQRSTUVWXYZ????? ???????????????0123456789??????";
    Fields f{};
    f.ca = f.fs = m[0] & 7;
    f.metype = m[4] >> 3;
    f.mesub = m[4] & 7;
    f.dr = m[1] >> 3;
    f.um = ((m[1] & 7) << 3) | (m[2] >> 5);
    const int a = ((m[3] & 0x80) >> 5) | (m[2] & 0x02) | ((m[2] & 0x08) >> 3);
    const int b = ((m[3] & 0x02) << 1) | ((m[3] & 0x08) >> 2) | ((m[3] & 0x20) >> 5);
    const int c = ((m[2] & 0x01) << 2) | ((m[2] & 0x04) >> 1) | ((m[2] & 0x10) >> 4);
    const int d = ((m[3] & 0x01) << 2) | ((m[3] & 0x04) >> 1) | ((m[3] & 0x10) >> 4);
    f.identity = a * 1000 + b * 100 + c * 10 + d;
    if (df == 0 || df == 4 || df == 16 || df == 20) f.altitude = altitude13(m, f.unit);
    if (df != 17) return f;
    if (f.metype >= 1 && f.metype <= 4) {
        uint64_t v = 0;
        for (int k = 5; k <= 10; ++k) v = (v << 8) | m[k];
        for (int k = 0; k < 8; ++k) f.flight[k] = ais[(v >> (42 - 6 * k)) & 63];
    } else if (f.metype >= 9 && f.metype <= 18) {
        f.fflag = m[6] & 4;
        f.altitude = altitude12(m, f.unit);
        f.raw_latitude = ((m[6] & 3) << 15) | (m[7] << 7) | (m[8] >> 1);
        f.raw_longitude = ((m[8] & 1) << 16) | (m[9] << 8) | m[10];
    } else if (f.metype == 19 && f.mesub >= 1 && f.mesub <= 4) {
        if (f.mesub <= 2) {
            f.ew_dir = (m[5] >> 2) & 1;
            f.ew_velocity = ((m[5] & 3) << 8) | m[6];
            f.ns_dir = m[7] >> 7;
            f.ns_velocity = ((m[7] & 0x7f) << 3) | (m[8] >> 5);
            f.velocity = int(std::sqrt(double(f.ns_velocity * f.ns_velocity + f.ew_velocity * f.ew_velocity)));
            if (f.velocity) {
                const int ewv = f.ew_dir ? -f.ew_velocity : f.ew_velocity, nsv = f.ns_dir ? -f.ns_velocity : f.ns_velocity;
                f.heading = int(std::atan2(double(ewv), double(nsv)) * 360 / (M_PI * 2));
                if (f.heading < 0) f.heading += 360;
            }
        } else {
            f.heading = int((360.0 / 128) * (((m[5] & 3) << 5) | (m[6] >> 3)));
        }
    }
    return f;
}

int cpr_nl(double lat)      /* the table of 1090-WP-9-14 as ADSB1090.cpp:993-1055 holds it */
{
    static const double edge[58] = {
        10.47047130, 14.82817437, 18.18626357, 21.02939493, 23.54504487, 25.82924707, 27.93898710, 29.91135686, 31.77209708,
        33.53993436, 35.22899598, 36.85025108, 38.41241892, 39.92256684, 41.38651832, 42.80914012, 44.19454951, 45.54626723,
        46.86733252, 48.16039128, 49.42776439, 50.67150166, 51.89342469, 53.09516153, 54.27817472, 55.44378444, 56.59318756,
        57.72747354, 58.84763776, 59.95459277, 61.04917774, 62.13216659, 63.20427479, 64.26616523, 65.31845310, 66.36171008,
        67.39646774, 68.42322022, 69.44242631, 70.45451075, 71.45986473, 72.45884545, 73.45177442, 74.43893416, 75.42056257,
        76.39684391, 77.36789461, 78.33374083, 79.29428225, 80.24923213, 81.19801349, 82.13956981, 83.07199445, 83.99173563,
        84.89166191, 85.75541621, 86.53536998, 87.00000000};
    if (lat < 0) lat = -lat;
    for (int k = 0; k < 58; ++k)
        if (lat < edge[k]) return 59 - k;
    return 1;
}
int cpr_mod(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; }
int32_t to_i32(double v)     /* the reference's static_cast<int32_t> as x86-64 executes it (cvttsd2si: out of range -> INT32_MIN) */
{
    return (v >= 2147483648.0 || v <= -2147483649.0 || v != v) ? INT32_MIN : int32_t(v);
}

int cpr_global(double lat0, double lon0, double lat1, double lon1, bool even_newer, int32_t out[2])   /* DecodeCpr :1079-1121 */
{
    const int j = int(std::floor(((59 * lat0 - 60 * lat1) / 131072) + 0.5));
    double rlat0 = (360.0 / 60) * (cpr_mod(j, 60) + lat0 / 131072);
    double rlat1 = (360.0 / 59) * (cpr_mod(j, 59) + lat1 / 131072);
    if (rlat0 >= 270) rlat0 -= 360;
    if (rlat1 >= 270) rlat1 -= 360;
    if (cpr_nl(rlat0) != cpr_nl(rlat1)) return 0;
    const double rlat = even_newer ? rlat0 : rlat1;
    const int nl = cpr_nl(rlat), ni = std::max(nl - (even_newer ? 0 : 1), 1);
    const int m = int(std::floor((((lon0 * (nl - 1)) - (lon1 * nl)) / 131072) + 0.5));
    double lon = (360.0 / ni) * (cpr_mod(m, ni) + (even_newer ? lon0 : lon1) / 131072) * 10000000;
    if (lon > 180 * 10000000.0) lon -= 3600000000.0;
    out[0] = to_i32(rlat * 10000000);
    out[1] = to_i32(lon);
    return 1;
}
}  // namespace

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
/* CRC stage (repairs DF11/17 single-bit errors) then every other Message field; out = adsbcu_fields_b as 20 int32 */
void oracle1090b_fields(oracle1090b *o, const uint8_t *msg, int32_t out[20])
{
    const Decoded d = o->decode(msg);
    const Fields f = fields_of(d.msg, d.msgtype);
    std::memcpy(out, &f, sizeof f);
}
int oracle1090b_cpr(int even_lat, int even_lon, int odd_lat, int odd_lon, int even_is_newer, int32_t out[2])
{
    out[0] = out[1] = INT32_MIN;
    return cpr_global(even_lat, even_lon, odd_lat, odd_lon, even_is_newer != 0, out);
}
/* many frames / pairs per call, for timing the CPU side without per-call binding overhead */
void oracle1090b_fields_bulk(const uint8_t *msgs14, const uint8_t *msgtype, size_t n, int32_t *out20)
{
    for (size_t k = 0; k < n; ++k) {      /* frames already past the CRC stage, like adsbcu_1090b_decode_fields */
        const Fields f = fields_of(msgs14 + 14 * k, msgtype[k]);
        std::memcpy(out20 + 20 * k, &f, sizeof f);
    }
}
void oracle1090b_cpr_bulk(const int32_t *pairs5, size_t n, int32_t *out3)
{
    for (size_t k = 0; k < n; ++k) {
        const int32_t *p = pairs5 + 5 * k;
        int32_t r[2] = {0, 0};
        out3[3 * k + 2] = cpr_global(p[0], p[1], p[2], p[3], p[4] != 0, r);
        out3[3 * k] = r[0]; out3[3 * k + 1] = r[1];
    }
}
size_t oracle1090b_cache(const oracle1090b *o, uint32_t *out, size_t cap)
{
    size_t k = 0;
    for (auto &kv : o->icao) { if (k < cap) out[k] = kv.first; ++k; }
    return k;
}

}  // extern "C"
