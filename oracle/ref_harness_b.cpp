/*
 * TEST INFRASTRUCTURE ONLY.  Compiles the reference's own ADSB1090.cpp where it lies
 * (unmodified; see oracle/Makefile) into oracle/_ref/libref1090b.so and exposes what
 * tests/test_oracle_b.py needs to pin oracle/oracle1090b.cpp against it:
 *   - HandleData on a buffer -> IListener::OnChanged callbacks, statistics counters, ICAO cache
 *   - ModesChecksum / FixSingleBitErrors / DecodeModesMessage on explicit frames
 * -D__ANDROID__ only selects the branch of RTLSDR.hpp:335-355 that compiles (the desktop
 * branch passes a bool where a pointer is expected); no Android API is involved.
 */
#include "ADSB1090.cpp"   /* found through -I$(REF) */

#include <cstring>

extern "C" {
/* librtlsdr is absent: never-called definitions for the linker */
uint32_t rtlsdr_get_device_count(void) { return 0; }
int rtlsdr_open(rtlsdr_dev_t **, uint32_t) { return -1; }
int rtlsdr_open_sys_dev(rtlsdr_dev_t **, int) { return -1; }
int rtlsdr_close(rtlsdr_dev_t *) { return 0; }
int rtlsdr_get_usb_strings(rtlsdr_dev_t *, char *, char *, char *) { return -1; }
int rtlsdr_get_device_usb_strings(uint32_t, char *, char *, char *) { return -1; }
int rtlsdr_set_tuner_gain_mode(rtlsdr_dev_t *, int) { return 0; }
int rtlsdr_get_tuner_gains(rtlsdr_dev_t *, int *) { return 0; }
int rtlsdr_set_tuner_gain(rtlsdr_dev_t *, int) { return 0; }
int rtlsdr_get_tuner_gain(rtlsdr_dev_t *) { return 0; }
int rtlsdr_set_freq_correction(rtlsdr_dev_t *, int) { return 0; }
int rtlsdr_set_agc_mode(rtlsdr_dev_t *, int) { return 0; }
int rtlsdr_set_center_freq(rtlsdr_dev_t *, uint32_t) { return 0; }
int rtlsdr_set_sample_rate(rtlsdr_dev_t *, uint32_t) { return 0; }
int rtlsdr_reset_buffer(rtlsdr_dev_t *) { return 0; }
int rtlsdr_read_async(rtlsdr_dev_t *, rtlsdr_read_async_cb_t, void *, uint32_t, uint32_t) { return -1; }
int rtlsdr_cancel_async(rtlsdr_dev_t *) { return 0; }
}

namespace {
#pragma pack(push, 1)
struct Callback {          /* one IListener::OnChanged */
    uint32_t addr;
    int32_t  altitude;
    uint32_t speed, track;
    int32_t  lat1e7, lon1e7;
    char     callsign[8];
};
#pragma pack(pop)

struct Tap : ADSB::IListener {
    std::vector<Callback> calls;
    void OnChanged(ADSB::IAirCraft const &a) override
    {
        Callback c{};
        c.addr = a.Addr(); c.altitude = a.Altitude(); c.speed = a.Speed(); c.track = a.Heading();
        c.lat1e7 = a.Lat1E7(); c.lon1e7 = a.Lon1E7();
        auto f = a.FlightNumber();
        std::memcpy(c.callsign, f.data(), f.size() < 8 ? f.size() : 8);
        calls.push_back(c);
    }
    void OnDeviceStatusChanged(ADSB::Source, bool) override {}
};
struct NoDevice : RTLSDR::IDeviceSelector {
    [[nodiscard]] bool SelectDevice(RTLSDR::DeviceInfo const &) const override { return false; }
};
struct Ref {
    Tap tap;
    NoDevice sel;
    std::shared_ptr<ADSB::TrafficManager> tm = std::make_shared<ADSB::TrafficManager>();
    std::unique_ptr<ADSB1090Handler> h;
    Ref()
    {
        tm->SetListener(&tap);
        h = std::make_unique<ADSB1090Handler>(tm, &sel, ADSB::Source::ADSB1090);
        h->listener = &tap;
    }
};
}  // namespace

extern "C" {
void *ref1090b_create(void) { return new Ref; }
void ref1090b_destroy(void *r) { delete static_cast<Ref *>(r); }
/* ADSB1090Handler::HandleData on one buffer; returns the number of OnChanged callbacks it caused */
size_t ref1090b_handle_data(void *rp, const uint8_t *iq, size_t nbytes, void *out, size_t cap)
{
    Ref *r = static_cast<Ref *>(rp);
    r->tap.calls.clear();
    r->h->HandleData(std::span<uint8_t const>(iq, nbytes));
    const size_t n = r->tap.calls.size();
    std::memcpy(out, r->tap.calls.data(), sizeof(Callback) * (n < cap ? n : cap));
    return n;
}
void ref1090b_stats(void *rp, uint64_t out[7])
{
    auto &h = *static_cast<Ref *>(rp)->h;
    out[0] = uint64_t(h.statValidPreamble); out[1] = uint64_t(h.statDemodulated); out[2] = uint64_t(h.statGoodcrc);
    out[3] = uint64_t(h.statBadcrc); out[4] = uint64_t(h.statFixed); out[5] = uint64_t(h.statSingleBitFix);
    out[6] = uint64_t(h.statOutOfPhase);
}
void ref1090b_lut(void *rp, uint16_t out[129 * 129])
{
    std::memcpy(out, static_cast<Ref *>(rp)->h->magnitudesLookupTable.data(), 129 * 129 * 2);
}
uint32_t ref1090b_checksum(const uint8_t *msg, int bits)
{
    std::array<uint8_t, 14> m{};
    std::memcpy(m.data(), msg, 14);
    return ModesChecksum(m, size_t(bits));
}
int ref1090b_fix_single(uint8_t *msg, int bits)
{
    std::array<uint8_t, 14> m{};
    std::memcpy(m.data(), msg, 14);
    const int r = FixSingleBitErrors(m, size_t(bits));
    std::memcpy(msg, m.data(), 14);
    return r;
}
void ref1090b_decode(void *rp, uint8_t *msg, int32_t out[6])
{
    std::array<uint8_t, 14> m{};
    std::memcpy(m.data(), msg, 14);
    const Message mm = static_cast<Ref *>(rp)->h->DecodeModesMessage(m);
    std::memcpy(msg, mm.msg.data(), 14);
    out[0] = int32_t(mm.msgbits); out[1] = mm.msgtype; out[2] = mm.crcok; out[3] = mm.errorbit; out[4] = int32_t(mm.crc);
    out[5] = (mm.aa1 << 16) | (mm.aa2 << 8) | mm.aa3;
}
/* every field DecodeModesMessage fills after the CRC stage (ADSB1090.cpp:525-672); layout = adsbcu_fields_b */
void ref1090b_fields(void *rp, const uint8_t *msg, int32_t out[20])
{
    std::array<uint8_t, 14> m{};
    std::memcpy(m.data(), msg, 14);
    const Message mm = static_cast<Ref *>(rp)->h->DecodeModesMessage(m);
    out[0] = mm.ca; out[1] = mm.metype; out[2] = mm.mesub; out[3] = mm.heading; out[4] = mm.fflag;
    out[5] = mm.rawLatitude; out[6] = mm.rawLongitude;
    std::memcpy(&out[7], mm.flight.data(), 8);
    out[9] = mm.ewDir; out[10] = mm.ewVelocity; out[11] = mm.nsDir; out[12] = mm.nsVelocity; out[13] = mm.velocity;
    out[14] = mm.fs; out[15] = mm.dr; out[16] = mm.um; out[17] = mm.identity; out[18] = mm.altitude; out[19] = int32_t(mm.unit);
}
/* DecodeCpr (ADSB1090.cpp:1079-1121) on one even/odd pair; returns 1 if the position was updated */
int ref1090b_cpr(int even_lat, int even_lon, int odd_lat, int odd_lon, int even_is_newer, int32_t out[2])
{
    int32_t got[2][2];
    for (int pass = 0; pass < 2; ++pass) {      /* two sentinels tell "not updated" from a value that equals one of them */
        ADSB::AirCraftImpl a{};
        a.cprEvenLat = even_lat; a.cprEvenLon = even_lon; a.cprOddLat = odd_lat; a.cprOddLon = odd_lon;
        a.cprEvenTime = ADSB::AirCraftImpl::time_point{} + std::chrono::seconds(even_is_newer ? 2 : 1);
        a.cprOddTime = ADSB::AirCraftImpl::time_point{} + std::chrono::seconds(even_is_newer ? 1 : 2);
        a.lat1E7 = a.lon1E7 = pass ? 12345 : INT32_MIN;
        DecodeCpr(a);
        got[pass][0] = a.lat1E7; got[pass][1] = a.lon1E7;
    }
    out[0] = got[0][0]; out[1] = got[0][1];
    return got[0][0] == got[1][0] && got[0][1] == got[1][1];
}
size_t ref1090b_cache(void *rp, uint32_t *out, size_t cap)
{
    size_t k = 0;
    for (auto &kv : static_cast<Ref *>(rp)->h->icaoTimestamps) { if (k < cap) out[k] = kv.first; ++k; }
    return k;
}
}
