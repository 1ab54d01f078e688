// Host-side mirror of the reference's handler interface for the three hot paths.
//
// The reference's capture thread hands every ring-buffer slot to
//     RTLSDR::IDataHandler::HandleData(std::span<uint8_t const> const&)        (RTLSDR.hpp:39-49, :518-539)
// and the handlers (ADSB1090Handler, UAT978Handler) demodulate on the calling thread and push each
// frame to the next layer.  The classes below keep that shape -- same method names, same argument
// meaning, one consumer thread per handler, sink called synchronously in stream order -- and do
// the demodulation through the C ABI of include/adsbcu.h instead of the CPU loops:
//
//   ModeS2400Handler   flightaware demodulate2400 at 2.4 MS/s (dump1090/demod_2400.c:99-444); the
//                      sink sees what useModesMessage() sees (mode_s.c:2274)
//   ADSB1090Handler    libadsb's own 2.0 MS/s demodulator (ADSB1090.cpp:158-175, :741-959); the sink
//                      sees what UseModesMessage(mm) sees with mm.crcok set (ADSB1090.cpp:937,968-971)
//   UAT978Handler      UAT978.cpp:43-60 + dump978 process_buffer; the sink has the signature of
//                      dump_raw_message(char updown, uint8_t* data, int len, int rs_errors)
//                      (uat2json-wrapper.cpp:14)
//
// Inside libadsb, define ADSBCU_WITH_LIBADSB before including this file (after RTLSDR.hpp): the
// classes then derive from RTLSDR::IDataHandler itself and can be passed to RTLSDR::Start().
// Errors: HandleData is void in the reference and only throws (RTLSDR.hpp:495); a failing ABI call
// throws std::runtime_error carrying adsbcu_last_error().  There is no CPU fallback.
#pragma once

#include "../../include/adsbcu.h"

#include <chrono>
#include <cstdint>
#include <functional>
#include <span>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace adsbcu_host {

#ifdef ADSBCU_WITH_LIBADSB
using IDataHandler = RTLSDR::IDataHandler;
#else
struct IDataHandler {
    virtual ~IDataHandler() = default;
    virtual void HandleData(std::span<uint8_t const> const &data) = 0;
    virtual void OnDeviceStatusChanged(bool available) = 0;
};
#endif

class Engine {
public:
    explicit Engine(adsbcu_config cfg)
    {
        cfg.struct_size = sizeof(cfg);
        const int rc = adsbcu_create(&cfg, &h_);
        if (rc != ADSBCU_OK) throw std::runtime_error(std::string("adsbcu_create: ") + adsbcu_last_error(nullptr));
    }
    ~Engine() { adsbcu_destroy(h_); }
    Engine(Engine const &) = delete;
    Engine &operator=(Engine const &) = delete;
    adsbcu_handle *get() const { return h_; }
    void check(int rc, char const *what) const
    {
        if (rc != ADSBCU_OK) throw std::runtime_error(std::string(what) + ": " + adsbcu_last_error(h_));
    }

private:
    adsbcu_handle *h_ = nullptr;
};

inline adsbcu_config make_config(int device, int nfix_crc, int fix_df, uint32_t max_feed_samples)
{
    adsbcu_config c{};
    c.struct_size = sizeof(c);
    c.device = device;
    c.nfix_crc = nfix_crc;
    c.fix_df = fix_df;
    c.max_feed_samples = max_feed_samples;
    return c;
}

// ---- Path A -------------------------------------------------------------------------------------
class ModeS2400Handler : public IDataHandler {
public:
    using Sink = std::function<void(adsbcu_msg const &)>;   // == useModesMessage(struct modesMessage*)
    // mode_ac: also run demodulate2400AC on every buffer, after the Mode S pass (dump1090 --modeac, dump1090.c:902-905);
    // one HandleData call is then treated as one mag_buf of the reference
    // live: the samples come from an SDR, not from a file -- the reference then flips its two ICAO filter tables every
    // MODES_ICAO_FILTER_TTL = 60 s of wall time (backgroundTasks -> icaoFilterExpire, dump1090.c:469-472,
    // icao_filter.c:112-127; the first call flips at once); file replays never do
    ModeS2400Handler(Sink sink, int nfix_crc = 0, int fix_df = 1, int device = 0, uint32_t max_feed_samples = 1u << 24,
                     bool mode_ac = false, bool live = false)
        : eng_(make_config(device, nfix_crc, fix_df, max_feed_samples)), sink_(std::move(sink)), out_(1u << 16), mode_ac_(mode_ac), live_(live)
    {
    }
    // The rest of decodeModesMessage for every accepted message (mode_s.c:626-796 with decodeExtendedSquitter and decodeCommB):
    // when set, each buffer's messages go through the three batched decode calls and the sink gets message + fields, which is
    // what useModesMessage(struct modesMessage*) sees in the reference.  Called after the plain sink, in the same order.
    using DecodedSink = std::function<void(adsbcu_msg const &, adsbcu_fields_a const &, adsbcu_fields_es const &, adsbcu_fields_commb const &)>;
    void SetDecodedSink(DecodedSink s) { decoded_sink_ = std::move(s); }

    void HandleData(std::span<uint8_t const> const &data) override
    {
        if (live_) {
            const auto now = std::chrono::steady_clock::now();
            if (now >= next_flip_) {
                eng_.check(adsbcu_1090_filter_expire(eng_.get()), "adsbcu_1090_filter_expire");
                next_flip_ = now + std::chrono::seconds(60);
            }
        }
        size_t n = 0;
        const size_t cap = out_.size();
        int rc = adsbcu_1090_feed(eng_.get(), data.data(), data.size() / 2, out_.data(), cap, &n);
        if (rc == ADSBCU_ERR_OVERFLOW && n > cap) {
            // the stream state already advanced past this buffer: the messages beyond cap are lost, like a
            // dropped mag_buf in the reference (fifo.c:148-152); grow so that it cannot happen twice
            out_.resize(n + n / 2);
            rc = ADSBCU_OK;
            n = cap;
        }
        eng_.check(rc, "adsbcu_1090_feed");
        if (decoded_sink_ && n) {
            hdr_.resize(n); es_.resize(n); mb_.resize(n);
            eng_.check(adsbcu_1090_decode_fields(eng_.get(), out_.data(), n, hdr_.data()), "adsbcu_1090_decode_fields");
            eng_.check(adsbcu_1090_decode_es(eng_.get(), out_.data(), n, es_.data()), "adsbcu_1090_decode_es");
            eng_.check(adsbcu_1090_decode_commb(eng_.get(), out_.data(), n, mb_.data()), "adsbcu_1090_decode_commb");
        }
        for (size_t k = 0; k < n; ++k) {
            if (sink_) sink_(out_[k]);
            if (decoded_sink_) decoded_sink_(out_[k], hdr_[k], es_[k], mb_[k]);
        }
        if (mode_ac_) {
            const size_t ns = data.size() / 2;
            size_t na = 0;
            // sampleTimestamp of the buffer (sdr_ifile.c:202): samples so far at 12 MHz / 2.4 MS/s
            const uint64_t ts = uint64_t(double(samples_seen_) * 12e6 / 2400000.0);
            eng_.check(adsbcu_1090_modeac_block(eng_.get(), data.data(), ns, tail_.data(), tail_.size() / 2, ts, out_.data(),
                                                out_.size(), &na), "adsbcu_1090_modeac_block");
            for (size_t k = 0; k < na; ++k) if (sink_) sink_(out_[k]);
            // keep the last 326 samples as the next buffer's overlap (fifo.c:187-196)
            std::vector<uint8_t> t(tail_);
            t.insert(t.end(), data.begin(), data.begin() + 2 * ns);
            if (t.size() > 2 * 326) t.erase(t.begin(), t.end() - 2 * 326);
            tail_.swap(t);
            samples_seen_ += ns;
        }
    }
    void OnDeviceStatusChanged(bool) override {}
    void Reset(bool clear_filter) { eng_.check(adsbcu_1090_reset(eng_.get(), clear_filter ? 1 : 0), "adsbcu_1090_reset"); }
    adsbcu_stats Stats() const
    {
        adsbcu_stats s{};
        adsbcu_get_stats(eng_.get(), &s);
        return s;
    }

private:
    Engine eng_;
    Sink sink_;
    std::vector<adsbcu_msg> out_;
    DecodedSink decoded_sink_;
    std::vector<adsbcu_fields_a> hdr_;
    std::vector<adsbcu_fields_es> es_;
    std::vector<adsbcu_fields_commb> mb_;
    bool mode_ac_ = false, live_ = false;
    std::chrono::steady_clock::time_point next_flip_{};
    std::vector<uint8_t> tail_;
    uint64_t samples_seen_ = 0;
};

// ---- Path B -------------------------------------------------------------------------------------
// What IListener::OnChanged(IAirCraft const&) sees after one message (ADSBListener.h, AircraftImpl.h:9-48)
struct AircraftState {
    uint32_t addr = 0;
    char     callsign[8] = {};
    int32_t  altitude = 0;
    uint32_t speed = 0, track = 0;
    int32_t  lat1E7 = 0, lon1E7 = 0;
};

class ADSB1090Handler : public IDataHandler {
public:
    using Sink = std::function<void(adsbcu_msg_b const &)>;   // == UseModesMessage(Message const&), crcok only
    using AircraftSink = std::function<void(AircraftState const &)>;   // == IListener::OnChanged
    explicit ADSB1090Handler(Sink sink, int device = 0)
        : eng_(make_config(device, 0, 1, 1u << 22)), sink_(std::move(sink)), out_(1u << 16)
    {
    }
    // With an aircraft sink the handler also does what follows UseModesMessage in the reference: field decode
    // (ADSB1090.cpp:597-672) and CPR (:1079-1121) in bulk on the device, the aircraft table of
    // InteractiveReceiveData (:1125-1175) here, one callback per message in stream order.
    ADSB1090Handler(Sink sink, AircraftSink aircraft, int device = 0) : ADSB1090Handler(std::move(sink), device)
    {
        aircraft_ = std::move(aircraft);
    }
    void HandleData(std::span<uint8_t const> const &data) override
    {
        size_t n = 0;
        eng_.check(adsbcu_1090b_feed(eng_.get(), data.data(), data.size(), out_.data(), out_.size(), &n), "adsbcu_1090b_feed");
        if (sink_) for (size_t k = 0; k < n; ++k) sink_(out_[k]);
        if (aircraft_ && n) track(n);
    }
    void OnDeviceStatusChanged(bool) override {}

private:
    struct Entry {
        AircraftState st;
        int32_t evenLat = 0, evenLon = 0, oddLat = 0, oddLon = 0;
        std::chrono::system_clock::time_point evenTime{}, oddTime{};
    };
    void track(size_t n)
    {
        fields_.resize(n);
        eng_.check(adsbcu_1090b_decode_fields(eng_.get(), out_.data(), n, fields_.data()), "adsbcu_1090b_decode_fields");
        // pass 1: the CPR pairs every position message will hand to DecodeCpr, in stream order
        pairs_.clear();
        pair_of_.assign(n, -1);
        for (size_t k = 0; k < n; ++k) {
            adsbcu_fields_b const &f = fields_[k];
            if (out_[k].msgtype != 17 || f.metype < 9 || f.metype > 18) continue;
            Entry &e = table_[out_[k].addr];
            const auto now = std::chrono::system_clock::now();
            if (f.fflag) { e.oddLat = f.raw_latitude; e.oddLon = f.raw_longitude; e.oddTime = now; }
            else { e.evenLat = f.raw_latitude; e.evenLon = f.raw_longitude; e.evenTime = now; }
            const auto apart = std::chrono::duration_cast<std::chrono::seconds>(e.evenTime - e.oddTime).count();
            if ((apart < 0 ? -apart : apart) > 10) continue;                       // ADSB1090.cpp:1161
            pair_of_[k] = int(pairs_.size());
            pairs_.push_back(adsbcu_cpr_pair{e.evenLat, e.evenLon, e.oddLat, e.oddLon, e.evenTime > e.oddTime ? 1 : 0});
        }
        fixes_.resize(pairs_.size());
        if (!pairs_.empty())
            eng_.check(adsbcu_1090b_cpr_decode(eng_.get(), pairs_.data(), pairs_.size(), fixes_.data()), "adsbcu_1090b_cpr_decode");
        // pass 2: InteractiveReceiveData's updates and the callback, message by message
        for (size_t k = 0; k < n; ++k) {
            adsbcu_fields_b const &f = fields_[k];
            const int df = out_[k].msgtype;
            Entry &e = table_[out_[k].addr];
            e.st.addr = out_[k].addr;
            if (df == 0 || df == 4 || df == 20) {
                e.st.altitude = f.altitude;
            } else if (df == 17) {
                if (f.metype >= 1 && f.metype <= 4) {
                    for (int j = 0; j < 8; ++j) e.st.callsign[j] = f.flight[j];
                } else if (f.metype >= 9 && f.metype <= 18) {
                    e.st.altitude = f.altitude;
                    if (pair_of_[k] >= 0 && fixes_[size_t(pair_of_[k])].valid) {
                        e.st.lat1E7 = fixes_[size_t(pair_of_[k])].lat1e7;
                        e.st.lon1E7 = fixes_[size_t(pair_of_[k])].lon1e7;
                    }
                } else if (f.metype == 19 && (f.mesub == 1 || f.mesub == 2)) {
                    e.st.speed = uint32_t(f.velocity);
                    e.st.track = uint32_t(f.heading);
                }
            }
            aircraft_(e.st);
        }
    }

    Engine eng_;
    Sink sink_;
    AircraftSink aircraft_;
    std::vector<adsbcu_msg_b> out_;
    std::vector<adsbcu_fields_b> fields_;
    std::vector<adsbcu_cpr_pair> pairs_;
    std::vector<adsbcu_cpr_fix> fixes_;
    std::vector<int> pair_of_;
    std::unordered_map<uint32_t, Entry> table_;
};

// ---- Path C -------------------------------------------------------------------------------------
class UAT978Handler : public IDataHandler {
public:
    using Sink = std::function<void(char updown, uint8_t *data, int len, int rs_errors)>;   // == dump_raw_message
    explicit UAT978Handler(Sink sink, int device = 0)
        : eng_(make_config(device, 0, 1, 1u << 22)), sink_(std::move(sink)), out_(1u << 12)
    {
    }
    void HandleData(std::span<uint8_t const> const &data) override
    {
        size_t n = 0;
        eng_.check(adsbcu_978_feed(eng_.get(), data.data(), data.size(), out_.data(), out_.size(), &n), "adsbcu_978_feed");
        for (size_t k = 0; k < n; ++k) sink_(char(out_[k].updown), out_[k].data, int(out_[k].len), int(out_[k].rs_errors));
    }
    void OnDeviceStatusChanged(bool) override {}

private:
    Engine eng_;
    Sink sink_;
    std::vector<adsbcu_uat_frame> out_;
};

}  // namespace adsbcu_host
