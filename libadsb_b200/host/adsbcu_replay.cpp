// File replay through the handler shim: the analogue of `dump1090 --ifile X --raw --mlat`
// (sdr_ifile.c:178-245 -> demodulate2400 -> displayModesMessage, mode_s.c:1897-1903), of
// tests/test_1090.cpp's TestEmbedded loop for libadsb's own handler, and of UAT978Handler fed from
// a file.  Output formats:
//   path a : "@<12 hex 12 MHz timestamp><hex msg>;"  one line per message (net_io.c:529-545), and with
//            --beast FILE the Beast binary stream of writeBeastMessage (net_io.c:440-487) on the verbatim bytes;
//            --decode adds one "# ..." line per message with what decodeModesMessage would have filled in (header fields,
//            extended squitter, Comm-B: the three batched decode calls of the handler)
//   path b : "*<hex msg>;", or with --aircraft one "addr,altitude,speed,track,lat1E7,lon1E7,callsign" line per
//            IListener::OnChanged callback of the reference (field decode + CPR on the device, aircraft table on the host)
//   path uat: "<updown><hex payload>;rs=<n>;"  (dump978 raw line format)
// usage: adsbcu_replay --path a|b|uat [--fix|--fix-2bit] [--no-fix-df] [--modeac] [--aircraft] [--decode] [--block BYTES] [--beast FILE] capture.bin
// (--modeac: every block is one mag_buf of the reference; use --block 262144 to mirror dump1090 --ifile)
#include "adsbcu_handlers.hpp"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>

using namespace adsbcu_host;

int main(int argc, char **argv)
{
    std::string path = "a", file, beast_file;
    int nfix = 0, fix_df = 1;
    bool mode_ac = false, aircraft = false, decode = false;
    size_t block = 262144;   // RTLSDR::BufferLength (RTLSDR.hpp:55)
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--path" && i + 1 < argc) path = argv[++i];
        else if (a == "--fix") nfix = 1;
        else if (a == "--fix-2bit") nfix = 2;
        else if (a == "--no-fix-df") fix_df = 0;
        else if (a == "--modeac") mode_ac = true;
        else if (a == "--aircraft") aircraft = true;
        else if (a == "--decode") decode = true;
        else if (a == "--block" && i + 1 < argc) block = std::strtoull(argv[++i], nullptr, 0);
        else if (a == "--beast" && i + 1 < argc) beast_file = argv[++i];
        else file = a;
    }
    if (file.empty() || block < 2) { std::fprintf(stderr, "usage: adsbcu_replay --path a|b|uat [options] capture.bin\n"); return 2; }
    std::FILE *fp = std::fopen(file.c_str(), "rb");
    if (!fp) { std::perror(file.c_str()); return 2; }
    std::vector<uint8_t> beast;
    size_t count = 0;
    try {
        std::unique_ptr<IDataHandler> h;
        if (path == "a") {
            h = std::make_unique<ModeS2400Handler>([&](adsbcu_msg const &m) {
                std::printf("@%012llX", (unsigned long long) m.timestamp);
                for (int j = 0; j < m.msgbits / 8; ++j) std::printf("%02x", m.msg[j]);
                std::printf(";\n");
                if (!beast_file.empty()) {          // the --net-verbatim Beast stream (writeBeastMessage, net_io.c:440-487), behind the C ABI
                    uint8_t tmp[64];
                    const size_t nb = adsbcu_1090_encode_beast(&m, 1, 1, tmp, sizeof(tmp));
                    beast.insert(beast.end(), tmp, tmp + nb);
                }
                ++count;
            }, nfix, fix_df, 0, uint32_t(std::max<size_t>(block / 2, 1u << 20)), mode_ac);
            if (decode)
                static_cast<ModeS2400Handler *>(h.get())->SetDecodedSink(
                    [](adsbcu_msg const &m, adsbcu_fields_a const &f, adsbcu_fields_es const &e, adsbcu_fields_commb const &b) {
                        std::printf("# DF%u airground=%u alt=%d/%u squawk=%04x/%u es=%u metype=%u.%u addr=%07X addrtype=%u valid=%08X callsign=%.8s "
                                    "cpr=%u,%u,%u alt_baro=%d alt_geom=%d gs=%.3f heading=%.3f commb=%u mbvalid=%05X mbcallsign=%.8s\n",
                                    unsigned(m.msgtype), unsigned(e.es ? e.airground : f.airground), f.altitude_baro, unsigned(f.altitude_baro_valid),
                                    unsigned(f.squawk), unsigned(f.squawk_valid), unsigned(e.es), unsigned(e.metype), unsigned(e.mesub), e.addr,
                                    unsigned(e.addrtype), e.valid, std::string(e.callsign, 8).c_str(), unsigned(e.cpr_odd), e.cpr_lat, e.cpr_lon,
                                    e.altitude_baro, e.altitude_geom, double(e.gs_selected), double(e.heading), unsigned(b.commb_format), b.valid,
                                    std::string(b.callsign, 8).c_str());
                    });
        } else if (path == "b") {
            auto raw = [&](adsbcu_msg_b const &m) {
                std::printf("*");
                for (int j = 0; j < m.msgbits / 8; ++j) std::printf("%02x", m.msg[j]);
                std::printf(";\n");
                ++count;
            };
            if (aircraft)   // one line per IListener::OnChanged: addr,altitude,speed,track,lat1E7,lon1E7,callsign
                h = std::make_unique<ADSB1090Handler>(nullptr, [&](AircraftState const &a) {
                    std::printf("%06X,%d,%u,%u,%d,%d,%.8s\n", a.addr, a.altitude, a.speed, a.track, a.lat1E7, a.lon1E7,
                                std::string(a.callsign, 8).c_str());
                    ++count;
                });
            else
                h = std::make_unique<ADSB1090Handler>(raw);
        } else {
            h = std::make_unique<UAT978Handler>([&](char updown, uint8_t *d, int len, int rs) {
                std::printf("%c", updown);
                for (int j = 0; j < len; ++j) std::printf("%02x", d[j]);
                std::printf(";rs=%d;\n", rs);
                ++count;
            });
        }
        std::vector<uint8_t> buf(block & ~size_t(1));
        for (;;) {
            const size_t got = std::fread(buf.data(), 1, buf.size(), fp);
            if (got < 2) break;
            h->HandleData(std::span<uint8_t const>(buf.data(), got & ~size_t(1)));
            if (got < buf.size()) break;
        }
    } catch (std::exception const &e) {
        std::fprintf(stderr, "adsbcu_replay: %s\n", e.what());
        std::fclose(fp);
        return 3;
    }
    std::fclose(fp);
    if (!beast_file.empty()) {
        std::FILE *bf = std::fopen(beast_file.c_str(), "wb");
        if (!bf) { std::perror(beast_file.c_str()); return 2; }
        std::fwrite(beast.data(), 1, beast.size(), bf);
        std::fclose(bf);
    }
    std::fprintf(stderr, "%zu frames\n", count);
    return 0;
}
