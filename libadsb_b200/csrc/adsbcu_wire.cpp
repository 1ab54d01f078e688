// Wire formats of the decoded set: what dump1090's net output writes per message.  Host code: the
// reference formats these on the CPU too (net_io.c), from the same fields adsbcu_msg carries.
#include "../../include/adsbcu.h"

#include <cmath>
#include <cstdio>

namespace {
struct Sink {
    uint8_t *out; size_t cap, n = 0;
    void put(uint8_t b) { if (n < cap) out[n] = b; ++n; }
    void put_escaped(uint8_t b) { put(b); if (b == 0x1A) put(b); }
};
}  // namespace

extern "C" {

// writeBeastMessage (net_io.c:440-487) for every message: <esc> '2'|'3', 6-byte 12 MHz timestamp, signal byte, frame bytes,
// every 0x1A in the payload doubled.  verbatim != 0: the bytes as sliced (modesSendBeastVerbatimOutput, net_io.c:405-412),
// else the corrected bytes without the messages the cooked output drops (2-bit repairs, net_io.c:423-438; the aircraft-
// reliability filter of that function belongs to the tracker and is not applied here).
size_t adsbcu_1090_encode_beast(const adsbcu_msg *msgs, size_t n, int verbatim, uint8_t *out, size_t cap)
{
    Sink s{out, out ? cap : 0};
    for (size_t k = 0; k < n; ++k) {
        const adsbcu_msg &m = msgs[k];
        if (m.msgtype == 32) continue;                       // Mode A/C replies have their own Beast type ('1'); not produced here
        if (!verbatim && m.correctedbits >= 2) continue;
        const int len = m.msgbits / 8;
        s.put(0x1A);
        s.put(len == 7 ? '2' : '3');
        for (int sh = 40; sh >= 0; sh -= 8) s.put_escaped(uint8_t(m.timestamp >> sh));
        int sig = int(std::round(std::sqrt(m.signal_level) * 255));
        if (m.signal_level > 0 && sig < 1) sig = 1;
        if (sig > 255) sig = 255;
        s.put_escaped(uint8_t(sig));
        const uint8_t *p = verbatim ? m.verbatim : m.msg;
        for (int j = 0; j < len; ++j) s.put_escaped(p[j]);
    }
    return s.n;
}

// modesSendRawOutput with timestamps (net_io.c:510-545, --mlat): "@" 12 hex digits of the timestamp, the corrected frame in
// hex, ";" and a newline per message
size_t adsbcu_1090_encode_avr(const adsbcu_msg *msgs, size_t n, uint8_t *out, size_t cap)
{
    Sink s{out, out ? cap : 0};
    static const char hex[] = "0123456789ABCDEF";
    for (size_t k = 0; k < n; ++k) {
        const adsbcu_msg &m = msgs[k];
        if (m.msgtype == 32) continue;
        s.put('@');
        for (int sh = 44; sh >= 0; sh -= 4) s.put(uint8_t(hex[(m.timestamp >> sh) & 15]));
        for (int j = 0; j < m.msgbits / 8; ++j) { s.put(uint8_t(hex[m.msg[j] >> 4])); s.put(uint8_t(hex[m.msg[j] & 15])); }
        s.put(';');
        s.put('\n');
    }
    return s.n;
}

}  // extern "C"
