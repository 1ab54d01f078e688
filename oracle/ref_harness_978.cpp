/*
 * TEST INFRASTRUCTURE ONLY.  Compiles the reference's own UAT978.cpp where it lies (unmodified; see
 * oracle/Makefile) into oracle/_ref/libref978.so.  The two stages of Path C whose source IS in the reference tree
 * -- InitATan2Table (UAT978.cpp:76-100) and HandleData's staging / carry loop (UAT978.cpp:43-60) -- run here as the
 * reference compiled them; dump978's process_buffer() / init_fec(), which the tree only declares (UAT978.cpp:9-10),
 * are supplied by the test through a callback (tests/test_oracle_978.py plugs in oracle/oracle978.c's restatement
 * and records every call: length, offset, return value).
 */
#include "UAT978.cpp"   /* found through -I$(REF) */

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

typedef int (*ref978_process_fn)(const uint16_t *phi, int len, uint64_t offset);
static ref978_process_fn g_process = nullptr;
static int g_init_fec_calls = 0;
/* the two symbols UAT978.cpp expects from dump978 */
void init_fec() { ++g_init_fec_calls; }
int process_buffer(uint16_t const *phi, int len, uint64_t offset) { return g_process ? g_process(phi, len, offset) : 0; }
}

namespace {
struct NoDevice : RTLSDR::IDeviceSelector {
    [[nodiscard]] bool SelectDevice(RTLSDR::DeviceInfo const &) const override { return false; }
};
struct Ref978 {
    NoDevice sel;
    std::shared_ptr<ADSB::TrafficManager> tm = std::make_shared<ADSB::TrafficManager>();
    std::unique_ptr<UAT978Handler> h = std::make_unique<UAT978Handler>(tm, &sel, ADSB::Source::UAT978);
};
}  // namespace

extern "C" {
void *ref978_create(void) { return new Ref978; }
void ref978_destroy(void *r) { delete static_cast<Ref978 *>(r); }
void ref978_set_process_buffer(ref978_process_fn fn) { g_process = fn; }
int ref978_init_fec_calls(void) { return g_init_fec_calls; }
/* the table InitATan2Table built in the handler's constructor */
void ref978_phase_table(void *r, uint16_t out[65536]) { std::memcpy(out, static_cast<Ref978 *>(r)->h->iqphase.data(), 65536 * 2); }
/* UAT978Handler::HandleData on one buffer */
void ref978_handle_data(void *r, const uint8_t *iq, size_t nbytes) { static_cast<Ref978 *>(r)->h->HandleData(std::span<uint8_t const>(iq, nbytes)); }
void ref978_state(void *r, uint64_t out[2])
{
    out[0] = static_cast<Ref978 *>(r)->h->used;
    out[1] = static_cast<Ref978 *>(r)->h->offset;
}
}
