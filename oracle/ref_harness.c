/*
 * TEST INFRASTRUCTURE ONLY -- not part of the shipped product path.
 *
 * Harness linked into oracle/_ref/dump1090_ref, the UNMODIFIED flightaware
 * dump1090 compiled from the sources where they lie under
 * /root/reference/dump1090 (see oracle/Makefile).  It adds exactly two things:
 *
 *  1. the four interactive* entry points declared in dump1090.h (the real
 *     interactive.c needs <curses.h>, absent in this image); all no-ops, never
 *     reached in --ifile batch mode.
 *  2. a link-time wrap of useModesMessage() (mode_s.c), the sink that
 *     demodulate2400() hands every accepted message to (demod_2400.c:422).
 *     When REF1090_OUT is set the wrap appends one packed binary record per
 *     message (the parity tuple: bytes, verbatim, crc, corrected bits,
 *     12 MHz timestamp, score, addr, RSSI) and then calls the real sink, so
 *     the reference's own behaviour (tracking, stats, printing) is unchanged.
 */
#include "dump1090.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void interactiveInit(void) {}
void interactiveShowData(void) {}
void interactiveCleanup(void) {}
void interactiveNoConnection(void) {}

void __real_useModesMessage(struct modesMessage *mm);

#pragma pack(push, 1)
struct ref_record {
    uint64_t timestamp;      /* mm->timestampMsg (12 MHz) */
    uint8_t  msg[14];        /* corrected bytes */
    uint8_t  verbatim[14];   /* as sliced */
    uint8_t  msgbits;        /* 56 / 112 */
    uint8_t  correctedbits;
    uint8_t  msgtype;        /* DF */
    uint8_t  score;
    uint32_t crc;
    uint32_t addr;
    double   signal_level;
};
#pragma pack(pop)

static FILE *out_fp;
static int   out_checked;

static void close_out(void)
{
    if (out_fp) fclose(out_fp);
    out_fp = NULL;
}

void __wrap_useModesMessage(struct modesMessage *mm)
{
    if (!out_checked) {
        const char *p = getenv("REF1090_OUT");
        out_checked = 1;
        if (p && *p) {
            out_fp = fopen(p, "wb");
            if (out_fp) {
                setvbuf(out_fp, NULL, _IOFBF, 1 << 20);
                atexit(close_out);
            }
        }
    }
    if (out_fp) {
        struct ref_record r;
        memset(&r, 0, sizeof(r));
        r.timestamp = mm->timestampMsg;
        memcpy(r.msg, mm->msg, 14);
        memcpy(r.verbatim, mm->verbatim, 14);
        r.msgbits = (uint8_t) mm->msgbits;
        r.correctedbits = (uint8_t) mm->correctedbits;
        r.msgtype = (uint8_t) mm->msgtype;
        r.score = (uint8_t) mm->score;
        r.crc = mm->crc;
        r.addr = mm->addr;
        r.signal_level = mm->signalLevel;
        fwrite(&r, sizeof(r), 1, out_fp);
    }
    {
        /* second tap (REF1090_FIELDS=path): the surveillance fields decodeModesMessage filled in (mode_s.c:626-779), in the
         * layout of adsbcu_fields_a, for tests/make_golden_fields.py */
        static FILE *ffp;
        static int fchecked;
        if (!fchecked) {
            const char *p = getenv("REF1090_FIELDS");
            fchecked = 1;
            if (p && *p) ffp = fopen(p, "wb");
        }
        if (ffp) {
            struct { uint16_t AC, ID; uint8_t CA, CC, CF, DR, FS, KE, ND, RI, SL, UM, VS, airground, alert_valid, alert, spi_valid, spi,
                     altitude_baro_valid, altitude_baro_unit, squawk_valid, es_pending; uint16_t squawk; uint16_t pad; int32_t altitude_baro; } f;
            memset(&f, 0, sizeof(f));
            f.AC = (uint16_t) mm->AC; f.ID = (uint16_t) mm->ID; f.CA = (uint8_t) mm->CA; f.CC = (uint8_t) mm->CC; f.CF = (uint8_t) mm->CF;
            f.DR = (uint8_t) mm->DR; f.FS = (uint8_t) mm->FS; f.KE = (uint8_t) mm->KE; f.ND = (uint8_t) mm->ND; f.RI = (uint8_t) mm->RI;
            f.SL = (uint8_t) mm->SL; f.UM = (uint8_t) mm->UM; f.VS = (uint8_t) mm->VS; f.airground = (uint8_t) mm->airground;
            f.alert_valid = mm->alert_valid; f.alert = mm->alert; f.spi_valid = mm->spi_valid; f.spi = mm->spi;
            f.altitude_baro_valid = mm->altitude_baro_valid; f.altitude_baro_unit = (uint8_t) mm->altitude_baro_unit;
            f.squawk_valid = mm->squawk_valid; f.es_pending = (mm->msgtype == 17 || mm->msgtype == 18);
            f.squawk = (uint16_t) mm->squawk; f.altitude_baro = mm->altitude_baro;
            fwrite(&f, sizeof(f), 1, ffp);
            fflush(ffp);
        }
    }
    __real_useModesMessage(mm);
}
