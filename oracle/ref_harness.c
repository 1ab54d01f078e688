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
    __real_useModesMessage(mm);
}
