/*
 * TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the
 * product path (libadsb_b200/).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it.
 *
 * oracle1090: plain-C CPU restatement of the reference's 1090 MHz hot path
 * (flightaware dump1090 as vendored under /root/reference/dump1090):
 * uint8 I/Q -> magnitude LUT -> demodulate2400 -> scoreModesMessage /
 * correctMessage / CRC-24 syndrome repair -> ICAO filter -> accepted messages.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_ref.py checks this restatement
 * message-for-message against oracle/_ref/dump1090_ref (the reference's own
 * code compiled from /root/reference, see oracle/Makefile) on synthetic
 * captures, and tests/golden/ holds reference-generated vectors for the GPU box.
 */
#ifndef ORACLE1090_H
#define ORACLE1090_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#pragma pack(push, 1)
/* Same layout as struct ref_record in oracle/ref_harness.c */
typedef struct oracle1090_msg {
    uint64_t timestamp;      /* 12 MHz clock: demod_2400.c:366 */
    uint8_t  msg[14];        /* after correction */
    uint8_t  verbatim[14];   /* as sliced */
    uint8_t  msgbits;        /* 56 or 112 */
    uint8_t  correctedbits;
    uint8_t  msgtype;        /* downlink format of the corrected message */
    uint8_t  score;          /* score_rank, mode_s.h:32-80 */
    uint32_t crc;
    uint32_t addr;
    double   signal_level;   /* demod_2400.c:382-393 */
} oracle1090_msg;
#pragma pack(pop)

typedef struct oracle1090 oracle1090;

/* nfix_crc: 0 (default), 1 (--fix), 2 (--fix-2bit); fix_df: 1 unless --no-fix-df.
 * block_samples: samples per mag_buf (reference: 131072, dump1090.h:92-94); 0 = default. */
oracle1090 *oracle1090_create(int nfix_crc, int fix_df, unsigned block_samples);
void oracle1090_destroy(oracle1090 *o);

/* --modeac: also run demodulate2400AC (demod_2400.c:539-726) on every mag_buf, after demodulate2400. */
void oracle1090_set_mode_ac(oracle1090 *o, int on);

/* --iformat: 0 UC8 (default), 1 SC16, 2 SC16Q11; for 1 and 2 `iq` of oracle1090_feed points at int16 pairs. */
void oracle1090_set_input_format(oracle1090 *o, int fmt);

/* Feed `nsamples` I/Q pairs, processed the way sdr_ifile.c:178-245 + fifo.c:173-210
 * + demod_2400.c:99-444 would process a file read: in blocks of block_samples
 * (the last one short).  Returns the number of messages accepted during this
 * call; the first `cap` of them are written to `out`. */
size_t oracle1090_feed(oracle1090 *o, const uint8_t *iq, size_t nsamples,
                       oracle1090_msg *out, size_t cap);

/* Stage-level entry points (used by per-stage GPU parity tests). */
void     oracle1090_mag_table(uint16_t out[65536]);            /* tables.c:9-41 */
uint32_t oracle1090_crc(const uint8_t *msg, int bits);         /* crc.c:65-81 */
int      oracle1090_score(oracle1090 *o, const uint8_t msg[14]); /* mode_s.c:350-466 */
/* syndrome table dump: returns entry count; writes up to cap entries as
 * (syndrome, nerr, bit0, bit1) uint32 quadruples. bits = 56 or 112. */
size_t   oracle1090_syndrome_table(oracle1090 *o, int bits, uint32_t *out, size_t cap);
void     oracle1090_filter_add(oracle1090 *o, uint32_t addr);
int      oracle1090_filter_test(oracle1090 *o, uint32_t addr);
void     oracle1090_filter_expire(oracle1090 *o);   /* one firing of icaoFilterExpire's timer (icao_filter.c:112-127) */
/* preamble test only: returns 1 if demodulate2400 would try phases at m[j] */
int      oracle1090_preamble(const uint16_t *m);
/* slice one phase (4..8) at m[j]: writes 14 bytes */
void     oracle1090_slice(const uint16_t *m, int try_phase, uint8_t out[14]);

#ifdef __cplusplus
}
#endif
#endif
