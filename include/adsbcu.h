/*
 * adsbcu.h -- C ABI of the B200 (sm_100a) Mode S demodulation engine.
 *
 * This is the thin extern "C" layer the reference's host C++ would call from
 * below its RTLSDR::IDataHandler::HandleData(span<uint8_t const>) entry
 * (reference RTLSDR.hpp:39-49; caller RTLSDR.hpp:518-539).  It replaces, for
 * the 1090 MHz band, the work the reference does between "raw uint8 I/Q
 * buffer" and "accepted Mode S message":
 *
 *   dump1090/convert.c:22-44 + dsp/impl/magnitude_uc8.c:30-63   (I/Q -> uint16 magnitude)
 *   dump1090/fifo.c:173-210                                      (326-sample overlap carry)
 *   dump1090/demod_2400.c:99-444  demodulate2400()               (preamble, 5-phase slicer)
 *   dump1090/mode_s.c:252-466     correctMessage/scoreModesMessage
 *   dump1090/crc.c:65-418         modesChecksum / syndrome tables
 *   dump1090/icao_filter.c:40-109 recently-seen address filter
 *   dump1090/mode_s.c:520-626,781-788  decodeModesMessage prefix + filter insertions
 *
 * and hands back what demodulate2400() passes to useModesMessage()
 * (mode_s.c:2274): message bytes, CRC, corrected-bit count, 12 MHz timestamp,
 * score and RSSI.  Plain pointers and sizes only; no exceptions cross it; every
 * function returns 0 on success or a negative adsbcu_status.
 *
 * Threading: one host thread per handle at a time (the reference's consumer
 * thread model).  Each handle owns one CUDA stream.
 */
#ifndef ADSBCU_H
#define ADSBCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADSBCU_ABI_VERSION 2   /* 2: adsbcu_stats fields renamed (round-2 kernels), entry points added; nothing removed */

typedef enum adsbcu_status {
    ADSBCU_OK            = 0,
    ADSBCU_ERR_ARG       = -1,   /* null pointer, bad size, misaligned device pointer */
    ADSBCU_ERR_CUDA      = -2,   /* a CUDA runtime call failed; see adsbcu_last_error() */
    ADSBCU_ERR_NO_DEVICE = -3,   /* no usable sm_100 device */
    ADSBCU_ERR_OVERFLOW  = -4,   /* out[] too small, or an internal candidate list overflowed */
    ADSBCU_ERR_NOMEM     = -5
} adsbcu_status;

/* Mirrors the parity-relevant knobs of the reference's global `Modes`
 * (dump1090/dump1090.c:108-150, :624-633). */
typedef struct adsbcu_config {
    uint32_t struct_size;      /* sizeof(adsbcu_config), for ABI evolution */
    int32_t  device;           /* CUDA device ordinal */
    int32_t  nfix_crc;         /* 0 = none (reference default), 1 = --fix, 2 = --fix-2bit */
    int32_t  fix_df;           /* 1 (default) = also repair the DF field; 0 = --no-fix-df */
    uint32_t block_samples;    /* samples per reference mag_buf; 0 -> 131072 (dump1090.h:92-94).
                                  Only the skip rule at block boundaries depends on it
                                  (demod_2400.c:432-443). */
    uint32_t max_feed_samples; /* largest nsamples accepted by one feed call; 0 -> 1<<24 */
} adsbcu_config;

#pragma pack(push, 1)
/* One accepted message == one useModesMessage() call of the reference. */
typedef struct adsbcu_msg {
    uint64_t timestamp;        /* 12 MHz: 5*sample + 768 + phase (demod_2400.c:366) */
    uint8_t  msg[14];          /* corrected bytes; only msgbits/8 are meaningful */
    uint8_t  verbatim[14];     /* bytes as sliced, before correction */
    uint8_t  msgbits;          /* 56 or 112 */
    uint8_t  correctedbits;    /* 0..2 */
    uint8_t  msgtype;          /* downlink format */
    uint8_t  score;            /* score_rank (mode_s.h:32-80) */
    uint32_t crc;              /* mm.crc (mode_s.c:541-549) */
    uint32_t addr;             /* mm.addr incl. MODES_NON_ICAO_ADDRESS flag (bit 24) */
    double   signal_level;     /* mean squared magnitude, full scale = 1.0 */
} adsbcu_msg;
#pragma pack(pop)

/* Counters and device timings of the most recent feed call. */
typedef struct adsbcu_stats {
    uint64_t samples;          /* samples in the feed */
    uint64_t preambles;        /* positions that passed the preamble detector */
    uint64_t phase_records;    /* (unused since the round-2 kernels: always 0) */
    uint64_t resolved_records; /* records that survived the known-address filter */
    uint64_t messages;         /* accepted messages */
    float    ms_h2d;           /* host->device copy (0 for device feeds) */
    float    ms_front;         /* front kernels (magnitude, slicer decisions, preambles), summed over the chunks */
    float    ms_messages;      /* message kernels (DF filter, CRC, repair, scoring) */
    float    ms_order;         /* order kernels (late records, position order) */
    float    ms_resolve;       /* resolve kernels (the ordered pass of demodulate2400, on the device) */
    uint32_t launches;         /* kernels launched by the feed */
    uint32_t chunks;           /* chunks the feed was cut into (= launches of each per-chunk kernel) */
} adsbcu_stats;

typedef struct adsbcu_handle adsbcu_handle;

int         adsbcu_abi_version(void);
const char *adsbcu_last_error(const adsbcu_handle *h);   /* h may be NULL: last create() error */

int  adsbcu_create(const adsbcu_config *cfg, adsbcu_handle **out);
void adsbcu_destroy(adsbcu_handle *h);

/* Forget stream state: overlap carry, last_message_end and sample clock, as a
 * MAGBUF_DISCONTINUOUS buffer does (demod_2400.c:112-115, fifo.c:187-190); if
 * clear_filter != 0 also icaoFilterInit() (icao_filter.c:64-69). */
int adsbcu_1090_reset(adsbcu_handle *h, int clear_filter);

/* Demodulate `nsamples` interleaved uint8 I/Q pairs that continue the stream.
 * iq is HOST memory; the copy to the device is part of the call.  Accepted
 * messages are appended to out[0..cap) in stream order, *n_out receives their
 * number.  Equivalent to the reference consuming the same bytes through
 * sdr_ifile.c:178-245 -> demodulate2400().
 * out[] may be ordinary host memory, pinned host memory (cudaHostAlloc / cudaHostRegister: the GPU writes the records
 * there itself, no copy at the end) or DEVICE memory of the handle's GPU (8-byte aligned): the message set then stays on
 * the device, complete when the call returns, ready for adsbcu_1090_decode_device.  The same holds for the two feeds below. */
int adsbcu_1090_feed(adsbcu_handle *h, const uint8_t *iq, size_t nsamples,
                     adsbcu_msg *out, size_t cap, size_t *n_out);

/* Same, for I/Q already resident in device memory (16-byte aligned). */
int adsbcu_1090_feed_device(adsbcu_handle *h, const uint8_t *iq_dev, size_t nsamples,
                            adsbcu_msg *out, size_t cap, size_t *n_out);

/* The other two input formats of the reference (convert.c:46-92, --iformat SC16 / SC16Q11): interleaved
 * little-endian int16 I/Q pairs (HOST memory), q11 != 0 for the 12-bit Q11 scaling.  Magnitudes follow
 * magnitude_sc16.c:27-45 / magnitude_sc16q11.c:29-47; everything downstream is the same path as
 * adsbcu_1090_feed and the stream may mix calls of both. */
int adsbcu_1090_feed_sc16(adsbcu_handle *h, const int16_t *iq, size_t nsamples, int q11,
                          adsbcu_msg *out, size_t cap, size_t *n_out);

/* Stage-level entry (replaces starch_magnitude_uc8, magnitude_uc8.c:30-63):
 * mag_out[k] = LUT[I,Q] for k < nsamples, both HOST pointers. */
int adsbcu_1090_magnitude(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, uint16_t *mag_out);

/* Filter access for host code that shares the reference's icaoFilterAdd/Test view. */
int adsbcu_1090_filter_add(adsbcu_handle *h, uint32_t addr);
int adsbcu_1090_filter_test(adsbcu_handle *h, uint32_t addr);   /* 1 = present, 0 = absent */
/* icaoFilterExpire (icao_filter.c:112-127): the reference's main loop calls it once a minute for live SDR input, never
 * for file input (dump1090.c:469-472).  The other table is cleared and becomes the active one, so an address stays
 * known for one to two periods after it was last added.  The caller keeps the clock (the handler shim calls this
 * every 60 s of wall time in live mode); between two calls the engine behaves like file mode. */
int adsbcu_1090_filter_expire(adsbcu_handle *h);

int adsbcu_get_stats(const adsbcu_handle *h, adsbcu_stats *out);

/* ---- Path A on a capture sharded over several GPUs -------------------------------------------------
 * demodulate2400() is sequential in three places only: the skip after an accepted message
 * (demod_2400.c:410-419,432-443), the ICAO filter (mode_s.c:370-446,781-788) and best-of-five-phases
 * (demod_2400.c:336).  The entry points below split the path there: every shard produces state-free
 * candidates on its own GPU, one handle replays the sequential rules over the merged candidates, so
 * the message set does not depend on the number of shards.  Per window of the capture:
 *
 *   every shard:  adsbcu_1090_seek -> adsbcu_1090_scan_candidates -> adsbcu_1090_known_export
 *   exchange   :  OR of the exported sets (host; a few MiB)   -> adsbcu_1090_known_import on every shard
 *   every shard:  adsbcu_1090_revive          (candidates that only became possible through another shard's addresses)
 *   one handle :  adsbcu_1090_resolve(candidates of all shards, ascending (pos, phase))
 */
#pragma pack(push, 1)
typedef struct adsbcu_cand {
    uint64_t pos;              /* start position: sample index in the capture (the padded mag_buf coordinate) */
    uint8_t  phase;            /* try_phase 4..8 */
    uint8_t  kind;             /* state-independent class; with the filter state it gives score_rank */
    int8_t   corrections;
    uint8_t  msgtype;
    uint8_t  verbatim[14];
    uint8_t  msg[14];
    uint32_t crc;
    uint32_t addr;
    uint32_t test_key;         /* argument of icaoFilterTest for this candidate */
    uint32_t add_key;          /* argument of icaoFilterAdd if accepted, 0xFFFFFFFF: none */
    double   signal_level;     /* mean squared magnitude over the signal window, full scale = 1.0 */
} adsbcu_cand;
#pragma pack(pop)
#define ADSBCU_KNOWN_WORDS (1u << 20)   /* 2 x 2^24 bits: addresses that may enter the filter [ICAO | DF18 NT] */

/* Position the stream: the next sample fed is sample `stream_pos` of the capture.  prev_iq (HOST)
 * holds the nprev <= 326 samples just before it (fewer only at the capture start, where the
 * reference's first overlap is zeros, fifo.c:45). */
int adsbcu_1090_seek(adsbcu_handle *h, uint64_t stream_pos, const uint8_t *prev_iq, size_t nprev);
/* Scan nsamples <= max_feed_samples (HOST memory, or device memory if on_device != 0) and return the
 * candidates in ascending (pos, phase) order.  Touches neither the ICAO filter nor the skip state.
 * A device buffer must stay valid until adsbcu_1090_revive has returned. */
int adsbcu_1090_scan_candidates(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, int on_device,
                                adsbcu_cand *out, size_t cap, size_t *n_out);
int adsbcu_1090_known_export(adsbcu_handle *h, uint32_t *words /* [ADSBCU_KNOWN_WORDS] */);
int adsbcu_1090_known_import(adsbcu_handle *h, const uint32_t *words);   /* OR into this handle's set */
/* Candidates of the last scan that were dropped only because their address was not in the local set
 * and is in the set now; ascending (pos, phase).  The list is consumed. */
int adsbcu_1090_revive(adsbcu_handle *h, adsbcu_cand *out, size_t cap, size_t *n_out);
/* The ordered pass of demodulate2400 over candidates in ascending (pos, phase) order: continues this
 * handle's ICAO filter and skip state.  Works on a handle created with cfg.device = -1 (a merge-only
 * handle: no GPU needed, every scan/feed entry point then returns ADSBCU_ERR_NO_DEVICE). */
int adsbcu_1090_resolve(adsbcu_handle *h, const adsbcu_cand *cands, size_t n,
                        adsbcu_msg *out, size_t cap, size_t *n_out);

/* ---- One long capture over several GPUs, every shard resolved on its own GPU (BASELINE config 4) -----
 * The reference's unbounded-stream mechanism is the 131072-sample block + 326-sample overlap + carried
 * last_message_end (fifo.c:173-210, demod_2400.c:106-115,432-443, dump1090.c:159-167): what one sample range
 * needs from everything in front of it is (a) the 326 samples before it, (b) the skip state the last accepted
 * message left behind and (c) the contents of the ICAO filter (file mode: entries are never removed,
 * icao_filter.c:58-83).  (b) and (c) are a few KiB:
 *
 *   adsbcu_1090_state_export   after a feed: the skip state + every address in the filter
 *   adsbcu_1090_state_import   after adsbcu_1090_seek, before the first feed of a range: start from that state
 *
 * adsbcu_1090_replay_sharded is the orchestrator on top of them.  Rank r of `world` holds ITS contiguous range of
 * the capture in device memory (a list of pieces in stream order) and gets the messages of that range; the only
 * exchange is the callback, an all-gather of at most ADSBCU_SHARD_WORDS 32-bit words per rank (host side; MPI,
 * torch.distributed, threads -- whatever the caller has).  Round 0: every rank demodulates its range as if
 * nothing preceded it.  Round k+1: a rank whose start state (skip state of rank r-1's end, union of the filters
 * of all ranks in front of it) differs from the one it used runs its range again.  When no start state changes
 * any more every rank has used exactly the state the sequential run would have had -- by induction over the
 * ranks the fixed point is unique -- so the message set equals the single-stream one whatever `world` is.
 * Typically rank 0 makes one pass and every other rank two (the second can only start when every rank in front has
 * finished its first), so `world` equal ranges take the time of two passes over one of them.
 * Not exact (ADSBCU_ERR_OVERFLOW) once the filter holds ADSBCU_SHARD_MAX_ADDRS addresses: from there on the
 * reference's table drops insertions and membership depends on the insertion order. */
typedef struct adsbcu_1090_state {
    uint64_t stream_pos;       /* samples consumed so far (information on export, ignored on import) */
    uint64_t last_a, last_lme; /* start and last_message_end of the last accepted message, stream coordinates */
    uint32_t has_last;
    uint32_t reserved;
} adsbcu_1090_state;
int adsbcu_1090_state_export(adsbcu_handle *h, adsbcu_1090_state *st, uint32_t *addrs, size_t cap, size_t *n_addrs);
int adsbcu_1090_state_import(adsbcu_handle *h, const adsbcu_1090_state *st, const uint32_t *addrs, size_t n_addrs);

#define ADSBCU_SHARD_MAX_ADDRS 4000u
#define ADSBCU_SHARD_WORDS     (ADSBCU_SHARD_MAX_ADDRS + 8u)
typedef struct adsbcu_piece { const uint8_t *iq; uint64_t nsamples; } adsbcu_piece;   /* device memory, 16-byte aligned (or host, see flags) */
/* Collective: contributes mine[0..n_mine) and fills all[q * ADSBCU_SHARD_WORDS + i] / counts[q] for every rank q. */
typedef int (*adsbcu_allgather_fn)(void *ctx, const uint32_t *mine, uint32_t n_mine, uint32_t *all, uint32_t *counts);
typedef struct adsbcu_shard_report {
    uint32_t passes;           /* times THIS rank demodulated its range */
    uint32_t rounds;           /* exchange rounds until no start state changed (same on every rank) */
    uint32_t start_addrs;      /* filter entries this rank started its final pass with */
    uint32_t chunks;           /* the adsbcu_stats counters below, summed over all feeds of all passes of this rank */
    uint64_t launches;
    double   ms_front, ms_messages, ms_order, ms_resolve, ms_h2d;
} adsbcu_shard_report;
enum { ADSBCU_SHARD_HOST_PIECES = 1,     /* flags: the pieces are HOST memory (uploaded chunk by chunk, on every pass) */
       ADSBCU_SHARD_HAS_TAIL = 2 };      /* _tail only, the same on every rank: the capture ends with a range that rank 0 holds */
int adsbcu_1090_replay_sharded(adsbcu_handle *h, int rank, int world, const adsbcu_piece *pieces, size_t npieces,
                               uint64_t range_start, const uint8_t *halo_iq /* HOST, <= 326 samples */, size_t n_halo,
                               adsbcu_allgather_fn allgather, void *ctx, unsigned flags,
                               adsbcu_msg *out, size_t cap, size_t *n_out, adsbcu_shard_report *report);
/* The same with world + 1 ranges: range r on rank r, and the LAST range of the capture (the "tail") on rank 0 as well.  The
 * first range never runs twice and nobody waits for the end state of the last one, so rank 0 demodulates its own range in
 * round 0 and the tail in round 1 while the other ranks make their two passes: two passes over 1 / (world + 1) of the capture.
 * Every rank sets ADSBCU_SHARD_HAS_TAIL; rank 0 passes the tail's pieces, start and halo (ignored elsewhere).  Rank 0's
 * out[] holds the messages of its own range, then the tail's: *n_first tells where the tail begins.  The capture's message
 * list is rank 0's first part, ranks 1 .. world-1, rank 0's tail part. */
int adsbcu_1090_replay_sharded_tail(adsbcu_handle *h, int rank, int world, const adsbcu_piece *pieces, size_t npieces,
                                    uint64_t range_start, const uint8_t *halo_iq, size_t n_halo,
                                    const adsbcu_piece *tail_pieces, size_t n_tail_pieces, uint64_t tail_start,
                                    const uint8_t *tail_halo_iq, size_t n_tail_halo,
                                    adsbcu_allgather_fn allgather, void *ctx, unsigned flags,
                                    adsbcu_msg *out, size_t cap, size_t *n_out, size_t *n_first, adsbcu_shard_report *report);

/* ---- Mode A/C on one mag_buf (dump1090 --modeac) ----------------------------------------------------
 * One call == one demodulate2400AC(struct mag_buf*) call of the reference (demod_2400.c:539-726), which
 * dump1090's main loop makes right after demodulate2400 on the same buffer (dump1090.c:902-905).  The
 * detector's noise level is the mean level / power of THIS buffer's new samples (demod_2400.c:550-551,
 * magnitude_power_uc8.c:48-90), so the unit is the reference's mag_buf: `iq` holds the buffer's nsamples new
 * samples (131072 in live mode), `overlap_iq` the n_overlap <= 326 samples that preceded it (fewer only at the
 * start of a stream, where the reference's overlap is zero magnitudes), both HOST memory.  block_timestamp is
 * mag->sampleTimestamp (12 MHz).  Replies come back as adsbcu_msg with msgtype 32, msgbits 16, msg[0..1] the
 * Mode A code in the reference's layout and addr = code | MODES_NON_ICAO_ADDRESS (mode_ac.c:123-158). */
int adsbcu_1090_modeac_block(adsbcu_handle *h, const uint8_t *iq, size_t nsamples,
                             const uint8_t *overlap_iq, size_t n_overlap, uint64_t block_timestamp,
                             adsbcu_msg *out, size_t cap, size_t *n_out);

/* The same Mode A/C block for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same detector, same kernels and same replies as adsbcu_1090_modeac_block, but
 *   - all scratch (magnitudes, level / power sums, candidate records) lives in ONE file-scope device array of the
 *     library, about 386 KiB, that every GPU which loads the library holds until the process exits: the call makes
 *     no cudaMalloc / cudaFree / cudaHostAlloc and no host<->device copy, and never waits for the GPU;
 *   - nsamples is limited to ADSBCU_MODEAC_STATIC_MAX_SAMPLES (the reference's mag_buf, MODES_MAG_BUF_SAMPLES);
 *     a larger block is rejected with ADSBCU_ERR_ARG and a message in adsbcu_last_error();
 *   - iq_dev (16-byte aligned), overlap_iq_dev, out_dev and result_dev are DEVICE pointers owned by the caller;
 *     the work is enqueued on cuda_stream (a cudaStream_t; NULL = the default stream) and the results are valid
 *     once that stream has reached this point: result_dev[0] = number of replies (if it exceeds cap only the
 *     first cap were written), result_dev[1] = ADSBCU_STATIC_* flags (0 = complete);
 *   - every call on a GPU uses the same array, so calls that overlap on that GPU (different streams, different
 *     handles or threads) race; calls on one stream, or otherwise ordered, are safe.  The array keeps the last
 *     call's contents; each call clears what must start at zero itself.
 * `h` supplies the device and its lookup tables; nothing in it is modified except the error string.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress): make that call outside stream capture. */
#define ADSBCU_MODEAC_STATIC_MAX_SAMPLES 131072
enum { ADSBCU_STATIC_CANDIDATE_OVERFLOW = 1,   /* more than MAX_SAMPLES/16 candidate positions: no replies written */
       ADSBCU_STATIC_OUT_OVERFLOW = 2 };       /* more replies than cap */
int adsbcu_1090_modeac_block_static(adsbcu_handle *h, const uint8_t *iq_dev, size_t nsamples,
                                    const uint8_t *overlap_iq_dev, size_t n_overlap, uint64_t block_timestamp,
                                    adsbcu_msg *out_dev, size_t cap, uint32_t *result_dev /* [2] */, void *cuda_stream);

/* ---- Path A, next layer: dump1090's surveillance field decode ---------------------
 * What decodeModesMessage does with an accepted frame outside the ME / MB payloads (mode_s.c:626-779): the header
 * fields of every downlink format (AC, CA, CC, CF, DR, FS, ID, KE, ND, RI, SL, UM, VS), the 13-bit altitude
 * (decodeAC13Field incl. the Gillham path, mode_s.c:122-150, mode_ac.c:61-122), the squawk (decodeID13Field,
 * mode_s.c:96-116), air/ground state and alert / SPI flags.  Many messages per call, host arrays in and out.
 * For DF17 / DF18 the reference's extended-squitter decoder (mode_s.c:798-1551, adsbcu_1090_decode_es below) may refine
 * airground, altitude, squawk, alert and spi afterwards: es_pending = 1 marks those records. */
typedef struct adsbcu_fields_a {
    uint16_t AC, ID;             /* raw 13-bit fields (0 where the format has none) */
    uint8_t  CA, CC, CF, DR, FS, KE, ND, RI, SL, UM, VS;
    uint8_t  airground;          /* airground_t: 0 AG_INVALID, 1 AG_GROUND, 2 AG_AIRBORNE, 3 AG_UNCERTAIN (dump1090.h:179-184) */
    uint8_t  alert_valid, alert, spi_valid, spi;
    uint8_t  altitude_baro_valid, altitude_baro_unit;   /* unit: 0 UNIT_FEET, 1 UNIT_METERS */
    uint8_t  squawk_valid, es_pending;
    uint16_t squawk;             /* hex-coded octal digits, as dump1090 prints it (%04x) */
    int32_t  altitude_baro;      /* feet; INVALID_ALTITUDE = -9999 */
} adsbcu_fields_a;
int adsbcu_1090_decode_fields(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_a *out);

/* ---- Path A, next layer: the extended-squitter payload (second half) -------------------------
 * Replaces decodeExtendedSquitter and the nine decodeES* functions behind it (mode_s.c:798-1551) for DF17 / DF18
 * records: one record per message with everything the reference's struct modesMessage (dump1090.h:444-655) holds
 * after decodeModesMessage returns, for the fields the ES decoder can touch.  The values that depend on the
 * surveillance header are folded in (airground from CA, mode_s.c:636-657; source / reliable, mode_s.c:580-585), so a
 * caller needs adsbcu_fields_a only for the other downlink formats.  Records of other DFs come back zeroed (es = 0).
 * valid is the OR of the ADSBCU_ES_* bits == the reference's *_valid bit-fields; enumerations carry the reference's
 * values (dump1090.h:133-242).  Integer fields are bit-exact; the four float fields of metype 19 subtypes 1/2
 * (sqrtf / atan2 of the velocity components, mode_s.c:894-902) are IEEE results of the same expressions - sqrtf is
 * correctly rounded on both sides, the heading goes through the device's double atan2 (tests/test_gpu_fields_es.py
 * walks every (E/W, N/S) pair against glibc).
 * Pinned against the reference binary's own decodeModesMessage (oracle/ref_harness.c, third tap;
 * tests/golden/fields_es.npz). */
enum {
    ADSBCU_ES_ALTITUDE_BARO = 1u << 0,  ADSBCU_ES_ALTITUDE_GEOM = 1u << 1,  ADSBCU_ES_HEADING = 1u << 2,      ADSBCU_ES_GS = 1u << 3,
    ADSBCU_ES_IAS = 1u << 4,            ADSBCU_ES_TAS = 1u << 5,            ADSBCU_ES_BARO_RATE = 1u << 6,    ADSBCU_ES_GEOM_RATE = 1u << 7,
    ADSBCU_ES_SQUAWK = 1u << 8,         ADSBCU_ES_CALLSIGN = 1u << 9,       ADSBCU_ES_CPR = 1u << 10,         ADSBCU_ES_CATEGORY = 1u << 11,
    ADSBCU_ES_GEOM_DELTA = 1u << 12,    ADSBCU_ES_SPI = 1u << 13,           ADSBCU_ES_ALERT = 1u << 14,       ADSBCU_ES_EMERGENCY = 1u << 15,
    ADSBCU_ES_NIC_A = 1u << 16,         ADSBCU_ES_NIC_B = 1u << 17,         ADSBCU_ES_NIC_C = 1u << 18,       ADSBCU_ES_NIC_BARO = 1u << 19,
    ADSBCU_ES_NAC_P = 1u << 20,         ADSBCU_ES_NAC_V = 1u << 21,         ADSBCU_ES_GVA = 1u << 22,         ADSBCU_ES_SDA = 1u << 23,
    ADSBCU_ES_NAV_HEADING = 1u << 24,   ADSBCU_ES_NAV_FMS_ALTITUDE = 1u << 25, ADSBCU_ES_NAV_MCP_ALTITUDE = 1u << 26, ADSBCU_ES_NAV_QNH = 1u << 27,
    ADSBCU_ES_NAV_MODES = 1u << 28,     ADSBCU_ES_OPSTATUS = 1u << 29,      ADSBCU_ES_CC_LW = 1u << 30
};
/* opstatus flag bits (the one-bit members of modesMessage.opstatus, dump1090.h:596-610) */
enum {
    ADSBCU_OPS_OM_ACAS_RA = 1u << 0, ADSBCU_OPS_OM_IDENT = 1u << 1, ADSBCU_OPS_OM_ATC = 1u << 2,  ADSBCU_OPS_OM_SAF = 1u << 3,
    ADSBCU_OPS_CC_ACAS = 1u << 4,    ADSBCU_OPS_CC_CDTI = 1u << 5,  ADSBCU_OPS_CC_1090_IN = 1u << 6, ADSBCU_OPS_CC_ARV = 1u << 7,
    ADSBCU_OPS_CC_TS = 1u << 8,      ADSBCU_OPS_CC_UAT_IN = 1u << 9, ADSBCU_OPS_CC_POA = 1u << 10,  ADSBCU_OPS_CC_B2_LOW = 1u << 11
};
typedef struct adsbcu_fields_es {
    uint32_t addr;               /* AA, with MODES_NON_ICAO_ADDRESS (1 << 24) where the reference sets it */
    uint32_t valid;              /* ADSBCU_ES_* */
    uint8_t  es, metype, mesub, addrtype;            /* es: 1 for DF17 / DF18; addrtype_t (dump1090.h:147-161) */
    uint8_t  source, reliable, airground, CA_CF;     /* datasource_t; CA (DF17) or CF (DF18) */
    uint8_t  cpr_type, cpr_odd, heading_type, nav_heading_type;
    uint8_t  alert, spi, emergency, category;
    uint8_t  altitude_baro_unit, altitude_geom_unit, nav_altitude_source, nav_modes;
    char     callsign[8];        /* as decoded, also when ADSBCU_ES_CALLSIGN is off (the reference keeps the bad one too) */
    int32_t  altitude_baro, altitude_geom, geom_delta, baro_rate, geom_rate;
    uint32_t cpr_lat, cpr_lon;
    uint16_t ias, tas, squawk, ops_flags;            /* ops_flags: ADSBCU_OPS_* */
    float    heading, gs_v0, gs_v2, gs_selected;
    float    nav_heading, nav_qnh;
    int32_t  nav_fms_altitude, nav_mcp_altitude;
    uint8_t  nic_a, nic_b, nic_c, nic_baro, nac_p, nac_v, sil, sil_type, gva, sda;
    uint8_t  ops_version, ops_tah, ops_hrd, ops_cc_tc, ops_cc_lw, ops_cc_antenna_offset;
} adsbcu_fields_es;              /* 120 bytes */
/* HOST arrays of n records; the decode runs on the handle's device (one thread per message). */
int adsbcu_1090_decode_es(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_es *out);

/* ---- Path A, next layer: the Comm-B payload of DF20 / DF21 (third part) -----------------------
 * Replaces decodeCommB (comm_b.c:50-86) and its ten register decoders (comm_b.c:88-985): the MB field is scored as
 * an empty response, BDS 1,0 / 2,0 / 3,0 / 1,7 / 4,0 / 5,0 / 6,0 / 4,4 / 0,5 in the reference's order, the single
 * best-scoring interpretation is decoded, a tie gives COMMB_AMBIGUOUS, no positive score COMMB_UNKNOWN, and replies
 * with DR != 0 or repaired bits are COMMB_NOT_DECODED.  (The reference also tests UM, but decodeModesMessage fills
 * UM in only after it has called decodeCommB - mode_s.c:731-734 against :757-760 - so that test never fires; the
 * same here.)  One record per message, zeroed with commb_format = 0 for other downlink formats.  commb_format carries
 * commb_format_t (dump1090.h:203-217); valid is the OR of ADSBCU_MB_*.  Floats are the reference's expressions
 * narrowed as it narrows them (mach: the float the reference widens into its double member).
 * Pinned against the reference binary's own decodeModesMessage (oracle/ref_harness.c, third tap;
 * tests/golden/fields_mb.npz). */
enum {
    ADSBCU_MB_CALLSIGN = 1u << 0,   ADSBCU_MB_ROLL = 1u << 1,        ADSBCU_MB_HEADING = 1u << 2,      ADSBCU_MB_GS = 1u << 3,
    ADSBCU_MB_TRACK_RATE = 1u << 4, ADSBCU_MB_TAS = 1u << 5,         ADSBCU_MB_IAS = 1u << 6,          ADSBCU_MB_MACH = 1u << 7,
    ADSBCU_MB_BARO_RATE = 1u << 8,  ADSBCU_MB_GEOM_RATE = 1u << 9,   ADSBCU_MB_NAV_MCP_ALTITUDE = 1u << 10, ADSBCU_MB_NAV_FMS_ALTITUDE = 1u << 11,
    ADSBCU_MB_NAV_QNH = 1u << 12,   ADSBCU_MB_NAV_MODES = 1u << 13,  ADSBCU_MB_MRAR_SOURCE = 1u << 14, ADSBCU_MB_WIND = 1u << 15,
    ADSBCU_MB_TEMPERATURE = 1u << 16, ADSBCU_MB_PRESSURE = 1u << 17, ADSBCU_MB_TURBULENCE = 1u << 18,  ADSBCU_MB_HUMIDITY = 1u << 19
};
typedef struct adsbcu_fields_commb {
    uint32_t valid;              /* ADSBCU_MB_* */
    uint8_t  commb_format, heading_type, nav_modes, nav_altitude_source;
    uint8_t  mrar_source, turbulence, reserved[2];
    char     callsign[8];
    float    roll, heading, track_rate, gs, mach, nav_qnh;
    float    wind_speed, wind_dir, temperature, pressure, humidity;
    int32_t  baro_rate, geom_rate, nav_mcp_altitude, nav_fms_altitude;
    uint16_t ias, tas;
} adsbcu_fields_commb;           /* 84 bytes */
/* HOST arrays of n records; the decode runs on the handle's device (one thread per message). */
int adsbcu_1090_decode_commb(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_commb *out);

/* The three decoders above in ONE launch for records that are already in device memory (the message set of a long
 * replay kept on the GPU, or records a caller's own kernel produced): msgs_dev and the output arrays are DEVICE
 * pointers of n records each, 4-byte aligned; any of the three outputs may be NULL (at least one is not).  Nothing is
 * allocated, copied or waited for: the kernel is enqueued on cuda_stream (a cudaStream_t; NULL = the default stream)
 * and the results are there when that stream reaches it.  Same records, bit for bit, as the host-array entry points
 * (decodeModesMessage, mode_s.c:626-779 + :798-1551 + comm_b.c:50-985). */
int adsbcu_1090_decode_device(adsbcu_handle *h, const adsbcu_msg *msgs_dev, size_t n, adsbcu_fields_a *fields_dev,
                              adsbcu_fields_es *es_dev, adsbcu_fields_commb *commb_dev, void *cuda_stream);

/* ---- Wire formats of the decoded set ---------------------------------------------
 * Host-side formatting of adsbcu_msg records exactly as dump1090's network outputs write them; no handle, no GPU.
 *   adsbcu_1090_encode_beast   writeBeastMessage (net_io.c:440-487) per message.  verbatim != 0: the bytes as sliced,
 *                              every message (the --net-verbatim Beast port, net_io.c:405-412); verbatim == 0: corrected
 *                              bytes, messages with two repaired bits left out (cooked port, net_io.c:423-438).
 *   adsbcu_1090_encode_avr     the raw port's "@<12 hex timestamp><frame hex>;\n" line (net_io.c:510-545 with --mlat).
 * Both return the number of bytes the output needs; at most cap bytes are written (call with out = NULL to size it).
 * tests/golden/beast_ref.npz holds the bytes the reference itself sent on its Beast port for the golden capture. */
size_t adsbcu_1090_encode_beast(const adsbcu_msg *msgs, size_t n, int verbatim, uint8_t *out, size_t cap);
size_t adsbcu_1090_encode_avr(const adsbcu_msg *msgs, size_t n, uint8_t *out, size_t cap);

/* ---- Path A, next layer: CPR position decode ----------------------------------
 * The three decoders of the flightaware tree's cpr.c, many records per call, results as the same IEEE
 * doubles the reference computes (bit for bit; dump1090's own cprtests.c vectors are in tests/golden):
 *   mode ADSBCU_CPR_AIRBORNE           decodeCPRairborne(even_lat, even_lon, odd_lat, odd_lon, fflag, ...)   cpr.c:162-214
 *   mode ADSBCU_CPR_SURFACE            decodeCPRsurface(reflat, reflon, even.., odd.., fflag, ...)            cpr.c:216-311
 *   mode ADSBCU_CPR_RELATIVE_AIRBORNE  decodeCPRrelative(reflat, reflon, cprlat, cprlon, fflag, 0, ...)      cpr.c:325-369
 *   mode ADSBCU_CPR_RELATIVE_SURFACE   decodeCPRrelative(reflat, reflon, cprlat, cprlon, fflag, 1, ...)
 * For the relative modes the single frame's raw values travel in even_lat / even_lon whatever its parity.
 * Which decoder to call, with which reference position and how to vet the answer stays with the tracker
 * (track.c:doGlobalCPR / doLocalCPR). */
enum { ADSBCU_CPR_AIRBORNE = 0, ADSBCU_CPR_SURFACE = 1, ADSBCU_CPR_RELATIVE_AIRBORNE = 2, ADSBCU_CPR_RELATIVE_SURFACE = 3 };
typedef struct adsbcu_cpr_in {
    double  reflat, reflon;             /* reference position, degrees (unused for ADSBCU_CPR_AIRBORNE) */
    int32_t even_lat, even_lon;         /* 17-bit raw CPR values of the even frame (or of the only frame) */
    int32_t odd_lat, odd_lon;           /* ... of the odd frame */
    int32_t fflag;                      /* non-zero: the odd frame is the latest / the only frame is odd */
    int32_t mode;                       /* ADSBCU_CPR_* */
} adsbcu_cpr_in;                        /* 40 bytes */
typedef struct adsbcu_cpr_out {
    double  lat, lon;                   /* *out_lat, *out_lon; 0 unless result == 0 */
    int32_t result;                     /* the reference's return value: 0 ok, -1 zone crossing / ambiguous, -2 bad data */
    int32_t reserved;
} adsbcu_cpr_out;                       /* 24 bytes */
/* out[k] = decoder(in[k]); HOST arrays of n records; an unknown mode is ADSBCU_ERR_ARG for the whole call. */
int adsbcu_1090_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_in *in, size_t n, adsbcu_cpr_out *out);

/* ---- Path B: the 1090 MHz demodulator libadsb itself compiles (2.0 MS/s) --------------------
 * Replaces the body of ADSB1090Handler::HandleData (reference ADSB1090.cpp:158-175) down to the
 * point where the reference calls UseModesMessage() with mm.crcok set (ADSB1090.cpp:741-959,
 * 491-596, 266-435): LUT magnitude, preamble test, plain and phase-corrected decode attempts, mean
 * delta filter, CRC-24 check, single-bit repair of DF11/DF17, Address/Parity recovery against the
 * 60 s cache of recently seen ICAO addresses.  Field decoding and CPR (ADSB1090.cpp:597-672, 1079-1121) have their
 * own bulk entry points below; the aircraft table and the IListener callback stay in the caller.  Fixed reference configuration
 * (ADSB1090.cpp:111-122): fixErrors, checkCRC, not aggressive. */
#pragma pack(push, 1)
typedef struct adsbcu_msg_b {
    uint32_t pos;              /* j: sample index of the preamble inside this buffer */
    uint32_t crc;              /* mm.crc */
    uint32_t addr;             /* aa1<<16 | aa2<<8 | aa3 (recovered address for Address/Parity formats) */
    uint8_t  msg[14];          /* mm.msg, after single-bit repair */
    uint8_t  msgbits;          /* 56 or 112 */
    uint8_t  msgtype;          /* DF as sliced (a repair does not update it, ADSB1090.cpp:501,513-519) */
    int8_t   errorbit;         /* repaired bit, or -1 */
    uint8_t  corrected;        /* 1: decoded on the phase-correction retry (ADSB1090.cpp:817-826) */
    uint16_t reserved;
} adsbcu_msg_b;
#pragma pack(pop)

/* One call == one HandleData(span) call: `iq` is nbytes of interleaved uint8 I/Q (HOST memory);
 * no sample state is carried between calls (the reference drops messages straddling buffers,
 * ADSB1090.cpp:772); only the ICAO cache persists.  Messages come back in buffer order. */
int adsbcu_1090b_feed(adsbcu_handle *h, const uint8_t *iq, size_t nbytes,
                      adsbcu_msg_b *out, size_t cap, size_t *n_out);
/* Forget the recently-seen ICAO cache (a fresh ADSB1090Handler). */
int adsbcu_1090b_reset(adsbcu_handle *h);

/* ---- Path B, next layer: bulk field decode and CPR ---------------------------
 * What DecodeModesMessage does after the CRC stage (ADSB1090.cpp:525-672) and what
 * InteractiveReceiveData hands to DecodeCpr (ADSB1090.cpp:986-1121), for many frames per call.  Both
 * are pure functions of their input: no handle state is read or written.  Field names follow
 * `struct Message` (ADSB1090.cpp:30-107); aa1..aa3 are adsbcu_msg_b::addr already. */
typedef struct adsbcu_fields_b {
    int32_t ca, metype, mesub;          /* mm.ca, mm.metype, mm.mesub */
    int32_t heading;                    /* mm.heading: DF17 type 19, degrees 0..359 */
    int32_t fflag;                      /* mm.fflag: msg[6] & 4, non-zero = odd CPR frame */
    int32_t raw_latitude, raw_longitude;/* 17-bit CPR values */
    char    flight[8];                  /* mm.flight, all zero unless DF17 type 1..4 */
    int32_t ew_dir, ew_velocity, ns_dir, ns_velocity, velocity;
    int32_t fs, dr, um, identity, altitude;
    int32_t unit;                       /* Message::ModeSUnit: 0 feet, 1 metres */
} adsbcu_fields_b;                      /* 80 bytes */

/* out[k] = the fields of msgs[k].msg (HOST arrays of n records).  The input frames are the ones
 * adsbcu_1090b_feed returned, i.e. after the CRC stage's repair. */
int adsbcu_1090b_decode_fields(adsbcu_handle *h, const adsbcu_msg_b *msgs, size_t n, adsbcu_fields_b *out);

typedef struct adsbcu_cpr_pair {
    int32_t even_lat, even_lon;         /* a.cprEvenLat / a.cprEvenLon */
    int32_t odd_lat, odd_lon;           /* a.cprOddLat / a.cprOddLon */
    int32_t even_is_newer;              /* a.cprEvenTime > a.cprOddTime */
} adsbcu_cpr_pair;
typedef struct adsbcu_cpr_fix {
    int32_t lat1e7, lon1e7;             /* a.lat1E7 / a.lon1E7; untouched (0) when !valid */
    int32_t valid;                      /* 0: the two frames lie in different latitude zones, DecodeCpr returned early */
} adsbcu_cpr_fix;
/* out[k] = DecodeCpr on pairs[k] (HOST arrays).  The caller keeps the reference's 10 s pairing rule
 * (ADSB1090.cpp:1161).  IEEE double arithmetic in the reference's operation order; a latitude outside the
 * int32 range converts to INT32_MIN like the reference's x86-64 build. */
int adsbcu_1090b_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_pair *pairs, size_t n, adsbcu_cpr_fix *out);


/* ---- Path C: 978 MHz UAT -----------------------------------------------------------------------
 * Replaces the body of UAT978Handler::HandleData (reference UAT978.cpp:43-60): the phase-table
 * mapping (UAT978.cpp:52,76-100), the 65536-sample staging buffer with carry-over (:48-59) and the
 * dump978 entry points it calls through its extern "C" FFI, `int process_buffer(uint16_t const*,
 * int len, uint64_t offset)` and `void init_fec()` (UAT978.cpp:9-10).  Every frame the reference
 * would pass to `dump_raw_message(char updown, uint8_t *data, int len, int rs_errors)`
 * (uat2json-wrapper.cpp:14) comes back as one adsbcu_uat_frame, in stream order.  dump978 itself
 * is an absent submodule of the reference: parity for this path is unpinned. */
#pragma pack(push, 1)
typedef struct adsbcu_uat_frame {
    uint64_t offset;           /* stream index of the sample holding the first sync bit */
    int16_t  rs_errors;        /* Reed-Solomon symbols corrected (sum over the six blocks for uplink) */
    uint16_t len;              /* 18 (basic ADS-B), 34 (long ADS-B) or 432 (ground uplink) payload bytes */
    uint8_t  updown;           /* '-' ADS-B downlink, '+' ground uplink */
    uint8_t  reserved[3];
    uint8_t  data[432];
} adsbcu_uat_frame;
#pragma pack(pop)

/* One call == one UAT978Handler::HandleData(span) call on `nbytes` of interleaved uint8 I/Q
 * sampled at 2.083334 MS/s (HOST memory).  The stream continues across calls. */
int adsbcu_978_feed(adsbcu_handle *h, const uint8_t *iq, size_t nbytes,
                    adsbcu_uat_frame *out, size_t cap, size_t *n_out);
/* The same with the samples already in DEVICE memory (resident replay; bench.py's `value`). */
int adsbcu_978_feed_device(adsbcu_handle *h, const uint8_t *iq_dev, size_t nbytes,
                           adsbcu_uat_frame *out, size_t cap, size_t *n_out);
typedef struct adsbcu_978_stats {
    float    ms_device;        /* device time of the last feed: upload + phase kernel + every process_buffer() call's kernels */
    uint32_t calls;            /* process_buffer() calls the feed made (UAT978.cpp:54) */
    uint32_t launches;         /* kernels launched */
    uint32_t reserved;
} adsbcu_978_stats;
int adsbcu_978_get_stats(const adsbcu_handle *h, adsbcu_978_stats *out);
/* A fresh UAT978Handler: forget the staged samples and the stream offset. */
int adsbcu_978_reset(adsbcu_handle *h);
/* Stage-level entry: phi_out[k] = iqphase[{I,Q}] (UAT978.cpp:52) for k < nsamples, HOST pointers. */
int adsbcu_978_phase(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, uint16_t *phi_out);

#ifdef __cplusplus
}
#endif
#endif
