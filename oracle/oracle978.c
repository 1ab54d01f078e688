/*
 * TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path
 * (libadsb_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it.
 *
 * oracle978: plain-C CPU restatement of the 978 MHz UAT path behind libadsb's
 * UAT978Handler::HandleData (reference UAT978.cpp:43-60,76-100).
 *
 * Parity status: stages C1 (InitATan2Table) and C2 (HandleData staging / carry) are PINNED to the reference's own
 * UAT978.cpp compiled where it lies (oracle/_ref/libref978.so, oracle/ref_harness_978.cpp; reference-generated
 * fixture tests/golden/uat_ref.npz): all 65536 table entries and the exact sequence of process_buffer(len, offset)
 * calls.  Stages C3-C5 (sync search, frame demodulation, Reed-Solomon) remain PARITY UNPINNED: their arithmetic
 * lives in a dependency
 * that is absent from /root/reference: the `dump978` submodule, pinned by
 * .SUBMODULES.json to github.com/ankurvdev/dump978 @ a461b30732162bfe1424d37ba4475e0db62b893b
 * (a fork of flightaware/dump978; the reference builds legacy/dump978.c, legacy/fec.c,
 * libs/fec/decode_rs_char.c, libs/fec/init_rs_char.c -- CMakeLists.txt:88-95).  What is in
 * the tree pins only the call sites: UAT978.cpp:9-10 (init_fec, process_buffer),
 * UAT978.cpp:43-60 (staging buffer, carry-over), UAT978.cpp:76-100 (phase table),
 * uat2json-wrapper.cpp:14 (dump_raw_message sink).  process_buffer, check_sync_word,
 * demod_*_frame, correct_*_frame and the Reed-Solomon codec below restate the published
 * dump978 "legacy" demodulator and Phil Karn's libfec char decoder; the reference's
 * recorded captures and expected outputs for 978 MHz have no inputs in the tree
 * (tests/testdata/TestEnv_rtlsdr_978_*.txt are aircraft-level text only).  The checks that
 * exist are: DO-282B frame geometry, RS encode->corrupt->decode round trips with exact
 * correction counts, and modulate->demodulate round trips (tests/test_oracle_978.py).
 */
#define _DEFAULT_SOURCE
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NN 255
#define A0 255
#define SYNC_BITS 36
#define ADSB_SYNC_WORD 0xEACDDA4E2ULL
#define UPLINK_SYNC_WORD 0x153225B1DULL
#define SHORT_FRAME_DATA_BYTES 18
#define SHORT_FRAME_BYTES 30
#define LONG_FRAME_DATA_BYTES 34
#define LONG_FRAME_BYTES 48
#define LONG_FRAME_BITS (LONG_FRAME_BYTES * 8)
#define SHORT_FRAME_BITS (SHORT_FRAME_BYTES * 8)
#define UPLINK_BLOCK_DATA_BYTES 72
#define UPLINK_BLOCK_BYTES 92
#define UPLINK_FRAME_BLOCKS 6
#define UPLINK_FRAME_DATA_BYTES (UPLINK_FRAME_BLOCKS * UPLINK_BLOCK_DATA_BYTES)
#define UPLINK_FRAME_BYTES (UPLINK_FRAME_BLOCKS * UPLINK_BLOCK_BYTES)
#define UPLINK_FRAME_BITS (UPLINK_FRAME_BYTES * 8)
#define MAX_SYNC_ERRORS 4

/* ---- Reed-Solomon over GF(256), libfec "char" codec semantics ------------------------------- */
typedef struct rs {
    uint8_t alpha_to[256], index_of[256], genpoly[33];
    int nroots, fcr, prim, iprim, pad;
} rs_t;

static int modnn(int x)
{
    while (x >= NN) { x -= NN; x = (x >> 8) + (x & NN); }
    return x;
}

static void rs_init(rs_t *rs, int gfpoly, int fcr, int prim, int nroots, int pad)
{
    int i, j, sr = 1, root, iprim;
    rs->index_of[0] = A0;
    rs->alpha_to[A0] = 0;
    for (i = 0; i < NN; i++) {
        rs->index_of[sr] = (uint8_t)i;
        rs->alpha_to[i] = (uint8_t)sr;
        sr <<= 1;
        if (sr & 256) sr ^= gfpoly;
        sr &= NN;
    }
    rs->fcr = fcr; rs->prim = prim; rs->nroots = nroots; rs->pad = pad;
    for (iprim = 1; (iprim % prim) != 0; iprim += NN) {}
    rs->iprim = iprim / prim;
    rs->genpoly[0] = 1;
    for (i = 0, root = fcr * prim; i < nroots; i++, root += prim) {
        rs->genpoly[i + 1] = 1;
        for (j = i; j > 0; j--) {
            if (rs->genpoly[j] != 0) rs->genpoly[j] = rs->genpoly[j - 1] ^ rs->alpha_to[modnn(rs->index_of[rs->genpoly[j]] + root)];
            else rs->genpoly[j] = rs->genpoly[j - 1];
        }
        rs->genpoly[0] = rs->alpha_to[modnn(rs->index_of[rs->genpoly[0]] + root)];
    }
    for (i = 0; i <= nroots; i++) rs->genpoly[i] = rs->index_of[rs->genpoly[i]];
}

static void rs_encode(const rs_t *rs, const uint8_t *data, uint8_t *bb)
{
    int i, j, nroots = rs->nroots;
    memset(bb, 0, (size_t)nroots);
    for (i = 0; i < NN - nroots - rs->pad; i++) {
        const int feedback = rs->index_of[data[i] ^ bb[0]];
        if (feedback != A0)
            for (j = 1; j < nroots; j++) bb[j] ^= rs->alpha_to[modnn(feedback + rs->genpoly[nroots - j])];
        memmove(&bb[0], &bb[1], (size_t)(nroots - 1));
        bb[nroots - 1] = feedback != A0 ? rs->alpha_to[modnn(feedback + rs->genpoly[0])] : 0;
    }
}

/* returns the number of corrected symbols, or -1 (data untouched) */
static int rs_decode(const rs_t *rs, uint8_t *data)
{
    const int nroots = rs->nroots, fcr = rs->fcr, prim = rs->prim, pad = rs->pad;
    int deg_lambda, el, deg_omega, i, j, r, k, q, tmp, num1, num2, den, discr_r, syn_error, count;
    uint8_t lambda[33], s[32], b[33], t[33], omega[33], root[32], reg[33], loc[32];

    for (i = 0; i < nroots; i++) s[i] = data[0];
    for (j = 1; j < NN - pad; j++)
        for (i = 0; i < nroots; i++) {
            if (s[i] == 0) s[i] = data[j];
            else s[i] = data[j] ^ rs->alpha_to[modnn(rs->index_of[s[i]] + (fcr + i) * prim)];
        }
    syn_error = 0;
    for (i = 0; i < nroots; i++) { syn_error |= s[i]; s[i] = rs->index_of[s[i]]; }
    if (!syn_error) return 0;

    memset(&lambda[1], 0, (size_t)nroots);
    lambda[0] = 1;
    for (i = 0; i < nroots + 1; i++) b[i] = rs->index_of[lambda[i]];
    r = 0; el = 0;
    while (++r <= nroots) {
        discr_r = 0;
        for (i = 0; i < r; i++)
            if (lambda[i] != 0 && s[r - i - 1] != A0) discr_r ^= rs->alpha_to[modnn(rs->index_of[lambda[i]] + s[r - i - 1])];
        discr_r = rs->index_of[discr_r];
        if (discr_r == A0) {
            memmove(&b[1], b, (size_t)nroots);
            b[0] = A0;
        } else {
            t[0] = lambda[0];
            for (i = 0; i < nroots; i++) {
                if (b[i] != A0) t[i + 1] = lambda[i + 1] ^ rs->alpha_to[modnn(discr_r + b[i])];
                else t[i + 1] = lambda[i + 1];
            }
            if (2 * el <= r - 1) {
                el = r - el;
                for (i = 0; i <= nroots; i++) b[i] = (lambda[i] == 0) ? A0 : (uint8_t)modnn(rs->index_of[lambda[i]] - discr_r + NN);
            } else {
                memmove(&b[1], b, (size_t)nroots);
                b[0] = A0;
            }
            memcpy(lambda, t, (size_t)(nroots + 1));
        }
    }
    deg_lambda = 0;
    for (i = 0; i < nroots + 1; i++) {
        lambda[i] = rs->index_of[lambda[i]];
        if (lambda[i] != A0) deg_lambda = i;
    }
    memcpy(&reg[1], &lambda[1], (size_t)nroots);
    count = 0;
    for (i = 1, k = rs->iprim - 1; i <= NN; i++, k = modnn(k + rs->iprim)) {
        q = 1;
        for (j = deg_lambda; j > 0; j--)
            if (reg[j] != A0) { reg[j] = (uint8_t)modnn(reg[j] + j); q ^= rs->alpha_to[reg[j]]; }
        if (q != 0) continue;
        root[count] = (uint8_t)i;
        loc[count] = (uint8_t)k;
        if (++count == deg_lambda) break;
    }
    if (deg_lambda != count) return -1;
    deg_omega = deg_lambda - 1;
    for (i = 0; i <= deg_omega; i++) {
        tmp = 0;
        for (j = i; j >= 0; j--)
            if (s[i - j] != A0 && lambda[j] != A0) tmp ^= rs->alpha_to[modnn(s[i - j] + lambda[j])];
        omega[i] = rs->index_of[tmp];
    }
    for (j = count - 1; j >= 0; j--) {
        num1 = 0;
        for (i = deg_omega; i >= 0; i--)
            if (omega[i] != A0) num1 ^= rs->alpha_to[modnn(omega[i] + i * root[j])];
        num2 = rs->alpha_to[modnn(root[j] * (fcr - 1) + NN)];
        den = 0;
        for (i = (deg_lambda < nroots - 1 ? deg_lambda : nroots - 1) & ~1; i >= 0; i -= 2)
            if (lambda[i + 1] != A0) den ^= rs->alpha_to[modnn(lambda[i + 1] + i * root[j])];
        if (num1 != 0 && loc[j] >= pad)
            data[loc[j] - pad] ^= rs->alpha_to[modnn(rs->index_of[num1] + rs->index_of[num2] + NN - rs->index_of[den])];
    }
    return count;
}

/* ---- fec.c of dump978: the three UAT codes (GF poly 0x187, fcr 120, prim 1) ---------------------- */
static rs_t rs_adsb_short, rs_adsb_long, rs_uplink;
static int fec_ready = 0;
static void init_fec(void)
{
    if (fec_ready) return;
    rs_init(&rs_adsb_short, 0x187, 120, 1, 12, 225);   /* RS(30,18) */
    rs_init(&rs_adsb_long, 0x187, 120, 1, 14, 207);    /* RS(48,34) */
    rs_init(&rs_uplink, 0x187, 120, 1, 20, 163);       /* RS(92,72) */
    fec_ready = 1;
}

/* returns 2 long, 1 basic, -1 failed.  Relies on rs_decode not touching the data on failure. */
static int correct_adsb_frame(uint8_t *to, int *rs_errors)
{
    int n = rs_decode(&rs_adsb_long, to);
    if (n >= 0 && n <= 7 && (to[0] >> 3) != 0) { *rs_errors = n; return 2; }
    n = rs_decode(&rs_adsb_short, to);
    if (n >= 0 && n <= 6 && (to[0] >> 3) == 0) { *rs_errors = n; return 1; }
    *rs_errors = 9999;
    return -1;
}

/* de-interleave (byte i of block b = from[i*6+b]) and correct the six blocks; `to` needs 432+20 bytes */
static int correct_uplink_frame(const uint8_t *from, uint8_t *to, int *rs_errors)
{
    int block, i, total = 0;
    for (block = 0; block < UPLINK_FRAME_BLOCKS; ++block) {
        uint8_t *blockdata = &to[block * UPLINK_BLOCK_DATA_BYTES];
        for (i = 0; i < UPLINK_BLOCK_BYTES; ++i) blockdata[i] = from[i * UPLINK_FRAME_BLOCKS + block];
        const int n = rs_decode(&rs_uplink, blockdata);
        if (n < 0 || n > 10) { *rs_errors = 9999; return -1; }
        total += n;   /* the next block overwrites this block's parity bytes */
    }
    *rs_errors = total;
    return 1;
}

/* ---- dump978.c demodulator --------------------------------------------------------------------- */
static int16_t phi_difference(uint16_t from, uint16_t to)
{
    int32_t d = (int32_t)to - (int32_t)from;
    if (d >= 32768) return (int16_t)(d - 65536);
    if (d < -32768) return (int16_t)(d + 65536);
    return (int16_t)d;
}

static int sync_word_fuzzy_compare(uint64_t word, uint64_t expected)
{
    uint64_t diff = word ^ expected;
    int k;
    for (k = 0; k < MAX_SYNC_ERRORS && diff; ++k) diff &= diff - 1;
    return diff == 0;      /* at most 4 bits differ */
}

static int check_sync_word(const uint16_t *phi, uint64_t pattern, int16_t *center)
{
    int i, zero_bits = 0, one_bits = 0, error_bits = 0;
    int32_t zero_total = 0, one_total = 0;
    for (i = 0; i < SYNC_BITS; ++i) {
        const int16_t dphi = phi_difference(phi[i * 2], phi[i * 2 + 1]);
        if (pattern & (1ULL << (35 - i))) { ++one_bits; one_total += dphi; }
        else { ++zero_bits; zero_total += dphi; }
    }
    zero_total /= zero_bits;
    one_total /= one_bits;
    *center = (int16_t)((one_total + zero_total) / 2);
    for (i = 0; i < SYNC_BITS; ++i) {
        const int16_t dphi = phi_difference(phi[i * 2], phi[i * 2 + 1]);
        if (pattern & (1ULL << (35 - i))) { if (dphi < *center) ++error_bits; }
        else { if (dphi >= *center) ++error_bits; }
    }
    return error_bits <= MAX_SYNC_ERRORS;
}

static void demod_frame(const uint16_t *phi, uint8_t *frame, int bytes, int16_t center)
{
    while (--bytes >= 0) {
        uint8_t b = 0;
        int k;
        for (k = 0; k < 8; ++k)
            if (phi_difference(phi[2 * k], phi[2 * k + 1]) > center) b |= (uint8_t)(0x80 >> k);
        *frame++ = b;
        phi += 16;
    }
}

/* bits consumed, 0 on failure */
static int demod_adsb_frame(const uint16_t *phi, uint8_t *to, int *rs_errors)
{
    int16_t center;
    if (!check_sync_word(phi, ADSB_SYNC_WORD, &center)) { *rs_errors = 9999; return 0; }
    demod_frame(phi + SYNC_BITS * 2, to, LONG_FRAME_BYTES, center);
    const int type = correct_adsb_frame(to, rs_errors);
    if (type == 2) return SYNC_BITS + LONG_FRAME_BITS;
    if (type == 1) return SYNC_BITS + SHORT_FRAME_BITS;
    return 0;
}

static int demod_uplink_frame(const uint16_t *phi, uint8_t *to, int *rs_errors)
{
    int16_t center;
    uint8_t interleaved[UPLINK_FRAME_BYTES];
    if (!check_sync_word(phi, UPLINK_SYNC_WORD, &center)) { *rs_errors = 9999; return 0; }
    demod_frame(phi + SYNC_BITS * 2, interleaved, UPLINK_FRAME_BYTES, center);
    if (correct_uplink_frame(interleaved, to, rs_errors) == 1) return UPLINK_FRAME_BITS + SYNC_BITS;
    return 0;
}

#pragma pack(push, 1)
typedef struct oracle978_frame {
    uint64_t offset;      /* stream index of the sample holding the first sync bit */
    int16_t  rs_errors;   /* symbols corrected (sum over the six blocks for uplink) */
    uint16_t len;         /* 18, 34 or 432 payload bytes: what dump_raw_message() is handed */
    uint8_t  updown;      /* '-' ADS-B (downlink), '+' ground uplink */
    uint8_t  pad[3];
    uint8_t  data[UPLINK_FRAME_DATA_BYTES];
} oracle978_frame;
#pragma pack(pop)

typedef struct oracle978 {
    uint16_t iqphase[65536];
    uint16_t buffer[65536];     /* UAT978.cpp:108 */
    size_t used;
    uint64_t offset;
    oracle978_frame *out;
    size_t cap, nout;
    int64_t calls[3 * 4096];    /* (len, offset, returned) of the process_buffer() calls HandleData made: compared with the
                                   calls the compiled reference makes (oracle/_ref/libref978.so) */
    size_t ncalls;
} oracle978;

static void emit(oracle978 *o, uint8_t updown, const uint8_t *data, int len, int rs_errors, uint64_t offset)
{
    if (o->nout < o->cap) {
        oracle978_frame *f = &o->out[o->nout];
        memset(f, 0, sizeof(*f));
        f->offset = offset; f->rs_errors = (int16_t)rs_errors; f->len = (uint16_t)len; f->updown = updown;
        memcpy(f->data, data, (size_t)len);
    }
    o->nout++;
}

/* dump978.c process_buffer(): scan `len` phase samples, two bit alignments, for the two sync words;
 * returns the samples consumed; the caller presents the rest again (UAT978.cpp:54-58). */
static int process_buffer(oracle978 *o, const uint16_t *phi, int len, uint64_t offset)
{
    uint64_t sync0 = 0, sync1 = 0;
    const uint64_t mask = (1ULL << SYNC_BITS) - 1;
    const int lenbits = len / 2 - (SYNC_BITS + UPLINK_FRAME_BITS);
    int bit;
    for (bit = 0; bit < lenbits; ++bit) {
        const int16_t dphi0 = phi_difference(phi[bit * 2], phi[bit * 2 + 1]);
        const int16_t dphi1 = phi_difference(phi[bit * 2 + 1], phi[bit * 2 + 2]);
        sync0 = ((sync0 << 1) | (dphi0 > 0 ? 1 : 0)) & mask;
        sync1 = ((sync1 << 1) | (dphi1 > 0 ? 1 : 0)) & mask;
        if (bit < SYNC_BITS) continue;
        const int a0 = sync_word_fuzzy_compare(sync0, ADSB_SYNC_WORD), a1 = sync_word_fuzzy_compare(sync1, ADSB_SYNC_WORD);
        if (a0 || a1) {
            const int startbit = bit - SYNC_BITS + 1, index = startbit * 2 + (a0 ? 0 : 1);
            uint8_t f0[LONG_FRAME_BYTES], f1[LONG_FRAME_BYTES];
            int rs0 = -1, rs1 = -1;
            const int skip0 = demod_adsb_frame(phi + index, f0, &rs0), skip1 = demod_adsb_frame(phi + index + 1, f1, &rs1);
            if (skip0 && rs0 <= rs1) {
                emit(o, '-', f0, (f0[0] >> 3) == 0 ? SHORT_FRAME_DATA_BYTES : LONG_FRAME_DATA_BYTES, rs0, offset + (uint64_t)index);
                bit = startbit + skip0;
                continue;
            } else if (skip1 && rs1 <= rs0) {
                emit(o, '-', f1, (f1[0] >> 3) == 0 ? SHORT_FRAME_DATA_BYTES : LONG_FRAME_DATA_BYTES, rs1, offset + (uint64_t)index + 1);
                bit = startbit + skip1;
                continue;
            }
        }
        const int u0 = sync_word_fuzzy_compare(sync0, UPLINK_SYNC_WORD), u1 = sync_word_fuzzy_compare(sync1, UPLINK_SYNC_WORD);
        if (u0 || u1) {
            const int startbit = bit - SYNC_BITS + 1, index = startbit * 2 + (u0 ? 0 : 1);
            uint8_t f0[UPLINK_FRAME_DATA_BYTES + 20], f1[UPLINK_FRAME_DATA_BYTES + 20];
            int rs0 = -1, rs1 = -1;
            const int skip0 = demod_uplink_frame(phi + index, f0, &rs0), skip1 = demod_uplink_frame(phi + index + 1, f1, &rs1);
            if (skip0 && rs0 <= rs1) {
                emit(o, '+', f0, UPLINK_FRAME_DATA_BYTES, rs0, offset + (uint64_t)index);
                bit = startbit + skip0;
                continue;
            } else if (skip1 && rs1 <= rs0) {
                emit(o, '+', f1, UPLINK_FRAME_DATA_BYTES, rs1, offset + (uint64_t)index + 1);
                bit = startbit + skip1;
                continue;
            }
        }
    }
    return (bit - SYNC_BITS) * 2;
}

/* ---- C ABI used by the tests --------------------------------------------------------------------- */
oracle978 *oracle978_create(void)
{
    oracle978 *o = calloc(1, sizeof(*o));
    unsigned i, q;
    if (!o) return NULL;
    init_fec();
    for (i = 0; i < 256; ++i) {            /* UAT978.cpp:76-100 */
        const double dI = i - 127.5;
        for (q = 0; q < 256; ++q) {
            const double dQ = q - 127.5;
            const double ang = atan2(dQ, dI) + M_PI;
            const double scaled = round(32768 * ang / M_PI);
            o->iqphase[i | (q << 8)] = (uint16_t)(scaled < 0 ? 0 : scaled > 65535 ? 65535 : scaled);
        }
    }
    return o;
}
void oracle978_destroy(oracle978 *o) { free(o); }
void oracle978_phase_table(const oracle978 *o, uint16_t out[65536]) { memcpy(out, o->iqphase, sizeof(o->iqphase)); }

/* UAT978Handler::HandleData (UAT978.cpp:43-60) on `nbytes` of interleaved uint8 I/Q */
size_t oracle978_handle_data(oracle978 *o, const uint8_t *iq, size_t nbytes, oracle978_frame *out, size_t cap)
{
    const size_t n = nbytes / 2;
    size_t j = 0, i;
    o->out = out; o->cap = cap; o->nout = 0;
    while (j < n) {
        for (i = o->used; i < 65536 && j < n; i++, j++) o->buffer[i] = o->iqphase[iq[2 * j] | (iq[2 * j + 1] << 8)];
        if ((int)i / 2 - (SYNC_BITS + UPLINK_FRAME_BITS) < SYNC_BITS) {   /* fewer than 36 scannable bits: see below */
            if (o->ncalls < 4096) { o->calls[3 * o->ncalls] = (int64_t)i; o->calls[3 * o->ncalls + 1] = (int64_t)o->offset; o->calls[3 * o->ncalls + 2] = 0; }
            o->ncalls++;
            o->used = i;
            continue;
        }   /* nothing scannable yet (the
            reference's buffers are 262144 bytes, so it never gets here; its return value would be negative) */
        const int done = process_buffer(o, o->buffer, (int)i, o->offset);
        if (o->ncalls < 4096) { o->calls[3 * o->ncalls] = (int64_t)i; o->calls[3 * o->ncalls + 1] = (int64_t)o->offset; o->calls[3 * o->ncalls + 2] = done; }
        o->ncalls++;
        o->offset += (uint64_t)done;
        /* UAT978.cpp:57 passes the number of remaining SAMPLES as memmove's BYTE count: only the first half of the
         * unconsumed samples moves to the front; indices [(i-done)/2, i-done) keep what they held before (for an odd
         * remainder one sample gets its low byte only).  The compiled reference does exactly this
         * (oracle/_ref/libref978.so, tests/golden/uat_ref.npz), so frames that reach into that stretch are seen
         * corrupted, and a drop-in has to see them the same way. */
        memmove(o->buffer, o->buffer + done, (size_t)((int)i - done));
        o->used = (size_t)((int)i - done);
    }
    return o->nout;
}

/* the calls logged so far: 3 int64 each; returns their number (the log keeps the first 4096) */
size_t oracle978_calls(const oracle978 *o, int64_t *out, size_t cap)
{
    const size_t n = o->ncalls < 4096 ? o->ncalls : 4096;
    memcpy(out, o->calls, 3 * sizeof(int64_t) * (n < cap ? n : cap));
    return o->ncalls;
}
void oracle978_state(const oracle978 *o, uint64_t out[2]) { out[0] = o->used; out[1] = o->offset; }

/* one process_buffer() call on explicit phase samples (the extern "C" entry of UAT978.cpp:10) */
int oracle978_process_buffer(oracle978 *o, const uint16_t *phi, int len, uint64_t offset, oracle978_frame *out, size_t cap, size_t *nout)
{
    o->out = out; o->cap = cap; o->nout = 0;
    const int done = len / 2 - (SYNC_BITS + UPLINK_FRAME_BITS) >= SYNC_BITS ? process_buffer(o, phi, len, offset) : 0;
    *nout = o->nout;
    return done;
}

/* codec entry points for the round-trip tests: kind 0 = RS(30,18), 1 = RS(48,34), 2 = RS(92,72) */
static const rs_t *code(int kind) { init_fec(); return kind == 0 ? &rs_adsb_short : kind == 1 ? &rs_adsb_long : &rs_uplink; }
void oracle978_rs_encode(int kind, const uint8_t *data, uint8_t *parity) { rs_encode(code(kind), data, parity); }
int oracle978_rs_decode(int kind, uint8_t *block) { return rs_decode(code(kind), block); }
int oracle978_correct_adsb(uint8_t frame[48], int *rs_errors) { init_fec(); return correct_adsb_frame(frame, rs_errors); }
int oracle978_correct_uplink(const uint8_t from[552], uint8_t to[452], int *rs_errors) { init_fec(); return correct_uplink_frame(from, to, rs_errors); }
