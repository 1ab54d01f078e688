/*
 * TEST INFRASTRUCTURE ONLY -- see oracle1090.h.  CPU restatement of the
 * reference's 1090 MHz path; every function cites the reference lines it
 * follows (paths relative to /root/reference).  Written from the behaviour of
 * that code, not copied from it.
 */
#define _DEFAULT_SOURCE            /* M_SQRT2 under -std=c11 */
#include "oracle1090.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define OVERLAP        326u      /* dump1090/dump1090.c:159 (8+112+16 us at 2.4 MS/s) */
#define DEFAULT_BLOCK  131072u   /* dump1090/dump1090.h:92-94 */
#define FILTER_SLOTS   4096u     /* dump1090/icao_filter.c:23 */
#define FILTER_EMPTY   0xFFFFFFFFu
#define NT_FLAG        (1u << 25) /* dump1090/icao_filter.h:24 */
#define NON_ICAO       (1u << 24) /* dump1090/dump1090.h:262 */
#define UNCHECKED      0xFFFFFFFFu

/* score ranks, dump1090/mode_s.h:32-80 */
enum {
    R_NOT_SET = 0, R_ALL_ZEROS, R_UNKNOWN_DF, R_UNCORRECTABLE, R_UNKNOWN_THRESHOLD,
    R_UNRELIABLE_UNKNOWN, R_DF11_IID_1ERR_UNK, R_DF11_ACQ_1ERR_UNK, R_DF11_IID_UNK,
    R_DF18_2ERR_UNK, R_DF17_2ERR_UNK, R_ACCEPT_THRESHOLD, R_UNRELIABLE_KNOWN,
    R_DF18_2ERR_KNOWN, R_DF17_2ERR_KNOWN, R_DF18_1ERR_UNK, R_DF17_1ERR_UNK,
    R_DF11_ACQ_UNK, R_DF11_IID_1ERR_KNOWN, R_DF11_ACQ_1ERR_KNOWN, R_DF11_IID_KNOWN,
    R_DF18_1ERR_KNOWN, R_DF17_1ERR_KNOWN, R_DF11_ACQ_KNOWN, R_DF18_UNK, R_DF17_UNK,
    R_DF18_KNOWN, R_DF17_KNOWN
};

typedef struct {
    uint32_t syndrome;
    int      nerr;       /* 1 or 2; -1 while flagged for removal */
    int8_t   bit[2];
} synd_entry;

struct oracle1090 {
    int nfix, fix_df;
    unsigned block;
    uint32_t crc_tab[256];
    uint32_t bit_synd[112];
    synd_entry *tab_short, *tab_long;
    int n_short, n_long;
    uint32_t valid_short, valid_long;   /* DF pre-filter bitsets */
    uint32_t filter2[2][FILTER_SLOTS];  /* icao_filter_a / icao_filter_b (icao_filter.c:28-30) */
    int filter_active;                  /* file mode: never flips (dump1090.c:469-472); oracle1090_filter_expire() is the live-mode flip */
    uint16_t lut[65536];
    uint16_t *mag;                      /* OVERLAP + block */
    uint16_t tail[OVERLAP];             /* fifo.c overlap_buffer */
    unsigned last_message_end;
    uint64_t sample_counter;
    uint8_t  scratch[2][14];            /* msg1/msg2 of demod_2400.c:103 (stale bytes persist) */
    int      scratch_sel;
    int      iformat;                   /* 0 UC8, 1 SC16, 2 SC16Q11 (convert.c:94-118) */
    int      mode_ac;                   /* --modeac: demodulate2400AC after demodulate2400 (dump1090.c:902-905) */
};

/* ---- magnitude table: dump1090/dsp/helpers/tables.c:9-41 ---------------------------- */
void oracle1090_mag_table(uint16_t out[65536])
{
    for (int i = 0; i < 256; i++) {
        for (int q = 0; q < 256; q++) {
            volatile float fi = (float) ((i - 127.4) / 128);   /* double expression narrowed to float */
            volatile float fq = (float) ((q - 127.4) / 128);
            volatile float ii = fi * fi;                       /* volatile: no FMA contraction */
            volatile float qq = fq * fq;
            volatile float sq = ii + qq;
            float mag = (float) round((double) (sqrtf(sq) * 65536.0f));
            if (mag > 65535) mag = 65535;
            out[(unsigned) i | ((unsigned) q << 8)] = (uint16_t) mag;  /* little-endian {I,Q}: dsp-types.h:6-14 */
        }
    }
}

/* ---- CRC-24: dump1090/crc.c:28-81 ---------------------------------------------------- */
static void crc_init(oracle1090 *o)
{
    for (unsigned v = 0; v < 256; v++) {
        uint32_t c = v << 16;
        for (int k = 0; k < 8; k++)
            c = (c & 0x800000) ? ((c << 1) ^ 0xFFF409u) : (c << 1);
        o->crc_tab[v] = c & 0xFFFFFF;
    }
}

static uint32_t crc24(const oracle1090 *o, const uint8_t *m, int bits)
{
    int n = bits / 8;
    uint32_t rem = 0;
    for (int i = 0; i < n - 3; i++)
        rem = ((rem << 8) ^ o->crc_tab[m[i] ^ ((rem >> 16) & 0xFF)]) & 0xFFFFFF;
    return rem ^ ((uint32_t) m[n - 3] << 16) ^ ((uint32_t) m[n - 2] << 8) ^ m[n - 1];
}

/* ---- syndrome tables: dump1090/crc.c:182-381 ----------------------------------------- */
static int cmp_synd(const void *a, const void *b)
{
    uint32_t x = ((const synd_entry *) a)->syndrome, y = ((const synd_entry *) b)->syndrome;
    return (x > y) - (x < y);
}

static synd_entry *find_synd(synd_entry *t, int n, uint32_t s)
{
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        int mid = (lo + hi) / 2;
        if (t[mid].syndrome == s) return &t[mid];
        if (t[mid].syndrome < s) lo = mid + 1; else hi = mid - 1;
    }
    return NULL;
}

/* flag entries whose syndrome is also produced by a (weight)-bit error: crc.c:152-177 */
static void flag_aliases(const oracle1090 *o, synd_entry *t, int n, int off, int bits,
                         int start, uint32_t acc, int depth, int lo_w, int hi_w)
{
    if (depth > hi_w) return;
    for (int i = start; i < bits; i++) {
        uint32_t s = acc ^ o->bit_synd[i + off];
        if (depth >= lo_w) {
            synd_entry *e = find_synd(t, n, s);
            if (e) e->nerr = -1;
        }
        flag_aliases(o, t, n, off, bits, i + 1, s, depth + 1, lo_w, hi_w);
    }
}

static synd_entry *build_table(const oracle1090 *o, int bits, int max_correct, int max_detect, int *n_out)
{
    *n_out = 0;
    if (max_correct <= 0) return NULL;
    int off = 112 - bits;
    int cap = bits + (max_correct >= 2 ? bits * (bits - 1) / 2 : 0);
    synd_entry *t = malloc((size_t) cap * sizeof(*t));
    int n = 0;
    for (int a = 0; a < bits; a++) {
        t[n++] = (synd_entry) { o->bit_synd[a + off], 1, { (int8_t) a, -1 } };
        if (max_correct >= 2)
            for (int b = a + 1; b < bits; b++)
                t[n++] = (synd_entry) { o->bit_synd[a + off] ^ o->bit_synd[b + off], 2, { (int8_t) a, (int8_t) b } };
    }
    qsort(t, (size_t) n, sizeof(*t), cmp_synd);
    /* drop every member of a group sharing one syndrome (crc.c:245-262) */
    int w = 0;
    for (int i = 0; i < n;) {
        int k = i + 1;
        while (k < n && t[k].syndrome == t[i].syndrome) k++;
        if (k == i + 1) t[w++] = t[i];
        i = k;
    }
    n = w;
    if (max_detect > max_correct) {           /* crc.c:265-292 */
        flag_aliases(o, t, n, off, bits, 0, 0, 1, max_correct + 1, max_detect);
        w = 0;
        for (int i = 0; i < n; i++)
            if (t[i].nerr != -1) t[w++] = t[i];
        n = w;
    }
    *n_out = n;
    return t;
}

/* ---- ICAO filter: dump1090/icao_filter.c:40-109 (two tables; expiry only when the test injects it) ----- */
static uint32_t filter_hash(uint32_t a)
{
    uint32_t h = 0;
    for (int k = 0; k < 3; k++) {       /* Jenkins one-at-a-time over the low three bytes */
        h += (a >> (8 * k)) & 0xFF;
        h += h << 10;
        h ^= h >> 6;
    }
    h += h << 3;
    h ^= h >> 11;
    h += h << 15;
    return h & (FILTER_SLOTS - 1);
}

void oracle1090_filter_add(oracle1090 *o, uint32_t addr)        /* icaoFilterAdd, icao_filter.c:58-84: the active table only */
{
    uint32_t *t = o->filter2[o->filter_active];
    uint32_t h0 = filter_hash(addr), h = h0;
    while (t[h] != FILTER_EMPTY && t[h] != addr) {
        h = (h + 1) & (FILTER_SLOTS - 1);
        if (h == h0) return;            /* table full: silently dropped */
    }
    if (t[h] == FILTER_EMPTY) t[h] = addr;
}

static int filter_probe(const uint32_t *t, uint32_t addr)
{
    uint32_t h0 = filter_hash(addr), h = h0;
    while (t[h] != FILTER_EMPTY && t[h] != addr) {
        h = (h + 1) & (FILTER_SLOTS - 1);
        if (h == h0) break;
    }
    return t[h] == addr;
}

int oracle1090_filter_test(oracle1090 *o, uint32_t addr)         /* icaoFilterTest, icao_filter.c:86-109: both tables */
{
    return filter_probe(o->filter2[0], addr) || filter_probe(o->filter2[1], addr);
}

/* icaoFilterExpire when its 60 s timer fires (icao_filter.c:112-127): the OTHER table is emptied and becomes the
 * active one; what was active stays readable until the next flip.  The reference calls it from backgroundTasks for
 * live SDR input only (dump1090.c:469-472); the tests inject the flips at chosen points of the stream. */
void oracle1090_filter_expire(oracle1090 *o)
{
    o->filter_active ^= 1;
    memset(o->filter2[o->filter_active], 0xFF, sizeof(o->filter2[0]));
}

/* ---- message correction: dump1090/mode_s.c:252-346 ----------------------------------- */
static const uint32_t k_fixable_short[3] = { 0x00000800u, 0x08008e08u, 0x08008e08u }; /* mode_s.c:260-262 */
static const uint32_t k_fixable_long[3]  = { 0x00060000u, 0x066f0006u, 0x6fff066fu }; /* mode_s.c:263-265 */

static unsigned df_of(const uint8_t *m) { return m[0] >> 3; }

static const synd_entry k_no_errors = { 0, 0, { -1, -1 } };

static const synd_entry *diagnose(const oracle1090 *o, uint32_t s, int bits)   /* crc.c:386-405 */
{
    if (s == 0) return &k_no_errors;
    synd_entry *t = bits == 56 ? o->tab_short : o->tab_long;
    int n = bits == 56 ? o->n_short : o->n_long;
    return t ? find_synd(t, n, s) : NULL;
}

static void apply_fix(uint8_t *m, const synd_entry *e)                        /* crc.c:409-418 */
{
    for (int i = 0; i < e->nerr; i++)
        m[e->bit[i] >> 3] ^= (uint8_t) (0x80 >> (e->bit[i] & 7));
}

static int correct_message(const oracle1090 *o, const uint8_t *in, uint8_t *out,
                           uint32_t *ss, uint32_t *ls)
{
    *ss = *ls = UNCHECKED;
    uint32_t dfbit = 1u << df_of(in);
    int sel = o->fix_df ? o->nfix : 0;
    const synd_entry *le = NULL, *se = NULL;

    if (k_fixable_long[sel] & dfbit) {
        *ls = crc24(o, in, 112);
        if ((df_of(in) == 17 || df_of(in) == 18) && *ls == 0) { memcpy(out, in, 14); return 0; }
        le = diagnose(o, *ls, 112);
    }
    if (k_fixable_short[sel] & dfbit) {
        *ss = crc24(o, in, 56);
        if (df_of(in) == 11 && (*ss & 0xFFFF80u) == 0) { memcpy(out, in, 7); return 0; }
        se = diagnose(o, *ss, 56);
    }
    unsigned lerr = le ? (unsigned) le->nerr : 999, serr = se ? (unsigned) se->nerr : 999;

    if (le && lerr <= serr) {                      /* fewer errors first; tie -> long first */
        memcpy(out, in, 14); apply_fix(out, le);
        if (df_of(out) == 17 || df_of(out) == 18) return (int) lerr;
    }
    if (se && serr == 1) {                         /* DF11: never more than one bit */
        memcpy(out, in, 7); apply_fix(out, se);
        if (df_of(out) == 11) return (int) serr;
    }
    if (le && lerr > serr) {
        memcpy(out, in, 14); apply_fix(out, le);
        if (df_of(out) == 17 || df_of(out) == 18) return (int) lerr;
    }
    memcpy(out, in, 14);
    return -1;
}

/* ---- scoring: dump1090/mode_s.c:350-466 ---------------------------------------------- */
static int score_message(oracle1090 *o, const uint8_t *raw)
{
    static const uint8_t zeros[7] = { 0 };
    if (!memcmp(raw, zeros, 7)) return R_ALL_ZEROS;

    uint8_t fixed[14];
    uint32_t ss, ls;
    int nc = correct_message(o, raw, fixed, &ss, &ls);
    uint32_t addr = ((uint32_t) fixed[1] << 16) | ((uint32_t) fixed[2] << 8) | fixed[3];
    int known;

    switch (df_of(fixed)) {
    case 0: case 4: case 5:
        if (ss == UNCHECKED) ss = crc24(o, fixed, 56);
        return oracle1090_filter_test(o, ss) ? R_UNRELIABLE_KNOWN : R_UNRELIABLE_UNKNOWN;
    case 16: case 20: case 21:
        if (ls == UNCHECKED) ls = crc24(o, fixed, 112);
        return oracle1090_filter_test(o, ls) ? R_UNRELIABLE_KNOWN : R_UNRELIABLE_UNKNOWN;
    case 24: case 25: case 26: case 27: case 28: case 29: case 30: case 31:
        return R_UNCORRECTABLE;                    /* enable_df24 is off */
    case 11: {
        if (ss == UNCHECKED) ss = crc24(o, fixed, 56);
        int acq = (ss & 0x7F) == 0;                /* IID taken from the syndrome of the *uncorrected* bytes */
        known = oracle1090_filter_test(o, addr);
        if (nc == 0) return acq ? (known ? R_DF11_ACQ_KNOWN : R_DF11_ACQ_UNK) : (known ? R_DF11_IID_KNOWN : R_DF11_IID_UNK);
        if (nc == 1) return acq ? (known ? R_DF11_ACQ_1ERR_KNOWN : R_DF11_ACQ_1ERR_UNK) : (known ? R_DF11_IID_1ERR_KNOWN : R_DF11_IID_1ERR_UNK);
        return R_UNCORRECTABLE;
    }
    case 17:
        known = oracle1090_filter_test(o, addr);
        if (nc == 0) return known ? R_DF17_KNOWN : R_DF17_UNK;
        if (nc == 1) return known ? R_DF17_1ERR_KNOWN : R_DF17_1ERR_UNK;
        if (nc == 2) return known ? R_DF17_2ERR_KNOWN : R_DF17_2ERR_UNK;
        return R_UNCORRECTABLE;
    case 18:
        known = oracle1090_filter_test(o, addr | NT_FLAG);
        if (nc == 0) return known ? R_DF18_KNOWN : R_DF18_UNK;
        if (nc == 1) return known ? R_DF18_1ERR_KNOWN : R_DF18_1ERR_UNK;
        if (nc == 2) return known ? R_DF18_2ERR_KNOWN : R_DF18_2ERR_UNK;
        return R_UNCORRECTABLE;
    default:
        return R_UNKNOWN_DF;
    }
}

int oracle1090_score(oracle1090 *o, const uint8_t msg[14]) { return score_message(o, msg); }
uint32_t oracle1090_crc(const uint8_t *msg, int bits)
{
    static oracle1090 tmp; static int init;
    if (!init) { crc_init(&tmp); init = 1; }
    return crc24(&tmp, msg, bits);
}

/* 1-based big-endian bit / field access, as util getbit/getbits in the reference */
static unsigned mbit(const uint8_t *d, unsigned n) { return (d[(n - 1) >> 3] >> (7 - ((n - 1) & 7))) & 1; }
static unsigned mbits(const uint8_t *d, unsigned a, unsigned b)
{
    unsigned v = 0;
    for (unsigned k = a; k <= b; k++) v = (v << 1) | mbit(d, k);
    return v;
}

/* Does decodeExtendedSquitter flag a DF18 address as non-ICAO?  Only this side effect
 * matters here because the flag ends up in the key handed to icaoFilterAdd
 * (mode_s.c:786-788).  Follows mode_s.c:1447-1548 with the IMF bit positions at
 * :869, :991, :1031-1033, :1126, :1137, :1333. */
static int df18_non_icao(const uint8_t *msg)
{
    const uint8_t *me = msg + 4;
    unsigned cf = msg[0] & 7, metype = me[0] >> 3, imf = 0;
    switch (cf) {
    case 0: break;
    case 1: case 5: return 1;   /* flag set, decoding continues but nothing clears it */
    case 2: case 6: imf = 1; break;
    case 3: return (int) mbit(me, 1);
    default: return 1;
    }
    if (!imf) return 0;
    switch (metype) {
    case 19: { unsigned sub = mbits(me, 6, 8); return (sub >= 1 && sub <= 4) ? (int) mbit(me, 9) : 0; }
    case 5: case 6: case 7: case 8: return (int) mbit(me, 21);
    case 0: case 9: case 10: case 11: case 12: case 13: case 14: case 15: case 16: case 17: case 18:
    case 20: case 21: case 22: return (int) mbit(me, 8);
    case 28: return (mbits(me, 6, 8) == 1) ? (int) mbit(me, 56) : 0;
    case 29: return (int) mbit(me, 51);
    case 31: return (int) mbit(me, 56);
    default: return 0;
    }
}

/* ---- preamble detector: dump1090/demod_2400.c:155-218 -------------------------------- */
int oracle1090_preamble(const uint16_t *p)
{
    if (!(p[0] < p[1] && p[12] > p[13])) return 0;
    unsigned hi, sig, noise;
    int up12 = p[11] < p[12];
    if (p[1] > p[2] && p[2] < p[3] && p[3] > p[4] && p[8] < p[9] && p[9] > p[10] && p[10] < p[11]) {
        hi = (p[1] + p[3] + p[9] + p[11] + p[12]) / 4;  sig = p[1] + p[3] + p[9];  noise = p[5] + p[6] + p[7];
    } else if (p[1] > p[2] && p[2] < p[3] && p[3] > p[4] && p[8] < p[9] && p[9] > p[10] && up12) {
        hi = (p[1] + p[3] + p[9] + p[12]) / 4;  sig = p[1] + p[3] + p[9] + p[12];  noise = p[5] + p[6] + p[7] + p[8];
    } else if (p[1] > p[2] && p[2] < p[3] && p[4] > p[5] && p[8] < p[9] && p[10] > p[11] && up12) {
        hi = (p[1] + p[3] + p[4] + p[9] + p[10] + p[12]) / 4;  sig = p[1] + p[12];  noise = p[6] + p[7];
    } else if (p[1] > p[2] && p[3] < p[4] && p[4] > p[5] && p[9] < p[10] && p[10] > p[11] && up12) {
        hi = (p[1] + p[4] + p[10] + p[12]) / 4;  sig = p[1] + p[4] + p[10] + p[12];  noise = p[5] + p[6] + p[7] + p[8];
    } else if (p[2] > p[3] && p[3] < p[4] && p[4] > p[5] && p[9] < p[10] && p[10] > p[11] && up12) {
        hi = (p[1] + p[2] + p[4] + p[10] + p[12]) / 4;  sig = p[4] + p[10] + p[12];  noise = p[6] + p[7] + p[8];
    } else {
        return 0;
    }
    if (sig * 2 < 3 * noise) return 0;
    static const int quiet[9] = { 5, 6, 7, 8, 14, 15, 16, 17, 18 };
    for (int k = 0; k < 9; k++)
        if (p[quiet[k]] >= hi) return 0;
    return 1;
}

/* ---- bit slicer: dump1090/demod_2400.c:45-59 and :223-313 in closed form --------------
 * Symbol k of try_phase p sits u = p + 12k fifths-of-a-sample after m[j+19]; the
 * correlator used is the one for sub-phase u mod 5 applied at sample floor(u/5). */
static int slice_bit(const uint16_t *s, int sub)
{
    switch (sub) {
    case 0: return 5 * s[0] - 3 * s[1] - 2 * s[2];
    case 1: return 4 * s[0] - s[1] - 3 * s[2];
    case 2: return 3 * s[0] + s[1] - 4 * s[2];
    case 3: return 2 * s[0] + 3 * s[1] - 5 * s[2];
    default: return s[0] + 5 * s[1] - 5 * s[2] - s[3];
    }
}

static void slice_bytes(const uint16_t *m, int try_phase, uint8_t *out, int first, int count)
{
    for (int b = first; b < first + count; b++) {
        unsigned v = 0;
        for (int k = 8 * b; k < 8 * b + 8; k++) {
            int u = try_phase + 12 * k;
            v = (v << 1) | (slice_bit(m + 19 + u / 5, u % 5) > 0);
        }
        out[b] = (uint8_t) v;
    }
}

void oracle1090_slice(const uint16_t *m, int try_phase, uint8_t out[14]) { slice_bytes(m, try_phase, out, 0, 14); }

static uint32_t damage_set(unsigned df, int dist)                   /* demod_2400.c:64-78 */
{
    uint32_t s = 1u << df;
    if (dist > 0)
        for (int b = 0; b < 5; b++) s |= damage_set(df ^ (1u << b), dist - 1);
    return s;
}

/* ---- one mag_buf: dump1090/demod_2400.c:99-444 ---------------------------------------- */
static size_t demod_block(oracle1090 *o, unsigned mlen, uint64_t block_ts,
                          oracle1090_msg *out, size_t cap, size_t have)
{
    const uint16_t *m = o->mag;
    size_t n = 0;
    if (o->last_message_end > mlen) o->last_message_end = mlen;

    for (unsigned j = o->last_message_end; j < mlen; j++) {
        if (!oracle1090_preamble(m + j)) continue;

        int best_score = R_NOT_SET, best_phase = -1;
        uint8_t *best = NULL;
        for (int ph = 4; ph <= 8; ph++) {
            uint8_t *buf = o->scratch[o->scratch_sel];
            slice_bytes(m + j, ph, buf, 0, 1);
            uint32_t dfbit = 1u << (buf[0] >> 3);
            int nbytes = (o->valid_long & dfbit) ? 14 : (o->valid_short & dfbit) ? 7 : 0;   /* :317-325 */
            if (!nbytes) continue;
            slice_bytes(m + j, ph, buf, 1, nbytes - 1);
            int sc = score_message(o, buf);
            if (sc > best_score) {            /* strict: lowest phase wins ties (:336) */
                best = buf; best_score = sc; best_phase = ph;
                o->scratch_sel ^= 1;
            }
        }
        if (best_score < R_ACCEPT_THRESHOLD) continue;

        int msglen = (best[0] & 0x80) ? 112 : 56;          /* from the *uncorrected* DF (:357) */
        oracle1090_msg mm;
        memset(&mm, 0, sizeof(mm));
        mm.timestamp = block_ts + (uint64_t) j * 5 + (8 + 56) * 12 + (uint64_t) best_phase;   /* :366 */
        mm.score = (uint8_t) best_score;

        /* decodeModesMessage prefix: mode_s.c:520-626 */
        memcpy(mm.verbatim, best, 14);
        uint32_t ss, ls;
        int nc = correct_message(o, best, mm.msg, &ss, &ls);
        mm.msgtype = (uint8_t) df_of(mm.msg);
        mm.msgbits = (mm.msgtype & 16) ? 112 : 56;
        if (mm.msgtype & 16) { if (ls == UNCHECKED) ls = crc24(o, mm.msg, 112); mm.crc = ls; }
        else                 { if (ss == UNCHECKED) ss = crc24(o, mm.msg, 56);  mm.crc = ss; }
        mm.correctedbits = (uint8_t) (nc > 0 ? nc : 0);
        uint32_t aa = ((uint32_t) mm.msg[1] << 16) | ((uint32_t) mm.msg[2] << 8) | mm.msg[3];
        switch (mm.msgtype) {
        case 11: case 17: mm.addr = aa; break;
        case 18: mm.addr = aa | (df18_non_icao(mm.msg) ? NON_ICAO : 0); break;
        default: mm.addr = mm.crc; break;                  /* Address/Parity formats */
        }
        /* filter insertions: mode_s.c:781-788 */
        if (!mm.correctedbits && (mm.msgtype == 17 || (mm.msgtype == 11 && (mm.crc & 0x7F) == 0)))
            oracle1090_filter_add(o, mm.addr);
        if (!mm.correctedbits && mm.msgtype == 18)
            oracle1090_filter_add(o, mm.addr | NT_FLAG);

        /* signal power: demod_2400.c:382-393 */
        int siglen = msglen * 12 / 5;
        uint64_t acc = 0;
        for (int k = 0; k < siglen; k++) { uint32_t v = m[j + 19 + k]; acc += (uint64_t) (v * v); }
        mm.signal_level = (acc / 65535.0 / 65535.0) / siglen;

        if (have + n < cap) out[have + n] = mm;
        n++;

        o->last_message_end = j + (unsigned) ((msglen + 8) * 12 / 5);    /* :410 */
        j = o->last_message_end - 8 * 12 / 5;                            /* :418; loop then adds 1 */
    }
    if (o->last_message_end < mlen) o->last_message_end = 0;            /* :432-443 */
    else o->last_message_end -= mlen;
    return n;
}

/* ---- Mode A/C on one mag_buf: dump1090/demod_2400.c:539-726 ----------------------------------
 * mean_level / mean_power are those of the block's NEW samples as magnitude_power_uc8 (lookup_unroll_4, the
 * generic flavour's first registered implementation; dsp/impl/magnitude_power_uc8.c:48-90) computes them. */
static size_t modeac_block(oracle1090 *o, unsigned mlen, uint64_t block_ts, double mean_level, double mean_power,
                           oracle1090_msg *out, size_t cap, size_t have)
{
    const uint16_t *m = o->mag;
    size_t n = 0;
    double noise_stddev = sqrt(mean_power - mean_level * mean_level);                     /* :550 */
    unsigned noise_level = (unsigned) ((mean_power + noise_stddev) * 65535 + 0.5);        /* :551 */

    for (unsigned f1_sample = 1; f1_sample < mlen; ++f1_sample) {
        if (!(m[f1_sample - 1] < m[f1_sample + 0])) continue;                              /* :606 */
        if (m[f1_sample + 2] > m[f1_sample + 0] || m[f1_sample + 2] > m[f1_sample + 1]) continue;
        unsigned f1_level = (m[f1_sample + 0] + m[f1_sample + 1]) / 2;
        if (noise_level * 2 > f1_level) continue;

        volatile float f1a_power = (float) m[f1_sample] * m[f1_sample];                   /* :622-625 */
        volatile float f1b_power = (float) m[f1_sample + 1] * m[f1_sample + 1];
        volatile float fsum = f1a_power + f1b_power;
        volatile float fraction = f1b_power / fsum;
        volatile float fr2 = fraction * fraction;
        volatile float fpos = f1_sample + fr2;
        volatile float f25 = 25 * fpos;
        unsigned f1_clock = (unsigned) (f25 + 0.5);

        unsigned f2_clock = f1_clock + (87 * 14);
        unsigned f2_sample = f2_clock / 25;
        if (!(m[f2_sample - 1] < m[f2_sample + 0])) continue;
        if (m[f2_sample + 2] > m[f2_sample + 0] || m[f2_sample + 2] > m[f2_sample + 1]) continue;
        unsigned f2_level = (m[f2_sample + 0] + m[f2_sample + 1]) / 2;
        if (noise_level * 2 > f2_level) continue;

        unsigned f1f2_level = (f1_level > f2_level ? f1_level : f2_level);
        float midpoint = sqrtf(noise_level * f1f2_level);                                  /* unsigned product -> float */
        unsigned signal_threshold = (unsigned) (midpoint * M_SQRT2 + 0.5);
        unsigned noise_threshold = (unsigned) (midpoint / M_SQRT2 + 0.5);

        unsigned uncertain_bits = 0, noisy_bits = 0, bits = 0, bit, clock;
        for (bit = 0, clock = f1_clock; bit < 20; ++bit, clock += 87) {
            unsigned sample = clock / 25;
            bits <<= 1; noisy_bits <<= 1; uncertain_bits <<= 1;
            if (m[sample + 2] >= signal_threshold) noisy_bits |= 1;
            if (m[sample + 0] >= signal_threshold || m[sample + 1] >= signal_threshold) bits |= 1;
            else if (m[sample + 0] > noise_threshold && m[sample + 1] > noise_threshold) uncertain_bits |= 1;
        }
        if ((bits & 0x80020) != 0x80020) continue;
        if ((bits & 0x0101B) != 0) continue;
        if (noisy_bits || uncertain_bits) continue;

        unsigned modeac =
            ((bits & 0x40000) ? 0x0010 : 0) | ((bits & 0x20000) ? 0x1000 : 0) | ((bits & 0x10000) ? 0x0020 : 0) |
            ((bits & 0x08000) ? 0x2000 : 0) | ((bits & 0x04000) ? 0x0040 : 0) | ((bits & 0x02000) ? 0x4000 : 0) |
            ((bits & 0x00800) ? 0x0100 : 0) | ((bits & 0x00400) ? 0x0001 : 0) | ((bits & 0x00200) ? 0x0200 : 0) |
            ((bits & 0x00100) ? 0x0002 : 0) | ((bits & 0x00080) ? 0x0400 : 0) | ((bits & 0x00040) ? 0x0004 : 0) |
            ((bits & 0x00004) ? 0x0080 : 0);

        oracle1090_msg mm;                                  /* decodeModeAMessage: mode_ac.c:123-158 */
        memset(&mm, 0, sizeof(mm));
        mm.timestamp = block_ts + f2_clock / 5;             /* :712 */
        mm.msgtype = 32;
        mm.msgbits = 16;
        mm.msg[0] = mm.verbatim[0] = (uint8_t) (modeac >> 8);
        mm.msg[1] = mm.verbatim[1] = (uint8_t) modeac;
        mm.addr = (modeac & 0x0000FF7F) | NON_ICAO;
        if (have + n < cap) out[have + n] = mm;
        n++;
        f1_sample += (20 * 87 / 25);                        /* :722 */
    }
    return n;
}

void oracle1090_set_mode_ac(oracle1090 *o, int on) { o->mode_ac = on; }
void oracle1090_set_input_format(oracle1090 *o, int fmt) { o->iformat = fmt; }

size_t oracle1090_feed(oracle1090 *o, const uint8_t *iq, size_t nsamples, oracle1090_msg *out, size_t cap)
{
    size_t total = 0, pos = 0;
    while (pos < nsamples) {
        size_t take = nsamples - pos < o->block ? nsamples - pos : o->block;
        memcpy(o->mag, o->tail, sizeof(o->tail));                                   /* fifo.c:187-193 */
        if (o->iformat) {       /* magnitude_sc16.c:27-45 / magnitude_sc16q11.c:29-47 ("exact_float") */
            const int16_t *s16 = (const int16_t *) (const void *) iq;
            const int scale = o->iformat == 2 ? 32 : 2;
            for (size_t k = 0; k < take; k++) {
                volatile float I = abs((int16_t) s16[2 * (pos + k)]) * scale;
                volatile float Q = abs((int16_t) s16[2 * (pos + k) + 1]) * scale;
                volatile float ii = I * I, qq = Q * Q;                              /* volatile: no FMA contraction */
                volatile float magsq = ii + qq;
                float mag = sqrtf(magsq);
                if (mag > 65535.0) mag = 65535.0;
                o->mag[OVERLAP + k] = (uint16_t) mag;
            }
        } else
        for (size_t k = 0; k < take; k++)                                           /* magnitude_uc8.c:30-63 */
            o->mag[OVERLAP + k] = o->lut[(unsigned) iq[2 * (pos + k)] | ((unsigned) iq[2 * (pos + k) + 1] << 8)];
        memcpy(o->tail, o->mag + take, sizeof(o->tail));                            /* fifo.c:196 */
        uint64_t ts = (uint64_t) ((double) o->sample_counter * 12e6 / 2400000.0);   /* sdr_ifile.c:202 */
        total += demod_block(o, (unsigned) take, ts, out, cap, total);
        if (o->mode_ac) {                                                           /* dump1090.c:902-905 */
            uint64_t sum_level = 0, sum_power = 0;                                  /* magnitude_power_uc8.c:55-89 */
            for (size_t k = 0; k < take; k++) { uint32_t v = o->mag[OVERLAP + k]; sum_level += v; sum_power += (uint64_t) (v * v); }
            double level = sum_level / 65536.0 / (unsigned) take, power = sum_power / 65536.0 / 65536.0 / (unsigned) take;
            total += modeac_block(o, (unsigned) take, ts, level, power, out, cap, total);
        }
        o->sample_counter += take;
        pos += take;
    }
    return total;
}

size_t oracle1090_syndrome_table(oracle1090 *o, int bits, uint32_t *out, size_t cap)
{
    synd_entry *t = bits == 56 ? o->tab_short : o->tab_long;
    size_t n = (size_t) (bits == 56 ? o->n_short : o->n_long);
    for (size_t i = 0; i < n && i < cap; i++) {
        out[4 * i] = t[i].syndrome; out[4 * i + 1] = (uint32_t) t[i].nerr;
        out[4 * i + 2] = (uint32_t) (int32_t) t[i].bit[0]; out[4 * i + 3] = (uint32_t) (int32_t) t[i].bit[1];
    }
    return n;
}

oracle1090 *oracle1090_create(int nfix_crc, int fix_df, unsigned block_samples)
{
    oracle1090 *o = calloc(1, sizeof(*o));
    if (!o) return NULL;
    o->nfix = nfix_crc < 0 ? 0 : nfix_crc > 2 ? 2 : nfix_crc;
    o->fix_df = fix_df != 0;
    o->block = block_samples ? block_samples : DEFAULT_BLOCK;
    crc_init(o);
    uint8_t z[14];
    for (int i = 0; i < 112; i++) {                     /* crc.c:57-62 */
        memset(z, 0, sizeof(z));
        z[i >> 3] = (uint8_t) (0x80 >> (i & 7));
        o->bit_synd[i] = crc24(o, z, 112);
    }
    if (o->nfix == 1) {                                 /* crc.c:355-381 */
        o->tab_short = build_table(o, 56, 1, 1, &o->n_short);
        o->tab_long = build_table(o, 112, 1, 1, &o->n_long);
    } else if (o->nfix == 2) {
        o->tab_short = build_table(o, 56, 2, 4, &o->n_short);
        o->tab_long = build_table(o, 112, 2, 4, &o->n_long);
    }
    o->valid_short = (1u << 0) | (1u << 4) | (1u << 5) | (1u << 11);           /* demod_2400.c:80-93 */
    o->valid_long = (1u << 16) | (1u << 17) | (1u << 18) | (1u << 20) | (1u << 21);
    if (o->fix_df && o->nfix) {
        o->valid_short |= damage_set(11, 1);
        o->valid_long |= damage_set(17, o->nfix) | damage_set(18, o->nfix);
    }
    memset(o->filter2, 0xFF, sizeof(o->filter2));
    o->filter_active = 0;
    oracle1090_mag_table(o->lut);
    o->mag = calloc(OVERLAP + o->block, sizeof(uint16_t));
    if (!o->mag) { oracle1090_destroy(o); return NULL; }
    return o;
}

void oracle1090_destroy(oracle1090 *o)
{
    if (!o) return;
    free(o->tab_short); free(o->tab_long); free(o->mag); free(o);
}

/* ---- surveillance fields of decodeModesMessage (mode_s.c:626-779), for adsbcu_1090_decode_fields ----------------
 * Pinned against the reference binary's own struct modesMessage through the second tap of oracle/ref_harness.c
 * (tests/golden/fields_a.npz).  Layout == adsbcu_fields_a (include/adsbcu.h). */
typedef struct oracle_fields_a {
    uint16_t AC, ID;
    uint8_t  CA, CC, CF, DR, FS, KE, ND, RI, SL, UM, VS, airground, alert_valid, alert, spi_valid, spi;
    uint8_t  altitude_baro_valid, altitude_baro_unit, squawk_valid, es_pending;
    uint16_t squawk, pad;
    int32_t  altitude_baro;
} oracle_fields_a;

static unsigned f_getbits(const uint8_t *m, unsigned first, unsigned last)      /* mode_s.h:104-150, bits numbered from 1 */
{
    unsigned v = 0, b;
    for (b = first; b <= last; ++b) v = (v << 1) | ((m[(b - 1) >> 3] >> (7 - ((b - 1) & 7))) & 1u);
    return v;
}

static int f_id13(int f)                                                          /* decodeID13Field, mode_s.c:96-116 */
{
    static const struct { int in, out; } map[12] = {
        {0x1000, 0x0010}, {0x0800, 0x1000}, {0x0400, 0x0020}, {0x0200, 0x2000}, {0x0100, 0x0040}, {0x0080, 0x4000},
        {0x0020, 0x0100}, {0x0010, 0x0001}, {0x0008, 0x0200}, {0x0004, 0x0002}, {0x0002, 0x0400}, {0x0001, 0x0004}};
    int g = 0, k;
    for (k = 0; k < 12; ++k) if (f & map[k].in) g |= map[k].out;
    return g;
}

static int f_mode_a_to_c(unsigned a)                                              /* modeAToModeC, mode_ac.c:61-122 via track.h:332-341 */
{
    unsigned one = 0, five = 0;
    a = ((a & 0x0007) | (a & 0x0070) | (a & 0x0700) | (a & 0x7000));           /* indexToModeA(modeAToIndex(a)) */
    if ((a & 0xFFFF8889u) != 0 || (a & 0xF0u) == 0) return -9999;
    if (a & 0x0010) one ^= 7;
    if (a & 0x0020) one ^= 3;
    if (a & 0x0040) one ^= 1;
    if ((one & 5) == 5) one ^= 2;
    if (one > 5) return -9999;
    if (a & 0x0002) five ^= 0x0FF;
    if (a & 0x0004) five ^= 0x07F;
    if (a & 0x1000) five ^= 0x03F;
    if (a & 0x2000) five ^= 0x01F;
    if (a & 0x4000) five ^= 0x00F;
    if (a & 0x0100) five ^= 0x007;
    if (a & 0x0200) five ^= 0x003;
    if (a & 0x0400) five ^= 0x001;
    if (five & 1) one = 6 - one;
    return (int)(five * 5 + one) - 13;
}

void oracle1090_decode_fields(const uint8_t *msg, int df, oracle_fields_a *f)
{
    memset(f, 0, sizeof(*f));
    if (df == 0 || df == 4 || df == 16 || df == 20) {
        f->AC = (uint16_t)f_getbits(msg, 20, 32);
        if (f->AC) {
            const int ac = f->AC;
            int alt;
            if (!(ac & 0x0040)) {                                                 /* decodeAC13Field, mode_s.c:122-150 */
                f->altitude_baro_unit = 0;
                if (ac & 0x0010) alt = (((ac & 0x1F80) >> 2) | ((ac & 0x0020) >> 1) | (ac & 0x000F)) * 25 - 1000;
                else { const int n = f_mode_a_to_c((unsigned)f_id13(ac)); alt = n < -12 ? -9999 : 100 * n; }
            } else { f->altitude_baro_unit = 1; alt = -9999; }
            f->altitude_baro = alt;
            if (alt != -9999) f->altitude_baro_valid = 1;
        }
    }
    if (df == 11 || df == 17) {
        f->CA = (uint8_t)f_getbits(msg, 6, 8);
        switch (f->CA) { case 0: case 6: case 7: f->airground = 3; break; case 4: f->airground = 1; break; case 5: f->airground = 2; break; default: break; }
    }
    if (df == 0) f->CC = (uint8_t)f_getbits(msg, 7, 7);
    if (df == 18) f->CF = (uint8_t)f_getbits(msg, 6, 8);
    if (df == 4 || df == 5 || df == 20 || df == 21) {
        f->DR = (uint8_t)f_getbits(msg, 9, 13);
        f->FS = (uint8_t)f_getbits(msg, 6, 8);
        f->alert_valid = f->spi_valid = 1;
        switch (f->FS) {
        case 0: f->airground = 3; break;
        case 1: f->airground = 1; break;
        case 2: f->airground = 3; f->alert = 1; break;
        case 3: f->airground = 1; f->alert = 1; break;
        case 4: f->airground = 3; f->alert = 1; f->spi = 1; break;
        case 5: f->airground = 3; f->spi = 1; break;
        default: f->spi_valid = f->alert_valid = 0; break;
        }
        f->UM = (uint8_t)f_getbits(msg, 14, 19);
    }
    if (df == 5 || df == 21) {
        f->ID = (uint16_t)f_getbits(msg, 20, 32);
        if (f->ID) { f->squawk = (uint16_t)f_id13(f->ID); f->squawk_valid = 1; }
    }
    if (df == 24) { f->KE = (uint8_t)f_getbits(msg, 4, 4); f->ND = (uint8_t)f_getbits(msg, 5, 8); }
    if (df == 0 || df == 16) {
        f->RI = (uint8_t)f_getbits(msg, 14, 17);
        f->SL = (uint8_t)f_getbits(msg, 9, 11);
        f->VS = (uint8_t)f_getbits(msg, 6, 6);
        f->airground = f->VS ? 1 : 3;
    }
    f->es_pending = (df == 17 || df == 18);
}

/* ---- extended-squitter payload of decodeModesMessage (mode_s.c:798-1551), for adsbcu_1090_decode_es ---------------
 * Pinned against the reference binary's own decodeModesMessage through the third tap of oracle/ref_harness.c
 * (tests/golden/fields_es.npz: frames in, struct modesMessage fields out).  Layout == adsbcu_fields_es. */
typedef struct oracle_fields_es {
    uint32_t addr, valid;
    uint8_t  es, metype, mesub, addrtype, source, reliable, airground, CA_CF;
    uint8_t  cpr_type, cpr_odd, heading_type, nav_heading_type, alert, spi, emergency, category;
    uint8_t  altitude_baro_unit, altitude_geom_unit, nav_altitude_source, nav_modes;
    char     callsign[8];
    int32_t  altitude_baro, altitude_geom, geom_delta, baro_rate, geom_rate;
    uint32_t cpr_lat, cpr_lon;
    uint16_t ias, tas, squawk, ops_flags;
    float    heading, gs_v0, gs_v2, gs_selected, nav_heading, nav_qnh;
    int32_t  nav_fms_altitude, nav_mcp_altitude;
    uint8_t  nic_a, nic_b, nic_c, nic_baro, nac_p, nac_v, sil, sil_type, gva, sda;
    uint8_t  ops_version, ops_tah, ops_hrd, ops_cc_tc, ops_cc_lw, ops_cc_antenna_offset;
} oracle_fields_es;

enum { V_ALT_BARO = 1u << 0, V_ALT_GEOM = 1u << 1, V_HEADING = 1u << 2, V_GS = 1u << 3, V_IAS = 1u << 4, V_TAS = 1u << 5, V_BARO_RATE = 1u << 6,
       V_GEOM_RATE = 1u << 7, V_SQUAWK = 1u << 8, V_CALLSIGN = 1u << 9, V_CPR = 1u << 10, V_CATEGORY = 1u << 11, V_GEOM_DELTA = 1u << 12,
       V_SPI = 1u << 13, V_ALERT = 1u << 14, V_EMERGENCY = 1u << 15, V_NIC_A = 1u << 16, V_NIC_B = 1u << 17, V_NIC_C = 1u << 18,
       V_NIC_BARO = 1u << 19, V_NAC_P = 1u << 20, V_NAC_V = 1u << 21, V_GVA = 1u << 22, V_SDA = 1u << 23, V_NAV_HEADING = 1u << 24,
       V_NAV_FMS = 1u << 25, V_NAV_MCP = 1u << 26, V_NAV_QNH = 1u << 27, V_NAV_MODES = 1u << 28, V_OPSTATUS = 1u << 29, V_CC_LW = 1u << 30 };
enum { O_OM_ACAS_RA = 1, O_OM_IDENT = 2, O_OM_ATC = 4, O_OM_SAF = 8, O_CC_ACAS = 16, O_CC_CDTI = 32, O_CC_1090_IN = 64, O_CC_ARV = 128,
       O_CC_TS = 256, O_CC_UAT_IN = 512, O_CC_POA = 1024, O_CC_B2_LOW = 2048 };
/* dump1090.h:147-161, 186-201, 219-242 */
enum { AT_ADSB_ICAO, AT_ADSB_ICAO_NT, AT_ADSR_ICAO, AT_TISB_ICAO, AT_ADSB_OTHER, AT_ADSR_OTHER, AT_TISB_TRACKFILE, AT_TISB_OTHER, AT_MODE_A, AT_UNKNOWN };
enum { H_INVALID, H_GROUND_TRACK, H_TRUE, H_MAGNETIC, H_MAGNETIC_OR_TRUE, H_TRACK_OR_HEADING };
enum { NM_AUTOPILOT = 1, NM_VNAV = 2, NM_ALT_HOLD = 4, NM_APPROACH = 8, NM_LNAV = 16, NM_TCAS = 32 };
enum { NA_INVALID, NA_UNKNOWN, NA_AIRCRAFT, NA_MCP, NA_FMS };
enum { SRC_TISB = 5, SRC_ADSR = 6, SRC_ADSB = 7 };
#define NON_ICAO (1u << 24)

static void es_set_imf(oracle_fields_es *e)                                      /* setIMF, mode_s.c:832-854 */
{
    e->addr |= NON_ICAO;
    if (e->addrtype == AT_ADSB_ICAO || e->addrtype == AT_ADSB_ICAO_NT) e->addrtype = AT_ADSB_OTHER;
    else if (e->addrtype == AT_TISB_ICAO) e->addrtype = AT_TISB_TRACKFILE;
    else if (e->addrtype == AT_ADSR_ICAO) e->addrtype = AT_ADSR_OTHER;
}

static int es_ac12(int f, uint8_t *unit)                                          /* decodeAC12Field, mode_s.c:156-178 */
{
    *unit = 0;
    if (f & 0x10) return ((((f & 0x0FE0) >> 1) | (f & 0x000F)) * 25) - 1000;
    {
        const int n = f_mode_a_to_c((unsigned)f_id13(((f & 0x0FC0) << 1) | (f & 0x003F)));
        return n < -12 ? -9999 : 100 * n;
    }
}

static float es_movement(unsigned m, int v2)                                      /* decodeMovementFieldV0 / V2, mode_s.c:184-230 */
{
    if (m >= 125) return 0;
    if (m == 124) return 180;
    if (m >= 109) return 100 + (m - 109 + 0.5) * 5;
    if (m >= 94) return 70 + (m - 94 + 0.5) * 2;
    if (m >= 39) return 15 + (m - 39 + 0.5) * 1;
    if (m >= 13) return 2 + (m - 13 + 0.5) * 0.50;
    if (m >= 9) return 1 + (m - 9 + 0.5) * 0.25;
    if (v2) {
        if (m >= 3) return 0.125 + (m - 3 + 0.5) * 0.875 / 6;
        if (m >= 2) return 0.125 / 2;
        return 0;
    }
    if (m >= 2) return 0.125 + (m - 2 + 0.5) * 0.125;
    return 0;
}

#define MEBITS(a, b) f_getbits(me, (a), (b))
#define MEBIT(a) f_getbits(me, (a), (a))

static void es_velocity(oracle_fields_es *e, const uint8_t *me, int check_imf)   /* decodeESAirborneVelocity, mode_s.c:856-963 */
{
    unsigned raw;
    e->mesub = (uint8_t)MEBITS(6, 8);
    if (e->mesub < 1 || e->mesub > 4) return;
    if (check_imf && MEBIT(9)) es_set_imf(e);
    e->valid |= V_NAC_V;
    e->nac_v = (uint8_t)MEBITS(11, 13);
    if (e->mesub <= 2) {
        const unsigned ew_raw = MEBITS(15, 24), ns_raw = MEBITS(26, 35);
        if (ew_raw && ns_raw) {
            const int mul = e->mesub == 2 ? 4 : 1;
            const int ew = (int)(ew_raw - 1) * (MEBIT(14) ? -1 : 1) * mul, ns = (int)(ns_raw - 1) * (MEBIT(25) ? -1 : 1) * mul;
            e->gs_v0 = e->gs_v2 = e->gs_selected = sqrtf((ns * ns) + (ew * ew) + 0.5);
            e->valid |= V_GS;
            if (e->gs_selected > 0) {
                float track = atan2(ew, ns) * 180.0 / M_PI;
                if (track < 0) track += 360;
                e->heading = track;
                e->heading_type = H_GROUND_TRACK;
                e->valid |= V_HEADING;
            }
        }
    } else {
        unsigned airspeed;
        if (MEBIT(14)) {
            e->valid |= V_HEADING;
            e->heading = MEBITS(15, 24) * 360.0 / 1024.0;
            e->heading_type = H_MAGNETIC_OR_TRUE;
        }
        airspeed = MEBITS(26, 35);
        if (airspeed) {
            const unsigned speed = (airspeed - 1) * (e->mesub == 4 ? 4 : 1);
            if (MEBIT(25)) { e->valid |= V_TAS; e->tas = (uint16_t)speed; } else { e->valid |= V_IAS; e->ias = (uint16_t)speed; }
        }
    }
    raw = MEBITS(38, 46);
    if (raw) {
        const int rate = (int)(raw - 1) * (MEBIT(37) ? -64 : 64);
        if (MEBIT(36)) { e->baro_rate = rate; e->valid |= V_BARO_RATE; } else { e->geom_rate = rate; e->valid |= V_GEOM_RATE; }
    }
    raw = MEBITS(50, 56);
    if (raw) { e->valid |= V_GEOM_DELTA; e->geom_delta = (int)(raw - 1) * (MEBIT(49) ? -25 : 25); }
}

static void es_surface(oracle_fields_es *e, const uint8_t *me, int check_imf)    /* decodeESSurfacePosition, mode_s.c:965-1001 */
{
    const unsigned movement = MEBITS(6, 12);
    e->airground = 1;
    e->valid |= V_CPR;
    e->cpr_type = 0;
    if (movement > 0 && movement < 125) {
        e->valid |= V_GS;
        e->gs_selected = e->gs_v0 = es_movement(movement, 0);
        e->gs_v2 = es_movement(movement, 1);
    }
    if (MEBIT(13)) { e->valid |= V_HEADING; e->heading = MEBITS(14, 20) * 360.0 / 128.0; e->heading_type = H_TRACK_OR_HEADING; }
    if (check_imf && MEBIT(21)) es_set_imf(e);
    e->cpr_odd = (uint8_t)MEBIT(22);
    e->cpr_lat = MEBITS(23, 39);
    e->cpr_lon = MEBITS(40, 56);
}

static void es_airborne(oracle_fields_es *e, const uint8_t *me, int check_imf)   /* decodeESAirbornePosition, mode_s.c:1003-1092 */
{
    const unsigned ac12 = MEBITS(9, 20);
    switch (MEBITS(6, 7)) {
    case 0: e->valid |= V_ALERT | V_SPI; e->alert = e->spi = 0; break;
    case 1: case 2: e->valid |= V_ALERT; e->alert = 1; break;
    default: e->valid |= V_ALERT | V_SPI; e->alert = 0; e->spi = 1; break;
    }
    if (check_imf) { if (MEBIT(8)) es_set_imf(e); }
    else { e->valid |= V_NIC_B; e->nic_b = (uint8_t)MEBIT(8); }
    if (e->metype != 0) {
        e->cpr_lat = MEBITS(23, 39);
        e->cpr_lon = MEBITS(40, 56);
        if (!(ac12 == 0 && e->cpr_lon == 0 && (e->cpr_lat & 0x0fff) == 0 && e->metype == 15)) {
            e->valid |= V_CPR;
            e->cpr_type = 1;
            e->cpr_odd = (uint8_t)MEBIT(22);
        }
    }
    if (ac12 && e->airground != 1) {
        uint8_t unit;
        const int alt = es_ac12((int)ac12, &unit);
        if (alt != -9999) {
            if (e->airground == 0) e->airground = 3;
            if (e->metype >= 20 && e->metype <= 22) { e->altitude_geom = alt; e->altitude_geom_unit = unit; e->valid |= V_ALT_GEOM; }
            else { e->altitude_baro = alt; e->altitude_baro_unit = unit; e->valid |= V_ALT_BARO; }
        }
    }
}

static void es_target(oracle_fields_es *e, const uint8_t *me, int check_imf)     /* decodeESTargetStatus, mode_s.c:1131-1324 */
{
    e->mesub = (uint8_t)MEBITS(6, 7);
    if (check_imf && MEBIT(51)) es_set_imf(e);
    if (e->mesub == 0 && MEBIT(11) == 0) {
        unsigned h_source;
        int alt;
        switch (MEBITS(8, 9)) { case 1: e->nav_altitude_source = NA_MCP; break; case 2: e->nav_altitude_source = NA_AIRCRAFT; break;
                                case 3: e->nav_altitude_source = NA_FMS; break; default: break; }
        switch (MEBITS(14, 15)) {
        case 1:
            e->valid |= V_NAV_MODES;
            e->nav_modes |= e->nav_altitude_source == NA_FMS ? NM_VNAV : NM_AUTOPILOT;
            break;
        case 2:
            e->valid |= V_NAV_MODES;
            e->nav_modes |= e->nav_altitude_source == NA_FMS ? NM_VNAV : e->nav_altitude_source == NA_AIRCRAFT ? NM_ALT_HOLD : NM_AUTOPILOT;
            break;
        default: break;
        }
        alt = -1000 + 100 * (int)MEBITS(16, 25);
        if (e->nav_altitude_source == NA_MCP) { e->valid |= V_NAV_MCP; e->nav_mcp_altitude = alt; }
        else if (e->nav_altitude_source == NA_FMS) { e->valid |= V_NAV_FMS; e->nav_fms_altitude = alt; }
        h_source = MEBITS(26, 27);
        if (h_source != 0) {
            e->valid |= V_NAV_HEADING;
            e->nav_heading = MEBITS(28, 36);
            e->nav_heading_type = MEBIT(37) ? H_GROUND_TRACK : H_MAGNETIC_OR_TRUE;
        }
        switch (MEBITS(38, 39)) {
        case 1: case 2: e->valid |= V_NAV_MODES; e->nav_modes |= h_source == 3 ? NM_LNAV : NM_AUTOPILOT; break;
        default: break;
        }
        e->valid |= V_NAC_P | V_NIC_BARO;
        e->nac_p = (uint8_t)MEBITS(40, 43);
        e->nic_baro = (uint8_t)MEBIT(44);
        e->sil = (uint8_t)MEBITS(45, 46);
        e->sil_type = 1;
        switch (MEBITS(52, 53)) {
        case 1: e->valid |= V_NAV_MODES; break;
        case 2: case 3: e->valid |= V_NAV_MODES; e->nav_modes |= NM_TCAS; break;
        default: e->nav_modes |= NM_TCAS; break;
        }
        e->valid |= V_EMERGENCY;
        e->emergency = (uint8_t)MEBITS(54, 56);
    } else if (e->mesub == 1) {
        const unsigned is_fms = MEBIT(9), alt_bits = MEBITS(10, 20), baro_bits = MEBITS(21, 29);
        if (alt_bits != 0) {
            if (is_fms) { e->valid |= V_NAV_FMS; e->nav_fms_altitude = (int)(alt_bits - 1) * 32; }
            else { e->valid |= V_NAV_MCP; e->nav_mcp_altitude = (int)(alt_bits - 1) * 32; }
        }
        if (baro_bits != 0) { e->valid |= V_NAV_QNH; e->nav_qnh = 800.0 + (baro_bits - 1) * 0.8; }
        if (MEBIT(30)) { e->valid |= V_NAV_HEADING; e->nav_heading = MEBITS(31, 39) * 180.0 / 256.0; e->nav_heading_type = H_MAGNETIC_OR_TRUE; }
        e->valid |= V_NAC_P | V_NIC_BARO;
        e->nac_p = (uint8_t)MEBITS(40, 43);
        e->nic_baro = (uint8_t)MEBIT(44);
        e->sil = (uint8_t)MEBITS(45, 46);
        e->sil_type = 1;
        if (MEBIT(47)) {
            e->valid |= V_NAV_MODES;
            e->nav_modes = (uint8_t)((MEBIT(48) ? NM_AUTOPILOT : 0) | (MEBIT(49) ? NM_VNAV : 0) | (MEBIT(50) ? NM_ALT_HOLD : 0) |
                                     (MEBIT(52) ? NM_APPROACH : 0) | (MEBIT(53) ? NM_TCAS : 0) | (MEBIT(54) ? NM_LNAV : 0));
        }
    }
}

static void es_opstatus(oracle_fields_es *e, const uint8_t *me, int check_imf)   /* decodeESOperationalStatus, mode_s.c:1326-1445 */
{
    unsigned o = 0;
    e->mesub = (uint8_t)MEBITS(6, 8);
    if (check_imf && MEBIT(56)) es_set_imf(e);
    if (e->mesub > 1) return;
    e->valid |= V_OPSTATUS;
    e->ops_version = (uint8_t)MEBITS(41, 43);
    switch (e->ops_version) {
    case 0:
        if (e->mesub == 0 && MEBITS(9, 10) == 0) { if (!MEBIT(12)) o |= O_CC_ACAS; if (MEBIT(13)) o |= O_CC_CDTI; }
        break;
    case 1:
        if (MEBITS(25, 26) == 0) { if (MEBIT(27)) o |= O_OM_ACAS_RA; if (MEBIT(28)) o |= O_OM_IDENT; if (MEBIT(29)) o |= O_OM_ATC; }
        if (e->mesub == 0 && MEBITS(9, 10) == 0 && MEBITS(13, 14) == 0) {
            if (!MEBIT(11)) o |= O_CC_ACAS;
            if (MEBIT(12)) o |= O_CC_CDTI;
            if (MEBIT(15)) o |= O_CC_ARV;
            if (MEBIT(16)) o |= O_CC_TS;
            e->ops_cc_tc = (uint8_t)MEBITS(17, 18);
        } else if (e->mesub == 1 && MEBITS(9, 10) == 0 && MEBITS(13, 14) == 0) {
            if (MEBIT(11)) o |= O_CC_POA;
            if (MEBIT(12)) o |= O_CC_CDTI;
            if (MEBIT(15)) o |= O_CC_B2_LOW;
            e->valid |= V_CC_LW;
            e->ops_cc_lw = (uint8_t)MEBITS(21, 24);
        }
        e->valid |= V_NIC_A | V_NAC_P;
        e->nic_a = (uint8_t)MEBIT(44);
        e->nac_p = (uint8_t)MEBITS(45, 48);
        e->sil_type = 1;
        e->sil = (uint8_t)MEBITS(51, 52);
        e->ops_hrd = MEBIT(54) ? H_MAGNETIC : H_TRUE;
        if (e->mesub == 0) { e->valid |= V_NIC_BARO; e->nic_baro = (uint8_t)MEBIT(53); }
        else e->ops_tah = MEBIT(53) ? e->ops_hrd : H_GROUND_TRACK;
        break;
    case 2:
        if (MEBITS(25, 26) == 0) {
            if (MEBIT(27)) o |= O_OM_ACAS_RA;
            if (MEBIT(28)) o |= O_OM_IDENT;
            if (MEBIT(29)) o |= O_OM_ATC;
            if (MEBIT(30)) o |= O_OM_SAF;
            e->valid |= V_SDA;
            e->sda = (uint8_t)MEBITS(31, 32);
        }
        if (e->mesub == 0 && MEBITS(9, 10) == 0) {
            if (MEBIT(11)) o |= O_CC_ACAS;
            if (MEBIT(12)) o |= O_CC_1090_IN;
            if (MEBIT(15)) o |= O_CC_ARV;
            if (MEBIT(16)) o |= O_CC_TS;
            e->ops_cc_tc = (uint8_t)MEBITS(17, 18);
            if (MEBIT(19)) o |= O_CC_UAT_IN;
        } else if (e->mesub == 1 && MEBITS(9, 10) == 0) {
            if (MEBIT(11)) o |= O_CC_POA;
            if (MEBIT(12)) o |= O_CC_1090_IN;
            if (MEBIT(15)) o |= O_CC_B2_LOW;
            if (MEBIT(16)) o |= O_CC_UAT_IN;
            e->valid |= V_NAC_V | V_NIC_C | V_CC_LW;
            e->nac_v = (uint8_t)MEBITS(17, 19);
            e->nic_c = (uint8_t)MEBIT(20);
            e->ops_cc_lw = (uint8_t)MEBITS(21, 24);
            e->ops_cc_antenna_offset = (uint8_t)MEBITS(33, 40);
        }
        e->valid |= V_NIC_A | V_NAC_P;
        e->nic_a = (uint8_t)MEBIT(44);
        e->nac_p = (uint8_t)MEBITS(45, 48);
        e->sil = (uint8_t)MEBITS(51, 52);
        e->sil_type = MEBIT(55) ? 2 : 3;
        e->ops_hrd = MEBIT(54) ? H_MAGNETIC : H_TRUE;
        if (e->mesub == 0) { e->valid |= V_GVA | V_NIC_BARO; e->gva = (uint8_t)MEBITS(49, 50); e->nic_baro = (uint8_t)MEBIT(53); }
        else e->ops_tah = MEBIT(53) ? e->ops_hrd : H_GROUND_TRACK;
        break;
    default: break;
    }
    e->ops_flags = (uint16_t)o;
}

/* msg: the 14 corrected bytes; df: the downlink format; corrected: bits repaired (mm->correctedbits). */
void oracle1090_decode_es(const uint8_t *msg, int df, int corrected, oracle_fields_es *e)
{
    static const char ais[65] = "@ABCDEFGHIJKLMNOPQRSTUVWXYZ[\\]^_ !\"#$%&'()*+,-./0123456789:;<=>?";   /* ais_charset.c:3 */
    const uint8_t *me = msg + 4;
    int check_imf = 0, k;
    memset(e, 0, sizeof(*e));
    if (df != 17 && df != 18) return;
    e->es = 1;
    e->source = SRC_ADSB;                                                         /* mode_s.c:580-585 */
    e->reliable = corrected == 0;
    e->addr = f_getbits(msg, 9, 32);                                              /* mode_s.c:624-626 */
    e->CA_CF = (uint8_t)f_getbits(msg, 6, 8);
    if (df == 17)                                                                 /* CA, mode_s.c:636-657 */
        switch (e->CA_CF) { case 0: case 6: case 7: e->airground = 3; break; case 4: e->airground = 1; break; case 5: e->airground = 2; break; default: break; }
    e->metype = (uint8_t)MEBITS(1, 5);                                            /* decodeExtendedSquitter, mode_s.c:1447-1551 */
    if (df == 18) {
        switch (e->CA_CF) {
        case 0: e->addrtype = AT_ADSB_ICAO_NT; break;
        case 1: e->addrtype = AT_ADSB_OTHER; e->addr |= NON_ICAO; break;
        case 2: e->source = SRC_TISB; e->addrtype = AT_TISB_ICAO; check_imf = 1; break;
        case 3: e->source = SRC_TISB; e->addrtype = AT_TISB_ICAO; if (MEBIT(1)) es_set_imf(e); return;
        case 5: e->addrtype = AT_TISB_OTHER; e->source = SRC_TISB; e->addr |= NON_ICAO; break;
        case 6: e->addrtype = AT_ADSR_ICAO; e->source = SRC_ADSR; check_imf = 1; break;
        default: e->addrtype = AT_UNKNOWN; e->addr |= NON_ICAO; return;
        }
    }
    switch (e->metype) {
    case 1: case 2: case 3: case 4:                                               /* decodeESIdentAndCategory, mode_s.c:798-829 */
        e->mesub = (uint8_t)MEBITS(6, 8);
        e->valid |= V_CALLSIGN | V_CATEGORY;
        for (k = 0; k < 8; ++k) {
            const char c = ais[MEBITS(9 + 6 * k, 14 + 6 * k)];
            e->callsign[k] = c;
        }
        for (k = 0; k < 8; ++k) {
            const char c = e->callsign[k];
            if (!((c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == ' ')) { e->valid &= ~(uint32_t)V_CALLSIGN; break; }
        }
        e->category = (uint8_t)(((0x0E - e->metype) << 4) | e->mesub);
        break;
    case 19: es_velocity(e, me, check_imf); break;
    case 5: case 6: case 7: case 8: es_surface(e, me, check_imf); break;
    case 0: case 9: case 10: case 11: case 12: case 13: case 14: case 15: case 16: case 17: case 18: case 20: case 21: case 22:
        es_airborne(e, me, check_imf);
        break;
    case 23:                                                                      /* decodeESTestMessage, mode_s.c:1094-1107 */
        e->mesub = (uint8_t)MEBITS(6, 8);
        if (e->mesub == 7) { const int id = (int)MEBITS(9, 21); if (id) { e->valid |= V_SQUAWK; e->squawk = (uint16_t)f_id13(id); } }
        break;
    case 28:                                                                      /* decodeESAircraftStatus, mode_s.c:1109-1129 */
        e->mesub = (uint8_t)MEBITS(6, 8);
        if (e->mesub == 1) {
            const int id = (int)MEBITS(12, 24);
            e->valid |= V_EMERGENCY;
            e->emergency = (uint8_t)MEBITS(9, 11);
            if (id) { e->valid |= V_SQUAWK; e->squawk = (uint16_t)f_id13(id); }
            if (check_imf && MEBIT(56)) es_set_imf(e);
        }
        break;
    case 29: es_target(e, me, check_imf); break;
    case 31: es_opstatus(e, me, check_imf); break;
    case 24: case 30: break;
    default: e->reliable = 0; break;                                              /* "Dubious", mode_s.c:1546-1549 */
    }
}

/* n frames of 14 corrected bytes each; corrected[k] = bits repaired (NULL: none); DF from the first five bits. */
void oracle1090_decode_es_batch(const uint8_t *frames, size_t n, const uint8_t *corrected, oracle_fields_es *out)
{
    size_t k;
    for (k = 0; k < n; ++k) oracle1090_decode_es(frames + 14 * k, frames[14 * k] >> 3, corrected ? corrected[k] : 0, out + k);
}

/* ---- Comm-B payload of DF20 / DF21 (decodeCommB and its register decoders, comm_b.c:50-985), for adsbcu_1090_decode_commb ----
 * Pinned against the reference binary's own decodeModesMessage through the third tap of oracle/ref_harness.c
 * (tests/golden/fields_mb.npz).  Layout == adsbcu_fields_commb. */
typedef struct oracle_fields_mb {
    uint32_t valid;
    uint8_t  commb_format, heading_type, nav_modes, nav_altitude_source, mrar_source, turbulence, reserved[2];
    char     callsign[8];
    float    roll, heading, track_rate, gs, mach, nav_qnh, wind_speed, wind_dir, temperature, pressure, humidity;
    int32_t  baro_rate, geom_rate, nav_mcp_altitude, nav_fms_altitude;
    uint16_t ias, tas;
} oracle_fields_mb;

enum { B_CALLSIGN = 1u << 0, B_ROLL = 1u << 1, B_HEADING = 1u << 2, B_GS = 1u << 3, B_TRACK_RATE = 1u << 4, B_TAS = 1u << 5, B_IAS = 1u << 6,
       B_MACH = 1u << 7, B_BARO_RATE = 1u << 8, B_GEOM_RATE = 1u << 9, B_NAV_MCP = 1u << 10, B_NAV_FMS = 1u << 11, B_NAV_QNH = 1u << 12,
       B_NAV_MODES = 1u << 13, B_MRAR_SOURCE = 1u << 14, B_WIND = 1u << 15, B_TEMPERATURE = 1u << 16, B_PRESSURE = 1u << 17,
       B_TURBULENCE = 1u << 18, B_HUMIDITY = 1u << 19 };
/* commb_format_t, dump1090.h:203-217 */
enum { CB_UNKNOWN, CB_AMBIGUOUS, CB_NOT_DECODED, CB_EMPTY_RESPONSE, CB_DATALINK_CAPS, CB_GICB_CAPS, CB_AIRCRAFT_IDENT, CB_ACAS_RA,
       CB_VERTICAL_INTENT, CB_TRACK_TURN, CB_HEADING_SPEED, CB_MRAR, CB_AIRBORNE_POSITION };

typedef struct mb_ctx { const uint8_t *mb; int df; unsigned ac; } mb_ctx;
#define MB(a, b) f_getbits(c->mb, (a), (b))
#define MB1(a) f_getbits(c->mb, (a), (a))

static int mb_empty(const mb_ctx *c, oracle_fields_mb *o)                        /* decodeEmptyResponse, comm_b.c:88-126 */
{
    int i;
    if (c->mb[0] != 0x00 && c->mb[0] != 0x40 && c->mb[0] != 0x50 && c->mb[0] != 0x60) return 0;
    for (i = 1; i < 7; ++i) if (c->mb[i]) return 0;
    if (o) o->commb_format = CB_EMPTY_RESPONSE;
    return 56;
}

static int mb_bds10(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:129-150 */
{
    if (c->mb[0] != 0x10 || MB(10, 14) != 0) return 0;
    if (o) o->commb_format = CB_DATALINK_CAPS;
    return 56;
}

static int mb_bds17(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:153-234 */
{
    int score = 0;
    const unsigned es = MB(1, 6);
    if (MB(25, 56) != 0) return 0;
    score += MB1(7) ? 1 : -2;
    if (MB1(10)) score -= 2;
    if (MB1(11)) score -= 2;
    if (MB1(12)) score -= 2;
    if (MB1(13)) score -= 1;
    if (MB1(14)) score -= 1;
    if (MB1(20)) score -= 2;
    if (MB1(21)) score -= 2;
    if (MB1(22)) score -= 2;
    if ((es >> 1) == 0x1F) score += 5 + (es & 1 ? 1 : 0);
    else if (es == 0) score += 1;
    else if ((es >> 1) == 0x07) score += 3;
    else score -= 12;
    if (MB1(16) && MB1(24)) score += 2 + (MB1(9) ? 1 : 0);
    else if (!MB1(16) && !MB1(24) && !MB1(9)) score += 1;
    else score -= 6;
    if (o) o->commb_format = CB_GICB_CAPS;
    return score;
}

static int mb_bds20(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:237-281 */
{
    static const char ais[65] = "@ABCDEFGHIJKLMNOPQRSTUVWXYZ[\\]^_ !\"#$%&'()*+,-./0123456789:;<=>?";
    char cs[8];
    int score = 8, valid = 1, k;
    if (c->mb[0] != 0x20) return 0;
    for (k = 0; k < 8; ++k) cs[k] = ais[MB(9 + 6 * k, 14 + 6 * k)];
    for (k = 0; k < 8; ++k) {
        if ((cs[k] >= 'A' && cs[k] <= 'Z') || (cs[k] >= '0' && cs[k] <= '9') || cs[k] == ' ') score += 6;
        else if (cs[k] == '@') valid = 0;
        else return 0;
    }
    if (o) {
        o->commb_format = CB_AIRCRAFT_IDENT;
        if (valid) { memcpy(o->callsign, cs, 8); o->valid |= B_CALLSIGN; }
    }
    return score;
}

static int mb_bds30(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:284-299 */
{
    if (c->mb[0] != 0x30) return 0;
    if (o) o->commb_format = CB_ACAS_RA;
    return 56;
}

static int mb_bds40(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:302-465 */
{
    const unsigned mcp_valid = MB1(1), mcp_raw = MB(2, 13), fms_valid = MB1(14), fms_raw = MB(15, 26), baro_valid = MB1(27), baro_raw = MB(28, 39);
    const unsigned mode_valid = MB1(48), mode_raw = MB(49, 51), source_valid = MB1(54), source_raw = MB(55, 56);
    unsigned mcp_alt = 0, fms_alt = 0;
    float baro = 0;
    int score = 0;
    if (!mcp_valid && !fms_valid && !baro_valid && !mode_valid && !source_valid) return 0;
    if (mcp_valid && mcp_raw != 0) { mcp_alt = mcp_raw * 16; if (mcp_alt >= 1000 && mcp_alt <= 50000) score += 13; else return 0; }
    else if (!mcp_valid && mcp_raw == 0) score += 1;
    else return 0;
    if (fms_valid && fms_raw != 0) { fms_alt = fms_raw * 16; if (fms_alt >= 1000 && fms_alt <= 50000) score += 13; else return 0; }
    else if (!fms_valid && fms_raw == 0) score += 1;
    else return 0;
    if (baro_valid && baro_raw != 0) { baro = 800 + baro_raw * 0.1; if (baro >= 900 && baro <= 1100) score += 13; else return 0; }
    else if (!baro_valid && baro_raw == 0) score += 1;
    else return 0;
    if (MB(40, 47) != 0) return 0;
    if (mode_valid) score += 4; else if (mode_raw == 0) score += 1; else return 0;
    if (MB(52, 53) != 0) return 0;
    if (source_valid) score += 3; else if (source_raw == 0) score += 1; else return 0;
    if (mcp_valid && fms_valid && mcp_alt != fms_alt) score -= 4;
    if (mcp_valid) { const unsigned r = mcp_alt % 500; if (!(r < 16 || r > 484)) score -= 4; }
    if (fms_valid) { const unsigned r = fms_alt % 500; if (!(r < 16 || r > 484)) score -= 4; }
    if (o) {
        o->commb_format = CB_VERTICAL_INTENT;
        if (mcp_valid) { o->valid |= B_NAV_MCP; o->nav_mcp_altitude = (int32_t)mcp_alt; }
        if (fms_valid) { o->valid |= B_NAV_FMS; o->nav_fms_altitude = (int32_t)fms_alt; }
        if (baro_valid) { o->valid |= B_NAV_QNH; o->nav_qnh = baro; }
        if (mode_valid) { o->valid |= B_NAV_MODES; o->nav_modes = (uint8_t)(((mode_raw & 4) ? 2 : 0) | ((mode_raw & 2) ? 4 : 0) | ((mode_raw & 1) ? 8 : 0)); }
        o->nav_altitude_source = source_valid ? (uint8_t)(source_raw + 1) : 0;    /* UNKNOWN, AIRCRAFT, MCP, FMS == 1..4 */
    }
    return score;
}

static int mb_bds50(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:468-623 */
{
    const unsigned roll_valid = MB1(1), roll_sign = MB1(2), roll_raw = MB(3, 11), track_valid = MB1(12), track_sign = MB1(13), track_raw = MB(14, 23);
    const unsigned gs_valid = MB1(24), gs_raw = MB(25, 34), rate_valid = MB1(35), rate_sign = MB1(36), rate_raw = MB(37, 45);
    const unsigned tas_valid = MB1(46), tas_raw = MB(47, 56);
    float roll, track, track_rate = 0;
    unsigned gs, tas;
    int score = 0;
    if (!roll_valid || !track_valid || !gs_valid || !tas_valid) return 0;
    roll = roll_raw * 45.0 / 256.0;
    if (roll_sign) roll -= 90.0;
    if (roll >= -40 && roll < 40) score += 11; else return 0;
    score += 12;
    track = track_raw * 90.0 / 512.0;
    if (track_sign) track += 180.0;
    if (gs_raw == 0) return 0;
    gs = gs_raw * 2;
    if (gs >= 50 && gs <= 700) score += 11; else return 0;
    if (rate_valid) {
        track_rate = rate_raw * 8.0 / 256.0;
        if (rate_sign) track_rate -= 16;
        if (track_rate >= -10.0 && track_rate <= 10.0) score += 11; else return 0;
    } else if (rate_raw == 0 && !rate_sign) score += 1;
    else return 0;
    if (tas_raw == 0) return 0;
    tas = tas_raw * 2;
    if (tas >= 50 && tas <= 700) score += 11; else return 0;
    /* comm_b.c:576-581 compares the two VALID bits (always 1 - 1), so its "inconsistent data" penalty never applies */
    if (rate_valid) {                                                             /* comm_b.c:584-590 */
        const double turn_rate = 68625 * tan(roll * M_PI / 180.0) / (tas * 20 * M_PI);
        if (fabs(turn_rate - track_rate) > 2.0) score -= 6;
    }
    if (o) {
        o->commb_format = CB_TRACK_TURN;
        o->valid |= B_ROLL | B_HEADING | B_GS | B_TAS;
        o->roll = roll;
        o->heading = track;
        o->heading_type = 1;
        o->gs = gs;
        if (rate_valid) { o->valid |= B_TRACK_RATE; o->track_rate = track_rate; }
        o->tas = (uint16_t)tas;
    }
    return score;
}

static int mb_bds60(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:626-775 */
{
    const unsigned hdg_valid = MB1(1), hdg_sign = MB1(2), hdg_raw = MB(3, 12), ias_valid = MB1(13), ias_raw = MB(14, 23);
    const unsigned mach_valid = MB1(24), mach_raw = MB(25, 34), baro_valid = MB1(35), baro_sign = MB1(36), baro_raw = MB(37, 45);
    const unsigned ins_valid = MB1(46), ins_sign = MB1(47), ins_raw = MB(48, 56);
    float heading, mach;
    int baro_rate = 0, ins_rate = 0, score = 0;
    if (!hdg_valid || !ias_valid || !mach_valid || (!baro_valid && !ins_valid)) return 0;
    heading = hdg_raw * 90.0 / 512.0;
    if (hdg_sign) heading += 180.0;
    score += 12;
    if (ias_raw == 0) return 0;
    if (ias_raw >= 50 && ias_raw <= 700) score += 11; else return 0;
    if (mach_raw == 0) return 0;
    mach = mach_raw * 2.048 / 512;
    if (mach >= 0.1 && mach <= 0.9) score += 11; else return 0;
    if (baro_valid) {
        baro_rate = (int)baro_raw * 32 - (baro_sign ? 16384 : 0);
        if (baro_rate >= -6000 && baro_rate <= 6000) score += 11; else return 0;
    } else if (baro_raw == 0) score += 1;
    else return 0;
    if (ins_valid) {
        ins_rate = (int)ins_raw * 32 - (ins_sign ? 16384 : 0);
        if (ins_rate >= -6000 && ins_rate <= 6000) score += 11; else return 0;
    } else if (ins_raw == 0) score += 1;
    else return 0;
    if (baro_valid && ins_valid && abs(baro_rate - ins_rate) > 2000) score -= 12;
    if (o) {
        o->commb_format = CB_HEADING_SPEED;
        o->valid |= B_HEADING | B_IAS | B_MACH;
        o->heading = heading;
        o->heading_type = 3;
        o->ias = (uint16_t)ias_raw;
        o->mach = mach;
        if (baro_valid) { o->valid |= B_BARO_RATE; o->baro_rate = baro_rate; }
        if (ins_valid) { o->valid |= B_GEOM_RATE; o->geom_rate = ins_rate; }
    }
    return score;
}

static int mb_bds44(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:778-937 */
{
    const unsigned source = MB(1, 4), wind_valid = MB1(5), windspeed_raw = MB(6, 14), winddir_raw = MB(15, 23);
    const unsigned sat_valid = MB1(24), sat_sign = MB1(25), sat_raw = MB(26, 34), asp_valid = MB1(35), asp_raw = MB(36, 46);
    const unsigned turb_valid = MB1(47), turb_raw = MB(48, 49), hum_valid = MB1(50), hum_raw = MB(51, 56);
    float sat;
    int score = 0;
    if (source == 0 || source >= 5) return 0;
    if (!wind_valid || !sat_valid) return 0;
    if (!asp_valid && asp_raw != 0) return 0;
    if (!turb_valid && turb_raw != 0) return 0;
    if (!hum_valid && hum_raw != 0) return 0;
    if (windspeed_raw == 0) score += 2; else if (windspeed_raw <= 250) score += 19; else return 0;
    sat = sat_raw * 0.25;
    if (sat_sign) sat -= 128;
    if (sat == 0) score += 2; else if (sat >= -80 && sat <= 60) score += 11; else return 0;
    if (asp_valid) { if (asp_raw >= 25 && asp_raw <= 1100) score += 12; else return 0; } else score += 1;
    score += turb_valid ? 3 : 1;
    score += hum_valid ? 7 : 1;
    if (source == 3 && score > 0) score = 1;                                      /* GICB look-alike, comm_b.c:889-902 */
    if (o) {
        o->commb_format = CB_MRAR;
        o->valid |= B_MRAR_SOURCE | B_WIND | B_TEMPERATURE;
        o->mrar_source = (uint8_t)source;
        o->wind_speed = windspeed_raw;
        o->wind_dir = winddir_raw * (180.0 / 256.0);
        o->temperature = sat;
        if (asp_valid) { o->valid |= B_PRESSURE; o->pressure = asp_raw; }
        if (turb_valid) { o->valid |= B_TURBULENCE; o->turbulence = (uint8_t)turb_raw; }
        if (hum_valid) { o->valid |= B_HUMIDITY; o->humidity = hum_raw * (100.0 / 64.0); }
    }
    return score;
}

static int mb_bds05(const mb_ctx *c, oracle_fields_mb *o)                        /* comm_b.c:944-985 */
{
    const unsigned typecode = MB(1, 5), ac12 = MB(9, 20);
    if (c->df != 20) return 0;
    if (typecode < 9 || typecode > 18) return 0;
    if (MB1(21)) return 0;
    if (!ac12) return 0;
    if (c->ac != (((ac12 & 0x0FC0) << 1) | (ac12 & 0x003F))) return 0;
    if (MB(23, 39) == 0 || MB(40, 56) == 0) return 0;
    if (o) o->commb_format = CB_AIRBORNE_POSITION;
    return 100;
}

void oracle1090_decode_commb(const uint8_t *msg, int df, int corrected, oracle_fields_mb *o)   /* decodeCommB, comm_b.c:50-86 */
{
    typedef int (*fn)(const mb_ctx *, oracle_fields_mb *);
    static const fn order[10] = {mb_empty, mb_bds10, mb_bds20, mb_bds30, mb_bds17, mb_bds40, mb_bds50, mb_bds60, mb_bds44, mb_bds05};
    mb_ctx c;
    int best = 0, ambiguous = 0, k;
    fn winner = NULL;
    memset(o, 0, sizeof(*o));
    if (df != 20 && df != 21) return;
    c.mb = msg + 4;
    c.df = df;
    c.ac = df == 20 ? f_getbits(msg, 20, 32) : 0;                                 /* mm->AC, mode_s.c:628-631 */
    /* DR (bits 9-13, mode_s.c:668-670); UM is still zero when the reference gets here (mode_s.c:731-734 precedes :757-760) */
    if (f_getbits(msg, 9, 13) != 0 || corrected > 0) { o->commb_format = CB_NOT_DECODED; return; }
    for (k = 0; k < 10; ++k) {
        const int score = order[k](&c, NULL);
        if (score > best) { best = score; winner = order[k]; ambiguous = 0; }
        else if (score == best) ambiguous = 1;
    }
    if (!winner) o->commb_format = CB_UNKNOWN;
    else if (ambiguous) o->commb_format = CB_AMBIGUOUS;
    else winner(&c, o);
}

void oracle1090_decode_commb_batch(const uint8_t *frames, size_t n, const uint8_t *corrected, oracle_fields_mb *out)
{
    size_t k;
    for (k = 0; k < n; ++k) oracle1090_decode_commb(frames + 14 * k, frames[14 * k] >> 3, corrected ? corrected[k] : 0, out + k);
}
