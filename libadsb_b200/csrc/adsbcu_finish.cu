// order_kernel + level_kernel -- the ordered tail of demodulate2400() on the device (sm_100a).
//
// What is sequential in the reference (demod_2400.c:222,336,349-355,410-419,432-443 and mode_s.c:781-788): pick the
// best-scoring phase per start position given the CURRENT ICAO filter, accept, skip over the accepted message, add the
// address to the filter, carry last_message_end across mag_buf boundaries.  Round 1 replayed this on one host thread
// (0.7 ms per 60 s capture, the slowest stage of the resident step).  Here it runs on the GPU, exactly:
//
//  order_kernel (one CTA per front-kernel segment)
//    * parked Address/Parity-style records whose address has been announced by now are evaluated from the symbol
//      planes like the message kernel does;
//    * the segment's records are rank-sorted by (position, phase);
//    * segments in index order are the chunk in stream order: every CTA publishes its count, sums the counts before
//      it and writes its sorted records (and a 16-byte view with the filter state folded in) to their final place;
//  resolve_kernel (one CTA of 1024 threads)
//    * the skip rule cannot reach across a gap of 288 samples (a message ends at most 288 samples after its start),
//      so the record stream splits into independent RUNS there; the only state shared between runs is the ICAO
//      filter.  Runs are resolved in parallel (contiguous blocks of runs per thread) against
//          known(key, pos) = key is in the reference's two 4096-slot tables  OR  first_add[key] < pos
//      where first_add (a small hash map in shared memory) holds the addresses that accepted messages of THIS chunk
//      add.  Iterating "resolve all runs -> collect first_add" reaches the sequential result: the decision at a
//      position only depends on adds at earlier positions, so the first wrong decision moves strictly later with
//      every round, and a round that reproduces its own map is the unique consistent answer.  In steady state (no
//      new address in the chunk) that is one round;
//    * accepted messages are written in stream order as adsbcu_msg records into a per-chunk device buffer, the filter
//      tables are updated verbatim (icao_filter.c:71-84, in stream order);
//  level_kernel (a second, wide launch)
//    * mean squared magnitude of every ACCEPTED message (demod_2400.c:382-403), one warp per message, and the final
//      record goes to mapped host memory or the caller's pinned buffer.
//    If the tables could fill up (4096 slots), the map overflows or the rounds do not settle, one thread replays
//    the chunk with the reference's own sequential loop: slow, exact, never seen on realistic traffic.
#include "adsbcu_device.cuh"
#include "../../include/adsbcu.h"

namespace adsbcu {
namespace {

constexpr int kXThreads = 256;
constexpr uint32_t kMapSlots = 2048, kMapMaxKeys = 1024, kMapEmpty = 0xFFFFFFFFu;
constexpr uint32_t kMapDropped = 0xFFFFFFFEu;     // first_add of an address that was announced when the table was already full
constexpr int kSortCap = 4096;                 // records of one segment that phase 1 can sort in shared memory

// ---- the reference's ICAO filter on the device (icao_filter.c:40-109), verbatim -------------------------------
__device__ __forceinline__ uint32_t icao_hash(uint32_t v)
{
    uint32_t h = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        h += (v >> (8 * k)) & 0xFFu;
        h += h << 10;
        h ^= h >> 6;
    }
    h += h << 3;
    h ^= h >> 11;
    h += h << 15;
    return h & (kIcaoSlots - 1);
}
__device__ __forceinline__ bool icao_probe(const uint32_t *t, uint32_t v)
{
    const uint32_t h0 = icao_hash(v);
    uint32_t h = h0;
    while (t[h] != kIcaoEmpty && t[h] != v) {
        h = (h + 1) & (kIcaoSlots - 1);
        if (h == h0) break;
    }
    return t[h] == v;
}
__device__ __forceinline__ bool icao_test(const StreamState *st, uint32_t v) { return icao_probe(st->table[0], v) || icao_probe(st->table[1], v); }
__device__ __forceinline__ void icao_add(StreamState *st, uint32_t v)
{
    uint32_t *t = st->table[st->active];
    const uint32_t h0 = icao_hash(v);
    uint32_t h = h0;
    while (t[h] != kIcaoEmpty && t[h] != v) {
        h = (h + 1) & (kIcaoSlots - 1);
        if (h == h0) return;                    // table full: the reference drops the address
    }
    if (t[h] == kIcaoEmpty) { t[h] = v; ++st->active_count; }
}

// ---- first_add map (shared memory): key -> position of the first accepted message of this chunk that adds it ----
struct AddMap {
    uint32_t key[kMapSlots];
    uint32_t pos[kMapSlots];
};
__device__ __forceinline__ uint32_t map_hash(uint32_t k) { return (k * 2654435761u) >> 21 & (kMapSlots - 1); }
__device__ __forceinline__ uint32_t map_lookup(const AddMap &m, uint32_t k)    // position, or 0xFFFFFFFF
{
    uint32_t h = map_hash(k);
    for (uint32_t n = 0; n < kMapSlots; ++n) {
        const uint32_t e = m.key[h];
        if (e == k) return m.pos[h];
        if (e == kMapEmpty) return kMapEmpty;
        h = (h + 1) & (kMapSlots - 1);
    }
    return kMapEmpty;
}
__device__ __forceinline__ void map_insert_min(AddMap &m, uint32_t k, uint32_t pos, uint32_t *nkeys)
{
    uint32_t h = map_hash(k);
    for (uint32_t n = 0; n < kMapSlots; ++n) {
        const uint32_t e = atomicCAS(&m.key[h], kMapEmpty, k);
        if (e == kMapEmpty) atomicAdd(nkeys, 1u);
        if (e == kMapEmpty || e == k) { atomicMin(&m.pos[h], pos); return; }
        h = (h + 1) & (kMapSlots - 1);
    }
    atomicAdd(nkeys, kMapSlots);                // cannot happen below the key limit; forces the sequential replay
}

// ---- one accepted message -> adsbcu_msg (what demodulate2400 hands to useModesMessage) ------------------------
__device__ __forceinline__ void write_msg(adsbcu_msg *out, const FullRec &r, uint64_t a, int score, uint32_t block, double level)
{
    union { adsbcu_msg m; uint64_t w[7]; } u;
    static_assert(sizeof(adsbcu_msg) == 56, "adsbcu_msg layout");
    // sampleTimestamp of the mag_buf (sdr_ifile.c:202) + 5*j + 768 + phase; with 12e6/2.4e6 == 5 this is 5*a + 768 + phase
    // up to the double rounding of the block's base, which is reproduced here
    const uint64_t bstart = (block & (block - 1u)) == 0u ? a & ~uint64_t(block - 1u) : a / block * block;
    // (below 2^53 / 12e6 samples the product is exact and so is the quotient 5 * bstart: no double arithmetic needed)
    const uint64_t ts_base = bstart < 750000000ull ? 5ull * bstart : uint64_t(__ddiv_rn(__dmul_rn(double(bstart), 12e6), 2400000.0));
    u.m.timestamp = ts_base + (a - bstart) * 5 + (8 + 56) * 12 + r.phase;
#pragma unroll
    for (int b = 0; b < 14; ++b) { u.m.msg[b] = r.msg[b]; u.m.verbatim[b] = r.verbatim[b]; }
    u.m.msgbits = (r.msgtype & 16) ? 112 : 56;
    u.m.correctedbits = uint8_t(r.corrections);
    u.m.msgtype = r.msgtype;
    u.m.score = uint8_t(score);
    u.m.crc = r.crc;
    u.m.addr = r.addr;
    u.m.signal_level = level;
    uint64_t *dst = reinterpret_cast<uint64_t *>(out);
#pragma unroll
    for (int k = 0; k < 7; ++k) dst[k] = u.w[k];
}

struct SkipState {
    uint32_t has_last;
    uint64_t last_a, last_lme;
};

// Compact view of a record for the ordered pass: x = feed position, y = key of icaoFilterTest, z = key of
// icaoFilterAdd (or 0xFFFFFFFF), w = kind | 112-bit flag << 8.  The resolve kernel adds 0x200 (the key is in the filter
// when the batch starts) and 0x400 (accepting the record adds an address the active table does not hold).
__device__ __forceinline__ uint4 view_of(const FullRec &r)
{
    return make_uint4(r.pos, r.test_key, r.add_key, uint32_t(r.kind) | ((r.verbatim[0] & 0x80u) ? 0x100u : 0u));
}

// demodulate2400's loop over the records [i0, i1) (ascending (pos, phase)): skip rule, best phase, accept.
//   known(key, pos)            -> filter state seen by the message at feed position pos
//   accept(index, score, a)    -> called for every accepted message, in order
template <class Known, class Accept>
__device__ __forceinline__ void resolve_range(const uint4 *view, uint32_t i0, uint32_t i1, uint64_t feed_base, uint32_t block,
                                              SkipState &ss, Known known, Accept accept)
{
    uint32_t i = i0;
    uint4 cur = i < i1 ? view[i] : make_uint4(0, 0, 0, 0);
    while (i < i1) {
        const uint32_t pos = cur.x;
        const uint64_t a = feed_base + pos;
        bool skip = false;
        if (ss.has_last) {
            // inside the mag_buf of the last accepted message scanning resumes 18 samples before its end
            // (j = last_message_end - 19, then ++j); a later mag_buf starts exactly at its end (demod_2400.c:410-443)
            const bool same_block = a / block == ss.last_a / block;
            skip = a < (same_block ? ss.last_lme - 18 : ss.last_lme);
        }
        int best = 0;                           // SR_NOT_SET
        uint32_t bi = i, last_key = kUnchecked, best_w = 0;
        bool last_known = false;
        do {                                    // phases ascending; strict > keeps the lowest phase on ties
            if (!skip) {
                if (cur.y != last_key) { last_key = cur.y; last_known = known(cur.y, pos); }
                const int sc = rank_of(cur.w & 0xFFu, last_known);
                if (sc > best) { best = sc; bi = i; best_w = cur.w; }
            }
            ++i;
            if (i < i1) cur = view[i];
        } while (i < i1 && cur.x == pos);
        if (skip || best < kAcceptThreshold) continue;
        accept(bi, best, a);
        ss.has_last = 1u;
        ss.last_a = a;
        ss.last_lme = a + uint64_t((((best_w & 0x100u) ? 112 : 56) + 8) * 12 / 5);
    }
}

struct SortSmem {
    uint32_t key[kSortCap];                    // (position << 3 | phase - 4) of record i
    uint32_t tkey[kSortCap];                   // its filter key (argument of icaoFilterTest)
    uint16_t by_rank[kSortCap];                // record index at sorted position r; 0xFFFF once the record is dropped
    uint8_t  kind[kSortCap];
    uint32_t warp_tot[kXThreads / 32];
    uint32_t nrec, offset, nkeep;
};

#ifdef ADSBCU_ORDER_TIMING            // debug: nanosecond marks of the order kernel in ChunkCounters::dbg (profiles/order_probe.py)
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define ORDER_MARK(k) do { if (threadIdx.x == 0) fa.c.ctr->dbg[k] = gtime(); } while (0)
#else
#define ORDER_MARK(k) do { } while (0)
#endif

__global__ void __launch_bounds__(kXThreads) order_kernel(FinishArgs fa)
{
    ADSBCU_DYN_SMEM(smem_raw);
    const ChunkArgs &a = fa.c;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t valid_long = a.tab->valid_long, valid_short = a.tab->valid_short, fixable_short = a.tab->fixable_short;
    const uint32_t *crcb = &a.tab->crc_byte[0][0];

    // ================= phase 1: this CTA's segment -- late records, order ========================================
    {
        SortSmem &ss = *reinterpret_cast<SortSmem *>(smem_raw);
        const uint32_t seg = blockIdx.x;
        if (seg == 0) ORDER_MARK(0);
        FullRec *rec = a.rec + size_t(seg) * a.rec_cap;
        const PhaseRec *pend = a.pend + size_t(seg) * a.pend_cap;
        const uint32_t n0 = min(a.rec_count[seg], a.rec_cap), npend = min(a.pend_count[seg], a.pend_cap);
        if (tid == 0) ss.nrec = n0;
        __syncthreads();
        for (uint32_t i = tid; i < npend; i += kXThreads) {
            const PhaseRec p = pend[i];
            const uint32_t key = p.info & 0x1FFFFFFu;
            if (!((a.bitmap[key >> 5] >> (key & 31)) & 1u)) {
                if (fa.deferred) {           // sharded captures: another shard may still announce the address
                    const uint32_t slot = atomicAdd(fa.n_deferred, 1u);
                    if (slot < fa.deferred_cap) fa.deferred[slot] = p; else a.ctr->overflow = 1u;
                }
                continue;
            }
            const uint32_t ph = 4u + ((p.info >> 25) & 7u), cpos = p.pos - a.pos_begin;
            const uint32_t t0 = 5u * (cpos + kFLead + 19u) + ph;
            const uint32_t c = t0 % 12u, n0b = t0 / 12u, f = n0b / 10u, r = n0b - 10u * f;
            const uint16_t *pf = a.planes + size_t(c) * a.plane_stride + f;
            const uint32_t w20 = (uint32_t(pf[0]) << 10) | pf[1];
            const uint32_t dfbit = 1u << ((w20 >> (15u - r)) & 31u);
            const bool is_long = (valid_long & dfbit) != 0;
            if (!is_long && !(valid_short & dfbit)) continue;             // cannot happen for a parked record
            uint32_t M[4], xs, xl;
            slice_planes(pf, r, is_long, crcb, fixable_short, M, xs, xl);
            const Frame fr = frame_from_words(M, is_long);
            const Classified cl = classify(fr, xs, xl, a.tab);
            const uint32_t slot = atomicAdd(&ss.nrec, 1u);
            if (slot < a.rec_cap) fill_rec(rec[slot], p.pos, ph, fr, cl); else a.ctr->overflow = 1u;
        }
        __syncthreads();
        if (seg == gridDim.x - 1) ORDER_MARK(1);
        const uint32_t n = min(ss.nrec, a.rec_cap);
        // rank sort by (position, phase): keys are unique
        if (n > uint32_t(kSortCap)) { if (tid == 0) a.ctr->overflow = 1u; }
        const uint32_t ns = min(n, uint32_t(kSortCap));
        for (uint32_t i = tid; i < ns; i += kXThreads) {
            ss.key[i] = ((rec[i].pos - a.pos_begin) << 3) | (rec[i].phase - 4u);
            ss.tkey[i] = rec[i].test_key;
            ss.kind[i] = rec[i].kind;
        }
        __syncthreads();
        for (uint32_t i = tid; i < ns; i += kXThreads) {
            const uint32_t me = ss.key[i];
            uint32_t rank = 0;
            for (uint32_t j = 0; j < ns; ++j) rank += ss.key[j] < me;
            ss.by_rank[rank] = uint16_t(i);
        }
        __syncthreads();
        // A record can never be the accepted phase of its position when another phase of the SAME position with the SAME filter
        // key beats it whether or not that key is known (higher score_rank, or equal rank and lower phase: demod_2400.c:336 keeps
        // the first best).  Most records are such losers -- a clean message decodes on three or four of the five phases -- and
        // dropping them here cuts the work of the one-CTA ordered pass about three times.  (Neighbours in sorted order.)
        bool drop[(kSortCap + kXThreads - 1) / kXThreads];
        uint32_t mykeep = 0;
        constexpr uint32_t kPer = kSortCap / kXThreads;           // a thread owns kPer consecutive sorted positions
        for (uint32_t q = 0; q < kPer; ++q) {
            const uint32_t r = tid * kPer + q;
            drop[q] = true;
            if (r >= ns) continue;
            const uint32_t i = ss.by_rank[r], me = ss.key[i], tk = ss.tkey[i], kd = ss.kind[i];
            const int m0 = rank_of(kd, false), m1 = rank_of(kd, true);
            bool dead = false;
            for (int dlt = -4; dlt <= 4 && !dead; ++dlt) {
                const int64_t r2 = int64_t(r) + dlt;
                if (dlt == 0 || r2 < 0 || r2 >= int64_t(ns)) continue;
                const uint32_t j = ss.by_rank[r2];
                if ((ss.key[j] >> 3) != (me >> 3) || ss.tkey[j] != tk) continue;
                const int o0 = rank_of(ss.kind[j], false), o1 = rank_of(ss.kind[j], true);
                const bool first = ss.key[j] < me;                    // lower phase
                dead = (o0 > m0 || (o0 == m0 && first)) && (o1 > m1 || (o1 == m1 && first));
            }
            drop[q] = dead;
            mykeep += !dead;
        }
        // exclusive scan of the per-thread keep counts, then the survivors move up in sorted order
        uint32_t incl = mykeep;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        if (lane == 31) ss.warp_tot[warp] = incl;
        __syncthreads();
        uint32_t base = incl - mykeep;
        for (int w2 = 0; w2 < warp; ++w2) base += ss.warp_tot[w2];
        uint32_t nkeep_all = 0;
        for (int w2 = 0; w2 < kXThreads / 32; ++w2) nkeep_all += ss.warp_tot[w2];
        uint16_t mine[kPer];
        for (uint32_t q = 0; q < kPer; ++q) mine[q] = (tid * kPer + q < ns) ? ss.by_rank[tid * kPer + q] : uint16_t(0xFFFF);
        __syncthreads();
        for (uint32_t q = 0, k = base; q < kPer; ++q) if (!drop[q]) ss.by_rank[k++] = mine[q];
        __syncthreads();
        const uint32_t nk = n > uint32_t(kSortCap) ? n : nkeep_all;      // (an oversized segment already raised the overflow flag)
        if (tid == 0) {
            a.rec_count[seg] = nk;
            *reinterpret_cast<volatile unsigned long long *>(&fa.seg_flag[seg]) = (unsigned long long)fa.epoch << 32 | nk;
            ss.offset = 0;
        }
        __syncthreads();
        uint32_t before = 0;
        for (uint32_t s2 = tid; s2 < seg; s2 += kXThreads) {
            unsigned long long v;
            do { v = *reinterpret_cast<volatile unsigned long long *>(&fa.seg_flag[s2]); } while (uint32_t(v >> 32) != fa.epoch);
            before += uint32_t(v);
        }
        before = __reduce_add_sync(0xFFFFFFFFu, before);
        if (lane == 0 && before) atomicAdd(&ss.offset, before);
        __syncthreads();
        if (seg == gridDim.x - 1) ORDER_MARK(2);
        const uint32_t offset = ss.offset;
        if (offset + nk > a.full_cap) { if (tid == 0) a.ctr->overflow = 1u; }
        else if (n <= uint32_t(kSortCap)) {
            for (uint32_t r = tid; r < nk; r += kXThreads) {
                const uint32_t i = ss.by_rank[r];
                const uint4 *src = reinterpret_cast<const uint4 *>(rec + i);
                uint4 *dst = reinterpret_cast<uint4 *>(a.ordered + offset + r);
                const uint4 q0 = src[0], q1 = src[1], q2 = src[2], q3 = src[3];
                dst[0] = q0; dst[1] = q1; dst[2] = q2; dst[3] = q3;
                fa.view[offset + r] = view_of(rec[i]);
            }
        }
    }
}

// ================= resolve_kernel: the ordered pass of demodulate2400 ===========================================
// One CTA of 1024 threads.  The chunk's records are taken in batches of up to kRBatch that end at run boundaries; a
// batch's 8-byte (position, flags) entries live in shared memory, so the per-run walks -- which are sequential by
// nature and would otherwise pay an uncoalesced global load per record -- run at shared-memory latency.  A batch is
// iterated to its fixed point (see the file header), its messages are written, its new addresses enter the filter
// table, then the next batch starts from that state: exactly the chunk algorithm at a finer grain.
constexpr int kRThreads = 1024;
constexpr uint32_t kRBatch = 15 * 1024;

struct ResolveSmem {
    uint32_t set[2 * kIcaoSlots];              // the contents of the reference's two filter tables as of the batch's start, as ONE
                                               // open-addressing set (entry = address | kInActive if the active table holds it): the
                                               // answers of icaoFilterTest without its probe chains, which are 4096 long once a
                                               // table is full (file mode never empties it)
    uint2 ent[kRBatch];                        // x: feed position, y: view flags (bits 0..15) | score of the accepted record << 16
    AddMap cur, next;
    uint16_t cand[kRBatch];                    // batch-relative indices of the records that can be accepted, ascending
    uint32_t warp_tot[kRThreads / 32 + 1];
    uint8_t scratch[kMapSlots];
    uint32_t cnt[kRBatch / kRThreads * 32];    // strided_compact: flags per (iteration, warp)
    uint32_t nkeys_cur, nkeys_next, changed, nruns, cut, last_acc, batch_msgs;
    uint8_t rank[24];                          // score_rank by (kind, known)
};

constexpr uint32_t kInActive = 1u << 30, kSetMask = 2u * kIcaoSlots - 1u;
__device__ __forceinline__ uint32_t set_hash(uint32_t v) { return (v * 2654435761u) >> 19 & kSetMask; }
__device__ __forceinline__ void set_insert(uint32_t *set, uint32_t v, uint32_t flag)
{
    uint32_t h = set_hash(v);
    for (uint32_t n = 0; n <= kSetMask; ++n) {
        const uint32_t e = atomicCAS(&set[h], kIcaoEmpty, v | flag);
        if (e == kIcaoEmpty) return;
        if ((e & ~kInActive) == v) { if (flag) atomicOr(&set[h], flag); return; }
        h = (h + 1) & kSetMask;
    }
}
__device__ __forceinline__ uint32_t set_find(const uint32_t *set, uint32_t v)     // the entry, or kIcaoEmpty
{
    uint32_t h = set_hash(v);
    for (uint32_t n = 0; n <= kSetMask; ++n) {
        const uint32_t e = set[h];
        if (e == kIcaoEmpty || (e & ~kInActive) == v) return e;
        h = (h + 1) & kSetMask;
    }
    return kIcaoEmpty;
}

// exclusive prefix of a per-thread flag over the CTA: ballot inside the warp, one shuffle scan of the 32 warp totals
__device__ __forceinline__ uint32_t block_flag_scan(bool f, uint32_t *warp_tot, uint32_t &total)
{
    static_assert(kRThreads == 1024, "one lane per warp total");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, f);
    if (lane == 0) warp_tot[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        const uint32_t c = warp_tot[lane];
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        warp_tot[lane] = incl - c;             // exclusive prefix of this warp's total
        if (lane == 31) warp_tot[32] = incl;
    }
    __syncthreads();
    const uint32_t before = warp_tot[warp];
    total = warp_tot[32];
    __syncthreads();
    return before + __popc(bal & ((1u << lane) - 1u));
}

// Compaction of up to 32 x 1024 flags held one bit per thread and iteration (bit `it` of `mask` belongs to element
// it * 1024 + tid): exclusive offsets in element order with three barriers in all, instead of one block scan per 1024
// elements.  emit(it, offset) is called for every set flag; returns the number of set flags.
template <class Emit>
__device__ __forceinline__ uint32_t strided_compact(uint32_t mask, uint32_t nit, uint32_t *cnt /* [32 * nit] */, uint32_t *warp_tot, Emit &&emit)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t it = 0; it < nit; ++it) {
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, (mask >> it) & 1u);
        if (lane == 0) cnt[it * 32u + warp] = __popc(b);
    }
    __syncthreads();
    // exclusive scan of the 32 * nit (<= 1024) counts, one per thread
    const uint32_t nval = 32u * nit;
    const uint32_t c = uint32_t(threadIdx.x) < nval ? cnt[threadIdx.x] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = warp_tot[lane];
        uint32_t wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, d); if (lane >= d) wi += t; }
        warp_tot[lane] = wi - w;
        if (lane == 31) warp_tot[32] = wi;
    }
    __syncthreads();
    if (uint32_t(threadIdx.x) < nval) cnt[threadIdx.x] = warp_tot[warp] + incl - c;
    const uint32_t total = warp_tot[32];
    __syncthreads();
    for (uint32_t it = 0; it < nit; ++it) {
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, (mask >> it) & 1u);
        if ((mask >> it) & 1u) emit(it, cnt[it * 32u + warp] + __popc(b & ((1u << lane) - 1u)));
    }
    __syncthreads();
    return total;
}

static_assert(sizeof(ResolveSmem) <= 227 * 1024, "resolve kernel shared memory exceeds the sm_100 per-CTA limit");

__global__ void __launch_bounds__(kRThreads, 1) resolve_kernel(FinishArgs fa)
{
    ADSBCU_DYN_SMEM(smem_raw);
    ResolveSmem &sm = *reinterpret_cast<ResolveSmem *>(smem_raw);
    const ChunkArgs &a = fa.c;
    StreamState *st = fa.st;
    const int tid = threadIdx.x;

    ORDER_MARK(3);
    // records of the chunk; a failed chunk (or a failed earlier chunk of the feed) leaves the stream state alone
    uint32_t n = 0;
    {
        uint32_t mine = 0;
        for (uint32_t s2 = tid; s2 < a.nseg; s2 += kRThreads) mine += a.rec_count[s2];
        mine = __reduce_add_sync(0xFFFFFFFFu, mine);
        if (tid == 0) sm.cut = 0;
        __syncthreads();
        if ((tid & 31) == 0 && mine) atomicAdd(&sm.cut, mine);
        __syncthreads();
        n = sm.cut;
        __syncthreads();
    }
    const bool failed = a.ctr->overflow != 0u || st->poison != 0u || n > a.full_cap;
    if (tid == 0) {
        a.ctr->n_full = failed ? 0u : n;
        if (failed) { st->poison = 1u; a.ctr->overflow = 1u; }                 // later chunks in flight must not touch the state
    }
    if (failed || fa.candidates_only) return;

    const FullRec *recs = a.ordered;
    const uint4 *view = fa.view;                  // 16-byte views, filter state as of the chunk's start folded in by the order kernel
    const uint32_t block = fa.block_samples;
    const int bshift = (block & (block - 1u)) == 0u ? 31 - __clz(int(block)) : -1;
    auto blk_of = [&](uint64_t x) { return bshift >= 0 ? x >> bshift : x / block; };
    const uint64_t feed_base = fa.feed_base;
    uint2 *out = fa.accepted;                     // (record index, score) of every accepted message: n entries at most
    const uint64_t feed_count = st->n_out;
    uint32_t *run_start = fa.run_start;
    const SkipState carried{st->has_last, st->last_a, st->last_lme};
    if (tid < 22) sm.rank[tid] = uint8_t(rank_of(uint32_t(tid) >> 1, (tid & 1) != 0));
    if (tid == 0) sm.last_acc = 0;
    const uint32_t active = st->active;
    uint32_t total = 0;                           // messages written so far (uniform)
    uint32_t ia = 0;
    bool sequential = fa.force_sequential != 0;
    SkipState entry = carried;                    // skip state in front of the batch: the last accepted message so far
    const uint32_t batch_max = min(kRBatch, max(fa.batch, 2u));
    uint32_t batch_cap = batch_max;               // halved when a batch announces more new addresses than the map holds

    while (ia < n && !sequential) {
        // ---- the batch's entries into shared memory (coalesced); later batches see the table the earlier ones left ----
        const uint32_t ib_t = min(n, ia + batch_cap);
        for (uint32_t i = tid; i < 2u * kIcaoSlots; i += kRThreads) sm.set[i] = kIcaoEmpty;
        __syncthreads();
        for (uint32_t i = tid; i < 2u * kIcaoSlots; i += kRThreads) {
            const uint32_t v = __ldcg(&st->table[0][i]);                 // (past L1: the batch before wrote them); both tables are contiguous
            if (v != kIcaoEmpty) set_insert(sm.set, v, (i >> 12) == active ? kInActive : 0u);
        }
        const bool table_full = __ldcg(&st->active_count) >= kIcaoSlots;  // every insertion would be dropped (icao_filter.c:71-84)
        __syncthreads();
#pragma unroll 4
        for (uint32_t i = ia + tid; i < ib_t; i += kRThreads) {
            const uint4 v = view[i];
            uint32_t w = v.w & 0x1FFu;
            if (set_find(sm.set, v.y) != kIcaoEmpty) w |= 0x200u;                                   // known when the batch starts
            if (v.z != kUnchecked && !table_full) {
                const uint32_t e = set_find(sm.set, v.z);
                if (e == kIcaoEmpty || !(e & kInActive)) w |= 0x400u;                               // accepting it adds a new address
            }
            sm.ent[i - ia] = make_uint2(v.x, w);
        }
        for (uint32_t i = tid; i < kMapSlots; i += kRThreads) { sm.cur.key[i] = kMapEmpty; sm.cur.pos[i] = kMapEmpty; sm.next.key[i] = kMapEmpty; sm.next.pos[i] = kMapEmpty; }
        if (tid == 0) { sm.nkeys_cur = 0; sm.nkeys_next = 0; sm.changed = 0; sm.nruns = 0; sm.cut = 0; sm.batch_msgs = 0; }
        __syncthreads();
        // ---- the batch ends between two start positions (records of one position are never split) ---------------------
        uint32_t ib = ib_t;
        if (ib_t < n) {
            uint32_t best = 0;
            for (uint32_t i = max(ia + 1, ib_t - min(ib_t - ia, 8u)) + tid; i < ib_t; i += kRThreads)
                if (sm.ent[i - ia].x != sm.ent[i - 1 - ia].x) best = i;
            best = __reduce_max_sync(0xFFFFFFFFu, best);
            if ((tid & 31) == 0 && best) atomicMax(&sm.cut, best);
            __syncthreads();
            ib = sm.cut;
            if (ib == 0) { sequential = true; break; }                   // (a batch of a handful of records at one position)
        }
        if (ia == 0) ORDER_MARK(4);

        // ---- rounds: scores and best-of-phases in parallel over the records, then the skip rule over the few records
        // that can be accepted at all ("candidates": the winner of a position whose score reaches the threshold).  Only an
        // accepted message changes the skip state, and it reaches at most 288 samples ahead, so the candidate list falls
        // apart into runs wherever two neighbours are 288 or more samples apart; every thread walks its runs (r = tid,
        // tid + 1024, ...) on its own against the batch's first_add map.  The gaps are between CANDIDATES, not between
        // records: dense noise records (--fix-2bit) that can never win do not glue runs together.
        bool retry = false;
        for (int round = 0;; ++round) {
            const bool use_map = sm.nkeys_cur != 0u;
            // scores (provisional, in the high half of the entry)
            for (uint32_t i = ia + tid; i < ib; i += kRThreads) {
                const uint2 e = sm.ent[i - ia];
                const uint32_t w = e.y & 0xFFFFu;
                const bool known = (w & 0x200u) || (use_map && map_lookup(sm.cur, view[i].y) < e.x);
                const uint32_t sc = sm.rank[2u * (w & 0xFFu) + (known ? 1u : 0u)];
                sm.ent[i - ia].y = w | (sc << 16);
            }
            __syncthreads();
            // winners: the first record of its position with the highest score (phases ascending; strict > keeps the lowest
            // phase on ties), if that score reaches the threshold.  Compacted in index order into sm.cand.
            const uint32_t nit = (ib - ia + kRThreads - 1u) / kRThreads;
            uint32_t wmask = 0;
            for (uint32_t it = 0; it < nit; ++it) {
                const uint32_t i = ia + it * kRThreads + tid;
                if (i >= ib) break;
                const uint2 e = sm.ent[i - ia];
                const uint32_t sc = e.y >> 16;
                bool win = sc >= uint32_t(kAcceptThreshold);
                for (uint32_t j = i; win && j > ia && sm.ent[j - 1 - ia].x == e.x; --j) win = (sm.ent[j - 1 - ia].y >> 16) < sc;
                for (uint32_t j = i + 1; win && j < ib && sm.ent[j - ia].x == e.x; ++j) win = (sm.ent[j - ia].y >> 16) <= sc;
                if (win) wmask |= 1u << it;
            }
            const uint32_t ncand = strided_compact(wmask, nit, sm.cnt, sm.warp_tot,
                                                   [&](uint32_t it, uint32_t off) { sm.cand[off] = uint16_t(it * kRThreads + tid); });
            __syncthreads();
            // every record that is not a winner loses its provisional score; runs of the candidate list
            for (uint32_t i = ia + tid; i < ib; i += kRThreads) sm.ent[i - ia].y &= 0xFFFFu;
            __syncthreads();
            const uint32_t nitc = (ncand + kRThreads - 1u) / kRThreads;
            uint32_t rmask = 0;
            for (uint32_t it = 0; it < nitc; ++it) {
                const uint32_t k = it * kRThreads + tid;
                if (k < ncand && (k == 0 || sm.ent[sm.cand[k]].x - sm.ent[sm.cand[k - 1]].x >= 288u)) rmask |= 1u << it;
            }
            const uint32_t R = strided_compact(rmask, nitc, sm.cnt, sm.warp_tot,
                                               [&](uint32_t it, uint32_t off) { run_start[off] = it * kRThreads + tid; });
            if (tid == 0) run_start[R] = ncand;
            __syncthreads();
            for (uint32_t r = tid; r < R; r += kRThreads) {
                SkipState ss = r == 0 ? entry : SkipState{0u, 0, 0};
                const uint32_t k1 = run_start[r + 1];
                for (uint32_t k = run_start[r]; k < k1; ++k) {
                    const uint32_t j = sm.cand[k];
                    const uint2 e = sm.ent[j];
                    const uint32_t pos = e.x, w = e.y & 0xFFFFu;
                    const uint64_t pa = feed_base + pos;
                    // inside the mag_buf of the last accepted message scanning resumes 18 samples before its end
                    // (j = last_message_end - 19, then ++j); a later mag_buf starts exactly at its end (demod_2400.c:410-443)
                    const bool skip = ss.has_last && pa < (blk_of(pa) == blk_of(ss.last_a) ? ss.last_lme - 18 : ss.last_lme);
                    if (skip) continue;
                    const bool known = (w & 0x200u) || (use_map && map_lookup(sm.cur, view[ia + j].y) < pos);
                    sm.ent[j].y = w | (uint32_t(sm.rank[2u * (w & 0xFFu) + (known ? 1u : 0u)]) << 16);
                    if (w & 0x400u) map_insert_min(sm.next, view[ia + j].z, pos, &sm.nkeys_next);
                    ss.has_last = 1u;
                    ss.last_a = pa;
                    ss.last_lme = pa + uint64_t((((w & 0x100u) ? 112 : 56) + 8) * 12 / 5);
                }
            }
            __syncthreads();
            // The table takes insertions until its last slot is used and drops them from then on (icao_filter.c:71-84): if this
            // batch's new addresses do not all fit, the ones announced first get in and the others count as never added.
            if (const uint32_t room = kIcaoSlots - min(kIcaoSlots, st->active_count); sm.nkeys_next > room) {
                for (uint32_t i = tid; i < kMapSlots; i += kRThreads) {
                    const uint32_t mine = sm.next.pos[i];
                    if (sm.next.key[i] == kMapEmpty) continue;
                    uint32_t rank = 0;
                    for (uint32_t j = 0; j < kMapSlots; ++j) rank += sm.next.key[j] != kMapEmpty && sm.next.pos[j] < mine;
                    if (rank >= room) sm.scratch[i] = 1u; else sm.scratch[i] = 0u;
                }
                __syncthreads();
                for (uint32_t i = tid; i < kMapSlots; i += kRThreads)
                    if (sm.next.key[i] != kMapEmpty && sm.scratch[i]) sm.next.pos[i] = kMapDropped;
                __syncthreads();
            }
            // did this round reproduce the map it was computed with?
            if (sm.nkeys_next != sm.nkeys_cur && tid == 0) sm.changed = 1u;
            for (uint32_t i = tid; i < kMapSlots; i += kRThreads)
                if (sm.next.key[i] != kMapEmpty && map_lookup(sm.cur, sm.next.key[i]) != sm.next.pos[i]) sm.changed = 1u;
            __syncthreads();
            const bool changed = sm.changed != 0u;
            const uint32_t nk = sm.nkeys_next;
            __syncthreads();
            if (nk > kMapMaxKeys) {                   // more new addresses than the map is sized for: the same records in smaller batches
                if (ib - ia > 64u) { batch_cap = (ib - ia) / 2u; retry = true; } else sequential = true;
                break;
            }
            if (round >= 12) { sequential = true; break; }
            if (!changed) break;
            for (uint32_t i = tid; i < kMapSlots; i += kRThreads) {
                sm.cur.key[i] = sm.next.key[i]; sm.cur.pos[i] = sm.next.pos[i];
                sm.next.key[i] = kMapEmpty; sm.next.pos[i] = kMapEmpty;
            }
            if (tid == 0) { sm.nkeys_cur = nk; sm.nkeys_next = 0; sm.changed = 0; }
            __syncthreads();
        }
        if (retry) { __syncthreads(); continue; }
        batch_cap = batch_max;
        if (sequential) break;
        if (ia == 0) ORDER_MARK(5);

        // ---- emission, parallel over the batch's records: every accepted record becomes one message, in stream order.
        // First the offsets (a block scan per 1024 records; the offset replaces the position in the entry), then all
        // the messages without a barrier in between, so that the record reads of the whole batch are in flight together.
        const uint32_t nit_e = (ib - ia + kRThreads - 1u) / kRThreads;
        uint32_t emask = 0;
        for (uint32_t it = 0; it < nit_e; ++it) {
            const uint32_t i = ia + it * kRThreads + tid;
            if (i < ib && (sm.ent[i - ia].y >> 16)) emask |= 1u << it;
        }
        const uint32_t batch_total = strided_compact(emask, nit_e, sm.cnt, sm.warp_tot,
                                                     [&](uint32_t it, uint32_t off) { sm.ent[it * kRThreads + tid].x = off; });
        uint32_t my_last = 0;
#pragma unroll 4
        for (uint32_t i = ia + tid; i < ib; i += kRThreads) {
            const uint2 e = sm.ent[i - ia];
            if (e.y >> 16) {
                out[total + e.x] = make_uint2(i, e.y >> 16);
                my_last = i + 1u;
            }
        }
        my_last = __reduce_max_sync(0xFFFFFFFFu, my_last);
        if ((tid & 31) == 0 && my_last) atomicMax(&sm.last_acc, my_last);
        total += batch_total;
        if (ia == 0) ORDER_MARK(6);
        // ---- new addresses enter the active table (icao_filter.c:71-84).  The reference inserts them in stream order;
        // which slot of a probe chain an address lands in changes no answer of icaoFilterTest while the table has room
        // (and a batch that could fill it takes the sequential replay), so the insertions run in parallel.
        {
            uint32_t *t = st->table[st->active];
            uint32_t added = 0;
            for (uint32_t i = tid; i < kMapSlots; i += kRThreads) {
                const uint32_t v = sm.cur.key[i];
                if (v == kMapEmpty || sm.cur.pos[i] == kMapDropped) continue;
                uint32_t h = icao_hash(v);
                for (uint32_t probes = 0; probes < kIcaoSlots; ++probes) {
                    const uint32_t e = atomicCAS(&t[h], kIcaoEmpty, v);
                    if (e == kIcaoEmpty) { ++added; break; }
                    if (e == v) break;
                    h = (h + 1) & (kIcaoSlots - 1);
                }
            }
            if (added) atomicAdd(&st->active_count, added);
        }
        __threadfence();
        __syncthreads();
        if (const uint32_t li = sm.last_acc; li > ia) { // (uniform) the next batch starts behind this batch's last accepted message
            const uint4 v = view[li - 1];
            const uint64_t pa = feed_base + v.x;
            entry = SkipState{1u, pa, pa + uint64_t((((v.w & 0x100u) ? 112 : 56) + 8) * 12 / 5)};
        }
        __syncthreads();
        ia = ib;
    }

    if (tid == 0) a.ctr->n_sequential = sequential && ia < n ? n - ia : 0u;
    if (sequential && ia < n) {
        // the reference's own loop from record `ia` on, one thread, on the tables in global memory: exact whatever they hold
        __syncthreads();
        if (tid == 0) {
            SkipState ss = entry;
            uint64_t k = total;
            uint32_t last = sm.last_acc;
            resolve_range(view, ia, n, feed_base, block, ss,
                          [&](uint32_t key, uint32_t) { return icao_test(st, key); },
                          [&](uint32_t idx, int score, uint64_t) {
                              out[k] = make_uint2(idx, uint32_t(score));
                              ++k;
                              last = idx + 1u;
                              if (recs[idx].add_key != kUnchecked) icao_add(st, recs[idx].add_key);
                          });
            sm.last_acc = last;
            sm.batch_msgs = uint32_t(k);
        }
        __syncthreads();
        total = sm.batch_msgs;
    }
    __syncthreads();
    if (tid == 0) {
        // the stream's skip state: the last accepted message of the chunk
        if (const uint32_t li = sm.last_acc) {
            const uint4 v = view[li - 1];
            const uint64_t pa = feed_base + v.x;
            st->has_last = 1u; st->last_a = pa;
            st->last_lme = pa + uint64_t((((v.w & 0x100u) ? 112 : 56) + 8) * 12 / 5);
        }
        st->n_out = feed_count + total;
        a.ctr->n_msgs = total;
        a.ctr->msg_base = feed_count;
        __threadfence_system();
    }
    ORDER_MARK(7);
}


// The accepted messages of one chunk, one warp per message: mm.signalLevel (demod_2400.c:382-403: mean squared
// magnitude over the signal window, IEEE double divisions as on the host), then the 56-byte adsbcu_msg goes to the
// feed's output array (mapped host memory or the caller's pinned buffer), where the host never touches it again.
__global__ void __launch_bounds__(256) level_kernel(FinishArgs fa)
{
    const ChunkArgs &a = fa.c;
    const int lane = threadIdx.x & 31;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5), gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint32_t n = a.ctr->n_msgs;
    const uint64_t base = a.ctr->msg_base;
    for (uint32_t k = gwarp; k < n; k += nwarps) {
        const uint2 acc = fa.accepted[k];
        const FullRec &r = a.ordered[acc.x];
        const double lvl = signal_level_warp(a, r.pos, uint32_t(r.verbatim[0]) >> 3, lane);
        if (lane == 0 && base + k < fa.out_cap) write_msg(fa.out + base + k, r, fa.feed_base + r.pos, int(acc.y), fa.block_samples, lvl);
    }
}

// Candidate scans (sharded captures) hand records, not messages, to the host: the same signal level, per record.
__global__ void __launch_bounds__(256) level_records_kernel(FinishArgs fa)
{
    const ChunkArgs &a = fa.c;
    const int lane = threadIdx.x & 31;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5), gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint32_t n = min(a.ctr->n_full, a.full_cap);
    for (uint32_t k = gwarp; k < n; k += nwarps) {
        const double lvl = signal_level_warp(a, a.ordered[k].pos, uint32_t(a.ordered[k].verbatim[0]) >> 3, lane);
        if (lane == 0) a.ordered[k].signal_level = lvl;
    }
}

// ---- small state kernels ------------------------------------------------------------------------------------------
__global__ void state_reset_kernel(StreamState *st, int clear_filter, int clear_stream)
{
    const int tid = threadIdx.x + blockIdx.x * blockDim.x;
    if (clear_filter)
        for (uint32_t i = tid; i < 2u * kIcaoSlots; i += gridDim.x * blockDim.x) st->table[0][i] = kIcaoEmpty;   // both tables are contiguous
    if (tid == 0) {
        if (clear_filter) { st->active = 0; st->active_count = 0; }
        if (clear_stream) { st->has_last = 0; st->last_a = 0; st->last_lme = 0; }
        st->poison = 0; st->n_out = 0;
    }
}
__global__ void state_begin_feed_kernel(StreamState *st, int reset_count) { if (reset_count) st->n_out = 0; st->poison = 0; }
__global__ void filter_add_kernel(StreamState *st, uint32_t *bitmap, uint32_t addr)
{
    icao_add(st, addr);
    if (!(addr & kNonIcao)) {
        const uint32_t bi = bitmap_index(addr);
        atomicOr(&bitmap[bi >> 5], 1u << (bi & 31));
    }
}
// adsbcu_1090_state_import: the start state of a sample range.  Only membership matters (file mode never removes an
// entry), so the addresses go in concurrently, each claiming its slot with a CAS as the resolve kernel's insertion does.
__global__ void __launch_bounds__(256) state_import_kernel(StreamState *st, uint32_t *bitmap, uint32_t has_last, unsigned long long last_a,
                                                            unsigned long long last_lme, const uint32_t *addrs, uint32_t n)
{
    if (threadIdx.x == 0) { st->has_last = has_last; st->last_a = last_a; st->last_lme = last_lme; }
    uint32_t *t = st->table[st->active];
    for (uint32_t k = threadIdx.x; k < n; k += blockDim.x) {
        const uint32_t v = addrs[k];
        uint32_t h = icao_hash(v);
        for (uint32_t probe = 0; probe < kIcaoSlots; ++probe) {
            const uint32_t e = atomicCAS(&t[h], kIcaoEmpty, v);
            if (e == kIcaoEmpty) { atomicAdd(&st->active_count, 1u); break; }
            if (e == v) break;
            h = (h + 1) & (kIcaoSlots - 1);
        }
        if (!(v & kNonIcao)) {
            const uint32_t bi = bitmap_index(v);
            atomicOr(&bitmap[bi >> 5], 1u << (bi & 31));
        }
    }
}
__global__ void filter_test_kernel(const StreamState *st, uint32_t addr, uint32_t *result) { *result = icao_test(st, addr) ? 1u : 0u; }
// icaoFilterExpire (icao_filter.c:112-127): the other table is cleared and becomes the active one; the scan's
// "may ever be in the filter" set is rebuilt from the table that survives
__global__ void filter_expire_kernel(StreamState *st, uint32_t *bitmap, uint32_t nwords)
{
    __shared__ uint32_t s_keep;
    const uint32_t keep = st->active;           // the table that stays readable
    if (threadIdx.x == 0) s_keep = keep;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nwords; i += blockDim.x) bitmap[i] = 0u;
    for (uint32_t i = threadIdx.x; i < kIcaoSlots; i += blockDim.x) st->table[s_keep ^ 1u][i] = kIcaoEmpty;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kIcaoSlots; i += blockDim.x) {
        const uint32_t v = st->table[s_keep][i];
        if (v != kIcaoEmpty && !(v & kNonIcao)) { const uint32_t bi = bitmap_index(v); atomicOr(&bitmap[bi >> 5], 1u << (bi & 31)); }
    }
    __syncthreads();
    if (threadIdx.x == 0) { st->active = s_keep ^ 1u; st->active_count = 0; }
}

}  // namespace

cudaError_t init_finish()
{
    return cudaFuncSetAttribute(resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(ResolveSmem)));
}

// order_kernel: one CTA per segment; resolve_kernel: one CTA; level kernels: `level_grid` CTAs of 8 warps.  The three
// only depend on each other in this order (and resolve on the resolve of the chunk before): the host gives each its own stream.
cudaError_t launch_order(const FinishArgs &fa, cudaStream_t s)
{
    if (fa.c.pos_end <= fa.c.pos_begin || !fa.c.nseg) return cudaSuccess;
    ADSBCU_LAUNCH(order_kernel, fa.c.nseg, kXThreads, sizeof(SortSmem), s, fa);
    return cudaGetLastError();
}
cudaError_t launch_resolve(const FinishArgs &fa, cudaStream_t s)
{
    if (fa.c.pos_end <= fa.c.pos_begin || !fa.c.nseg) return cudaSuccess;
    ADSBCU_LAUNCH(resolve_kernel, 1, kRThreads, sizeof(ResolveSmem), s, fa);
    return cudaGetLastError();
}
cudaError_t launch_level(const FinishArgs &fa, uint32_t level_grid, cudaStream_t s)
{
    if (fa.c.pos_end <= fa.c.pos_begin || !fa.c.nseg) return cudaSuccess;
    if (fa.candidates_only) ADSBCU_LAUNCH(level_records_kernel, level_grid, 256, 0, s, fa);
    else ADSBCU_LAUNCH(level_kernel, level_grid, 256, 0, s, fa);
    return cudaGetLastError();
}
cudaError_t launch_state_reset(StreamState *st, int clear_filter, int clear_stream, cudaStream_t s)
{
    ADSBCU_LAUNCH(state_reset_kernel, 8, 256, 0, s, st, clear_filter, clear_stream);
    return cudaGetLastError();
}
cudaError_t launch_state_begin_feed(StreamState *st, int reset_count, cudaStream_t s)
{
    ADSBCU_LAUNCH(state_begin_feed_kernel, 1, 1, 0, s, st, reset_count);
    return cudaGetLastError();
}
cudaError_t launch_filter_add(StreamState *st, uint32_t *bitmap, uint32_t addr, cudaStream_t s)
{
    ADSBCU_LAUNCH(filter_add_kernel, 1, 1, 0, s, st, bitmap, addr);
    return cudaGetLastError();
}
cudaError_t launch_state_import(StreamState *st, uint32_t *bitmap, uint32_t has_last, uint64_t last_a, uint64_t last_lme,
                                const uint32_t *addrs_dev, uint32_t n, cudaStream_t s)
{
    ADSBCU_LAUNCH(state_import_kernel, 1, 256, 0, s, st, bitmap, has_last, (unsigned long long)last_a, (unsigned long long)last_lme, addrs_dev, n);
    return cudaGetLastError();
}
cudaError_t launch_filter_test(const StreamState *st, uint32_t addr, uint32_t *result_dev, cudaStream_t s)
{
    ADSBCU_LAUNCH(filter_test_kernel, 1, 1, 0, s, st, addr, result_dev);
    return cudaGetLastError();
}
cudaError_t launch_filter_expire(StreamState *st, uint32_t *bitmap, uint32_t nwords, cudaStream_t s)
{
    ADSBCU_LAUNCH(filter_expire_kernel, 1, 1024, 0, s, st, bitmap, nwords);
    return cudaGetLastError();
}

}  // namespace adsbcu
