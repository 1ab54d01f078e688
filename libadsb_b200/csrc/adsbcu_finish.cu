// order_kernel + level_kernel -- the ordered tail of demodulate2400() on the device (sm_100a).
//
// What is sequential in the reference (demod_2400.c:222,336,349-355,410-419,432-443 and mode_s.c:781-788): pick the
// best-scoring phase per start position given the CURRENT ICAO filter, accept, skip over the accepted message, add the
// address to the filter, carry last_message_end across mag_buf boundaries.  Round 1 replayed this on one host thread
// (0.7 ms per 60 s capture, the slowest stage of the resident step).  Here it runs on the GPU, exactly:
//
//  phase 1 (one CTA per front-kernel segment)
//    * parked Address/Parity-style records whose address has been announced by now are evaluated from the symbol
//      planes like the message kernel does;
//    * the segment's records are rank-sorted by (position, phase);
//  phase 2 (the CTA that finishes phase 1 last)
//    * segments in index order are the chunk in stream order: prefix sum, compaction into one array;
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
__device__ __forceinline__ void write_msg(adsbcu_msg *out, const FullRec &r, uint64_t a, int score, uint32_t block)
{
    union { adsbcu_msg m; uint64_t w[7]; } u;
    static_assert(sizeof(adsbcu_msg) == 56, "adsbcu_msg layout");
    // sampleTimestamp of the mag_buf (sdr_ifile.c:202) + 5*j + 768 + phase; with 12e6/2.4e6 == 5 this is 5*a + 768 + phase
    // up to the double rounding of the block's base, which is reproduced here
    const uint64_t blk = a / block, bstart = blk * block;
    const uint64_t ts_base = uint64_t(__ddiv_rn(__dmul_rn(double(bstart), 12e6), 2400000.0));
    u.m.timestamp = ts_base + (a - bstart) * 5 + (8 + 56) * 12 + r.phase;
#pragma unroll
    for (int b = 0; b < 14; ++b) { u.m.msg[b] = r.msg[b]; u.m.verbatim[b] = r.verbatim[b]; }
    u.m.msgbits = (r.msgtype & 16) ? 112 : 56;
    u.m.correctedbits = uint8_t(r.corrections);
    u.m.msgtype = r.msgtype;
    u.m.score = uint8_t(score);
    u.m.crc = r.crc;
    u.m.addr = r.addr;
    u.w[6] = uint64_t(r.pos);                  // the level kernel turns the feed position into mm.signalLevel
    uint64_t *dst = reinterpret_cast<uint64_t *>(out);
#pragma unroll
    for (int k = 0; k < 7; ++k) dst[k] = u.w[k];
}

struct SkipState {
    uint32_t has_last;
    uint64_t last_a, last_lme;
};

// Compact view of a record for the ordered pass: x = feed position, y = key of icaoFilterTest, z = key of
// icaoFilterAdd (or 0xFFFFFFFF), w = kind | 112-bit flag << 8 (| 0x200: the key is in the filter when the chunk starts,
// | 0x400: accepting the record adds an address the active table does not hold; both set by phase 1).  16 bytes, so a
// thread walking its runs stays in L1.
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

struct FinishSmem {
    AddMap cur, next;
    uint32_t scan[kXThreads];
    uint32_t nkeys_cur, nkeys_next, changed;
    uint32_t total;
    uint8_t rank[24];                          // score_rank by (kind, known)
};
struct SortSmem {
    uint32_t key[kSortCap];
    uint32_t nrec, is_last, offset;
};

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *scratch, uint32_t &total)
{
    const int tid = threadIdx.x;
    scratch[tid] = v;
    __syncthreads();
    for (int d = 1; d < kXThreads; d <<= 1) {
        const uint32_t t = tid >= d ? scratch[tid - d] : 0u;
        __syncthreads();
        scratch[tid] += t;
        __syncthreads();
    }
    total = scratch[kXThreads - 1];
    const uint32_t incl = scratch[tid];
    __syncthreads();
    return incl - v;
}

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
        const uint32_t n = min(ss.nrec, a.rec_cap);
        // rank sort by (position, phase): keys are unique
        if (n > uint32_t(kSortCap)) { if (tid == 0) a.ctr->overflow = 1u; }
        for (uint32_t i = tid; i < n && i < uint32_t(kSortCap); i += kXThreads) ss.key[i] = ((rec[i].pos - a.pos_begin) << 3) | (rec[i].phase - 4u);
        // Segments in index order are the chunk in stream order: publish this segment's count, sum the counts of the
        // segments before it (CTAs start in index order, so the ones waited for are running or done), and write the
        // sorted records straight to their place in the chunk's ordered array.
        if (tid == 0) {
            a.rec_count[seg] = n;
            *reinterpret_cast<volatile unsigned long long *>(&fa.seg_flag[seg]) = (unsigned long long)fa.epoch << 32 | n;
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
        const uint32_t offset = ss.offset;
        if (offset + n > a.full_cap) { if (tid == 0) a.ctr->overflow = 1u; }
        else if (n <= uint32_t(kSortCap)) {
            for (uint32_t i = tid; i < n; i += kXThreads) {
                const uint32_t me = ss.key[i];
                uint32_t rank = 0;
                for (uint32_t j = 0; j < n; ++j) rank += ss.key[j] < me;
                const uint4 *src = reinterpret_cast<const uint4 *>(rec + i);
                uint4 *dst = reinterpret_cast<uint4 *>(a.ordered + offset + rank);
                const uint4 q0 = src[0], q1 = src[1], q2 = src[2], q3 = src[3];
                dst[0] = q0; dst[1] = q1; dst[2] = q2; dst[3] = q3;
                uint4 v = view_of(rec[i]);
                if (icao_test(fa.st, v.y)) v.w |= 0x200u;                                              // known when the chunk starts
                if (v.z != kUnchecked && !icao_probe(fa.st->table[fa.st->active], v.z)) v.w |= 0x400u;    // accepting it adds a new address
                fa.view[offset + rank] = v;
            }
        }
        __syncthreads();
        __threadfence();
        if (tid == 0) ss.is_last = atomicAdd(fa.ticket, 1u) == gridDim.x - 1;
        __syncthreads();
        if (!ss.is_last) return;
        __threadfence();
    }

    // ================= phase 2: the last CTA -- the ordered pass ===============================================
    FinishSmem &sm = *reinterpret_cast<FinishSmem *>(smem_raw);
    StreamState *st = fa.st;
    __syncthreads();
    if (tid == 0) *fa.ticket = 0;                                             // ready for the next chunk
    uint32_t n_full = 0;
    {
        uint32_t mine = 0;
        for (uint32_t s2 = tid; s2 < a.nseg; s2 += kXThreads) mine += a.rec_count[s2];
        block_exclusive_scan(mine, sm.scan, n_full);
        if (n_full > a.full_cap) n_full = 0;                                   // (the overflow flag is set)
    }
    const bool failed = a.ctr->overflow != 0u || st->poison != 0u;
    if (tid == 0) {
        a.ctr->n_full = n_full;
        if (failed) st->poison = 1u;                                           // later chunks in flight must not touch the state
    }
    if (failed || fa.candidates_only) {
        __syncthreads();
        if (tid == 0) { __threadfence_system(); }
        return;
    }

    const FullRec *recs = a.ordered;
    const uint32_t n = n_full;
    const uint4 *view = fa.view;                  // 16-byte views, filter state as of the chunk's start folded in by phase 1
    uint8_t *dec = fa.dec;                        // score of every accepted record, 0 for the others
    const uint32_t block = fa.block_samples;
    const int bshift = (block & (block - 1u)) == 0u ? 31 - __clz(int(block)) : -1;
    auto blk_of = [&](uint64_t x) { return bshift >= 0 ? x >> bshift : x / block; };
    const uint64_t feed_base = fa.feed_base;
    adsbcu_msg *out = fa.chunk_msgs;              // n_full entries at most: every accepted message has its own record
    const uint64_t feed_count = st->n_out;
    uint32_t *run_start = fa.run_start;
    if (tid < 22) sm.rank[tid] = uint8_t(rank_of(uint32_t(tid) >> 1, (tid & 1) != 0));
    for (uint32_t i = tid; i < kMapSlots; i += kXThreads) { sm.cur.key[i] = kMapEmpty; sm.cur.pos[i] = kMapEmpty; sm.next.key[i] = kMapEmpty; sm.next.pos[i] = kMapEmpty; }
    if (tid == 0) { sm.nkeys_cur = 0; sm.nkeys_next = 0; sm.changed = 0; sm.total = 0; }

    // runs: a record starts a run when it is 288 or more samples after its predecessor
    uint32_t R = 0;
    {
        const uint32_t per = (n + kXThreads - 1) / kXThreads;
        const uint32_t i0 = min(n, tid * per), i1 = min(n, i0 + per);
        uint32_t mine = 0;
        for (uint32_t i = i0; i < i1; ++i) mine += (i == 0 || view[i].x - view[i - 1].x >= 288u);
        uint32_t off = block_exclusive_scan(mine, sm.scan, R);
        for (uint32_t i = i0; i < i1; ++i)
            if (i == 0 || view[i].x - view[i - 1].x >= 288u) run_start[off++] = i;
        if (tid == 0) run_start[R] = n;
        __syncthreads();
    }
    const uint32_t rper = (R + kXThreads - 1) / kXThreads;
    const uint32_t r0 = min(R, tid * rper), r1 = min(R, r0 + rper);
    const SkipState carried{st->has_last, st->last_a, st->last_lme};

    bool sequential = fa.force_sequential != 0;
    for (int round = 0; !sequential; ++round) {
        const bool use_map = sm.nkeys_cur != 0u;
        for (uint32_t r = r0; r < r1; ++r) {
            SkipState ss = r == 0 ? carried : SkipState{0u, 0, 0};
            uint32_t i = run_start[r];
            const uint32_t i1 = run_start[r + 1];
            uint4 cur = view[i];
            while (i < i1) {
                const uint32_t pos = cur.x;
                const uint64_t pa = feed_base + pos;
                // inside the mag_buf of the last accepted message scanning resumes 18 samples before its end
                // (j = last_message_end - 19, then ++j); a later mag_buf starts exactly at its end (demod_2400.c:410-443)
                const bool skip = ss.has_last && pa < (blk_of(pa) == blk_of(ss.last_a) ? ss.last_lme - 18 : ss.last_lme);
                uint32_t best = 0, bi = i, best_w = 0, best_add = kUnchecked;     // SR_NOT_SET
                do {                                // phases ascending; strict > keeps the lowest phase on ties
                    const bool known = (cur.w & 0x200u) || (use_map && map_lookup(sm.cur, cur.y) < pos);
                    const uint32_t sc = sm.rank[2u * (cur.w & 0xFFu) + (known ? 1u : 0u)];
                    if (sc > best) { best = sc; bi = i; best_w = cur.w; best_add = cur.z; }
                    dec[i] = 0;
                    ++i;
                    if (i < i1) cur = view[i];
                } while (i < i1 && cur.x == pos);
                if (skip || best < uint32_t(kAcceptThreshold)) continue;
                dec[bi] = uint8_t(best);
                if (best_w & 0x400u) map_insert_min(sm.next, best_add, pos, &sm.nkeys_next);
                ss.has_last = 1u;
                ss.last_a = pa;
                ss.last_lme = pa + uint64_t((((best_w & 0x100u) ? 112 : 56) + 8) * 12 / 5);
            }
        }
        __syncthreads();
        // did this round reproduce the map it was computed with?
        if (sm.nkeys_next != sm.nkeys_cur && tid == 0) sm.changed = 1u;
        for (uint32_t i = tid; i < kMapSlots; i += kXThreads)
            if (sm.next.key[i] != kMapEmpty && map_lookup(sm.cur, sm.next.key[i]) != sm.next.pos[i]) sm.changed = 1u;
        __syncthreads();
        const bool changed = sm.changed != 0u;
        const uint32_t nk = sm.nkeys_next;
        __syncthreads();
        if (nk > kMapMaxKeys || st->active_count + nk >= kIcaoSlots - 1 || round >= 12) { sequential = true; break; }
        if (!changed) break;
        for (uint32_t i = tid; i < kMapSlots; i += kXThreads) {
            sm.cur.key[i] = sm.next.key[i]; sm.cur.pos[i] = sm.next.pos[i];
            sm.next.key[i] = kMapEmpty; sm.next.pos[i] = kMapEmpty;
        }
        if (tid == 0) { sm.nkeys_cur = nk; sm.nkeys_next = 0; sm.changed = 0; }
        __syncthreads();
    }

    if (sequential) {
        // the reference's own loop, one thread, on the tables in global memory: exact whatever they hold
        if (tid == 0) {
            SkipState ss = carried;
            uint64_t k = 0;
            resolve_range(view, 0, n, feed_base, block, ss,
                          [&](uint32_t key, uint32_t) { return icao_test(st, key); },
                          [&](uint32_t idx, int score, uint64_t pa) {
                              write_msg(out + k, recs[idx], pa, score, block);
                              ++k;
                              if (recs[idx].add_key != kUnchecked) icao_add(st, recs[idx].add_key);
                          });
            st->has_last = ss.has_last; st->last_a = ss.last_a; st->last_lme = ss.last_lme;
            st->n_out = feed_count + k;
            a.ctr->n_msgs = uint32_t(k);
            a.ctr->msg_base = feed_count;
            __threadfence_system();
        }
        return;
    }

    // emission, parallel over the records: every accepted record becomes one message, in stream order
    {
        const uint32_t per = (n + kXThreads - 1) / kXThreads;
        const uint32_t i0 = min(n, tid * per), i1 = min(n, i0 + per);
        uint32_t mine = 0, last = 0;                                       // last: index + 1 of my last accepted record
        for (uint32_t i = i0; i < i1; ++i) if (dec[i]) { ++mine; last = i + 1; }
        uint32_t total = 0;
        uint32_t off = block_exclusive_scan(mine, sm.scan, total);
        for (uint32_t i = i0; i < i1; ++i)
            if (const uint32_t sc = dec[i]) write_msg(out + off++, recs[i], feed_base + recs[i].pos, int(sc), block);
        // the stream's skip state: the last accepted message of the chunk
        sm.scan[tid] = last;
        __syncthreads();
        for (int d = kXThreads / 2; d; d >>= 1) {
            if (tid < d) sm.scan[tid] = max(sm.scan[tid], sm.scan[tid + d]);
            __syncthreads();
        }
        if (tid == 0) {
            if (const uint32_t li = sm.scan[0]) {
                const uint4 v = view[li - 1];
                const uint64_t pa = feed_base + v.x;
                st->has_last = 1u; st->last_a = pa;
                st->last_lme = pa + uint64_t((((v.w & 0x100u) ? 112 : 56) + 8) * 12 / 5);
            }
            st->n_out = feed_count + total;
            a.ctr->n_msgs = total;
            a.ctr->msg_base = feed_count;
        }
    }
    // New addresses enter the active table (icao_filter.c:71-84).  The reference inserts them in stream order; which
    // slot of a probe chain an address lands in changes no answer of icaoFilterTest while the table has room (and a
    // chunk that could fill it took the sequential replay above), so the insertions run in parallel.
    {
        uint32_t *t = st->table[st->active];
        uint32_t added = 0;
        for (uint32_t i = tid; i < kMapSlots; i += kXThreads) {
            const uint32_t v = sm.cur.key[i];
            if (v == kMapEmpty) continue;
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
    __syncthreads();
    if (tid == 0) __threadfence_system();
}

// mm.signalLevel of the accepted messages of one chunk (demod_2400.c:382-403): mean squared magnitude over the
// signal window, IEEE double divisions as on the host.  One warp per message; the finished 56-byte record goes to the
// feed's output array (mapped host memory or the caller's pinned buffer), where the host never touches it again.
__global__ void __launch_bounds__(256) level_kernel(FinishArgs fa)
{
    const ChunkArgs &a = fa.c;
    const int lane = threadIdx.x & 31;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5), gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint32_t n = a.ctr->n_msgs;
    const uint64_t base = a.ctr->msg_base;
    for (uint32_t k = gwarp; k < n; k += nwarps) {
        const uint64_t *src = reinterpret_cast<const uint64_t *>(fa.chunk_msgs + k);
        const uint64_t w = lane < 7 ? src[lane] : 0ull;
        const uint32_t pos = uint32_t(__shfl_sync(0xFFFFFFFFu, w, 6));
        const uint32_t vb0 = uint32_t(__shfl_sync(0xFFFFFFFFu, w, 2) >> 48) & 0xFFu;       // verbatim[0]: byte 22 of the record
        const double lvl = signal_level_warp(a, pos, vb0 >> 3, lane);
        if (base + k < fa.out_cap && lane < 7) {
            uint64_t *dst = reinterpret_cast<uint64_t *>(fa.out + base + k);
            dst[lane] = lane == 6 ? uint64_t(__double_as_longlong(lvl)) : w;
        }
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

size_t finish_smem_bytes() { return sizeof(FinishSmem) > sizeof(SortSmem) ? sizeof(FinishSmem) : sizeof(SortSmem); }

cudaError_t init_finish()
{
    return cudaFuncSetAttribute(order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(finish_smem_bytes()));
}

// order_kernel: one CTA per segment; level_kernel: `level_grid` CTAs of 8 warps (skipped for candidate scans)
cudaError_t launch_finish(const FinishArgs &fa, uint32_t level_grid, cudaStream_t s)
{
    if (fa.c.pos_end <= fa.c.pos_begin || !fa.c.nseg) return cudaSuccess;
    ADSBCU_LAUNCH(order_kernel, fa.c.nseg, kXThreads, finish_smem_bytes(), s, fa);
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
