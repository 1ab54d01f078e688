// adsbcu_1090_replay_sharded: one capture over several GPUs, every shard resolved on its own GPU (BASELINE config 4,
// SURVEY 8e).  Host code on top of the public entry points only (include/adsbcu.h): reset / seek / state_import /
// feed_device / state_export.  No data-path collective: the callback carries a few KiB per rank and round.
//
// What a sample range needs from everything in front of it (fifo.c:173-210, demod_2400.c:106-115,432-443,
// icao_filter.c:58-83): the 326 samples before it (the caller's halo), the skip state of the last accepted message
// and the contents of the ICAO filter.  The last two are found by iteration: round 0 assumes "nothing in front",
// every later round gives rank r the end skip state of rank r-1 and the union of the end filters of the ranks < r,
// and a rank runs its range again only if that differs from what it used.  A state in which nothing changes is the
// sequential run's state on every rank (induction over the ranks), so the result does not depend on `world`.
//
// Who waits for whom: the first range never runs twice (nothing precedes it) and nobody needs the end state of the LAST
// range, which therefore only runs once the others have made their first pass.  adsbcu_1090_replay_sharded_tail puts
// both on rank 0: the capture is cut into world + 1 ranges, rank 0 demodulates the first in round 0 and the last in
// round 1, every other rank its own range in both rounds -- two passes over 1 / (world + 1) of the capture instead of
// two passes over 1 / world.
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"      // handle_set_error only: everything else goes through the public entry points

#include <algorithm>
#include <cstring>
#include <iterator>
#include <new>
#include <vector>

namespace {

struct StartState {
    adsbcu_1090_state skip{};
    std::vector<uint32_t> addrs;                       // sorted, unique
    bool operator==(const StartState &o) const
    {
        return skip.has_last == o.skip.has_last && (!skip.has_last || (skip.last_a == o.skip.last_a && skip.last_lme == o.skip.last_lme)) &&
               addrs == o.addrs;
    }
};

// end state of a rank as it travels through the callback: [has_last, last_a lo, hi, last_lme lo, hi, n, addrs...]
constexpr uint32_t kHead = 6;

}  // namespace

extern "C" int adsbcu_1090_replay_sharded_tail(adsbcu_handle *h, int rank, int world, const adsbcu_piece *pieces, size_t npieces,
                                               uint64_t range_start, const uint8_t *halo_iq, size_t n_halo,
                                               const adsbcu_piece *tail_pieces, size_t n_tail_pieces, uint64_t tail_start,
                                               const uint8_t *tail_halo_iq, size_t n_tail_halo, adsbcu_allgather_fn allgather, void *ctx,
                                               unsigned flags, adsbcu_msg *out, size_t cap, size_t *n_out, size_t *n_first,
                                               adsbcu_shard_report *report)
{
    if (n_out) *n_out = 0;
    if (n_first) *n_first = 0;
    const bool has_tail = (flags & ADSBCU_SHARD_HAS_TAIL) != 0 && world > 1;
    if (!h || world < 1 || rank < 0 || rank >= world || (!pieces && npieces) || (!out && cap) || (world > 1 && !allgather) ||
        (n_halo && !halo_iq) || n_halo > 326 || (has_tail && rank == 0 && ((!tail_pieces && n_tail_pieces) || (n_tail_halo && !tail_halo_iq) || n_tail_halo > 326)))
        return ADSBCU_ERR_ARG;
    try {
        const size_t nvirt = size_t(world) + (has_tail ? 1u : 0u);          // ranges of the capture: one per rank (+ the tail, on rank 0)
        std::vector<StartState> used(nvirt), next(nvirt);                   // start states of ALL ranges (every rank derives the same)
        const size_t nranks = size_t(world);
        std::vector<uint32_t> mine(ADSBCU_SHARD_WORDS), all(nranks * ADSBCU_SHARD_WORDS), counts(nranks);
        std::vector<uint32_t> end_addrs(2 * 4096);
        size_t n_own = 0, n_tail_msgs = 0;
        uint32_t passes = 0, rounds = 0;
        adsbcu_shard_report sum{};
        // one pass over a range from a start state: messages to out[at ...], end state packed into `packed` (if wanted)
        auto run_range = [&](const adsbcu_piece *pc, size_t npc, uint64_t start, const uint8_t *halo, size_t nh, const StartState &st, size_t at,
                             size_t &n_msgs, uint32_t *packed) -> int {
            int rc = adsbcu_1090_reset(h, 1);
            if (rc == ADSBCU_OK) rc = adsbcu_1090_seek(h, start, halo, nh);
            if (rc == ADSBCU_OK) rc = adsbcu_1090_state_import(h, &st.skip, st.addrs.data(), st.addrs.size());
            n_msgs = 0;
            bool overflow = false;
            for (size_t p = 0; p < npc && rc == ADSBCU_OK; ++p) {
                size_t got = 0;
                const size_t room = cap > at + n_msgs ? cap - at - n_msgs : 0;
                rc = (flags & ADSBCU_SHARD_HOST_PIECES)
                         ? adsbcu_1090_feed(h, pc[p].iq, size_t(pc[p].nsamples), room ? out + at + n_msgs : nullptr, room, &got)
                         : adsbcu_1090_feed_device(h, pc[p].iq, size_t(pc[p].nsamples), room ? out + at + n_msgs : nullptr, room, &got);
                adsbcu_stats fs{};
                if (adsbcu_get_stats(h, &fs) == ADSBCU_OK) {
                    sum.chunks += fs.chunks; sum.launches += fs.launches; sum.ms_front += fs.ms_front; sum.ms_messages += fs.ms_messages;
                    sum.ms_order += fs.ms_order; sum.ms_resolve += fs.ms_resolve; sum.ms_h2d += fs.ms_h2d;
                }
                if (rc == ADSBCU_ERR_OVERFLOW && got > room) { overflow = true; rc = ADSBCU_OK; }   // keep counting, report at the end
                n_msgs += got;
            }
            if (rc != ADSBCU_OK) return rc;
            adsbcu_1090_state end{};
            size_t n_end = 0;
            rc = adsbcu_1090_state_export(h, &end, end_addrs.data(), end_addrs.size(), &n_end);
            if (rc != ADSBCU_OK) return rc;
            ++passes;
            if (overflow) { adsbcu::handle_set_error(h, "out[] too small for the accepted messages of this range"); return ADSBCU_ERR_OVERFLOW; }
            if (world > 1 && n_end > ADSBCU_SHARD_MAX_ADDRS) {                // saturating filter: membership would depend on insertion order
                adsbcu::handle_set_error(h, "the ICAO filter holds more than ADSBCU_SHARD_MAX_ADDRS addresses: a sharded replay would not be exact");
                return ADSBCU_ERR_OVERFLOW;
            }
            if (packed) {
                n_end = std::min<size_t>(n_end, ADSBCU_SHARD_MAX_ADDRS);
                std::sort(end_addrs.begin(), end_addrs.begin() + long(n_end));
                packed[0] = end.has_last;
                packed[1] = uint32_t(end.last_a); packed[2] = uint32_t(end.last_a >> 32);
                packed[3] = uint32_t(end.last_lme); packed[4] = uint32_t(end.last_lme >> 32);
                packed[5] = uint32_t(n_end);
                std::copy(end_addrs.begin(), end_addrs.begin() + long(n_end), packed + kHead);
            }
            return ADSBCU_OK;
        };
        bool run_own = true, run_tail = false;
        for (;;) {
            if (run_own) {
                const int rc = run_range(pieces, npieces, range_start, halo_iq, n_halo, used[size_t(rank)], 0, n_own, mine.data());
                if (n_out) *n_out = n_own;
                if (n_first) *n_first = n_own;
                if (rc != ADSBCU_OK) return rc;
            }
            if (run_tail) {           // rank 0 only; its own (first) range never runs again, so the tail's messages go behind it
                const int rc = run_range(tail_pieces, n_tail_pieces, tail_start, tail_halo_iq, n_tail_halo, used[size_t(world)], n_own, n_tail_msgs, nullptr);
                if (n_out) *n_out = n_own + n_tail_msgs;
                if (rc != ADSBCU_OK) return rc;
            }
            if (world == 1) { rounds = 1; break; }
            // every rank contributes the latest end state of its own range (unchanged if it did not run this round)
            if (allgather(ctx, mine.data(), kHead + mine[5], all.data(), counts.data()) != 0) return ADSBCU_ERR_ARG;
            ++rounds;
            std::vector<uint32_t> acc;                                          // union of the end filters of the ranges in front
            for (size_t v = 0; v < nvirt; ++v) {
                StartState &s = next[v];
                s = StartState{};
                if (v > 0) {
                    const uint32_t *prev = &all[(v - 1) * ADSBCU_SHARD_WORDS];
                    if (counts[v - 1] < kHead || counts[v - 1] != kHead + prev[5]) return ADSBCU_ERR_ARG;
                    s.skip.has_last = prev[0];
                    s.skip.last_a = uint64_t(prev[1]) | (uint64_t(prev[2]) << 32);
                    s.skip.last_lme = uint64_t(prev[3]) | (uint64_t(prev[4]) << 32);
                    std::vector<uint32_t> merged;
                    std::set_union(acc.begin(), acc.end(), prev + kHead, prev + kHead + prev[5], std::back_inserter(merged));
                    acc.swap(merged);
                    s.addrs = acc;
                }
            }
            const bool tail_ran = rounds >= 2;                                  // (the tail's first pass is in round 1)
            bool any = has_tail && !tail_ran;
            for (size_t v = 0; v < nvirt; ++v) any = any || !(next[v] == used[v]);
            if (!any) break;                                                    // fixed point: same decision on every rank
            run_own = !(next[size_t(rank)] == used[size_t(rank)]);
            run_tail = has_tail && rank == 0 && (!tail_ran || !(next[size_t(world)] == used[size_t(world)]));
            used = next;
            if (rounds > uint32_t(world) + 3) return ADSBCU_ERR_ARG;            // cannot happen: range k is exact after round k
        }
        sum.passes = passes; sum.rounds = rounds; sum.start_addrs = uint32_t(used[size_t(rank)].addrs.size());
        sum.launches += 3ull * passes;                 // reset, seek (carry) and state_import kernels of every pass
        if (report) *report = sum;
        return ADSBCU_OK;
    } catch (const std::bad_alloc &) {
        return ADSBCU_ERR_NOMEM;
    }
}

extern "C" int adsbcu_1090_replay_sharded(adsbcu_handle *h, int rank, int world, const adsbcu_piece *pieces, size_t npieces,
                                          uint64_t range_start, const uint8_t *halo_iq, size_t n_halo, adsbcu_allgather_fn allgather,
                                          void *ctx, unsigned flags, adsbcu_msg *out, size_t cap, size_t *n_out, adsbcu_shard_report *report)
{
    return adsbcu_1090_replay_sharded_tail(h, rank, world, pieces, npieces, range_start, halo_iq, n_halo, nullptr, 0, 0, nullptr, 0, allgather, ctx,
                                           flags & ~unsigned(ADSBCU_SHARD_HAS_TAIL), out, cap, n_out, nullptr, report);
}
