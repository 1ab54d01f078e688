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

extern "C" int adsbcu_1090_replay_sharded(adsbcu_handle *h, int rank, int world, const adsbcu_piece *pieces, size_t npieces,
                                          uint64_t range_start, const uint8_t *halo_iq, size_t n_halo, adsbcu_allgather_fn allgather,
                                          void *ctx, unsigned flags, adsbcu_msg *out, size_t cap, size_t *n_out, adsbcu_shard_report *report)
{
    if (n_out) *n_out = 0;
    if (!h || world < 1 || rank < 0 || rank >= world || (!pieces && npieces) || (!out && cap) || (world > 1 && !allgather) ||
        (n_halo && !halo_iq) || n_halo > 326)
        return ADSBCU_ERR_ARG;
    try {
        const size_t nranks = size_t(world);
        std::vector<StartState> used(nranks), next(nranks);      // start states of ALL ranks (every rank derives the same)
        std::vector<uint32_t> mine(ADSBCU_SHARD_WORDS), all(nranks * ADSBCU_SHARD_WORDS), counts(nranks);
        std::vector<uint32_t> end_addrs(2 * 4096);
        size_t n_msgs = 0;
        uint32_t passes = 0, rounds = 0;
        adsbcu_shard_report sum{};
        bool run = true;
        for (;;) {
            if (run) {
                const StartState &st = used[size_t(rank)];
                int rc = adsbcu_1090_reset(h, 1);
                if (rc == ADSBCU_OK) rc = adsbcu_1090_seek(h, range_start, halo_iq, n_halo);
                if (rc == ADSBCU_OK) rc = adsbcu_1090_state_import(h, &st.skip, st.addrs.data(), st.addrs.size());
                n_msgs = 0;
                bool overflow = false;
                for (size_t p = 0; p < npieces && rc == ADSBCU_OK; ++p) {
                    size_t got = 0;
                    const size_t room = cap > n_msgs ? cap - n_msgs : 0;
                    rc = (flags & ADSBCU_SHARD_HOST_PIECES)
                             ? adsbcu_1090_feed(h, pieces[p].iq, size_t(pieces[p].nsamples), room ? out + n_msgs : nullptr, room, &got)
                             : adsbcu_1090_feed_device(h, pieces[p].iq, size_t(pieces[p].nsamples), room ? out + n_msgs : nullptr, room, &got);
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
                if (n_out) *n_out = n_msgs;
                if (overflow) { adsbcu::handle_set_error(h, "out[] too small for the accepted messages of this range"); return ADSBCU_ERR_OVERFLOW; }
                if (world > 1 && n_end > ADSBCU_SHARD_MAX_ADDRS) {                // saturating filter: membership would depend on insertion order
                    adsbcu::handle_set_error(h, "the ICAO filter holds more than ADSBCU_SHARD_MAX_ADDRS addresses: a sharded replay would not be exact");
                    return ADSBCU_ERR_OVERFLOW;
                }
                n_end = std::min<size_t>(n_end, ADSBCU_SHARD_MAX_ADDRS);
                std::sort(end_addrs.begin(), end_addrs.begin() + long(n_end));
                mine[0] = end.has_last;
                mine[1] = uint32_t(end.last_a); mine[2] = uint32_t(end.last_a >> 32);
                mine[3] = uint32_t(end.last_lme); mine[4] = uint32_t(end.last_lme >> 32);
                mine[5] = uint32_t(n_end);
                std::copy(end_addrs.begin(), end_addrs.begin() + long(n_end), mine.begin() + kHead);
            }
            if (world == 1) { rounds = 1; break; }
            // every rank contributes its latest end state (unchanged if it did not run this round)
            if (allgather(ctx, mine.data(), kHead + mine[5], all.data(), counts.data()) != 0) return ADSBCU_ERR_ARG;
            ++rounds;
            std::vector<uint32_t> acc;                                          // union of the end filters of the ranks in front
            for (int r = 0; r < world; ++r) {
                StartState &s = next[size_t(r)];
                s = StartState{};
                if (r > 0) {
                    const uint32_t *prev = &all[size_t(r - 1) * ADSBCU_SHARD_WORDS];
                    if (counts[size_t(r - 1)] < kHead || counts[size_t(r - 1)] != kHead + prev[5]) return ADSBCU_ERR_ARG;
                    s.skip.has_last = prev[0];
                    s.skip.last_a = uint64_t(prev[1]) | (uint64_t(prev[2]) << 32);
                    s.skip.last_lme = uint64_t(prev[3]) | (uint64_t(prev[4]) << 32);
                    std::vector<uint32_t> merged;
                    std::set_union(acc.begin(), acc.end(), prev + kHead, prev + kHead + prev[5], std::back_inserter(merged));
                    acc.swap(merged);
                    s.addrs = acc;
                }
            }
            bool any = false;
            for (int r = 0; r < world; ++r) any = any || !(next[size_t(r)] == used[size_t(r)]);
            if (!any) break;                                                    // fixed point: same decision on every rank
            run = !(next[size_t(rank)] == used[size_t(rank)]);
            used = next;
            if (rounds > uint32_t(world) + 2) return ADSBCU_ERR_ARG;            // cannot happen: rank k is exact after round k
        }
        sum.passes = passes; sum.rounds = rounds; sum.start_addrs = uint32_t(used[size_t(rank)].addrs.size());
        sum.launches += 3ull * passes;                 // reset, seek (carry) and state_import kernels of every pass
        if (report) *report = sum;
        return ADSBCU_OK;
    } catch (const std::bad_alloc &) {
        return ADSBCU_ERR_NOMEM;
    }
}
