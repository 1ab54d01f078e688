// C ABI (include/adsbcu.h): handle life cycle and the chunk pipeline of the 1090 MHz path.
//
// Per chunk of a feed:   [upload]  ->  front_kernel  ->  message_kernel                  (stream `stream`)
//                                                        -> order_kernel -> level_kernel  (stream `post_stream`, beside the next chunk)
// The front kernel runs on all but a few SMs, so the order / level kernels of chunk c (late records, order, the ordered
// pass of demodulate2400, signal levels, final adsbcu_msg records) have somewhere to run while chunk c+1 is scanned.  The host only
// submits, waits for a slot before reusing it, and copies the finished messages out (or nothing at all when the
// caller's out[] is pinned memory the GPU can write itself).  No step of the path runs on the CPU.
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

using namespace adsbcu;

namespace {

thread_local std::string g_create_error;

// Reference ICAO filter (icao_filter.c:23-127) on the host: only adsbcu_1090_resolve uses it (merged candidates of
// sharded captures; works on handles without a GPU).  Feeds keep their filter in device memory (StreamState).
struct IcaoFilter {
    static constexpr uint32_t kSlots = 4096, kEmpty = 0xFFFFFFFFu;
    uint32_t t[2][kSlots];
    int active = 0;
    IcaoFilter() { clear(); }
    void clear()
    {
        std::memset(t, 0xFF, sizeof(t));
        active = 0;
    }
    static uint32_t hash(uint32_t v)
    {
        uint32_t h = 0;
        for (int k = 0; k < 3; ++k) {
            h += (v >> (8 * k)) & 0xFF;
            h += h << 10;
            h ^= h >> 6;
        }
        h += h << 3;
        h ^= h >> 11;
        h += h << 15;
        return h & (kSlots - 1);
    }
    void add(uint32_t v)
    {
        uint32_t *a = t[active];
        const uint32_t h0 = hash(v);
        uint32_t h = h0;
        while (a[h] != kEmpty && a[h] != v) {
            h = (h + 1) & (kSlots - 1);
            if (h == h0) return;
        }
        if (a[h] == kEmpty) a[h] = v;
    }
    static bool probe(const uint32_t *a, uint32_t v)
    {
        const uint32_t h0 = hash(v);
        uint32_t h = h0;
        while (a[h] != kEmpty && a[h] != v) {
            h = (h + 1) & (kSlots - 1);
            if (h == h0) break;
        }
        return a[h] == v;
    }
    bool test(uint32_t v) const { return probe(t[0], v) || probe(t[1], v); }
    void expire()                       // icaoFilterExpire (icao_filter.c:112-127)
    {
        active ^= 1;
        std::memset(t[active], 0xFF, sizeof(t[active]));
    }
};

}  // namespace

constexpr int kSlots = 4;            // chunks in flight

// Everything one in-flight chunk owns.
struct Slot {
    uint16_t *d_planes = nullptr;
    uint32_t *d_pre = nullptr, *d_pre_count = nullptr;
    PhaseRec *d_pend = nullptr;
    uint32_t *d_pend_count = nullptr;
    FullRec *d_rec = nullptr;
    uint32_t *d_rec_count = nullptr;
    FullRec *d_ordered = nullptr;
    FullRec *h_cand = nullptr, *d_cand = nullptr;     // sharded scans: the ordered array in mapped host memory (allocated on first use)
    unsigned long long *d_seg_flag = nullptr;
    uint32_t *d_run_start = nullptr;
    uint4 *d_view = nullptr;
    uint2 *d_accepted = nullptr;      // accepted records of the chunk: the level kernel turns them into messages
    ChunkCounters *d_ctr = nullptr;   // = d_zero
    unsigned char *d_zero = nullptr;  // what a chunk must find zeroed, one block: counters | rec_count[nseg] | pend_count[nseg]
    size_t zero_bytes = 0;
    ChunkCounters *h_ctr = nullptr;   // pinned
    cudaEvent_t ev[9]{};              // front begin, front end, messages end | resolve begin, resolve end, counters on the host | order end, order begin | messages begin
};

struct adsbcu_handle {
    adsbcu_config cfg{};
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;        // front + message kernels
    cudaStream_t copy_stream = nullptr;   // host->device uploads of the feed, chunk by chunk
    cudaStream_t msg_stream = nullptr;    // message kernels when they run beside the next chunk's front kernel (msg_overlap)
    bool msg_overlap = true;
    SmPartition part;                     // dense kernels / post kernels on disjoint SM sets (green contexts), if available
    cudaStream_t post_stream = nullptr;   // resolve kernels, in chunk order (high priority, like the next two)
    cudaStream_t order_stream = nullptr;  // order kernels
    cudaStream_t level_stream = nullptr;  // level kernels + the chunk counters to the host
    cudaEvent_t ev_h2d[2]{};
    cudaEvent_t ev_up[kSlots]{};      // upload of the chunk using that slot has finished
    std::string error;

    DevTables *d_tab = nullptr;
    uint32_t *d_bitmap = nullptr;
    uint16_t *d_carry = nullptr;
    uint8_t *d_iq = nullptr;          // staging for host feeds
    int16_t *d_sc16 = nullptr;        // staging for SC16 / SC16Q11 host feeds (allocated on first use)
    StreamState *d_state = nullptr;
    uint32_t *h_word = nullptr, *d_word = nullptr;    // one mapped word for small device->host answers
    uint32_t *d_import = nullptr;     // adsbcu_1090_state_import: the address list on the device (allocated on first use)
    adsbcu_msg *h_msgs = nullptr, *d_msgs = nullptr;  // accepted messages of the running feed (pinned + mapped), grown on demand
    size_t msgs_cap = 0;

    // geometry
    uint32_t nseg = 0, msg_grid = 0, level_grid = 128;
    uint32_t unit_positions = 0, max_tps = 1, host_tps = 4, res_tps = 8;   // chunk = unit_positions * (tiles per segment)
    uint32_t ntiles_max = 0, plane_stride = 0, pre_cap = 0, pend_cap = 0, rec_cap = 0, full_cap = 0;
    uint32_t chunk_epoch = 0;         // sequence number of the last submitted chunk (never 0 in a launch)
    uint32_t grow = 1;                // list capacities are `grow` times their base size (doubled after an overflow)
    uint32_t resolve_batch = 1u << 30; // ADSBCU_RESOLVE_BATCH: records per batch of the ordered pass (tests)
    bool force_sequential = false;    // ADSBCU_RESOLVE_SEQ=1: the finish kernel always takes its one-thread replay (tests)
    Slot slot[kSlots];
    uint16_t *d_mag = nullptr;        // stage-level magnitude output
    size_t d_mag_cap = 0;

    // stream state that lives on the host
    uint64_t stream_pos = 0;          // samples consumed so far == padded coordinate of the next feed's first start
    adsbcu_stats stats{};
    PathB *pathb = nullptr;
    Fields *fields = nullptr;
    PathC *pathc = nullptr;

    // adsbcu_1090_resolve (merged candidates): host-side ordered pass with its own filter and skip state
    bool host_only = false;           // cfg.device == -1: merge-only handle, no CUDA at all
    bool has_last = false;
    uint64_t last_a = 0, last_lme = 0;
    IcaoFilter filter;

    // sharded captures (adsbcu_1090_scan_candidates / _revive)
    uint16_t *d_carry_prev = nullptr; // the carry the last scan started from (revive re-slices with it)
    PhaseRec *d_deferred = nullptr;
    uint32_t *d_n_deferred = nullptr, *d_known_tmp = nullptr;
    uint32_t deferred_cap = 0;
    const uint8_t *last_iq = nullptr; // device samples of the last candidate scan
    uint32_t last_n = 0;
    uint64_t last_base = 0;
    // Mode A/C (adsbcu_1090_modeac_block)
    uint8_t *d_ac_ov = nullptr;
    unsigned long long *d_ac_sums = nullptr;      // [2] sums, then the record counter
    AcRec *d_ac_rec = nullptr, *h_ac_rec = nullptr;
    uint32_t ac_cap = 0;
};

namespace adsbcu {
HandleView handle_view(const adsbcu_handle *h) { return HandleView{h->device, h->sm_count, h->d_tab, h->host_only}; }
void handle_set_error(adsbcu_handle *h, const char *what) { h->error = what; }
}  // namespace adsbcu

namespace {

bool cuda_ok(adsbcu_handle *h, cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true;
    char buf[256];
    std::snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
    if (h) h->error = buf; else g_create_error = buf;
    return false;
}
#define CK(call) do { if (!cuda_ok(h, (call), #call)) return ADSBCU_ERR_CUDA; } while (0)

void free_slots(adsbcu_handle *h)
{
    for (Slot &s : h->slot) {
        cudaFree(s.d_planes); cudaFree(s.d_pre); cudaFree(s.d_pre_count); cudaFree(s.d_pend);
        cudaFree(s.d_rec); cudaFree(s.d_ordered); cudaFree(s.d_seg_flag);
        cudaFree(s.d_run_start); cudaFree(s.d_zero); cudaFree(s.d_accepted); cudaFree(s.d_view);
        if (s.h_cand) cudaFreeHost(s.h_cand);
        if (s.h_ctr) cudaFreeHost(s.h_ctr);
        cudaEvent_t ev[9];
        std::memcpy(ev, s.ev, sizeof(ev));
        s = Slot{};
        std::memcpy(s.ev, ev, sizeof(ev));
    }
}

// Slot buffers for chunks of up to max_tps tiles per segment; list capacities scale with h->grow.
int alloc_slots(adsbcu_handle *h)
{
    const size_t chunk = size_t(h->unit_positions) * h->max_tps;
    h->ntiles_max = uint32_t((chunk + kFNeed + kWTile - 1) / kWTile);
    h->plane_stride = h->ntiles_max * 32 + 32;
    const size_t seg_pos = size_t((h->ntiles_max + h->nseg - 1) / h->nseg) * kWTile;
    // Gaussian noise confirms 1 preamble per 93 positions and parks 1 record per 90 (with --fix); full-scale uniform
    // noise about 4x that.  Overflow is detected, the lists are doubled and the chunk is run again.
    h->pre_cap = uint32_t(std::min<size_t>(seg_pos, std::max<size_t>(1024, seg_pos / 8 * h->grow)));
    h->pend_cap = uint32_t(std::min<size_t>(5 * seg_pos, std::max<size_t>(1024, seg_pos / 8 * h->grow)));
    h->rec_cap = uint32_t(std::min<size_t>(4096, std::max<size_t>(256, seg_pos / 16 * h->grow)));
    h->full_cap = uint32_t(std::min<size_t>(size_t(h->nseg) * h->rec_cap, std::max<size_t>(1u << 13, chunk / 64 * h->grow)));
    for (Slot &s : h->slot) {
        CK(cudaMalloc(&s.d_planes, size_t(12) * h->plane_stride * sizeof(uint16_t)));
        CK(cudaMalloc(&s.d_pre, size_t(h->nseg) * h->pre_cap * sizeof(uint32_t)));
        CK(cudaMalloc(&s.d_pre_count, size_t(h->nseg) * sizeof(uint32_t)));
        CK(cudaMalloc(&s.d_pend, size_t(h->nseg) * h->pend_cap * sizeof(PhaseRec)));
        CK(cudaMalloc(&s.d_rec, size_t(h->nseg) * h->rec_cap * sizeof(FullRec)));
        CK(cudaMalloc(&s.d_ordered, size_t(h->full_cap) * sizeof(FullRec)));
        CK(cudaMalloc(&s.d_seg_flag, size_t(h->nseg) * sizeof(unsigned long long)));
        CK(cudaMemset(s.d_seg_flag, 0, size_t(h->nseg) * sizeof(unsigned long long)));
        CK(cudaMalloc(&s.d_run_start, (size_t(h->full_cap) + 1) * sizeof(uint32_t)));
        CK(cudaMalloc(&s.d_accepted, size_t(h->full_cap) * sizeof(uint2)));
        CK(cudaMalloc(&s.d_view, size_t(h->full_cap) * sizeof(uint4)));
        static_assert(sizeof(ChunkCounters) <= 256, "zero block layout");
        s.zero_bytes = 256 + 2 * size_t(h->nseg) * sizeof(uint32_t);
        CK(cudaMalloc(&s.d_zero, s.zero_bytes));
        CK(cudaMemset(s.d_zero, 0, s.zero_bytes));
        s.d_ctr = reinterpret_cast<ChunkCounters *>(s.d_zero);
        s.d_rec_count = reinterpret_cast<uint32_t *>(s.d_zero + 256);
        s.d_pend_count = s.d_rec_count + h->nseg;
        CK(cudaHostAlloc(&s.h_ctr, sizeof(ChunkCounters), cudaHostAllocDefault));
        std::memset(s.h_ctr, 0, sizeof(ChunkCounters));
    }
    return ADSBCU_OK;
}

void free_all(adsbcu_handle *h)
{
    if (h->host_only) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->post_stream) cudaStreamSynchronize(h->post_stream);
    if (h->order_stream) cudaStreamSynchronize(h->order_stream);
    if (h->level_stream) cudaStreamSynchronize(h->level_stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    if (h->msg_stream) cudaStreamSynchronize(h->msg_stream);
    cudaFree(h->d_sc16); cudaFree(h->d_import);
    cudaFree(h->d_ac_ov); cudaFree(h->d_ac_sums); cudaFree(h->d_ac_rec);
    if (h->h_ac_rec) cudaFreeHost(h->h_ac_rec);
    cudaFree(h->d_carry_prev); cudaFree(h->d_deferred); cudaFree(h->d_n_deferred); cudaFree(h->d_known_tmp);
    pathb_free(h->pathb);
    h->pathb = nullptr;
    fields_free(h->fields);
    h->fields = nullptr;
    pathc_free(h->pathc);
    h->pathc = nullptr;
    cudaFree(h->d_tab); cudaFree(h->d_bitmap); cudaFree(h->d_carry); cudaFree(h->d_iq); cudaFree(h->d_mag); cudaFree(h->d_state);
    if (h->h_word) cudaFreeHost(h->h_word);
    if (h->h_msgs) cudaFreeHost(h->h_msgs);
    free_slots(h);
    for (Slot &s : h->slot)
        for (auto &e : s.ev) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_h2d) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_up) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->msg_stream) cudaStreamDestroy(h->msg_stream);
    if (h->post_stream) cudaStreamDestroy(h->post_stream);
    if (h->order_stream) cudaStreamDestroy(h->order_stream);
    if (h->level_stream) cudaStreamDestroy(h->level_stream);
}

// ---- ordered pass on the host: adsbcu_1090_resolve only ---------------------------------------------------
struct Sink {
    adsbcu_msg *out;
    size_t cap, n;
    bool overflow;
    adsbcu_cand *cands = nullptr;     // candidate mode: records are copied out instead of resolved
};

static_assert(sizeof(adsbcu_cand) == 64, "adsbcu_cand layout");

void copy_candidates(const FullRec *recs, size_t nrec, uint64_t feed_base, Sink &sink)
{
    for (size_t i = 0; i < nrec; ++i, ++sink.n) {
        if (sink.n >= sink.cap) { sink.overflow = true; continue; }
        const FullRec &r = recs[i];
        adsbcu_cand &c = sink.cands[sink.n];
        c.pos = feed_base + r.pos;
        c.phase = r.phase; c.kind = r.kind; c.corrections = r.corrections; c.msgtype = r.msgtype;
        std::memcpy(c.verbatim, r.verbatim, 14);
        std::memcpy(c.msg, r.msg, 14);
        c.crc = r.crc; c.addr = r.addr; c.test_key = r.test_key; c.add_key = r.add_key; c.signal_level = r.signal_level;
    }
}

// demodulate2400's sequential rules over candidates in ascending (pos, phase) order (demod_2400.c:336,349-355,
// 410-443; mode_s.c:781-788).  The device twin of this loop is resolve_range() in adsbcu_finish.cu.
void resolve_host(adsbcu_handle *h, const adsbcu_cand *recs, size_t nrec, Sink &sink)
{
    const uint64_t block = h->cfg.block_samples;
    auto blk_of = [&](uint64_t x) { return x / block; };
    size_t i = 0;
    while (i < nrec) {
        const uint64_t a = recs[i].pos;
        size_t e = i + 1;
        while (e < nrec && recs[e].pos == a) ++e;
        if (h->has_last) {
            const bool same_block = blk_of(a) == blk_of(h->last_a);
            if (a < (same_block ? h->last_lme - 18 : h->last_lme)) { i = e; continue; }
        }
        int best = 0;                       // SR_NOT_SET
        const adsbcu_cand *bp = nullptr;
        uint32_t last_key = kUnchecked;
        bool last_known = false;
        for (size_t k = i; k < e; ++k) {    // phases ascending; strict > keeps the lowest phase on ties
            if (const uint32_t key = recs[k].test_key; key != last_key) { last_key = key; last_known = h->filter.test(key); }
            const int sc = rank_of(recs[k].kind, last_known);
            if (sc > best) { best = sc; bp = &recs[k]; }
        }
        i = e;
        if (best < kAcceptThreshold) continue;
        const adsbcu_cand &r = *bp;
        if (sink.n < sink.cap) {
            adsbcu_msg &m = sink.out[sink.n];
            const uint64_t bstart = blk_of(a) * block;
            m.timestamp = uint64_t(double(bstart) * 12e6 / 2400000.0) + (a - bstart) * 5 + (8 + 56) * 12 + r.phase;
            std::memcpy(m.msg, r.msg, 14);
            std::memcpy(m.verbatim, r.verbatim, 14);
            m.msgbits = (r.msgtype & 16) ? 112 : 56;
            m.correctedbits = uint8_t(r.corrections);
            m.msgtype = r.msgtype;
            m.score = uint8_t(best);
            m.crc = r.crc;
            m.addr = r.addr;
            m.signal_level = r.signal_level;
        } else {
            sink.overflow = true;
        }
        ++sink.n;
        if (r.add_key != kUnchecked) h->filter.add(r.add_key);
        h->has_last = true;
        h->last_a = a;
        h->last_lme = a + uint64_t((((r.verbatim[0] & 0x80) ? 112 : 56) + 8) * 12 / 5);
    }
}

// ---- one feed of at most max_feed_samples: a pipeline of chunks ---------------------------------------------
// iq_host != nullptr: the samples are uploaded into d_iq chunk by chunk on copy_stream (a chunk only looks at samples
// before its own end, see ChunkArgs::nsamples).  sc16_host != nullptr (then iq_host must be null): SC16 / SC16Q11
// samples are uploaded into d_sc16, converted to uint16 magnitudes into d_iq (2 bytes per sample, like uc8 I/Q) and
// every kernel runs in magnitudes-in mode (convert.c:46-92).
// Resolve mode (sink.cands == nullptr): accepted messages go to out_dev[0 .. out_cap) on the device side, the count
// is StreamState::n_out.  Candidate mode: the ordered records of every chunk are copied to sink.cands.
constexpr int kOverflowRetry = -100;

int run_chunks(adsbcu_handle *h, const uint8_t *d_iq, const uint8_t *iq_host, uint32_t n, Sink &sink, adsbcu_msg *out_dev, uint64_t out_cap,
               const int16_t *sc16_host, int q11, const std::vector<uint32_t> &bounds, uint32_t first_chunk, uint32_t &failed_chunk,
               bool already_uploaded)
{
    adsbcu_stats &st = h->stats;
    const uint32_t nchunks = uint32_t(bounds.size() - 1);
    const bool upload = (iq_host || sc16_host) && !already_uploaded;
    failed_chunk = nchunks;

    auto submit = [&](uint32_t c) -> int {
        const int si = int(c % kSlots);
        Slot &s = h->slot[si];
        ChunkArgs a{};
        a.iq = d_iq;
        a.carry_mag = h->d_carry;
        a.pos_begin = bounds[c];
        a.pos_end = bounds[c + 1];
        a.nsamples = (iq_host || sc16_host) ? a.pos_end : n;      // host feeds: nothing past the chunk has been uploaded
        a.tab = h->d_tab;
        a.bitmap = h->d_bitmap;
        a.nseg = h->nseg;
        a.planes = s.d_planes; a.plane_stride = h->plane_stride;
        a.pre = s.d_pre; a.pre_cap = h->pre_cap; a.pre_count = s.d_pre_count;
        a.pend = s.d_pend; a.pend_cap = h->pend_cap; a.pend_count = s.d_pend_count;
        a.rec = s.d_rec; a.rec_cap = h->rec_cap; a.rec_count = s.d_rec_count;
        a.ordered = sink.cands ? s.d_cand : s.d_ordered; a.full_cap = h->full_cap;
        a.ctr = s.d_ctr;
        if (sc16_host) {
            a.iq = nullptr;
            a.mag_in = reinterpret_cast<const uint16_t *>(h->d_iq);
        }
        if (upload) {
            if (sc16_host)
                CK(cudaMemcpyAsync(h->d_sc16 + 2 * size_t(a.pos_begin), sc16_host + 2 * size_t(a.pos_begin),
                                   4 * size_t(a.pos_end - a.pos_begin), cudaMemcpyHostToDevice, h->copy_stream));
            else
                CK(cudaMemcpyAsync(h->d_iq + 2 * size_t(a.pos_begin), iq_host + 2 * size_t(a.pos_begin),
                                   2 * size_t(a.pos_end - a.pos_begin), cudaMemcpyHostToDevice, h->copy_stream));
            CK(cudaEventRecord(h->ev_up[si], h->copy_stream));
            if (c + 1 == nchunks) CK(cudaEventRecord(h->ev_h2d[1], h->copy_stream));
            CK(cudaStreamWaitEvent(h->stream, h->ev_up[si], 0));
            if (sc16_host)
                CK(launch_convert_sc16(h->d_sc16 + 2 * size_t(a.pos_begin), a.pos_end - a.pos_begin,
                                       reinterpret_cast<uint16_t *>(h->d_iq) + a.pos_begin, q11, h->stream));
        }
        // the slot's counters are zero: cleared when the slot was allocated and after every use (below)
        CK(cudaEventRecord(s.ev[0], h->stream));
        CK(launch_front(a, h->stream));
        CK(cudaEventRecord(s.ev[1], h->stream));
        cudaStream_t ms = h->msg_overlap ? h->msg_stream : h->stream;
        if (h->msg_overlap) CK(cudaStreamWaitEvent(ms, s.ev[1], 0));
        CK(cudaEventRecord(s.ev[8], ms));
        CK(launch_messages(a, h->msg_grid, ms));
        CK(cudaEventRecord(s.ev[2], ms));
        // the finish kernel only touches this chunk's slot, the stream state and (reading) the address set, where a
        // superset merely keeps more records for the exact ordered pass: it runs beside the next chunk's scan
        FinishArgs fa{};
        fa.c = a;
        fa.st = h->d_state;
        fa.out = out_dev; fa.out_cap = out_cap;
        fa.accepted = s.d_accepted;
        fa.feed_base = h->stream_pos;
        fa.block_samples = h->cfg.block_samples;
        fa.candidates_only = sink.cands ? 1u : 0u;
        fa.force_sequential = h->force_sequential ? 1u : 0u;
        fa.batch = h->resolve_batch;
        fa.seg_flag = s.d_seg_flag; fa.epoch = ++h->chunk_epoch ? h->chunk_epoch : ++h->chunk_epoch; fa.run_start = s.d_run_start; fa.view = s.d_view;
        if (sink.cands) { fa.deferred = h->d_deferred; fa.deferred_cap = h->deferred_cap; fa.n_deferred = h->d_n_deferred; }
        // order(c) needs messages(c); resolve(c) needs order(c) and resolve(c-1); level(c) needs resolve(c): three
        // streams, so that the one-CTA ordered pass of a chunk overlaps the wide kernels of its neighbours
        CK(cudaStreamWaitEvent(h->order_stream, s.ev[2], 0));
        CK(cudaEventRecord(s.ev[7], h->order_stream));
        CK(launch_order(fa, h->order_stream));
        CK(cudaEventRecord(s.ev[6], h->order_stream));
        CK(cudaStreamWaitEvent(h->post_stream, s.ev[6], 0));
        CK(cudaEventRecord(s.ev[3], h->post_stream));
        CK(launch_resolve(fa, h->post_stream));
        CK(cudaEventRecord(s.ev[4], h->post_stream));
        CK(cudaStreamWaitEvent(h->level_stream, s.ev[4], 0));
        CK(launch_level(fa, h->level_grid, h->level_stream));
        CK(cudaMemcpyAsync(s.h_ctr, s.d_ctr, sizeof(ChunkCounters), cudaMemcpyDeviceToHost, h->level_stream));
        CK(cudaMemsetAsync(s.d_zero, 0, s.zero_bytes, h->level_stream));      // ready for the slot's next chunk (submitted after ev[5])
        CK(cudaEventRecord(s.ev[5], h->level_stream));
        st.launches += 5 + (upload && sc16_host ? 1 : 0);      // front, messages, order, resolve, level (+ SC16 converter)
        st.chunks += 1;
        return ADSBCU_OK;
    };
    // host half of a chunk: its counters are on the host; candidate mode also copies the records out
    auto retire = [&](uint32_t c) -> int {
        Slot &s = h->slot[c % kSlots];
        CK(cudaEventSynchronize(s.ev[5]));
        const ChunkCounters k = *s.h_ctr;
        if (std::getenv("ADSBCU_ORDER_TIMING_PRINT")) {
            static unsigned long long t_base = 0;
            if (c == 0) t_base = k.dbg[0];
            std::fprintf(stderr, "chunk %u: order starts @%.1f us (%.1f), resolve @%.1f..%.1f (load+runs %.1f, rounds %.1f, emit %.1f, end %.1f; n_full %u)\n", c,
                         (k.dbg[0] - t_base) * 1e-3, (k.dbg[2] - k.dbg[0]) * 1e-3, (k.dbg[3] - t_base) * 1e-3, (k.dbg[7] - t_base) * 1e-3,
                         (k.dbg[4] - k.dbg[3]) * 1e-3, (k.dbg[5] - k.dbg[4]) * 1e-3, (k.dbg[6] - k.dbg[5]) * 1e-3, (k.dbg[7] - k.dbg[6]) * 1e-3, k.n_full);
        }
        if (k.n_sequential && std::getenv("ADSBCU_DEBUG_CHUNKS"))
            std::fprintf(stderr, "chunk %u: %u of %u records went through the one-thread replay\n", c, k.n_sequential, k.n_full);
        if (k.overflow) { if (failed_chunk == nchunks) failed_chunk = c; return ADSBCU_OK; }
        if (failed_chunk != nchunks) return ADSBCU_OK;                 // chunks after a failed one did not touch the state
        if (sink.cands) copy_candidates(s.h_cand, k.n_full, h->stream_pos, sink);
        st.preambles += k.preambles; st.phase_records += k.n_cand; st.resolved_records += k.n_full; st.messages += k.n_msgs;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]); st.ms_front += ms;
        cudaEventElapsedTime(&ms, s.ev[8], s.ev[2]); st.ms_messages += ms;
        cudaEventElapsedTime(&ms, s.ev[7], s.ev[6]); st.ms_order += ms;
        cudaEventElapsedTime(&ms, s.ev[3], s.ev[4]); st.ms_resolve += ms;
        return ADSBCU_OK;
    };

    int rc = ADSBCU_OK;
    uint32_t submitted = first_chunk, retired = first_chunk;
    while (retired < nchunks) {
        while (submitted < nchunks && submitted < retired + kSlots && failed_chunk == nchunks) {
            rc = submit(submitted);
            if (rc != ADSBCU_OK) break;
            ++submitted;
        }
        if (rc != ADSBCU_OK || retired >= submitted) break;
        rc = retire(retired);
        if (rc != ADSBCU_OK) break;
        ++retired;
    }
    if (rc != ADSBCU_OK) {
        cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->msg_stream); cudaStreamSynchronize(h->order_stream); cudaStreamSynchronize(h->post_stream); cudaStreamSynchronize(h->level_stream); cudaStreamSynchronize(h->copy_stream);
        for (Slot &s : h->slot) if (s.d_zero) cudaMemset(s.d_zero, 0, s.zero_bytes);
        return rc;
    }
    return failed_chunk != nchunks ? kOverflowRetry : ADSBCU_OK;
}

int run_feed(adsbcu_handle *h, const uint8_t *d_iq, const uint8_t *iq_host, uint32_t n, Sink &sink, adsbcu_msg *out_dev, uint64_t out_cap,
             const int16_t *sc16_host = nullptr, int q11 = 0)
{
    const bool from_host = iq_host || sc16_host;
    // chunk boundaries: the first chunks are short so that the pipeline fills quickly, later ones long so that
    // per-launch costs (LUT load, tails) are amortised; host feeds are bound by the upload and keep short chunks
    // throughout, so that little work is left when the last byte has arrived
    std::vector<uint32_t> bounds{0};
    const uint32_t tps_cap = std::min(h->max_tps, from_host ? h->host_tps : h->res_tps);
    for (uint32_t tps = std::min(2u, tps_cap); bounds.back() < n; tps = std::min(2 * tps, tps_cap))
        bounds.push_back(uint32_t(std::min<uint64_t>(n, uint64_t(bounds.back()) + uint64_t(h->unit_positions) * tps)));
    const uint32_t nchunks = uint32_t(bounds.size() - 1);
    if (std::getenv("ADSBCU_DEBUG_CHUNKS")) {            // profiling runs: which launch covers how many samples
        std::fprintf(stderr, "feed of %u samples, %u chunks:", n, nchunks);
        for (uint32_t c = 0; c < nchunks; ++c) std::fprintf(stderr, " %u", bounds[c + 1] - bounds[c]);
        std::fprintf(stderr, "\n");
    }
    if (from_host) CK(cudaEventRecord(h->ev_h2d[0], h->copy_stream));

    uint32_t first = 0, failed = nchunks;
    bool uploaded = false;
    int rc = ADSBCU_OK;
    for (int attempt = 0;; ++attempt) {
        rc = run_chunks(h, d_iq, iq_host, n, sink, out_dev, out_cap, sc16_host, q11, bounds, first, failed, uploaded);
        if (rc != kOverflowRetry) break;
        // a list of chunk `failed` overflowed: nothing from that chunk on has touched the stream state.  Double the
        // lists and run again from there (the samples are on the device already).
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaStreamSynchronize(h->msg_stream));
        CK(cudaStreamSynchronize(h->order_stream));
        CK(cudaStreamSynchronize(h->post_stream));
        CK(cudaStreamSynchronize(h->level_stream));
        CK(cudaStreamSynchronize(h->copy_stream));
        if (attempt >= 5) {
            h->error = "internal candidate list overflow (traffic denser than the lists can be grown for)";
            return ADSBCU_ERR_OVERFLOW;
        }
        h->grow *= 2;
        free_slots(h);
        if (const int r = alloc_slots(h); r != ADSBCU_OK) return r;
        if (sink.cands)
            for (Slot &s : h->slot) {
                CK(cudaHostAlloc(&s.h_cand, size_t(h->full_cap) * sizeof(FullRec), cudaHostAllocMapped));
                CK(cudaHostGetDevicePointer(&s.d_cand, s.h_cand, 0));
            }
        CK(launch_state_begin_feed(h->d_state, 0, h->post_stream));  // clears the poison mark, keeps the message count
        first = failed;
        uploaded = true;
    }
    if (rc != ADSBCU_OK) return rc;
    if (sink.cands) CK(cudaMemcpyAsync(h->d_carry_prev, h->d_carry, kOverlap * sizeof(uint16_t), cudaMemcpyDeviceToDevice, h->stream));
    if (sc16_host) CK(launch_carry_mag(reinterpret_cast<const uint16_t *>(h->d_iq), n, h->d_carry, h->stream));
    else CK(launch_carry(d_iq, n, h->d_carry, h->d_tab, h->stream));
    h->stats.launches += 1;               // carry kernel
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->order_stream));
    CK(cudaStreamSynchronize(h->post_stream));
    CK(cudaStreamSynchronize(h->level_stream));
    if (from_host && nchunks) { float ms = 0.f; cudaEventElapsedTime(&ms, h->ev_h2d[0], h->ev_h2d[1]); h->stats.ms_h2d += ms; }
    h->stats.samples += n;
    h->stream_pos += n;
    return ADSBCU_OK;
}

// accepted messages of a feed: where the finish kernel writes them
struct OutTarget {
    adsbcu_msg *dev = nullptr;        // device-visible address
    bool direct = false;              // the caller's out[] itself (pinned memory): nothing to copy afterwards
};

int prepare_out(adsbcu_handle *h, adsbcu_msg *out, size_t cap, OutTarget &t)
{
    t = OutTarget{};
    if (!cap) return ADSBCU_OK;
    // out[] in pinned (cudaHostAlloc / cudaHostRegister) memory: the GPU writes the messages there itself
    cudaPointerAttributes at{};
    const bool known = cudaPointerGetAttributes(&at, out) == cudaSuccess;
    if (known && at.type == cudaMemoryTypeHost && at.devicePointer && (reinterpret_cast<uintptr_t>(at.devicePointer) & 7u) == 0) {
        t.dev = static_cast<adsbcu_msg *>(at.devicePointer);
        t.direct = true;
        return ADSBCU_OK;
    }
    // out[] in device memory: the message set stays on the GPU (for adsbcu_1090_decode_device or the caller's own kernels)
    if (known && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)) {
        if (at.type == cudaMemoryTypeDevice && at.device != h->device) { h->error = "bad argument (out[] is memory of another device)"; return ADSBCU_ERR_ARG; }
        if (reinterpret_cast<uintptr_t>(out) & 7u) { h->error = "bad argument (device out[] must be 8-byte aligned)"; return ADSBCU_ERR_ARG; }
        t.dev = out;
        t.direct = true;
        return ADSBCU_OK;
    }
    cudaGetLastError();
    if (h->msgs_cap < cap) {
        if (h->h_msgs) { cudaFreeHost(h->h_msgs); h->h_msgs = nullptr; h->msgs_cap = 0; }
        CK(cudaHostAlloc(&h->h_msgs, cap * sizeof(adsbcu_msg), cudaHostAllocMapped));
        CK(cudaHostGetDevicePointer(&h->d_msgs, h->h_msgs, 0));
        h->msgs_cap = cap;
    }
    t.dev = h->d_msgs;
    return ADSBCU_OK;
}

// Demodulate nsamples (device memory if on_device, else host memory; int16 pairs if sc16) that continue the stream.
int feed_common(adsbcu_handle *h, const uint8_t *iq, bool on_device, const int16_t *sc16, int q11, size_t nsamples, adsbcu_msg *out,
                size_t cap, size_t *n_out)
{
    CK(cudaSetDevice(h->device));
    h->stats = adsbcu_stats{};
    OutTarget target;
    if (const int r = prepare_out(h, out, cap, target); r != ADSBCU_OK) return r;
    CK(launch_state_begin_feed(h->d_state, 1, h->post_stream));
    h->stats.launches += 1;
    Sink sink{out, cap, 0, false};
    // a feed larger than max_feed_samples is cut into takes (multiples of 8 samples keep device pointers 16 B aligned)
    const size_t max_take = h->cfg.max_feed_samples & ~size_t(7);
    for (size_t done = 0; done < nsamples;) {
        const size_t take = std::min<size_t>(nsamples - done, max_take);
        const int r = sc16        ? run_feed(h, h->d_iq, nullptr, uint32_t(take), sink, target.dev, cap, sc16 + 2 * done, q11)
                      : on_device ? run_feed(h, iq + 2 * done, nullptr, uint32_t(take), sink, target.dev, cap)
                                  : run_feed(h, h->d_iq, iq + 2 * done, uint32_t(take), sink, target.dev, cap);
        if (r != ADSBCU_OK) return r;
        done += take;
    }
    const size_t n = size_t(h->stats.messages);
    if (!target.direct && n) std::memcpy(out, h->h_msgs, std::min(n, cap) * sizeof(adsbcu_msg));
    if (n_out) *n_out = n;
    if (n > cap) { h->error = "out[] too small for the accepted messages"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

}  // namespace

// No C++ exception crosses the C ABI: allocation failures inside the library become ADSBCU_ERR_NOMEM
#define ABI_GUARD_BEGIN try {
#define ABI_GUARD_END(h) } catch (const std::bad_alloc &) { if (h) (h)->error = "out of host memory"; return ADSBCU_ERR_NOMEM; } \
                         catch (...) { if (h) (h)->error = "unexpected C++ exception"; return ADSBCU_ERR_NOMEM; }

extern "C" {

int adsbcu_abi_version(void) { return ADSBCU_ABI_VERSION; }

const char *adsbcu_last_error(const adsbcu_handle *h) { return h ? h->error.c_str() : g_create_error.c_str(); }

int adsbcu_create(const adsbcu_config *cfg, adsbcu_handle **out)
{
    if (!out) return ADSBCU_ERR_ARG;
    *out = nullptr;
    adsbcu_handle *h = nullptr;
    ABI_GUARD_BEGIN
    adsbcu_config c{};
    c.struct_size = sizeof(c); c.fix_df = 1;
    if (cfg) {
        if (cfg->struct_size < sizeof(uint32_t) * 2 || cfg->struct_size > sizeof(c)) { g_create_error = "bad struct_size"; return ADSBCU_ERR_ARG; }
        std::memcpy(&c, cfg, cfg->struct_size);
    }
    if (c.nfix_crc < 0 || c.nfix_crc > 2) { g_create_error = "nfix_crc must be 0..2"; return ADSBCU_ERR_ARG; }
    if (!c.block_samples) c.block_samples = 131072;
    if (!c.max_feed_samples) c.max_feed_samples = 1u << 24;
    if (c.block_samples < 1024) { g_create_error = "block_samples too small"; return ADSBCU_ERR_ARG; }
    if (c.max_feed_samples < 4096) { g_create_error = "max_feed_samples must be at least 4096"; return ADSBCU_ERR_ARG; }
    c.max_feed_samples = (c.max_feed_samples + 7u) & ~7u;

    if (c.device == -1) {              // merge-only handle: ICAO filter + skip state for adsbcu_1090_resolve, no CUDA
        adsbcu_handle *m = new (std::nothrow) adsbcu_handle;
        if (!m) return ADSBCU_ERR_NOMEM;
        m->cfg = c; m->device = -1; m->host_only = true;
        *out = m;
        return ADSBCU_OK;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || c.device < 0 || c.device >= ndev) {
        cudaGetLastError();
        g_create_error = "no usable CUDA device (this library has no CPU fallback)";
        return ADSBCU_ERR_NO_DEVICE;
    }
    h = new (std::nothrow) adsbcu_handle;
    if (!h) return ADSBCU_ERR_NOMEM;
    h->cfg = c;
    h->device = c.device;
    auto fail = [&](int code) { g_create_error = h->error; free_all(h); delete h; h = nullptr; return code; };
#define CKC(call) do { if (!cuda_ok(h, (call), #call)) return fail(ADSBCU_ERR_CUDA); } while (0)
    CKC(cudaSetDevice(h->device));
    cudaDeviceProp prop{};
    CKC(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major < 10) { h->error = "device is not sm_100 class"; return fail(ADSBCU_ERR_NO_DEVICE); }
    h->sm_count = prop.multiProcessorCount;
    CKC(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    if (const char *v = std::getenv("ADSBCU_MSG_OVERLAP")) h->msg_overlap = std::atoi(v) != 0;
    {
        // a few SMs of their own for the order / resolve / level kernels (green contexts), else priorities on one partition
        int post_sms = 0;                                   // off by default: measured slower (profiles/r02_partition_sweep.txt)
        if (const char *v = std::getenv("ADSBCU_SM_PARTITION")) post_sms = std::atoi(v);
        CKC(cudaFree(nullptr));                             // the primary context exists before the driver API is used
        if (post_sms > 0 && h->sm_count >= 64 && make_partition(h->device, post_sms, h->part)) {
            h->stream = h->part.main; h->msg_stream = h->part.msg;
            h->post_stream = h->part.resolve;
            if (std::getenv("ADSBCU_PARTITION_RESOLVE_ONLY")) {      // order / level kernels stay with the dense kernels (many CTAs)
                int lo = 0, hi = 0;
                CKC(cudaDeviceGetStreamPriorityRange(&lo, &hi));
                CKC(cudaStreamCreateWithPriority(&h->order_stream, cudaStreamNonBlocking, hi));
                CKC(cudaStreamCreateWithPriority(&h->level_stream, cudaStreamNonBlocking, hi));
                cudaStreamDestroy(h->part.order); cudaStreamDestroy(h->part.level);
            } else {
                h->order_stream = h->part.order; h->level_stream = h->part.level;
            }
        } else {
            int lo = 0, hi = 0;
            CKC(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            CKC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
            CKC(cudaStreamCreateWithFlags(&h->msg_stream, cudaStreamNonBlocking));
            CKC(cudaStreamCreateWithPriority(&h->post_stream, cudaStreamNonBlocking, hi));
            CKC(cudaStreamCreateWithPriority(&h->order_stream, cudaStreamNonBlocking, hi));
            CKC(cudaStreamCreateWithPriority(&h->level_stream, cudaStreamNonBlocking, hi));
        }
    }
    for (auto &e : h->ev_h2d) CKC(cudaEventCreate(&e));
    for (auto &e : h->ev_up) CKC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (Slot &s : h->slot)
        for (auto &e : s.ev) CKC(cudaEventCreate(&e));
    CKC(init_kernels());
    CKC(init_front());
    CKC(init_messages());
    CKC(init_finish());

    DevTables *tab = new (std::nothrow) DevTables;
    if (!tab) return fail(ADSBCU_ERR_NOMEM);
    build_tables(*tab, c.nfix_crc, c.fix_df);
    CKC(cudaMalloc(&h->d_tab, sizeof(DevTables)));
    cudaError_t e = cudaMemcpy(h->d_tab, tab, sizeof(DevTables), cudaMemcpyHostToDevice);
    delete tab;
    CKC(e);
    CKC(launch_build_lut(h->d_tab, h->stream));
    CKC(launch_build_lutf(h->d_tab, h->stream));

    // geometry: the front kernel leaves `spare` SMs to the order / level kernels of the previous chunk; every front
    // warp (kFWarps per SM) owns up to max_tps warp-tiles of a chunk
    uint32_t tps = 16, spare = h->sm_count > 16 ? 4 : 0;
    if (const char *v = std::getenv("ADSBCU_TILES_PER_SEG")) tps = uint32_t(std::max(1, std::atoi(v)));
    if (const char *v = std::getenv("ADSBCU_SPARE_SMS")) spare = uint32_t(std::max(0, std::min(h->sm_count - 1, std::atoi(v))));
    if (const char *v = std::getenv("ADSBCU_RES_TPS")) h->res_tps = uint32_t(std::max(1, std::atoi(v)));
    if (const char *v = std::getenv("ADSBCU_HOST_TPS")) h->host_tps = uint32_t(std::max(1, std::atoi(v)));
    if (const char *v = std::getenv("ADSBCU_LEVEL_GRID")) h->level_grid = uint32_t(std::max(1, std::atoi(v)));
    if (const char *v = std::getenv("ADSBCU_RESOLVE_SEQ")) h->force_sequential = std::atoi(v) != 0;
    if (const char *v = std::getenv("ADSBCU_RESOLVE_BATCH")) h->resolve_batch = uint32_t(std::max(2, std::atoi(v)));
    const size_t nmax = c.max_feed_samples;
    h->nseg = h->part.ok ? uint32_t(h->part.sm_main) : uint32_t(h->sm_count) - spare;
    h->unit_positions = h->nseg * uint32_t(kFWarps) * uint32_t(kWTile);
    h->msg_grid = (h->part.ok ? uint32_t(h->part.sm_main) : uint32_t(h->sm_count)) * 6;
    if (const char *v = std::getenv("ADSBCU_MSG_GRID")) h->msg_grid = uint32_t(std::max(1, std::atoi(v)));
    h->max_tps = uint32_t(std::max<size_t>(1, std::min<size_t>(tps, (nmax + h->unit_positions - 1) / h->unit_positions)));
    h->max_tps = std::max(h->max_tps, std::min(std::max(h->res_tps, h->host_tps), uint32_t((nmax + h->unit_positions - 1) / h->unit_positions)));
    CKC(cudaMalloc(&h->d_bitmap, (size_t(1) << 25) / 8));
    CKC(cudaMemset(h->d_bitmap, 0, (size_t(1) << 25) / 8));
    CKC(cudaMalloc(&h->d_carry, kOverlap * sizeof(uint16_t) + 32));
    CKC(cudaMemset(h->d_carry, 0, kOverlap * sizeof(uint16_t) + 32));
    CKC(cudaMalloc(&h->d_iq, nmax * 2 + 64));
    CKC(cudaMalloc(&h->d_state, sizeof(StreamState)));
    CKC(launch_state_reset(h->d_state, 1, 1, h->stream));
    CKC(cudaHostAlloc(&h->h_word, 64, cudaHostAllocMapped));
    CKC(cudaHostGetDevicePointer(&h->d_word, h->h_word, 0));
    if (alloc_slots(h) != ADSBCU_OK) return fail(ADSBCU_ERR_CUDA);
    CKC(cudaStreamSynchronize(h->stream));
#undef CKC
    *out = h;
    return ADSBCU_OK;
    } catch (...) {
        if (h) { free_all(h); delete h; }
        g_create_error = "out of host memory";
        return ADSBCU_ERR_NOMEM;
    }
}

void adsbcu_destroy(adsbcu_handle *h)
{
    if (!h) return;
    free_all(h);
    delete h;
}

#define NEED_DEVICE(h) do { if ((h)->host_only) { (h)->error = "merge-only handle (device = -1): no GPU work possible"; return ADSBCU_ERR_NO_DEVICE; } } while (0)

int adsbcu_1090_reset(adsbcu_handle *h, int clear_filter)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (clear_filter) h->filter.clear();
    h->stream_pos = 0; h->has_last = false; h->last_a = h->last_lme = 0;
    if (h->host_only) return ADSBCU_OK;
    CK(cudaSetDevice(h->device));
    CK(cudaMemsetAsync(h->d_carry, 0, kOverlap * sizeof(uint16_t), h->stream));
    if (clear_filter) CK(cudaMemsetAsync(h->d_bitmap, 0, (size_t(1) << 25) / 8, h->stream));
    CK(launch_state_reset(h->d_state, clear_filter, 1, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_feed_device(adsbcu_handle *h, const uint8_t *iq_dev, size_t nsamples, adsbcu_msg *out, size_t cap,
                            size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq_dev && nsamples) || (!out && cap) || (reinterpret_cast<uintptr_t>(iq_dev) & 15u)) {
        h->error = "bad argument (null or misaligned pointer)";
        return ADSBCU_ERR_ARG;
    }
    ABI_GUARD_BEGIN
    return feed_common(h, iq_dev, true, nullptr, 0, nsamples, out, cap, n_out);
    ABI_GUARD_END(h)
}

int adsbcu_1090_feed(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    ABI_GUARD_BEGIN
    return feed_common(h, iq, false, nullptr, 0, nsamples, out, cap, n_out);
    ABI_GUARD_END(h)
}

int adsbcu_1090_feed_sc16(adsbcu_handle *h, const int16_t *iq, size_t nsamples, int q11, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    ABI_GUARD_BEGIN
    CK(cudaSetDevice(h->device));
    if (!h->d_sc16) CK(cudaMalloc(&h->d_sc16, size_t(h->cfg.max_feed_samples) * 4 + 64));
    return feed_common(h, nullptr, false, iq, q11 ? 1 : 0, nsamples, out, cap, n_out);
    ABI_GUARD_END(h)
}

int adsbcu_1090_magnitude(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, uint16_t *mag_out)
{
    if (!h || (!iq && nsamples) || (!mag_out && nsamples)) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (nsamples > h->cfg.max_feed_samples) { h->error = "nsamples exceeds max_feed_samples"; return ADSBCU_ERR_ARG; }
    if (!nsamples) return ADSBCU_OK;
    CK(cudaSetDevice(h->device));
    if (h->d_mag_cap < nsamples) {
        cudaFree(h->d_mag); h->d_mag = nullptr; h->d_mag_cap = 0;
        CK(cudaMalloc(&h->d_mag, nsamples * sizeof(uint16_t) + 64));
        h->d_mag_cap = nsamples;
    }
    CK(cudaMemcpyAsync(h->d_iq, iq, nsamples * 2, cudaMemcpyHostToDevice, h->stream));
    CK(launch_magnitude(h->d_iq, uint32_t(nsamples), h->d_mag, h->d_tab, h->stream));
    CK(cudaMemcpyAsync(mag_out, h->d_mag, nsamples * sizeof(uint16_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_filter_add(adsbcu_handle *h, uint32_t addr)
{
    if (!h) return ADSBCU_ERR_ARG;
    h->filter.add(addr);
    if (h->host_only) return ADSBCU_OK;
    CK(cudaSetDevice(h->device));
    CK(launch_filter_add(h->d_state, h->d_bitmap, addr, h->stream));      // table insert + the scan's address set, one thread
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_filter_test(adsbcu_handle *h, uint32_t addr)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (h->host_only) return h->filter.test(addr) ? 1 : 0;
    CK(cudaSetDevice(h->device));
    CK(launch_filter_test(h->d_state, addr, h->d_word, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return *reinterpret_cast<volatile uint32_t *>(h->h_word) ? 1 : 0;
}

int adsbcu_1090_filter_expire(adsbcu_handle *h)
{
    if (!h) return ADSBCU_ERR_ARG;
    h->filter.expire();
    if (h->host_only) return ADSBCU_OK;
    CK(cudaSetDevice(h->device));
    CK(launch_filter_expire(h->d_state, h->d_bitmap, ADSBCU_KNOWN_WORDS, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_state_export(adsbcu_handle *h, adsbcu_1090_state *st, uint32_t *addrs, size_t cap, size_t *n_addrs)
{
    if (!h || !st || (!addrs && cap)) return ADSBCU_ERR_ARG;
    if (n_addrs) *n_addrs = 0;
    NEED_DEVICE(h);
    ABI_GUARD_BEGIN
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->post_stream));
    std::vector<unsigned char> raw(sizeof(StreamState));
    CK(cudaMemcpy(raw.data(), h->d_state, sizeof(StreamState), cudaMemcpyDeviceToHost));
    const StreamState &ss = *reinterpret_cast<const StreamState *>(raw.data());
    *st = adsbcu_1090_state{h->stream_pos, ss.last_a, ss.last_lme, ss.has_last, 0u};
    size_t n = 0;
    for (int t = 0; t < 2; ++t)
        for (uint32_t i = 0; i < kIcaoSlots; ++i)
            if (const uint32_t v = ss.table[t][i]; v != kIcaoEmpty) {
                if (n < cap) addrs[n] = v;
                ++n;
            }
    if (n_addrs) *n_addrs = n;
    if (n > cap) { h->error = "addrs[] too small for the filter's entries"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
    ABI_GUARD_END(h)
}

int adsbcu_1090_state_import(adsbcu_handle *h, const adsbcu_1090_state *st, const uint32_t *addrs, size_t n_addrs)
{
    if (!h || !st || (!addrs && n_addrs) || n_addrs > 2 * size_t(kIcaoSlots)) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    for (size_t k = 0; k < n_addrs; ++k) h->filter.add(addrs[k]);       // (the merge-only resolver's copy, as adsbcu_1090_filter_add does)
    CK(cudaSetDevice(h->device));
    if (n_addrs && !h->d_import) CK(cudaMalloc(&h->d_import, 2 * size_t(kIcaoSlots) * sizeof(uint32_t)));
    if (n_addrs) CK(cudaMemcpyAsync(h->d_import, addrs, n_addrs * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    CK(launch_state_import(h->d_state, h->d_bitmap, st->has_last, st->last_a, st->last_lme, h->d_import, uint32_t(n_addrs), h->stream));
    CK(cudaStreamSynchronize(h->stream));           // addrs[] may be pageable: it is free again when the call returns
    return ADSBCU_OK;
}

int adsbcu_1090b_feed(adsbcu_handle *h, const uint8_t *iq, size_t nbytes, adsbcu_msg_b *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nbytes) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    return pathb_feed(h->pathb, h->d_tab, h->sm_count, h->stream, iq, nbytes, out, cap, n_out, h->error);
}

int adsbcu_1090b_reset(adsbcu_handle *h)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    pathb_reset(h->pathb);
    return ADSBCU_OK;
}

int adsbcu_1090b_decode_fields(adsbcu_handle *h, const adsbcu_msg_b *msgs, size_t n, adsbcu_fields_b *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!msgs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return fields_decode(h->fields, h->stream, msgs, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090b_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_pair *pairs, size_t n, adsbcu_cpr_fix *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!pairs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return cpr_decode(h->fields, h->stream, pairs, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_in *in, size_t n, adsbcu_cpr_out *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!in || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return cpr_a_decode(h->fields, h->stream, in, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090_decode_fields(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_a *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!msgs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return fields_a_decode(h->fields, h->stream, msgs, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090_decode_es(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_es *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!msgs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 24)) { h->error = "bad argument (more than 2^24 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return fields_es_decode(h->fields, h->stream, msgs, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090_decode_commb(adsbcu_handle *h, const adsbcu_msg *msgs, size_t n, adsbcu_fields_commb *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!msgs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 24)) { h->error = "bad argument (more than 2^24 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return fields_commb_decode(h->fields, h->stream, msgs, n, out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_1090_decode_device(adsbcu_handle *h, const adsbcu_msg *msgs_dev, size_t n, adsbcu_fields_a *fields_dev, adsbcu_fields_es *es_dev,
                              adsbcu_fields_commb *commb_dev, void *cuda_stream)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!msgs_dev || (!fields_dev && !es_dev && !commb_dev))) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(msgs_dev) | reinterpret_cast<uintptr_t>(fields_dev) | reinterpret_cast<uintptr_t>(es_dev) |
         reinterpret_cast<uintptr_t>(commb_dev)) & 3u) { h->error = "bad argument (device arrays must be 4-byte aligned)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return fields_all_device(h->sm_count, static_cast<cudaStream_t>(cuda_stream), msgs_dev, n, fields_dev, es_dev, commb_dev, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_978_feed(adsbcu_handle *h, const uint8_t *iq, size_t nbytes, adsbcu_uat_frame *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nbytes) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return pathc_feed(h->pathc, h->sm_count, h->stream, iq, nbytes, false, out, cap, n_out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_978_feed_device(adsbcu_handle *h, const uint8_t *iq_dev, size_t nbytes, adsbcu_uat_frame *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq_dev && nbytes) || (!out && cap) || (reinterpret_cast<uintptr_t>(iq_dev) & 1u)) { h->error = "bad argument (null or odd device pointer)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    ABI_GUARD_BEGIN
    return pathc_feed(h->pathc, h->sm_count, h->stream, iq_dev, nbytes, true, out, cap, n_out, h->error);
    ABI_GUARD_END(h)
}

int adsbcu_978_get_stats(const adsbcu_handle *h, adsbcu_978_stats *out)
{
    if (!h || !out) return ADSBCU_ERR_ARG;
    *out = adsbcu_978_stats{};
    pathc_stats(h->pathc, &out->ms_device, &out->calls, &out->launches);
    return ADSBCU_OK;
}

int adsbcu_978_reset(adsbcu_handle *h)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    pathc_reset(h->pathc);
    return ADSBCU_OK;
}

int adsbcu_978_phase(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, uint16_t *phi_out)
{
    if (!h || (!iq && nsamples) || (!phi_out && nsamples)) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    CK(cudaSetDevice(h->device));
    return pathc_phase(h->pathc, h->sm_count, h->stream, iq, nsamples, phi_out, h->error);
}

// ---- sharded captures ---------------------------------------------------------------------------

int adsbcu_1090_seek(adsbcu_handle *h, uint64_t stream_pos, const uint8_t *prev_iq, size_t nprev)
{
    if (!h || (nprev && !prev_iq) || nprev > size_t(kOverlap)) return ADSBCU_ERR_ARG;
    if (h->host_only) { h->stream_pos = stream_pos; return ADSBCU_OK; }
    CK(cudaSetDevice(h->device));
    CK(cudaMemsetAsync(h->d_carry, 0, kOverlap * sizeof(uint16_t), h->stream));
    if (nprev) {
        CK(cudaMemcpyAsync(h->d_iq, prev_iq, 2 * nprev, cudaMemcpyHostToDevice, h->stream));
        CK(launch_carry(h->d_iq, uint32_t(nprev), h->d_carry, h->d_tab, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    h->stream_pos = stream_pos;
    return ADSBCU_OK;
}

int adsbcu_1090_scan_candidates(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, int on_device, adsbcu_cand *out,
                                size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap) || (on_device && (reinterpret_cast<uintptr_t>(iq) & 15u)) ||
        nsamples > h->cfg.max_feed_samples) {
        h->error = "bad argument (null or misaligned pointer, or nsamples > max_feed_samples)";
        return ADSBCU_ERR_ARG;
    }
    ABI_GUARD_BEGIN
    CK(cudaSetDevice(h->device));
    if (!h->d_deferred) {
        // noise leaves ~1 unknown-address record per 100 positions with --fix; dense traffic more
        h->deferred_cap = uint32_t(std::max<size_t>(1u << 16, size_t(h->cfg.max_feed_samples) / 12));
        CK(cudaMalloc(&h->d_deferred, size_t(h->deferred_cap) * sizeof(PhaseRec)));
        CK(cudaMalloc(&h->d_n_deferred, sizeof(uint32_t)));
        CK(cudaMalloc(&h->d_carry_prev, kOverlap * sizeof(uint16_t) + 32));
        CK(cudaMalloc(&h->d_known_tmp, size_t(ADSBCU_KNOWN_WORDS) * sizeof(uint32_t)));
    }
    for (Slot &s : h->slot)
        if (!s.h_cand) {
            CK(cudaHostAlloc(&s.h_cand, size_t(h->full_cap) * sizeof(FullRec), cudaHostAllocMapped));
            CK(cudaHostGetDevicePointer(&s.d_cand, s.h_cand, 0));
        }
    CK(cudaMemsetAsync(h->d_n_deferred, 0, sizeof(uint32_t), h->stream));     // (the order kernels wait for this stream)
    CK(launch_state_begin_feed(h->d_state, 1, h->post_stream));
    h->stats = adsbcu_stats{};
    Sink sink{nullptr, cap, 0, false, out};
    h->last_base = h->stream_pos;
    h->last_n = uint32_t(nsamples);
    h->last_iq = on_device ? iq : h->d_iq;
    if (nsamples) {
        const int r = on_device ? run_feed(h, iq, nullptr, uint32_t(nsamples), sink, nullptr, 0)
                                : run_feed(h, h->d_iq, iq, uint32_t(nsamples), sink, nullptr, 0);
        if (r != ADSBCU_OK) return r;
    }
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the candidates"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
    ABI_GUARD_END(h)
}

int adsbcu_1090_known_export(adsbcu_handle *h, uint32_t *words)
{
    if (!h || !words) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(words, h->d_bitmap, size_t(ADSBCU_KNOWN_WORDS) * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_known_import(adsbcu_handle *h, const uint32_t *words)
{
    if (!h || !words) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    CK(cudaSetDevice(h->device));
    if (!h->d_known_tmp) CK(cudaMalloc(&h->d_known_tmp, size_t(ADSBCU_KNOWN_WORDS) * sizeof(uint32_t)));
    CK(cudaMemcpyAsync(h->d_known_tmp, words, size_t(ADSBCU_KNOWN_WORDS) * 4, cudaMemcpyHostToDevice, h->stream));
    CK(launch_or_words(h->d_bitmap, h->d_known_tmp, ADSBCU_KNOWN_WORDS, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ADSBCU_OK;
}

int adsbcu_1090_revive(adsbcu_handle *h, adsbcu_cand *out, size_t cap, size_t *n_out)
{
    if (!h || (!out && cap)) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if (!h->d_deferred || !h->last_n) return ADSBCU_OK;
    ABI_GUARD_BEGIN
    CK(cudaSetDevice(h->device));
    Slot &s = h->slot[0];
    // the parked records whose address another shard announced are sliced again from the samples (the planes of
    // their chunk are gone): evaluate_kernel, one warp per record
    ScanArgs a{};
    a.iq = h->last_iq; a.carry_mag = h->d_carry_prev; a.nsamples = h->last_n;
    a.pos_begin = 0; a.pos_end = h->last_n;
    a.tab = h->d_tab; a.bitmap = h->d_bitmap;
    a.seg = s.d_pend; a.seg_cap = h->pend_cap; a.nseg = h->nseg; a.seg_count = s.d_pend_count;
    a.full = s.d_cand; a.full_cap = h->full_cap; a.ctr = s.d_ctr;
    a.deferred = h->d_deferred; a.deferred_cap = h->deferred_cap; a.n_deferred = h->d_n_deferred;
    CK(cudaMemsetAsync(s.d_ctr, 0, sizeof(ChunkCounters), h->stream));
    CK(launch_revive(a, h->stream));
    CK(cudaMemcpyAsync(s.h_ctr, s.d_ctr, sizeof(ChunkCounters), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemsetAsync(h->d_n_deferred, 0, sizeof(uint32_t), h->stream));   // consumed
    CK(cudaMemsetAsync(s.d_zero, 0, s.zero_bytes, h->stream));              // the slot goes back to the feeds zeroed
    CK(cudaStreamSynchronize(h->stream));
    const ChunkCounters k = *s.h_ctr;
    if (k.overflow || k.n_full > h->full_cap) { h->error = "revive: candidate list overflow"; return ADSBCU_ERR_OVERFLOW; }
    std::vector<FullRec> v(s.h_cand, s.h_cand + k.n_full);
    std::sort(v.begin(), v.end(), [](const FullRec &x, const FullRec &y) { return x.pos != y.pos ? x.pos < y.pos : x.phase < y.phase; });
    Sink sink{nullptr, cap, 0, false, out};
    copy_candidates(v.data(), v.size(), h->last_base, sink);
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the revived candidates"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
    ABI_GUARD_END(h)
}

int adsbcu_1090_resolve(adsbcu_handle *h, const adsbcu_cand *cands, size_t n, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h || (!cands && n) || (!out && cap)) return ADSBCU_ERR_ARG;
    Sink sink{out, cap, 0, false};
    resolve_host(h, cands, n, sink);
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the accepted messages"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

int adsbcu_1090_modeac_block(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, const uint8_t *overlap_iq, size_t n_overlap,
                             uint64_t block_timestamp, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap) || (n_overlap && !overlap_iq) || n_overlap > size_t(kOverlap) ||
        nsamples > h->cfg.max_feed_samples) {
        h->error = "bad argument (null pointer, more than 326 overlap samples, or nsamples > max_feed_samples)";
        return ADSBCU_ERR_ARG;
    }
    if (nsamples < 2) return ADSBCU_OK;
    CK(cudaSetDevice(h->device));
    const uint32_t n = uint32_t(nsamples);
    const uint32_t want = std::max<uint32_t>(4096u, n / 16);
    if (h->ac_cap < want) {
        cudaFree(h->d_ac_rec); h->d_ac_rec = nullptr;
        if (h->h_ac_rec) { cudaFreeHost(h->h_ac_rec); h->h_ac_rec = nullptr; }
        h->ac_cap = 0;
        CK(cudaMalloc(&h->d_ac_rec, size_t(want) * sizeof(AcRec)));
        CK(cudaHostAlloc(&h->h_ac_rec, size_t(want) * sizeof(AcRec), cudaHostAllocDefault));
        h->ac_cap = want;
    }
    if (!h->d_ac_ov) {
        CK(cudaMalloc(&h->d_ac_ov, 2 * kOverlap + 64));
        CK(cudaMalloc(&h->d_ac_sums, 4 * sizeof(unsigned long long)));
    }
    const size_t need_mag = nsamples + kOverlap + kAcLead + 64;
    if (h->d_mag_cap < need_mag) {
        cudaFree(h->d_mag); h->d_mag = nullptr; h->d_mag_cap = 0;
        CK(cudaMalloc(&h->d_mag, need_mag * sizeof(uint16_t) + 64));
        h->d_mag_cap = need_mag;
    }
    uint32_t *d_count = reinterpret_cast<uint32_t *>(h->d_ac_sums + 2);
    if (n_overlap) CK(cudaMemcpyAsync(h->d_ac_ov, overlap_iq, 2 * n_overlap, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->d_iq, iq, 2 * nsamples, cudaMemcpyHostToDevice, h->stream));
    CK(launch_modeac(h->d_tab, h->sm_count, h->stream, h->d_iq, n, h->d_ac_ov, uint32_t(n_overlap), h->d_mag, h->d_ac_sums,
                     h->d_ac_rec, h->ac_cap, d_count));
    uint32_t count = 0;
    CK(cudaMemcpyAsync(&count, d_count, sizeof(count), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (count > h->ac_cap) { h->error = "Mode A/C candidate list overflow"; return ADSBCU_ERR_OVERFLOW; }
    if (count) {
        CK(cudaMemcpyAsync(h->h_ac_rec, h->d_ac_rec, size_t(count) * sizeof(AcRec), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    std::sort(h->h_ac_rec, h->h_ac_rec + count, [](const AcRec &x, const AcRec &y) { return x.f1 < y.f1; });
    size_t nmsg = 0;
    bool overflow = false;
    uint32_t next_ok = 0;
    for (uint32_t k = 0; k < count; ++k) {
        const AcRec &r = h->h_ac_rec[k];
        if (r.f1 < next_ok) continue;                       // f1_sample += 20*87/25 after a decoded reply (demod_2400.c:722)
        next_ok = r.f1 + 20 * 87 / 25 + 1;
        if (nmsg < cap) {
            adsbcu_msg &m = out[nmsg];                      // decodeModeAMessage (mode_ac.c:123-158)
            std::memset(&m, 0, sizeof(m));
            m.timestamp = block_timestamp + r.f2_clock / 5; // 60 MHz -> 12 MHz, reported at F2 (demod_2400.c:712)
            m.msgtype = 32;
            m.msgbits = 16;
            m.msg[0] = m.verbatim[0] = uint8_t(r.modeac >> 8);
            m.msg[1] = m.verbatim[1] = uint8_t(r.modeac);
            m.addr = (r.modeac & 0x0000FF7Fu) | kNonIcao;
        } else {
            overflow = true;
        }
        ++nmsg;
    }
    if (n_out) *n_out = nmsg;
    if (overflow) { h->error = "out[] too small for the Mode A/C replies"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

int adsbcu_get_stats(const adsbcu_handle *h, adsbcu_stats *out)
{
    if (!h || !out) return ADSBCU_ERR_ARG;
    *out = h->stats;
    return ADSBCU_OK;
}

}  // extern "C"
