// C ABI (include/adsbcu.h) and the ordered host pass.
//
// The GPU evaluates every start position and every phase without any stream state.  What is
// left for the host is exactly the part of demodulate2400() that is sequential by definition
// (demod_2400.c:222,336,349-355,410-419,432-443 and mode_s.c:781-788): pick the best-scoring
// phase per position given the *current* ICAO filter, accept, skip over the message, add the
// address to the filter, carry last_message_end across mag_buf boundaries.
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <new>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace adsbcu;

namespace {

thread_local std::string g_create_error;

// Reference ICAO filter (icao_filter.c:23-109): two 4096-slot open-addressed tables, one
// active; the flip (icaoFilterExpire) only happens for live SDR input (dump1090.c:469-472).
struct IcaoFilter {
    static constexpr uint32_t kSlots = 4096, kEmpty = 0xFFFFFFFFu;
    uint32_t a[kSlots], b[kSlots];
    uint32_t *active;
    IcaoFilter() { clear(); }
    void clear()
    {
        std::memset(a, 0xFF, sizeof(a));
        std::memset(b, 0xFF, sizeof(b));
        active = a;
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
        const uint32_t h0 = hash(v);
        uint32_t h = h0;
        while (active[h] != kEmpty && active[h] != v) {
            h = (h + 1) & (kSlots - 1);
            if (h == h0) return;
        }
        if (active[h] == kEmpty) active[h] = v;
    }
    static bool probe(const uint32_t *t, uint32_t v)
    {
        const uint32_t h0 = hash(v);
        uint32_t h = h0;
        while (t[h] != kEmpty && t[h] != v) {
            h = (h + 1) & (kSlots - 1);
            if (h == h0) break;
        }
        return t[h] == v;
    }
    // table b only fills after an expiry flip (live SDR input, dump1090.c:469-472), which this engine never does
    bool test(uint32_t v) const { return probe(a, v); }
};

}  // namespace

constexpr int kSlots = 3;            // chunks in flight: GPU works on chunk c+1, c+2 while the host pass replays chunk c

// One long-lived helper thread per handle for the ordered host pass of long feeds (a thread spawned per feed starts
// on the submitting thread's core and steals the time the submissions need).
struct Resolver {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<void()> job;
    bool has_job = false, quit = false;
    std::atomic<bool> idle{true};
    void post(std::function<void()> fn)
    {
        if (!th.joinable()) th = std::thread([this] { loop(); });
        { std::lock_guard<std::mutex> g(m); job = std::move(fn); has_job = true; idle.store(false, std::memory_order_relaxed); }
        cv.notify_one();
    }
    void wait() { while (!idle.load(std::memory_order_acquire)) std::this_thread::yield(); }
    void loop()
    {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> g(m);
                cv.wait(g, [this] { return has_job || quit; });
                if (quit) return;
                fn = std::move(job); has_job = false;
            }
            fn();
            idle.store(true, std::memory_order_release);
        }
    }
    ~Resolver()
    {
        if (th.joinable()) {
            { std::lock_guard<std::mutex> g(m); quit = true; }
            cv.notify_one();
            th.join();
        }
    }
};

// Everything one in-flight chunk owns.
struct Slot {
    PhaseRec *d_seg = nullptr;        // nseg segments of seg_cap records
    uint32_t *d_seg_count = nullptr;
    ChunkCounters *d_ctr = nullptr;
    ChunkCounters *h_ctr = nullptr;   // pinned + mapped
    ChunkCounters *h_ctr_dev = nullptr;   // device alias of h_ctr
    FullRec *h_full = nullptr;        // pinned + mapped: the evaluate kernel writes it directly
    FullRec *d_full = nullptr;        // device alias of h_full
    cudaEvent_t ev[4]{};              // scan begin, scan end, evaluate end, done (counters on the host)
    uint32_t seq = 0;                 // sequence number of the chunk the slot holds (threaded feeds: h_ctr->pad once it is done)
};

struct adsbcu_handle {
    adsbcu_config cfg{};
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;    // kernels + counter downloads
    cudaStream_t copy_stream = nullptr;   // host->device uploads of the feed, chunk by chunk
    cudaStream_t post_stream = nullptr;   // ADSBCU_OVERLAP=1: filter/evaluate/signal of chunk c run here, beside the scan of chunk c+1
    bool overlap = false;
    cudaEvent_t ev_h2d[2]{};
    cudaEvent_t ev_up[kSlots]{};      // upload of the chunk using that slot has finished
    std::string error;

    DevTables *d_tab = nullptr;
    uint32_t *d_bitmap = nullptr;
    uint16_t *d_carry = nullptr;
    uint8_t *d_iq = nullptr;          // staging for host feeds
    int16_t *d_sc16 = nullptr;        // staging for SC16 / SC16Q11 host feeds (allocated on first use)
    uint32_t nseg = 0, seg_cap = 0, full_cap = 0;
    uint32_t skew = 0, skew_ns = 2500;   // measured on B200: no skew 0.5065 ms, sleep skew 0.506-0.516, short first tile 0.529 (profiles/sweep_skew.sh)
    bool resolver_thread = true;       // ADSBCU_RESOLVER_THREAD=0: the ordered host pass stays on the calling thread
    uint32_t chunk_seq = 0;
    Resolver resolver;
    int wait_policy = 50;              // ADSBCU_WAIT: microseconds the submitter sleeps while all slots are busy (0: both threads yield-spin)
    uint32_t thread_tps = 8;
    uint32_t unit_positions = 0, max_tps = 1, host_tps = 4;   // chunk = unit_positions * (tiles per segment), at most max_tps
    Slot slot[kSlots];
    uint16_t *d_mag = nullptr;        // stage-level magnitude output
    size_t d_mag_cap = 0;

    // stream state
    uint64_t stream_pos = 0;          // samples consumed so far == padded coordinate of the next feed's first start
    bool has_last = false;            // last accepted message: its start and last_message_end, both in
    uint64_t last_a = 0, last_lme = 0; // stream coordinates counted from the first (zero) overlap sample
    IcaoFilter filter;
    adsbcu_stats stats{};
    PathB *pathb = nullptr;
    Fields *fields = nullptr;
    PathC *pathc = nullptr;

    // sharded captures (adsbcu_1090_scan_candidates / _revive)
    bool host_only = false;           // cfg.device == -1: merge-only handle, no CUDA at all
    uint16_t *d_carry_prev = nullptr; // the carry the last scan started from (revive re-slices with it)
    PhaseRec *d_deferred = nullptr;
    uint32_t *d_n_deferred = nullptr, *d_known_tmp = nullptr;
    uint32_t deferred_cap = 0;
    // Mode A/C (adsbcu_1090_modeac_block)
    uint8_t *d_ac_ov = nullptr;
    unsigned long long *d_ac_sums = nullptr;      // [2] sums, then the record counter
    AcRec *d_ac_rec = nullptr, *h_ac_rec = nullptr;
    uint32_t ac_cap = 0;
    const uint8_t *last_iq = nullptr; // device samples of the last candidate scan
    uint32_t last_n = 0;
    uint64_t last_base = 0;
};

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

void free_all(adsbcu_handle *h)
{
    if (h->host_only) return;
    cudaSetDevice(h->device);
    cudaFree(h->d_sc16);
    cudaFree(h->d_ac_ov); cudaFree(h->d_ac_sums); cudaFree(h->d_ac_rec);
    if (h->h_ac_rec) cudaFreeHost(h->h_ac_rec);
    cudaFree(h->d_carry_prev); cudaFree(h->d_deferred); cudaFree(h->d_n_deferred); cudaFree(h->d_known_tmp);
    if (h->stream) cudaStreamSynchronize(h->stream);
    pathb_free(h->pathb);
    h->pathb = nullptr;
    fields_free(h->fields);
    h->fields = nullptr;
    pathc_free(h->pathc);
    h->pathc = nullptr;
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    cudaFree(h->d_tab); cudaFree(h->d_bitmap); cudaFree(h->d_carry); cudaFree(h->d_iq); cudaFree(h->d_mag);
    for (Slot &s : h->slot) {
        cudaFree(s.d_seg); cudaFree(s.d_seg_count); cudaFree(s.d_ctr);
        if (s.h_ctr) cudaFreeHost(s.h_ctr);
        if (s.h_full) cudaFreeHost(s.h_full);
        for (auto &e : s.ev) if (e) cudaEventDestroy(e);
    }
    for (auto &e : h->ev_h2d) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_up) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->post_stream) cudaStreamDestroy(h->post_stream);
}

// ---- ordered host pass ------------------------------------------------------------------
// recs: the surviving (position, phase) records of one chunk in (pos, phase) order.
// feed_base: padded-stream coordinate of the feed's first start position.
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

template <class Rec>
void resolve(adsbcu_handle *h, const Rec *recs, size_t nrec, uint64_t feed_base, Sink &sink)
{
    const uint64_t block = h->cfg.block_samples;
    // mag_buf index of a stream coordinate; block_samples is a power of two unless a test shrinks it oddly
    const int bshift = (block & (block - 1)) == 0 ? __builtin_ctzll(block) : -1;
    auto blk_of = [&](uint64_t x) { return bshift >= 0 ? x >> bshift : x / block; };
    uint64_t last_blk = h->has_last ? blk_of(h->last_a) : 0;
    uint64_t ts_blk = ~0ull, ts_bstart = 0, ts_base = 0;     // timestamp base of the mag_buf being replayed
    size_t i = 0;
    while (i < nrec) {
        // the records were written by the GPU into mapped host memory: keep a few cache lines in flight
        __builtin_prefetch(reinterpret_cast<const char *>(recs + i) + 1024);
        __builtin_prefetch(reinterpret_cast<const char *>(recs + i) + 1088);
        const auto pos = recs[i].pos;
        const uint64_t a = feed_base + pos;
        size_t e = i + 1;
        while (e < nrec && recs[e].pos == pos) ++e;
        if (h->has_last) {
            // inside the mag_buf of the last accepted message scanning resumes 18 samples before its
            // end (j = last_message_end - 19, then ++j); a later mag_buf starts exactly at its end
            const bool same_block = blk_of(a) == last_blk;
            if (a < (same_block ? h->last_lme - 18 : h->last_lme)) { i = e; continue; }
        }

        int best = 0;                       // SR_NOT_SET
        const Rec *bp = nullptr;
        uint32_t last_key = kUnchecked, best_key = kUnchecked;
        bool last_known = false, best_known = false;
        for (size_t k = i; k < e; ++k) {    // phases ascending; strict > keeps the lowest phase on ties
            if (const uint32_t key = recs[k].test_key; key != last_key) { last_key = key; last_known = h->filter.test(key); }
            const int sc = rank_of(recs[k].kind, last_known);
            if (sc > best) { best = sc; bp = &recs[k]; best_key = last_key; best_known = last_known; }
        }
        i = e;
        if (best < kAcceptThreshold) continue;

        const Rec &r = *bp;
        if (sink.n < sink.cap) {
            adsbcu_msg &m = sink.out[sink.n];
            // sampleTimestamp of the block (sdr_ifile.c:202) + 5*j + 768 + phase: with 12e6/2.4e6 == 5 this
            // is 5*a + 768 + phase where a counts from the first overlap sample of the stream
            const uint64_t blk = blk_of(a);
            if (blk != ts_blk) { ts_blk = blk; ts_bstart = blk * block; ts_base = uint64_t(double(ts_bstart) * 12e6 / 2400000.0); }
            m.timestamp = ts_base + (a - ts_bstart) * 5 + (8 + 56) * 12 + r.phase;
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
        // icaoFilterAdd of an address the filter just reported as present is a no-op: skip its hash + probe
        if (const uint32_t add_key = r.add_key; add_key != kUnchecked && !(add_key == best_key && best_known)) h->filter.add(add_key);

        // skip rule (demod_2400.c:410-419, :432-443): remember where this message ends
        h->has_last = true;
        h->last_a = a;
        last_blk = blk_of(a);
        h->last_lme = a + uint64_t((((r.verbatim[0] & 0x80) ? 112 : 56) + 8) * 12 / 5);
    }
}

// ---- one feed of at most max_feed_samples: a pipeline of chunks ---------------------------------
// For every chunk: [upload its samples] -> scan -> filter -> evaluate -> counters to the host.
// Everything is queued without host synchronisation; the host pass of chunk c runs while the GPU
// works on the following chunks.  iq_host != nullptr: the samples are uploaded into d_iq chunk by
// chunk on copy_stream (a chunk only looks at samples at or before its own end, see ScanArgs).
// sc16_host != nullptr (then iq_host must be null): SC16 / SC16Q11 samples; they are uploaded chunk by chunk into
// d_sc16, converted to uint16 magnitudes into d_iq (2 bytes per sample, like uc8 I/Q) and every kernel runs
// in magnitudes-in mode (convert.c:46-92).
int run_feed(adsbcu_handle *h, const uint8_t *d_iq, const uint8_t *iq_host, uint32_t n, Sink &sink,
             const int16_t *sc16_host = nullptr, int q11 = 0)
{
    const bool from_host = iq_host || sc16_host;
    adsbcu_stats &st = h->stats;
    // chunk boundaries: the first chunks are short so that the pipeline (upload -> kernels -> host pass)
    // fills quickly, later ones long so that per-launch costs (LUT load, tail) are amortised
    std::vector<uint32_t> bounds{0};
    // host feeds are bound by the upload: short chunks throughout keep the work left after the last byte arrived small
    // (with the host pass on its own thread, 8 tiles per segment and chunk balance kernel launch overheads against the
    // host work left when the GPU finishes: profiles/r01_host_pass_thread_sweep.txt)
    const uint32_t tps_cap = from_host ? std::min(h->host_tps, h->max_tps) : h->resolver_thread ? std::min(h->max_tps, h->thread_tps) : h->max_tps;
    for (uint32_t tps = std::min(4u, tps_cap); bounds.back() < n; tps = std::min(2 * tps, tps_cap))
        bounds.push_back(uint32_t(std::min<uint64_t>(n, uint64_t(bounds.back()) + uint64_t(h->unit_positions) * tps)));
    const uint32_t nchunks = uint32_t(bounds.size() - 1);
    const bool threaded = h->resolver_thread && nchunks >= 3;
    const bool post = h->overlap && !sink.cands && nchunks >= 2;
    float ms_h2d = 0.f;
    if (from_host) CK(cudaEventRecord(h->ev_h2d[0], h->copy_stream));

    auto submit = [&](uint32_t c) -> int {
        Slot &s = h->slot[c % kSlots];
        ScanArgs a{};
        a.iq = d_iq;
        a.carry_mag = h->d_carry;
        a.nsamples = n;
        a.pos_begin = bounds[c];
        a.pos_end = bounds[c + 1];
        a.tab = h->d_tab;
        a.bitmap = h->d_bitmap;
        a.seg = s.d_seg; a.seg_cap = h->seg_cap; a.nseg = h->nseg;
        a.seg_count = s.d_seg_count;
        a.full = s.d_full; a.full_cap = h->full_cap;
        a.ctr = s.d_ctr;
        a.skew = h->skew; a.skew_ns = h->skew_ns;
        if (sink.cands) { a.deferred = h->d_deferred; a.deferred_cap = h->deferred_cap; a.n_deferred = h->d_n_deferred; }
        if (sc16_host) {
            a.iq = nullptr;
            a.mag_in = reinterpret_cast<const uint16_t *>(h->d_iq);
            CK(cudaMemcpyAsync(h->d_sc16 + 2 * size_t(a.pos_begin), sc16_host + 2 * size_t(a.pos_begin),
                               4 * size_t(a.pos_end - a.pos_begin), cudaMemcpyHostToDevice, h->copy_stream));
            CK(cudaEventRecord(h->ev_up[c % kSlots], h->copy_stream));
            if (c + 1 == nchunks) CK(cudaEventRecord(h->ev_h2d[1], h->copy_stream));
            CK(cudaStreamWaitEvent(h->stream, h->ev_up[c % kSlots], 0));
            CK(launch_convert_sc16(h->d_sc16 + 2 * size_t(a.pos_begin), a.pos_end - a.pos_begin,
                                   reinterpret_cast<uint16_t *>(h->d_iq) + a.pos_begin, q11, h->stream));
        }
        if (iq_host) {
            CK(cudaMemcpyAsync(h->d_iq + 2 * size_t(a.pos_begin), iq_host + 2 * size_t(a.pos_begin),
                               2 * size_t(a.pos_end - a.pos_begin), cudaMemcpyHostToDevice, h->copy_stream));
            CK(cudaEventRecord(h->ev_up[c % kSlots], h->copy_stream));
            if (c + 1 == nchunks) CK(cudaEventRecord(h->ev_h2d[1], h->copy_stream));
            CK(cudaStreamWaitEvent(h->stream, h->ev_up[c % kSlots], 0));
        }
        CK(cudaMemsetAsync(s.d_ctr, 0, sizeof(ChunkCounters), h->stream));
        CK(cudaEventRecord(s.ev[0], h->stream));
        CK(launch_scan(a, h->stream));
        CK(cudaEventRecord(s.ev[1], h->stream));
        // the sparse kernels of this chunk only touch the chunk's own slot (and read the address bitmap, where a
        // superset keeps more records for the exact host pass): they may run beside the next chunk's scan
        const cudaStream_t ps = post ? h->post_stream : h->stream;
        if (post) CK(cudaStreamWaitEvent(ps, s.ev[1], 0));
        CK(launch_filter(a, ps));
        CK(launch_evaluate(a, ps));
        CK(cudaEventRecord(s.ev[2], ps));
        if (threaded) {
            s.seq = ++h->chunk_seq;
            CK(launch_signal(s.d_ctr, s.h_ctr_dev, s.seq, ps));
        } else {
            CK(cudaMemcpyAsync(s.h_ctr, s.d_ctr, sizeof(ChunkCounters), cudaMemcpyDeviceToHost, ps));
        }
        CK(cudaEventRecord(s.ev[3], ps));
        return ADSBCU_OK;
    };
    // host half of a chunk: counters are on the host, the records in mapped memory
    auto host_pass = [&](uint32_t c) -> int {
        Slot &s = h->slot[c % kSlots];
        const ChunkCounters k = *s.h_ctr;
        if (k.overflow || k.n_full > h->full_cap) {
            h->error = "internal candidate list overflow (traffic denser than the lists are sized for)";
            return ADSBCU_ERR_OVERFLOW;
        }
        const auto t0 = std::chrono::steady_clock::now();
        if (sink.cands) copy_candidates(s.h_full, k.n_full, h->stream_pos, sink);
        else resolve(h, s.h_full, k.n_full, h->stream_pos, sink);
        const auto t1 = std::chrono::steady_clock::now();
        st.preambles += k.preambles; st.phase_records += k.n_cand; st.resolved_records += k.n_full;
        st.ms_resolve += std::chrono::duration<float, std::milli>(t1 - t0).count();
        return ADSBCU_OK;
    };
    auto kernel_times = [&](uint32_t c) {
        Slot &s = h->slot[c % kSlots];
        float ms = 0.f;
        cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]); st.ms_scan += ms;
        cudaEventElapsedTime(&ms, s.ev[1], s.ev[2]); st.ms_finalize += ms;
        st.scan_launches += threaded ? 4 : 3;        // scan, filter, evaluate (+ signal)
    };
    auto retire = [&](uint32_t c) -> int {
        CK(cudaEventSynchronize(h->slot[c % kSlots].ev[3]));
        const int r = host_pass(c);
        if (r == ADSBCU_OK) kernel_times(c);
        return r;
    };

    int rc = ADSBCU_OK;
    if (threaded) {
        // Long feeds: the ordered host pass runs on its own thread, which makes no CUDA call at all (it polls the
        // sequence number the signal kernel writes into mapped host memory), so the ~10 runtime calls of every
        // submit and the host pass do not queue behind each other on one core.
        std::atomic<uint32_t> retired{0};
        std::atomic<int> worker_rc{ADSBCU_OK};
        std::atomic<bool> stop{false};
        const uint32_t first_seq = h->chunk_seq + 1;
        uint32_t spins = 0;
        h->resolver.post([&] {
            for (uint32_t c = 0; c < nchunks; ++c) {
                Slot &s = h->slot[c % kSlots];
                while (reinterpret_cast<const volatile ChunkCounters *>(s.h_ctr)->pad != first_seq + c) {
                    if (stop.load(std::memory_order_relaxed)) return;
                    if (h->wait_policy == 0) std::this_thread::yield(); else __builtin_ia32_pause();
                }
                std::atomic_thread_fence(std::memory_order_acquire);
                const int r = host_pass(c);
                if (r != ADSBCU_OK) { worker_rc.store(r); return; }
                retired.store(c + 1, std::memory_order_release);
            }
        });
        for (uint32_t c = 0; c < nchunks && rc == ADSBCU_OK; ++c) {
            while (c >= retired.load(std::memory_order_acquire) + kSlots && worker_rc.load() == ADSBCU_OK) {
                if (h->wait_policy == 0) std::this_thread::yield();
                else std::this_thread::sleep_for(std::chrono::microseconds(h->wait_policy));
                if ((++spins & 63u) != 0) continue;         // every few milliseconds: has the pipeline died?
                if (const cudaError_t q = cudaStreamQuery(h->stream); q != cudaSuccess && q != cudaErrorNotReady) {
                    cuda_ok(h, q, "scan pipeline");
                    rc = ADSBCU_ERR_CUDA;
                    break;
                }
            }
            if (rc != ADSBCU_OK) break;
            if (worker_rc.load() != ADSBCU_OK) break;
            if (c >= kSlots) kernel_times(c - kSlots);          // the slot's previous chunk is done: read its events before reuse
            rc = submit(c);
        }
        if (rc != ADSBCU_OK) { cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->post_stream); stop.store(true); }
        // wait for the resolver; a failed kernel would never write its sequence number, so watch the stream for errors
        while (!h->resolver.idle.load(std::memory_order_acquire)) {
            std::this_thread::yield();                  // (a sleep here would add its timer slack to every feed)
            if ((++spins & 4095u) != 0 || stop.load()) continue;
            if (const cudaError_t q = cudaStreamQuery(h->stream); q != cudaSuccess && q != cudaErrorNotReady) {
                cuda_ok(h, q, "scan pipeline");
                rc = ADSBCU_ERR_CUDA;
                stop.store(true);
            }
        }
        if (rc == ADSBCU_OK) rc = worker_rc.load();
        if (rc == ADSBCU_OK)
            for (uint32_t c = nchunks > kSlots ? nchunks - kSlots : 0; c < nchunks; ++c) kernel_times(c);
    } else {
        uint32_t submitted = 0, retired = 0;
        while (retired < nchunks) {
            while (submitted < nchunks && submitted < retired + kSlots) {
                rc = submit(submitted);
                if (rc != ADSBCU_OK) break;
                ++submitted;
            }
            if (rc != ADSBCU_OK) break;
            rc = retire(retired);
            if (rc != ADSBCU_OK) break;
            ++retired;
        }
    }
    if (rc != ADSBCU_OK) { cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->post_stream); cudaStreamSynchronize(h->copy_stream); return rc; }
    if (post) CK(cudaStreamWaitEvent(h->stream, h->slot[(nchunks - 1) % kSlots].ev[3], 0));
    if (sink.cands) CK(cudaMemcpyAsync(h->d_carry_prev, h->d_carry, kOverlap * sizeof(uint16_t), cudaMemcpyDeviceToDevice, h->stream));
    if (sc16_host) CK(launch_carry_mag(reinterpret_cast<const uint16_t *>(h->d_iq), n, h->d_carry, h->stream));
    else CK(launch_carry(d_iq, n, h->d_carry, h->d_tab, h->stream));
    st.scan_launches += 1;
    CK(cudaStreamSynchronize(h->stream));
    if (from_host && nchunks) { cudaEventElapsedTime(&ms_h2d, h->ev_h2d[0], h->ev_h2d[1]); st.ms_h2d += ms_h2d; }
    st.samples += n;
    h->stream_pos += n;
    return ADSBCU_OK;
}

int feed_common(adsbcu_handle *h, const uint8_t *iq, bool on_device, size_t nsamples, adsbcu_msg *out, size_t cap,
                size_t *n_out)
{
    CK(cudaSetDevice(h->device));
    h->stats = adsbcu_stats{};
    Sink sink{out, cap, 0, false};
    size_t done = 0;
    // a feed larger than max_feed_samples is cut into takes (multiples of 8 samples keep device pointers 16 B aligned)
    const size_t max_take = h->cfg.max_feed_samples & ~size_t(7);
    while (done < nsamples) {
        const size_t take = std::min<size_t>(nsamples - done, max_take);
        const int r = on_device ? run_feed(h, iq + 2 * done, nullptr, uint32_t(take), sink)
                                : run_feed(h, h->d_iq, iq + 2 * done, uint32_t(take), sink);
        if (r != ADSBCU_OK) return r;
        done += take;
    }
    h->stats.messages = sink.n;
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the accepted messages"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

}  // namespace

extern "C" {

int adsbcu_abi_version(void) { return ADSBCU_ABI_VERSION; }

const char *adsbcu_last_error(const adsbcu_handle *h) { return h ? h->error.c_str() : g_create_error.c_str(); }

int adsbcu_create(const adsbcu_config *cfg, adsbcu_handle **out)
{
    if (!out) return ADSBCU_ERR_ARG;
    *out = nullptr;
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
    adsbcu_handle *h = new (std::nothrow) adsbcu_handle;
    if (!h) return ADSBCU_ERR_NOMEM;
    h->cfg = c;
    h->device = c.device;
    auto fail = [&](int code) { g_create_error = h->error; free_all(h); delete h; return code; };
#define CKC(call) do { if (!cuda_ok(h, (call), #call)) return fail(ADSBCU_ERR_CUDA); } while (0)
    CKC(cudaSetDevice(h->device));
    cudaDeviceProp prop{};
    CKC(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major < 10) { h->error = "device is not sm_100 class"; return fail(ADSBCU_ERR_NO_DEVICE); }
    h->sm_count = prop.multiProcessorCount;
    CKC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    {
        int lo = 0, hi = 0;
        CKC(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CKC(cudaStreamCreateWithPriority(&h->post_stream, cudaStreamNonBlocking, hi));
    }
    for (auto &e : h->ev_h2d) CKC(cudaEventCreate(&e));
    for (auto &e : h->ev_up) CKC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CKC(init_kernels());

    DevTables *tab = new (std::nothrow) DevTables;
    if (!tab) return fail(ADSBCU_ERR_NOMEM);
    build_tables(*tab, c.nfix_crc, c.fix_df);
    CKC(cudaMalloc(&h->d_tab, sizeof(DevTables)));
    cudaError_t e = cudaMemcpy(h->d_tab, tab, sizeof(DevTables), cudaMemcpyHostToDevice);
    delete tab;
    CKC(e);
    CKC(launch_build_lut(h->d_tab, h->stream));

    // chunk geometry: every scan group (kGroups per SM) owns up to max_tps consecutive tiles of a chunk
    uint32_t tps = 16;
    if (const char *e = std::getenv("ADSBCU_TILES_PER_SEG")) tps = uint32_t(std::max(1, std::atoi(e)));
    if (const char *e = std::getenv("ADSBCU_SKEW")) h->skew = uint32_t(std::atoi(e));
    if (const char *e = std::getenv("ADSBCU_THREAD_TPS")) h->thread_tps = uint32_t(std::max(1, std::atoi(e)));
    if (const char *e = std::getenv("ADSBCU_WAIT")) h->wait_policy = std::atoi(e);
    if (const char *e = std::getenv("ADSBCU_OVERLAP")) h->overlap = std::atoi(e) != 0;
    if (const char *e = std::getenv("ADSBCU_RESOLVER_THREAD")) h->resolver_thread = std::atoi(e) != 0;
    if (const char *e = std::getenv("ADSBCU_HOST_TPS")) h->host_tps = uint32_t(std::max(1, std::atoi(e)));
    if (const char *e = std::getenv("ADSBCU_SKEW_NS")) h->skew_ns = uint32_t(std::max(0, std::atoi(e)));
    const size_t nmax = c.max_feed_samples;
    h->nseg = uint32_t(h->sm_count) * uint32_t(kGroups);
    h->unit_positions = h->nseg * uint32_t(kTile);
    h->max_tps = uint32_t(std::max<size_t>(1, std::min<size_t>(tps, (nmax + h->unit_positions - 1) / h->unit_positions)));
    const size_t chunk = size_t(h->unit_positions) * h->max_tps;
    // list sizes: 1 (position, phase) record per 8 positions of a segment (Gaussian noise keeps 1 per 90),
    // 1 evaluated record per 96 positions of a chunk (real traffic needs 1 per ~800); overflow is reported, never silent
    h->seg_cap = uint32_t(std::max<size_t>(2048, chunk / h->nseg / 8));
    h->full_cap = uint32_t(std::max<size_t>(1u << 13, chunk / 96));
    CKC(cudaMalloc(&h->d_bitmap, (size_t(1) << 25) / 8));
    CKC(cudaMemset(h->d_bitmap, 0, (size_t(1) << 25) / 8));
    CKC(cudaMalloc(&h->d_carry, kOverlap * sizeof(uint16_t) + 32));
    CKC(cudaMemset(h->d_carry, 0, kOverlap * sizeof(uint16_t) + 32));
    CKC(cudaMalloc(&h->d_iq, nmax * 2 + 64));
    for (Slot &s : h->slot) {
        CKC(cudaMalloc(&s.d_seg, size_t(h->nseg) * h->seg_cap * sizeof(PhaseRec)));
        CKC(cudaMalloc(&s.d_seg_count, size_t(h->nseg) * sizeof(uint32_t)));
        CKC(cudaMalloc(&s.d_ctr, sizeof(ChunkCounters)));
        CKC(cudaHostAlloc(&s.h_ctr, sizeof(ChunkCounters), cudaHostAllocMapped));
        std::memset(s.h_ctr, 0, sizeof(ChunkCounters));
        CKC(cudaHostGetDevicePointer(reinterpret_cast<void **>(&s.h_ctr_dev), s.h_ctr, 0));
        CKC(cudaHostAlloc(&s.h_full, size_t(h->full_cap) * sizeof(FullRec), cudaHostAllocMapped));
        CKC(cudaHostGetDevicePointer(&s.d_full, s.h_full, 0));
        for (int k = 0; k < 4; ++k) CKC(cudaEventCreate(&s.ev[k]));
    }
#undef CKC
    *out = h;
    return ADSBCU_OK;
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
    if (h->host_only) {
        if (clear_filter) h->filter.clear();
        h->stream_pos = 0; h->has_last = false; h->last_a = h->last_lme = 0;
        return ADSBCU_OK;
    }
    CK(cudaSetDevice(h->device));
    CK(cudaMemsetAsync(h->d_carry, 0, kOverlap * sizeof(uint16_t), h->stream));
    if (clear_filter) {
        h->filter.clear();
        CK(cudaMemsetAsync(h->d_bitmap, 0, (size_t(1) << 25) / 8, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    h->stream_pos = 0;
    h->has_last = false;
    h->last_a = h->last_lme = 0;
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
    return feed_common(h, iq_dev, true, nsamples, out, cap, n_out);
}

int adsbcu_1090_feed(adsbcu_handle *h, const uint8_t *iq, size_t nsamples, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    return feed_common(h, iq, false, nsamples, out, cap, n_out);
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
    // keep the device-side superset consistent: queries use the 24-bit address (+ NT table bit)
    if (!(addr & kNonIcao)) {
        const uint32_t bi = (addr & 0xFFFFFFu) | ((addr & kNtFlag) ? (1u << 24) : 0u);
        uint32_t word = 0;
        CK(cudaSetDevice(h->device));
        CK(cudaMemcpy(&word, h->d_bitmap + (bi >> 5), 4, cudaMemcpyDeviceToHost));
        word |= 1u << (bi & 31);
        CK(cudaMemcpy(h->d_bitmap + (bi >> 5), &word, 4, cudaMemcpyHostToDevice));
    }
    return ADSBCU_OK;
}

int adsbcu_1090_filter_test(adsbcu_handle *h, uint32_t addr)
{
    if (!h) return ADSBCU_ERR_ARG;
    return h->filter.test(addr) ? 1 : 0;
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
    return fields_decode(h->fields, h->stream, msgs, n, out, h->error);
}

int adsbcu_1090b_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_pair *pairs, size_t n, adsbcu_cpr_fix *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!pairs || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    return cpr_decode(h->fields, h->stream, pairs, n, out, h->error);
}

int adsbcu_1090_cpr_decode(adsbcu_handle *h, const adsbcu_cpr_in *in, size_t n, adsbcu_cpr_out *out)
{
    if (!h) return ADSBCU_ERR_ARG;
    NEED_DEVICE(h);
    if (n && (!in || !out)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    if (n > (size_t(1) << 28)) { h->error = "bad argument (more than 2^28 records in one call)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    return cpr_a_decode(h->fields, h->stream, in, n, out, h->error);
}

int adsbcu_978_feed(adsbcu_handle *h, const uint8_t *iq, size_t nbytes, adsbcu_uat_frame *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nbytes) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    return pathc_feed(h->pathc, h->sm_count, h->stream, iq, nbytes, out, cap, n_out, h->error);
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
    CK(cudaSetDevice(h->device));
    if (!h->d_deferred) {
        // noise leaves ~1 unknown-address record per 100 positions with --fix; dense traffic more
        h->deferred_cap = uint32_t(std::max<size_t>(1u << 16, size_t(h->cfg.max_feed_samples) / 12));
        CK(cudaMalloc(&h->d_deferred, size_t(h->deferred_cap) * sizeof(PhaseRec)));
        CK(cudaMalloc(&h->d_n_deferred, sizeof(uint32_t)));
        CK(cudaMalloc(&h->d_carry_prev, kOverlap * sizeof(uint16_t) + 32));
        CK(cudaMalloc(&h->d_known_tmp, size_t(ADSBCU_KNOWN_WORDS) * sizeof(uint32_t)));
    }
    CK(cudaMemsetAsync(h->d_n_deferred, 0, sizeof(uint32_t), h->stream));
    h->stats = adsbcu_stats{};
    Sink sink{nullptr, cap, 0, false, out};
    h->last_base = h->stream_pos;
    h->last_n = uint32_t(nsamples);
    h->last_iq = on_device ? iq : h->d_iq;
    if (nsamples) {
        const int r = on_device ? run_feed(h, iq, nullptr, uint32_t(nsamples), sink) : run_feed(h, h->d_iq, iq, uint32_t(nsamples), sink);
        if (r != ADSBCU_OK) return r;
    }
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the candidates"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
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
    CK(cudaSetDevice(h->device));
    Slot &s = h->slot[0];
    ScanArgs a{};
    a.iq = h->last_iq; a.carry_mag = h->d_carry_prev; a.nsamples = h->last_n;
    a.pos_begin = 0; a.pos_end = h->last_n;
    a.tab = h->d_tab; a.bitmap = h->d_bitmap;
    a.seg = s.d_seg; a.seg_cap = h->seg_cap; a.nseg = h->nseg; a.seg_count = s.d_seg_count;
    a.full = s.d_full; a.full_cap = h->full_cap; a.ctr = s.d_ctr;
    a.deferred = h->d_deferred; a.deferred_cap = h->deferred_cap; a.n_deferred = h->d_n_deferred;
    CK(cudaMemsetAsync(s.d_ctr, 0, sizeof(ChunkCounters), h->stream));
    CK(launch_revive(a, h->stream));
    CK(cudaMemcpyAsync(s.h_ctr, s.d_ctr, sizeof(ChunkCounters), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemsetAsync(h->d_n_deferred, 0, sizeof(uint32_t), h->stream));   // consumed
    CK(cudaStreamSynchronize(h->stream));
    const ChunkCounters k = *s.h_ctr;
    if (k.overflow || k.n_full > h->full_cap) { h->error = "revive: candidate list overflow"; return ADSBCU_ERR_OVERFLOW; }
    std::vector<FullRec> v(s.h_full, s.h_full + k.n_full);
    std::sort(v.begin(), v.end(), [](const FullRec &x, const FullRec &y) { return x.pos != y.pos ? x.pos < y.pos : x.phase < y.phase; });
    Sink sink{nullptr, cap, 0, false, out};
    copy_candidates(v.data(), v.size(), h->last_base, sink);
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the revived candidates"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

int adsbcu_1090_resolve(adsbcu_handle *h, const adsbcu_cand *cands, size_t n, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h || (!cands && n) || (!out && cap)) return ADSBCU_ERR_ARG;
    Sink sink{out, cap, 0, false};
    resolve(h, cands, n, 0, sink);
    if (n_out) *n_out = sink.n;
    if (sink.overflow) { h->error = "out[] too small for the accepted messages"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

int adsbcu_1090_feed_sc16(adsbcu_handle *h, const int16_t *iq, size_t nsamples, int q11, adsbcu_msg *out, size_t cap, size_t *n_out)
{
    if (!h) return ADSBCU_ERR_ARG;
    if (n_out) *n_out = 0;
    NEED_DEVICE(h);
    if ((!iq && nsamples) || (!out && cap)) { h->error = "bad argument (null pointer)"; return ADSBCU_ERR_ARG; }
    CK(cudaSetDevice(h->device));
    if (!h->d_sc16) CK(cudaMalloc(&h->d_sc16, size_t(h->cfg.max_feed_samples) * 4 + 64));
    h->stats = adsbcu_stats{};
    Sink sink{out, cap, 0, false};
    const size_t max_take = h->cfg.max_feed_samples & ~size_t(7);
    for (size_t done = 0; done < nsamples;) {
        const size_t take = std::min<size_t>(nsamples - done, max_take);
        const int r = run_feed(h, h->d_iq, nullptr, uint32_t(take), sink, iq + 2 * done, q11 ? 1 : 0);
        if (r != ADSBCU_OK) return r;
        done += take;
    }
    h->stats.messages = sink.n;
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
